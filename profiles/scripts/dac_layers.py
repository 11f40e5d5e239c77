"""Per-layer timing of the fast DAC decoder (44.1 kHz architecture): every C-ABI call is bracketed with CUDA events and
aggregated by (entry point, Cin, Cout, taps/stride, dilation), with the tensor-core rate each layer reaches."""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import dac_oracle  # noqa: E402  (random weights only)
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.dac import DacDecoderEngine  # noqa: E402

B, F = int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 860
sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=2)
codes = torch.randint(0, 1024, (B, 9, F)).cuda()
eng = DacDecoderEngine(sd, "cuda", impl=cabi.IMPL_FAST)
eng.decode(codes[:1, :, :64])
torch.cuda.synchronize()
events, orig = [], cabi.call


def timed(name, *args):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    orig(name, *args)
    e1.record()
    if name == "ptts_dac_conv1d_nwc":
        key, macs = (name, args[3], args[7], args[8], args[9]), args[1] * args[2] * args[3] * args[7] * args[8]
    elif name == "ptts_dac_convtr1d_nwc":
        key, macs = (name, args[3], args[7], args[8], 0), args[1] * args[2] * args[3] * args[7] * 2 * args[8]
    elif name == "ptts_dac_conv1d_px":
        mode = ("planes" if args[1] is not None else "fp32") + "->" + ("planes" if args[13] is not None else "fp32")
        key, macs = (name + " " + mode, args[4], args[7], args[8], args[9]), args[2] * args[3] * args[4] * args[7] * args[8]
    elif name == "ptts_dac_convtr1d_px":
        key, macs = (name, args[4], args[7], args[8], 0), args[2] * args[3] * args[4] * args[7] * 2 * args[8]
    else:
        key, macs = (name, 0, 0, 0, 0), 0
    events.append((key, macs, e0, e1))


cabi.call = timed
import parler_tts_b200.dac as dacmod  # noqa: E402

dacmod.cabi.call = timed
eng.decode(codes)
torch.cuda.synchronize()
agg = collections.OrderedDict()
for key, macs, e0, e1 in events:
    a = agg.setdefault(key, [0, 0.0, 0])
    a[0] += 1
    a[1] += e0.elapsed_time(e1)
    a[2] += macs
tot = sum(a[1] for a in agg.values())
print(f"B={B} F={F}: total {tot:.1f} ms")
for key, (n, ms, macs) in agg.items():
    print(f"{key[0][9:]:32s} Cin={key[1]:5d} Cout={key[2]:5d} k/s={key[3]:2d} dil={key[4]:2d} n={n:3d} {ms:8.2f} ms "
          f"({100 * ms / tot:4.1f}%)  {2 * macs / ms / 1e9 if ms else 0:7.1f} TFLOP/s fp32-eq")
