"""Times the DAC decode (44.1 kHz architecture, random weights) of B x F frames with both kernel families."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import dac_oracle  # noqa: E402  (random weights only)
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.dac import DacDecoderEngine  # noqa: E402

B, F = int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 860
sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=2)
codes = torch.randint(0, 1024, (B, 9, F)).cuda()
for name, impl in (("fast", cabi.IMPL_FAST), ("simple", cabi.IMPL_SIMPLE)):
    eng = DacDecoderEngine(sd, "cuda", impl=impl)
    eng.decode(codes[:1, :, :64])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = eng.decode(codes)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"dac {name}: {B}x{F} frames in {dt * 1e3:.1f} ms -> {B * F / dt:.0f} frames/s, "
          f"{B * F * 1.608e9 / dt / 1e12:.1f} TFLOP/s (fp32-equivalent)")
