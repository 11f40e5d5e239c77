"""DAC decode of B=64 x 860 frames: batch chunking of DacDecoderEngine.decode (max_frames_per_call) - 4 x 16, 2 x 32
(default) or 1 x 64 utterances per pass.  CUDA-event time of the whole decode, second call of each setting."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import dac_oracle  # noqa: E402  (random weights only)
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.dac import DacDecoderEngine  # noqa: E402

B, F = 64, 860
sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=2)
codes = torch.randint(0, 1024, (B, 9, F)).cuda()
eng = DacDecoderEngine(sd, "cuda", impl=cabi.IMPL_FAST)
ref = None
for per_call in (16 * 1024, 32 * 1024, 64 * 1024, 16 * 1024, 32 * 1024, 64 * 1024):
    torch.cuda.empty_cache()
    eng.decode(codes, max_frames_per_call=per_call)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = eng.decode(codes, max_frames_per_call=per_call)
    b.record()
    torch.cuda.synchronize()
    if ref is None:
        ref = out.clone()
    print(f"max_frames_per_call {per_call:6d} ({max(1, per_call // F):3d} utterances per pass): {a.elapsed_time(b):7.1f} ms, "
          f"max |diff| vs first setting {float((out - ref).abs().max()):.1e}, peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
    del out
