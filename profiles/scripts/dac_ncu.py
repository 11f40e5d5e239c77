"""One DAC decode (44.1 kHz architecture, B=8 x 860 frames) between cudaProfilerStart/Stop for
   ncu --profile-from-start off --set full -k regex:dac_conv_ws --launch-skip N --launch-count M ...
(29 dac_conv_ws launches per decode: conv_in, then per block the transposed conv and three (k7, k1) residual units)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import dac_oracle  # noqa: E402  (random weights only)
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.dac import DacDecoderEngine  # noqa: E402

B, F = int(sys.argv[1]) if len(sys.argv) > 1 else 8, 860
sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=2)
codes = torch.randint(0, 1024, (B, 9, F)).cuda()
eng = DacDecoderEngine(sd, "cuda", impl=cabi.IMPL_FAST)
eng.decode(codes)
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.decode(codes)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("dac_ncu done")
