"""One EAGER decode step of the benched configuration (Mini v0.1, bf16, B=64, mean context 466) between
cudaProfilerStart/Stop, for `ncu --profile-from-start off`:

  # step-level DRAM traffic and per-launch times
  ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/step_bytes.csv \
      --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum \
      python profiles/scripts/ncu_step.py [bf16|fp8] [layers]
  # --set full of the step's kernels on a 2-layer model of the same widths (same per-kernel shapes)
  ncu --profile-from-start off --clock-control none --set full -o gpurun_out/step_full python profiles/scripts/ncu_step.py bf16 2

The cache rows are whatever the allocator left there (zeros): traffic and timing do not depend on the values.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.engine import DecoderEngine, GenState  # noqa: E402

kv = sys.argv[1] if len(sys.argv) > 1 else "bf16"
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 24
dac = len(sys.argv) > 3 and sys.argv[3] == "dac"
bench.MINI["num_hidden_layers"] = layers
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
model = bench.build_models(dev, torch.bfloat16)
w = model.decoder.packed_weights()
eng = DecoderEngine(w)
B, S, P, K, M = 64, 48, 32, 9, 869
g = torch.Generator().manual_seed(1234)
ids = torch.randint(2, 32100, (B, S), generator=g).to(dev)
prm = torch.randint(2, 32100, (B, P), generator=g).to(dev)
with torch.no_grad():
    enc = model.project_encoder(model.encode_text(ids, torch.ones_like(ids)), torch.ones_like(ids))
st = GenState(w, B, S, P, 1, M, suppress_from=1024, kv_dtype="fp8" if kv == "fp8" else None)
st.reset(torch.full((B * K, 1), 1025, dtype=torch.int64, device=dev), seed=0)
eng.cross_kv(st, enc)
eng.prefill(st, model.prompt_table(), prm)
st.ctrl[0] = M // 2  # jump to the mean context (434 columns -> 466 cache rows)
eng.step(st)         # loads every kernel
eng.step(st)
torch.cuda.synchronize()
codes = torch.randint(0, 1024, (2, 9, 86), device=dev)
dac_eng = model.audio_encoder.engine() if dac else None
if dac:
    dac_eng.decode(codes)
    torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.step(st)
if dac:
    dac_eng.decode(codes)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ncu_step done:", kv, layers, "launches per step", int(cabi.load().ptts_launch_count()))
