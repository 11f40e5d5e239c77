"""One pass of T5 encoder + cross K/V + prompt prefill at the benched size (Mini v0.1, B=64, S=48, P=32, bf16) between
cudaProfilerStart / Stop, for the ncu launch list (serialised per-launch durations: kernels launched with programmatic
dependent launch overlap their predecessor, so CUPTI activity durations of the chain include waiting time):
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv \
        python profiles/scripts/prefill_ncu.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200.engine import DecoderEngine  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P = 64, 48, 32
model = bench.build_models(dev, torch.bfloat16)
model.generation_config.max_length = 40
g = torch.Generator().manual_seed(0)
ids = torch.randint(2, 32100, (B, S), generator=g).to(dev)
pids = torch.randint(2, 32100, (B, P), generator=g).to(dev)
procs = [ptts.SuppressTokensAtOrAbove(1024)]
for _ in range(2):
    model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs, seed=0, return_codes=True)
st = model.decoder._state_cache[1]
eng = DecoderEngine(model.decoder.packed_weights())
am = torch.ones_like(ids)
table = model.prompt_table()
start = torch.full((B * 9, 1), 1025, dtype=torch.int64, device=dev)
pids64 = pids.to(torch.int64).contiguous()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
enc = model.project_encoder(model.encode_text(ids, am), am)
st.reset(start, 0)
eng.cross_kv(st, enc)
eng.prefill(st, table, pids64, None)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("done")
