"""Kernel table (torch.profiler / CUPTI) of the two once-per-call phases in front of the decode loop at the benched size
(Mini v0.1, B=64, S=48, P=32, bf16): the T5 description encoder and the prompt prefill."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

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


def phase(name, fn):
    fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    print(f"==== {name}")
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=70))


phase("T5 encoder", lambda: model.encode_text(ids, am))
enc = model.project_encoder(model.encode_text(ids, am), am)


def prefill():
    st.reset(start, 0)
    eng.cross_kv(st, enc)
    eng.prefill(st, table, pids.to(torch.int64).contiguous(), None)


phase("cross K/V + prefill", prefill)
