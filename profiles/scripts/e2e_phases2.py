"""Finer split of the end-to-end generate() call at the benched size (Mini v0.1, B=64, 860 frames, bf16): CUDA-event
time of every phase in front of and behind the decode loop (H2D ids, T5 encoder, enc_to_dec_proj, cross K/V, prefill,
loop, de-delay, DAC decode, D2H audio), re-issued by hand through the same engine calls generate() makes."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.engine import DecoderEngine  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, 860
model = bench.build_models(dev, torch.bfloat16)
model.generation_config.max_length = N + 9
g = torch.Generator().manual_seed(0)
ids = torch.randint(2, 32100, (B, S), generator=g).pin_memory()
pids = torch.randint(2, 32100, (B, P), generator=g).pin_memory()
procs = [ptts.SuppressTokensAtOrAbove(1024)]
# warm: builds the state, captures the graph
for _ in range(2):
    model.generate(input_ids=ids.to(dev), prompt_input_ids=pids.to(dev), do_sample=False, logits_processor=procs, seed=0)
st = model.decoder._state_cache[1]
eng = DecoderEngine(model.decoder.packed_weights())
dac = model.audio_encoder.engine(model.decoder.impl)
K = 9


def ev():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


for it in range(3):
    torch.cuda.synchronize()
    marks = [("start", ev())]
    d_ids, d_pids = ids.to(dev, non_blocking=True), pids.to(dev, non_blocking=True)
    marks.append(("H2D ids", ev()))
    am = torch.ones_like(d_ids)
    hs = model.encode_text(d_ids, am)
    marks.append(("T5 encoder", ev()))
    enc = model.project_encoder(hs, am)
    table = model.prompt_table()
    marks.append(("enc_to_dec_proj", ev()))
    start = torch.full((B * K, 1), 1025, dtype=torch.int64, device=dev)
    st.reset(start, 0)
    marks.append(("state reset + delay mask", ev()))
    eng.cross_kv(st, enc)
    marks.append(("cross K/V (24 layers)", ev()))
    eng.prefill(st, table, d_pids.to(torch.int64).contiguous(), None)
    marks.append(("prefill (P + 1 positions)", ev()))
    ctrl = eng.run_loop(st, use_graph=True)
    marks.append(("decode loop", ev()))
    Lout = ctrl["stop_len"]
    codes = torch.empty(B, K, Lout - K, device=dev, dtype=torch.int64)
    cabi.call("ptts_dedelay", cabi.ptr(st.ids), st.max_len, cabi.ptr(st.mask), st.max_len, B, K, Lout, cabi.ptr(codes),
              cabi.stream_ptr())
    marks.append(("de-delay", ev()))
    audio = dac.decode_generated(codes, 1025, 1024, 1024)
    marks.append(("DAC decode (incl. special-id check)", ev()))
    host = audio.cpu()
    marks.append(("D2H audio", ev()))
    torch.cuda.synchronize()
    total = marks[0][1].elapsed_time(marks[-1][1])
    print(f"run {it}: total {total:8.1f} ms")
    for (_, a), (name, b) in zip(marks, marks[1:]):
        print(f"   {name:38s} {a.elapsed_time(b):8.2f} ms")
