"""The B=64 decode loop as n independent sub-batches (64 / n utterances each: own state, own step graph, own stream),
graph replays issued round-robin: while one sub-batch is in its latency-bound GEMM chain another streams its KV cache.
Device time of the 867-step loop for n = 1, 2, 4 and the tokens against the n = 1 run.
    python profiles/scripts/two_streams.py [kv_dtype]"""
import hashlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.engine import DecoderEngine, GenState  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, 860
K = 9
M = N + K
kvd = sys.argv[1] if len(sys.argv) > 1 else None
model = bench.build_models(dev, torch.bfloat16)
model.decoder.kv_dtype = kvd
w = model.decoder.packed_weights()
eng = DecoderEngine(w)
g = torch.Generator().manual_seed(1234)
ids = torch.randint(2, 32100, (B, S), generator=g).to(dev)
pids = torch.randint(2, 32100, (B, P), generator=g).to(dev)
with torch.no_grad():
    enc = model.project_encoder(model.encode_text(ids, torch.ones_like(ids)), torch.ones_like(ids))
table = model.prompt_table()
want = None
for n in (1, 2, 4, 1, 2, 4):
    b = B // n
    sts = [GenState(w, b, S, P, 1, M, impl=cabi.IMPL_FAST, suppress_from=1024, row_offset=i * b * K, kv_dtype=kvd)
           for i in range(n)]
    streams = [torch.cuda.Stream() for _ in range(n)]
    start = torch.full((b * K, 1), 1025, dtype=torch.int64, device=dev)

    def prepare():
        for i, st in enumerate(sts):
            st.reset(start, seed=0)
            eng.cross_kv(st, enc[i * b:(i + 1) * b].contiguous())
            eng.prefill(st, table, pids[i * b:(i + 1) * b].contiguous())

    prepare()
    graphs = []
    for st in sts:
        eng.step(st)
        graphs.append(eng.capture(st))
    best = None
    for rep in range(3):
        prepare()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in streams:
            s.wait_stream(torch.cuda.current_stream())
        for _ in range(M - 2):
            for s, gr in zip(streams, graphs):
                with torch.cuda.stream(s):
                    gr.replay()
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    seq = torch.cat([st.ids for st in sts], dim=0)
    sha = hashlib.sha1(seq.cpu().numpy().tobytes()).hexdigest()[:16]
    if want is None:
        want = seq.clone()
    same = float((seq == want).float().mean())
    print(f"n = {n} sub-batches of {b:2d}: loop {best:8.1f} ms, {best / (M - 2):.4f} ms per step of 64 utterances, "
          f"{B * N / best * 1e3:8.0f} frames/s, tokens equal to n = 1: {same:.4f} ({sha})", flush=True)
    del sts, graphs
    torch.cuda.empty_cache()
