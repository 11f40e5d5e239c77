"""Does splitting the batch into independent micro-batches on forked graph branches hide the dependent-kernel chain?
B=64 is run as n micro-batches of 64/n utterances (own GenState each, shared weights), every branch one whole decode
step, captured into ONE CUDA graph with fork/join; branch i is delayed by i*delay_us (spin kernel) so that one
branch's attention (HBM-bound) overlaps another branch's GEMM / LayerNorm chain (latency-bound).
Usage: python profiles/scripts/microbatch.py [frames_before_timing]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parler_tts_b200.engine import DecoderEngine, GenState  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, int(sys.argv[1]) if len(sys.argv) > 1 else 434
K, M = 9, N + 9
model = bench.build_models(dev, torch.bfloat16)
w = model.decoder.packed_weights()
eng = DecoderEngine(w)
table = model.prompt_table()
CYC_PER_US = 1965


def make(n):
    Bm = B // n
    sts = []
    for i in range(n):
        st = GenState(w, Bm, S, P, 1, M + 400, suppress_from=1024, row_offset=i * Bm * K)
        enc = (torch.randn(Bm, S, 1024, device=dev) * 0.3).to(torch.bfloat16)
        prompt = torch.randint(2, 32100, (Bm, P), device=dev)
        st.reset(torch.full((Bm * K, 1), 1025, dtype=torch.int64, device=dev), seed=0)
        eng.cross_kv(st, enc)
        eng.prefill(st, table, prompt)
        eng.run_loop(st, use_graph=True, poll_every=64, max_steps=N)
        sts.append(st)
    torch.cuda.synchronize()
    return sts


def timed(sts, delay_us, reps=60):
    n = len(sts)
    side = [torch.cuda.Stream() for _ in range(n - 1)]
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        main = torch.cuda.current_stream()
        for s in side:
            s.wait_stream(main)
        eng.step(sts[0])
        for i, (s, st) in enumerate(zip(side, sts[1:]), start=1):
            with torch.cuda.stream(s):
                if delay_us > 0:
                    torch.cuda._sleep(int(delay_us * i * CYC_PER_US))
                eng.step(st)
        for s in side:
            main.wait_stream(s)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for n in (1, 2, 4):
    sts = make(n)
    for delay in ((0,) if n == 1 else (0, 15, 30, 45)):
        ms = timed(sts, delay)
        ctx = P + sts[0].read_ctrl()["len"]
        print(f"micro-batches {n} x B={B // n}  delay {delay:3d} us/branch  ctx ~{ctx}: {ms:.4f} ms per step of B={B}",
              flush=True)
    del sts
    torch.cuda.empty_cache()
