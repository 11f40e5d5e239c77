"""Does splitting the batch into independent lanes (each its own PDL chain on its own stream, all forked inside ONE
CUDA graph) shorten the decode step?  The chain of ~290 dependent kernels is latency-bound for two thirds of its
length; two or four lanes let one lane's HBM-bound attention overlap another lane's latency-bound GEMM / LayerNorm
hand-offs.  Mini v0.1, B=64 in total, bf16, context ~470."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parler_tts_b200.engine import DecoderEngine, GenState  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
Btot, S, P, N = 64, 48, 32, int(sys.argv[1]) if len(sys.argv) > 1 else 440
K, M = 9, N + 9
model = bench.build_models(dev, torch.bfloat16)
w = model.decoder.packed_weights()
eng = DecoderEngine(w)


def make_lane(B, row0):
    st = GenState(w, B, S, P, 1, M + 80, suppress_from=1024, row_offset=row0 * K)
    g = torch.Generator().manual_seed(row0)
    enc = (torch.randn(B, S, 1024, generator=g) * 0.3).to(dev, torch.bfloat16)
    prompt = torch.randint(2, 32100, (B, P), generator=g).to(dev)
    st.reset(torch.full((B * K, 1), 1025, dtype=torch.int64, device=dev), seed=0)
    eng.cross_kv(st, enc)
    eng.prefill(st, model.prompt_table(), prompt)
    eng.run_loop(st, use_graph=True, poll_every=64, max_steps=N)
    return st


cfgs = [tuple(int(v) for v in a.split('x')) for a in sys.argv[2:]] or [(1, 64), (2, 32), (4, 16)]
for lanes, B in cfgs:
    states = [make_lane(B, i * B) for i in range(lanes)]
    torch.cuda.synchronize()
    main = torch.cuda.Stream()
    streams = [torch.cuda.Stream() for _ in range(lanes)]
    fork = torch.cuda.Event()
    joins = [torch.cuda.Event() for _ in range(lanes)]
    dummy = torch.zeros(lanes + 1, device=dev)
    g = torch.cuda.CUDAGraph()
    main.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=main):
        dummy[lanes:].add_(1)
        fork.record(main)
        for i, (st, s) in enumerate(zip(states, streams)):
            s.wait_event(fork)
            with torch.cuda.stream(s):
                dummy[i : i + 1].add_(1)  # a plain kernel first: a PDL launch must not directly follow a cross-stream edge
                eng.step(st)
                joins[i].record(s)
        for j in joins:
            main.wait_event(j)
        dummy[lanes:].add_(1)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"{lanes} lane(s) x B={B}: {e0.elapsed_time(e1) / reps:.3f} ms per decode step of the whole batch "
          f"(context {states[0].read_ctrl()['len'] + P})")
    del states, g
    torch.cuda.empty_cache()
