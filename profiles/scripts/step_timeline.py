"""Timeline of ONE replay of the decode-step CUDA graph (Mini v0.1, B=64): for every fast-family launch the first CTA
start, first / last CTA past griddepcontrol.wait and last CTA end (globaltimer, PTTS_TIMELINE=1), printed for one
decoder layer plus per-kernel-class aggregates: how long each kernel really is, how early PDL makes it resident and
how long the hand-over between dependent kernels takes."""
import ctypes as C
import os
import sys

os.environ["PTTS_TIMELINE"] = "1"
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.engine import DecoderEngine, GenState  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, int(sys.argv[1]) if len(sys.argv) > 1 else 440
K, M = 9, N + 9
model = bench.build_models(dev, torch.bfloat16)
w = model.decoder.packed_weights()
eng = DecoderEngine(w)
st = GenState(w, B, S, P, 1, M + 40, suppress_from=1024)
enc = (torch.randn(B, S, 1024, device=dev) * 0.3).to(torch.bfloat16)
prompt = torch.randint(2, 32100, (B, P), device=dev)
st.reset(torch.full((B * K, 1), 1025, dtype=torch.int64, device=dev), seed=0)
eng.cross_kv(st, enc)
eng.prefill(st, model.prompt_table(), prompt)
eng.run_loop(st, use_graph=True, poll_every=64, max_steps=N)
torch.cuda.synchronize()
lib = cabi.load()
lib.ptts_diag_timeline(0, None, None, 0)
st.graph.replay()
torch.cuda.synchronize()
buf = np.zeros(4096 * 4, dtype=np.uint64)
kinds = np.zeros(4096, dtype=np.int32)
n = lib.ptts_diag_timeline(1, buf.ctypes.data_as(C.c_void_p), kinds.ctypes.data_as(C.c_void_p), 4096)
t = buf.reshape(4096, 4)[:n].astype(np.float64)
ok = t[:, 3] > 0
t, kinds = t[ok], kinds[:n][ok]
order = np.argsort(t[:, 1])
t, kinds = t[order], kinds[order]
t0 = t[0, 0]
names = {1: "layernorm", 2: "self-attn", 3: "cross-attn", 4: "reduce-act", 5: "cross-block", 10: "gemm partial", 11: "gemm store", 12: "gemm scatter"}
print(f"context ~{P + N}; {len(t)} instrumented launches in one replay; step span {(t[:, 3].max() - t0) / 1e3:.1f} us")
print("  kernel          resident-before-dep  dep-spread  run(after dep)  handover(prev end -> dep)")
for i in range(12, 24):
    r = t[i]
    print(f"  {names.get(int(kinds[i]), kinds[i]):14s} {(r[1] - r[0]) / 1e3:10.2f} us {(r[2] - r[1]) / 1e3:10.2f} us "
          f"{(r[3] - r[1]) / 1e3:10.2f} us {(r[1] - t[i - 1, 3]) / 1e3:12.2f} us")
print("per class (median over the step): run after dep | handover from previous kernel end | resident before dep")
for k, nm in names.items():
    idx = np.where(kinds == k)[0]
    idx = idx[idx > 0]
    if len(idx) == 0:
        continue
    run = np.median(t[idx, 3] - t[idx, 1]) / 1e3
    hand = np.median(t[idx, 1] - t[idx - 1, 3]) / 1e3
    res = np.median(t[idx, 1] - t[idx, 0]) / 1e3
    print(f"  {nm:14s} n={len(idx):3d} run {run:6.2f} us  handover {hand:6.2f} us  resident-early {res:6.2f} us  "
          f"sum(run+handover) {np.sum(t[idx, 3] - t[idx - 1, 3]) / 1e3:8.1f} us")
