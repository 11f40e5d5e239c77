"""Per-phase SM-clock cycles of the fused cross-attention block (fast_xblock.cu) inside the replayed decode-step graph
(Mini v0.1, B=64): mean over all CTAs and 24 layers x replays of the cycles since the CTA's griddepcontrol.wait
returned.  Run: PTTS_XB_DBG=1 is set here; python profiles/scripts/xblock_stamps.py [frames]"""
import ctypes as C
import os
import sys

os.environ["PTTS_XB_DBG"] = "1"
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
lib.ptts_diag_xblock(None, 1)
for _ in range(10):
    st.graph.replay()
torch.cuda.synchronize()
buf = np.zeros(8, dtype=np.uint64)
lib.ptts_diag_xblock(buf.ctypes.data_as(C.c_void_p), 0)
n = float(buf[6])
names = ["LayerNorm done", "Wq slab landed", "q done", "attention done", "Wo slab landed", "end"]
print(f"context ~{P + N}; {int(n)} CTAs; cycles after the dependency wait (mean), 1965 MHz")
prev = 0.0
for i, nm in enumerate(names):
    c = float(buf[i]) / n
    print(f"  {nm:16s} {c:8.0f} cyc  {c / 1965:6.2f} us   (+{(c - prev) / 1965:5.2f} us)")
    prev = c
print(f"  max end          {float(buf[7]):8.0f} cyc  {float(buf[7]) / 1965:6.2f} us")
