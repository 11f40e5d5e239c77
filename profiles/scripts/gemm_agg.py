"""Per-CTA timing of the tcgen05 GEMMs inside the replayed decode-step graph (Mini v0.1, B=64), PTTS_GEMM_DBG=3:
for every GEMM launch of the step, mean / max over all CTAs (and replays) of SM-clock cycles since the CTA's
griddepcontrol.wait returned until: W slab landed, X slab landed, MMA done, CTA end."""
import ctypes as C
import os
import sys

os.environ["PTTS_GEMM_DBG"] = "3"
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
buf = np.zeros(256 * 8, dtype=np.uint64)
lib.ptts_diag_gemm_stamps(buf.ctypes.data_as(C.c_void_p), -2)
for _ in range(8):
    st.graph.replay()
torch.cuda.synchronize()
lib.ptts_diag_gemm_stamps(buf.ctypes.data_as(C.c_void_p), -1)
a = buf.reshape(256, 8).astype(np.float64)
nz = np.where(a[:, 4] > 0)[0]
start = [i for i in nz if a[(i - 1) & 255, 4] == 0]
start = start[0] if start else nz[0]
order = [(start + i) & 255 for i in range(256) if a[(start + i) & 255, 4] > 0]
# 6 split-K GEMMs per layer on the unfused chain, 4 when the cross-attention block is one launch (fast_xblock.cu)
per_layer = 4 if (len(order) - 1) % 24 == 0 and (len(order) - 1) // 24 == 4 else 6
names = ["qkv", "out_proj", "fc1", "fc2"] if per_layer == 4 else ["qkv", "out_proj", "q_cross", "out_cross", "fc1", "fc2"]
print(f"context ~{P + N}; {len(order)} GEMM launches per step; cycles after the CTA's dependency wait (mean | max)")
rows = np.array([[a[i, 0] / a[i, 4], a[i, 1] / a[i, 4], a[i, 2] / a[i, 4], a[i, 3] / a[i, 4], a[i, 4] / 8, a[i, 5],
                  a[i, 7], a[i, 6]] for i in order])
for j, nm in enumerate(names):
    r = rows[j:24 * per_layer:per_layer]
    m = np.median(r, axis=0)
    print(f"{nm:10s} CTAs {int(m[4]):4d} | mean: W landed {int(m[0]):6d}  X landed {int(m[1]):6d}  MMA done {int(m[2]):6d}  "
          f"end {int(m[3]):6d} | max: W {int(m[5]):6d}  X {int(m[6]):6d}  end {int(m[7]):6d}")
if len(order) > 24 * per_layer:
    r = rows[24 * per_layer]
    print(f"{'heads':10s} CTAs {int(r[4]):4d} | mean: W landed {int(r[0]):6d}  X landed {int(r[1]):6d}  MMA done {int(r[2]):6d}  "
          f"end {int(r[3]):6d} | max: W {int(r[5]):6d}  X {int(r[6]):6d}  end {int(r[7]):6d}")
