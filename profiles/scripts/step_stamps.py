"""In-kernel globaltimer stamps of the tcgen05 GEMMs inside the REAL replayed decode-step graph (Mini v0.1, B=64):
prints, per GEMM of a layer, the phase durations and the gap to the previous GEMM's end (= everything in between:
LayerNorm / attention kernels and launch boundaries)."""
import ctypes as C
import os
import sys

os.environ.setdefault("PTTS_GEMM_DBG", "1")  # 2 = SM clock cycles instead of globaltimer ns
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
st = GenState(w, B, S, P, 1, M, suppress_from=1024)
enc = (torch.randn(B, S, 1024, device=dev) * 0.3).to(torch.bfloat16)
prompt = torch.randint(2, 32100, (B, P), device=dev)
st.reset(torch.full((B * K, 1), 1025, dtype=torch.int64, device=dev), seed=0)
eng.cross_kv(st, enc)
eng.prefill(st, model.prompt_table(), prompt)
eng.run_loop(st, use_graph=True, poll_every=64)
torch.cuda.synchronize()
buf = np.zeros(256 * 8, dtype=np.uint64)
n = cabi.load().ptts_diag_gemm_stamps(buf.ctypes.data_as(C.c_void_p), 256)
order = [(n - 1 - i) & 255 for i in range(255)][::-1]
stm = buf.reshape(256, 8)[order].astype(np.int64)
names = ["qkv", "out_proj", "q_cross", "out_cross", "fc1", "fc2"]
# the step ends with the heads GEMM: find it as the launch followed by the largest gap... simply align on period 6 + 1
per_step = 24 * 6 + 1
tail = stm[-per_step:]           # one whole step: 144 layer GEMMs + heads
print("context at the sampled step ~", P + M - 1)
for j, nm in enumerate(names):
    rows = tail[j:144:6]
    d = np.diff(rows[:, :7], axis=1)
    prev_end = stm[-per_step - 1 + j:-1 - (144 - j) + 1:1]
    idx = np.arange(len(stm) - per_step + j, len(stm) - 1, 6)
    gap = stm[idx, 1] - stm[idx - 1, 6]
    print(f"{nm:10s} resident-before-dep {int(np.median(d[:, 0])):6d} | load {int(np.median(d[:, 1])):5d} mma {int(np.median(d[:, 2])):5d} "
          f"tmem/sync {int(np.median(d[:, 3])):5d} epi {int(np.median(d[:, 4])):5d} exit {int(np.median(d[:, 5])):5d} | "
          f"dep-resolved minus previous GEMM end {int(np.median(gap)):6d} ns | total after dep {int(np.median(rows[:, 6] - rows[:, 1])):6d}")
print("layer period (qkv start-to-start) ns:", int(np.median(np.diff(tail[0:144:6, 1]))))
