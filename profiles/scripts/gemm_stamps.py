"""Where does the time go inside the tcgen05 decode GEMM?  PTTS_GEMM_DBG=1 makes CTA 0 of every launch record
globaltimer stamps; this prints per-phase deltas for a PDL chain of launches of one shape in a CUDA graph."""
import ctypes as C
import os
import sys

os.environ["PTTS_GEMM_DBG"] = "1"
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from parler_tts_b200 import cabi  # noqa: E402

torch.cuda.set_device(0)
M, L = 64, 24
for name, N, K, res in (("out_proj", 1024, 1024, True), ("fc1", 4096, 1024, False), ("fc2", 1024, 4096, True)):
    w = (torch.randn(L, N, K, device="cuda") * 0.02).to(torch.bfloat16)
    x = (torch.randn(M, K, device="cuda") * 0.5).to(torch.bfloat16)
    y = torch.zeros(M, N, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        def chain():
            for l in range(L):
                cabi.call("ptts_linear", cabi.ptr(x), 1, cabi.ptr(w[l]), 1, None, M, N, K, 1.0, 0,
                          cabi.ptr(y) if res else None, None, cabi.ptr(y), 0, cabi.IMPL_FAST, cabi.stream_ptr())
        chain()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            chain()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
    buf = np.zeros(256 * 8, dtype=np.uint64)
    n = cabi.load().ptts_diag_gemm_stamps(buf.ctypes.data_as(C.c_void_p), 256)
    last = [(n - 1 - i) & 255 for i in range(L)][::-1]  # the last replay's launches, in order
    st = buf.reshape(256, 8)[last].astype(np.int64)
    d = np.diff(st[:, :7], axis=1)
    print(name, "per-phase ns (median over 24 launches): wait_pdl, load_done, mma_done, tmem+csync, epilogue, csync2 =",
          np.median(d, axis=0).astype(int).tolist(), " start-to-start:", int(np.median(np.diff(st[:, 0]))),
          " prev-end-to-this-pdl-return:", int(np.median(st[1:, 1] - st[:-1, 6])))
