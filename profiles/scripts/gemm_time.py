"""Times the decode-step GEMM shapes in isolation (M = 64 rows, 24 distinct weight matrices per shape so that weights
come from HBM, 24 PDL-chained launches captured in one CUDA graph).  Env: PTTS_GEMM=ring|cluster, PTTS_CS=<n>."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from parler_tts_b200 import cabi  # noqa: E402

torch.cuda.set_device(0)
M = int(os.environ.get("M", "64"))
tag = f"gemm={os.environ.get('PTTS_GEMM', 'cluster')} cs={os.environ.get('PTTS_CS', 'auto')}"
for name, N, K, act, res, ydt in (("qkv-like", 3072, 1024, 0, False, torch.bfloat16),
                                  ("out_proj", 1024, 1024, 0, True, torch.float32),
                                  ("fc1", 4096, 1024, 1, False, torch.bfloat16),
                                  ("fc2", 1024, 4096, 0, True, torch.float32),
                                  ("heads", 9792, 1024, 0, False, torch.float32)):
    L = 24
    w = (torch.randn(L, N, K, device="cuda") * 0.02).to(torch.bfloat16)
    x = (torch.randn(M, K, device="cuda") * 0.5).to(torch.bfloat16)
    y = torch.zeros(M, N, device="cuda", dtype=ydt)
    r = y if res else None

    def chain():
        for l in range(L):
            cabi.call("ptts_linear", cabi.ptr(x), 1, cabi.ptr(w[l]), 1, None, M, N, K, 1.0, act, cabi.ptr(r), None,
                      cabi.ptr(y), cabi.dtype_enum(ydt), cabi.IMPL_FAST, cabi.stream_ptr())

    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        chain()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            chain()
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20 / L
    print(f"{tag} {name:9s} N={N} K={K}: {us:6.2f} us/launch  {N * K * 2 / us / 1e3:7.1f} GB/s of weights")
