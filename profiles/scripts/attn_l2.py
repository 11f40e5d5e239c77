"""Is the flash-decode kernel limited by HBM or by itself?  Times ptts_attention (B=64, 16 heads, d=64, bf16) over
(a) 24 distinct caches (every launch streams from HBM) and (b) ONE cache launched back to back (served from L2 when
K+V fit), for several context lengths."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from parler_tts_b200 import cabi  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, heads, d, L, cmax = 64, 16, 64, 24, 512
kv = (torch.randn(L, 2, B, heads, cmax, d, device=dev) * 0.5).to(torch.bfloat16)
q = (torch.randn(B, heads * d, device=dev) * 0.1).to(torch.bfloat16)
out = torch.empty_like(q)
scratch = torch.empty(1 << 24, device=dev)


def run(layers, ctx, reps):
    ctx_t = torch.zeros(1, dtype=torch.int32, device=dev)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        def chain():
            for l in layers:
                cabi.call("ptts_attention", cabi.ptr(q), cabi.ptr(kv[l, 0]), cabi.ptr(kv[l, 1]), cabi.PTTS_BF16, B, heads, d, 1,
                          cmax, ctx, cabi.ptr(ctx_t), 0, None, 0, cabi.ptr(out), cabi.ptr(scratch), cabi.IMPL_FAST,
                          cabi.stream_ptr())
        chain()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            chain()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * len(layers))


for ctx in (64, 118, 236, 472):
    mb = B * heads * 2 * ctx * d * 2 / 1e6
    t_hbm = run(list(range(L)), ctx, 5)
    t_l2 = run([0] * L, ctx, 5)
    print(f"ctx {ctx:4d}  K+V {mb:6.1f} MB | distinct caches {t_hbm:6.2f} us = {mb / t_hbm / 1e3 * 1e3:6.0f} GB/s | "
          f"same cache {t_l2:6.2f} us = {mb / t_l2 / 1e3 * 1e3:6.0f} GB/s")
