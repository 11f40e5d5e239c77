"""Short program for the `ncu --set full` capture of the dominant kernel: the flash-decode self-attention kernel
(ptts_attention, B=64, 16 heads, d=64, bf16, ctx=466 = the mean context of the 860-frame run) launched over 24
distinct layer caches, twice.  Same arguments as the roofline leg of bench.py.
  python profiles/scripts/attn_ncu.py && ncu --set full --clock-control none --import-source on \
      -k regex:attn_decode -s 24 -c 3 -o gpurun_out/attn_full python profiles/scripts/attn_ncu.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from parler_tts_b200 import cabi  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, heads, d, L, P, M = 64, 16, 64, 24, 32, 869
cmax, ctx = P + M, P + M // 2
kv = torch.empty(L, 2, B, heads, cmax, d, device=dev, dtype=torch.bfloat16).normal_(0, 0.5)
q = (torch.randn(B, heads * d, device=dev) * 0.1).to(torch.bfloat16)
out = torch.empty_like(q)
scratch = torch.zeros(1 << 24, device=dev)
ctx_t = torch.tensor([ctx - P], dtype=torch.int32, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(2):
    e0.record()
    for l in range(L):
        cabi.call("ptts_attention", cabi.ptr(q), cabi.ptr(kv[l, 0]), cabi.ptr(kv[l, 1]), cabi.PTTS_BF16, B, heads, d, 1,
                  cmax, P, cabi.ptr(ctx_t), 0, None, 0, cabi.ptr(out), cabi.ptr(scratch), cabi.IMPL_FAST,
                  cabi.stream_ptr())
    e1.record()
    torch.cuda.synchronize()
    nbytes = B * heads * (2 * ctx * d * 2 + 2 * d * 2)
    us = e0.elapsed_time(e1) * 1e3 / L
    print(f"pass {rep}: {us:.2f} us per launch, {nbytes / us / 1e3:.0f} GB/s algorithmic ({nbytes} B per launch)")
