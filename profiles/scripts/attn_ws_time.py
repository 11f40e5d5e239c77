"""Isolated timing of the decode-step self-attention kernels over 24 distinct layer caches (nothing served from L2):
persistent warp-specialised kernel (ptts_attention_ws, bf16 / fp8) against the per-item kernels (ptts_attention /
ptts_attention_fp8), B=64, 16 heads, c_stride 901.  python profiles/scripts/attn_ws_time.py [n_old ...]"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from parler_tts_b200 import cabi  # noqa: E402

dev = "cuda"
B, HEADS, D, L, CS = 64, 16, 64, 24, 901
H = HEADS * D
ctxs = [int(a) for a in sys.argv[1:]] or [465]
part = torch.randn(2, B, 3 * H, device=dev) * 0.5
q = torch.randn(B, H, device=dev).to(torch.bfloat16)
out = torch.empty(B, H, device=dev, dtype=torch.bfloat16)
scratch = torch.zeros(1 << 20, device=dev)
kb = torch.randn(L, 2, B, HEADS, CS, D, device=dev).to(torch.bfloat16)
k8 = torch.randint(0, 120, (L, 2, B, HEADS, CS, D), device=dev, dtype=torch.uint8)
sc = torch.full((HEADS,), 0.02, device=dev)
scale = 1 / math.sqrt(D)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / (reps * L) * 1e3  # us per launch


for n_old in ctxs:
    ctx_t = torch.tensor([n_old + 1 - 32], dtype=torch.int32, device=dev)

    def ws(kv):
        c = kb if kv == cabi.PTTS_BF16 else k8
        for l in range(L):
            cabi.call("ptts_attention_ws", cabi.ptr(part), 2, B * 3 * H, 3 * H, scale, n_old, None, cabi.ptr(c[l, 0]),
                      cabi.ptr(c[l, 1]), kv, cabi.ptr(sc), cabi.ptr(sc), B, HEADS, D, CS, None, 0, cabi.ptr(out),
                      cabi.stream_ptr())

    def item_bf16():
        for l in range(L):
            cabi.call("ptts_attention", cabi.ptr(q), cabi.ptr(kb[l, 0]), cabi.ptr(kb[l, 1]), cabi.PTTS_BF16, B, HEADS, D, 1,
                      CS, 32, cabi.ptr(ctx_t), 0, None, 0, cabi.ptr(out), cabi.ptr(scratch), cabi.IMPL_FAST,
                      cabi.stream_ptr())

    def item_fp8():
        for l in range(L):
            cabi.call("ptts_attention_fp8", cabi.ptr(q), cabi.ptr(k8[l, 0]), cabi.ptr(k8[l, 1]), cabi.ptr(sc), cabi.ptr(sc),
                      B, HEADS, D, CS, 32, cabi.ptr(ctx_t), None, 0, cabi.ptr(out), cabi.ptr(scratch), cabi.stream_ptr())

    nbytes = B * HEADS * 2 * (n_old + 1) * D
    for name, fn, es in (("per-item bf16", item_bf16, 2), ("persistent bf16", lambda: ws(cabi.PTTS_BF16), 2),
                         ("per-item fp8", item_fp8, 1), ("persistent fp8", lambda: ws(cabi.PTTS_FP8), 1)):
        us = timed(fn)
        print(f"ctx {n_old + 1:4d}  {name:16s} {us:7.2f} us / launch   {nbytes * es / us / 1e3:7.0f} GB/s")
