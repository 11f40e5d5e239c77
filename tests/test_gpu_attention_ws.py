"""GPU: the warp-specialised persistent flash-decode kernel (fast_attn_ws.cu: TMA-staged tiles, both products on
tcgen05) against a plain torch fp32 reference of the same op - fused prologue (split-K partial sums -> q, new K / V row
appended), bf16 and fp8 caches, contexts around the 128-row tile boundaries, prompt masks, uneven items per CTA."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from parler_tts_b200 import cabi  # noqa: E402

DEV = "cuda"
HEADS, D = 16, 64
H = HEADS * D


def _reference(q, k_old, v_old, k_new, v_new, mask, n_old):
    """q (B, heads, d) scaled; k_old / v_old (B, heads, n_old, d); new row (B, heads, d); mask (B, mask_len) or None."""
    k = torch.cat([k_old, k_new[:, :, None]], dim=2)
    v = torch.cat([v_old, v_new[:, :, None]], dim=2)
    s = torch.einsum("bhd,bhjd->bhj", q, k)
    if mask is not None:
        ml = min(mask.shape[1], n_old)
        s[:, :, :ml] = torch.where(mask[:, None, :ml] != 0, s[:, :, :ml], torch.finfo(torch.bfloat16).min)
    return torch.einsum("bhj,bhjd->bhd", torch.softmax(s, dim=-1), v).reshape(q.shape[0], H)


@pytest.mark.parametrize("kv", ["bf16", "fp8"])
@pytest.mark.parametrize("B,c_stride,n_old,masked", [
    (64, 901, 465, False), (64, 901, 128, False), (64, 901, 127, False), (64, 901, 129, False), (64, 901, 0, False),
    (64, 901, 1, False), (64, 901, 900, False), (20, 300, 290, True), (37, 515, 256, False)])
def test_attention_ws_vs_torch(kv, B, c_stride, n_old, masked):
    g = torch.Generator(device=DEV).manual_seed(n_old * 7 + B)
    S = 2
    part = torch.randn(S, B, 3 * H, device=DEV, generator=g) * 0.7
    head_gain = torch.linspace(0.4, 2.5, HEADS, device=DEV).view(1, HEADS, 1, 1)
    k_cache = torch.randn(B, HEADS, c_stride, D, device=DEV, generator=g) * head_gain
    v_cache = torch.randn(B, HEADS, c_stride, D, device=DEV, generator=g) * head_gain
    scale = 1.0 / math.sqrt(D)
    qkv = part.sum(0)
    q = (qkv[:, :H] * scale).view(B, HEADS, D)
    k_new, v_new = qkv[:, H:2 * H].view(B, HEADS, D), qkv[:, 2 * H:].view(B, HEADS, D)
    mask = None
    if masked:
        mask = (torch.rand(B, 40, device=DEV, generator=g) > 0.3).to(torch.int32)
    if kv == "bf16":
        ck, cv = k_cache.to(torch.bfloat16), v_cache.to(torch.bfloat16)
        ks = vs = None
        kq, vq = ck.float(), cv.float()
        want_k, want_v = k_new.to(torch.bfloat16), v_new.to(torch.bfloat16)
        k_new_f, v_new_f = want_k.float(), want_v.float()
        kv_enum = cabi.PTTS_BF16
    else:
        ks = (k_cache.abs().amax(dim=(0, 2, 3)) * 4 / 448).contiguous()
        vs = (v_cache.abs().amax(dim=(0, 2, 3)) * 4 / 448).contiguous()
        ck = (k_cache / ks.view(1, HEADS, 1, 1)).to(torch.float8_e4m3fn).view(torch.uint8).contiguous()
        cv = (v_cache / vs.view(1, HEADS, 1, 1)).to(torch.float8_e4m3fn).view(torch.uint8).contiguous()
        kq = ck.view(torch.float8_e4m3fn).float() * ks.view(1, HEADS, 1, 1)
        vq = cv.view(torch.float8_e4m3fn).float() * vs.view(1, HEADS, 1, 1)
        want_k = (k_new / ks.view(1, HEADS, 1)).to(torch.float8_e4m3fn).view(torch.uint8)
        want_v = (v_new / vs.view(1, HEADS, 1)).to(torch.float8_e4m3fn).view(torch.uint8)
        k_new_f = want_k.view(torch.float8_e4m3fn).float() * ks.view(1, HEADS, 1)
        v_new_f = want_v.view(torch.float8_e4m3fn).float() * vs.view(1, HEADS, 1)
        kv_enum = cabi.PTTS_FP8
    ref = _reference(q, kq[:, :, :n_old], vq[:, :, :n_old], k_new_f, v_new_f, mask, n_old)
    out = torch.full((B, H), float("nan"), device=DEV, dtype=torch.bfloat16)
    cabi.call("ptts_attention_ws", cabi.ptr(part), S, B * 3 * H, 3 * H, scale, n_old, None, cabi.ptr(ck), cabi.ptr(cv),
              kv_enum, cabi.ptr(ks), cabi.ptr(vs), B, HEADS, D, c_stride, cabi.ptr(mask), 0 if mask is None else 40,
              cabi.ptr(out), cabi.stream_ptr())
    torch.cuda.synchronize()
    assert torch.equal(ck[:, :, n_old], want_k) and torch.equal(cv[:, :, n_old], want_v), "appended row"
    err = float((out.float() - ref).abs().max() / ref.abs().max())
    assert err < 8e-3, err


def test_attention_ws_rejects_small_batches():
    part = torch.zeros(1, 2, 3 * H, device=DEV)
    ck = torch.zeros(2, HEADS, 64, D, device=DEV, dtype=torch.bfloat16)
    out = torch.zeros(2, H, device=DEV, dtype=torch.bfloat16)
    rc = cabi.load().ptts_attention_ws(cabi.ptr(part), 1, 2 * 3 * H, 3 * H, 0.125, 3, None, cabi.ptr(ck), cabi.ptr(ck),
                                       cabi.PTTS_BF16, None, None, 2, HEADS, D, 64, None, 0, cabi.ptr(out),
                                       cabi.stream_ptr())
    assert rc == 3  # PTTS_E_UNSUPPORTED: the per-item split-KV kernels serve small batches
