"""GPU: ptts_attention_noalloc (static __device__ split-KV scratch, no allocation)
against the existing ptts_attention with caller scratch - same kernels, so the outputs must agree to the tolerance of
the other attention tests (they are expected to be identical) - on two calls in a row, the second with other inputs and
another batch so that the array's stale contents would show.  Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from parler_tts_b200 import cabi  # noqa: E402

DEV = "cuda"
HEADS, D = 16, 64


def _run_both(B, c_stride, ctx, dtype, seed, masked):
    g = torch.Generator(device=DEV).manual_seed(seed)
    q = (torch.randn(B, HEADS * D, device=DEV, generator=g) * 0.3).to(dtype)
    ck = torch.randn(B, HEADS, c_stride, D, device=DEV, generator=g).to(dtype)
    cv = torch.randn(B, HEADS, c_stride, D, device=DEV, generator=g).to(dtype)
    mask = (torch.rand(B, 30, device=DEV, generator=g) > 0.3).to(torch.int32) if masked else None
    ctx_t = torch.tensor([ctx - 7], dtype=torch.int32, device=DEV)
    want = torch.full((B, HEADS * D), float("nan"), device=DEV, dtype=dtype)
    got = torch.full_like(want, float("nan"))
    scratch = torch.zeros(1 << 20, device=DEV)
    cabi.call("ptts_attention", cabi.ptr(q), cabi.ptr(ck), cabi.ptr(cv), cabi.dtype_enum(dtype), B, HEADS, D, 1, c_stride,
              7, cabi.ptr(ctx_t), 0, cabi.ptr(mask), 0 if mask is None else 30, cabi.ptr(want), cabi.ptr(scratch),
              cabi.IMPL_FAST, cabi.stream_ptr())
    cabi.call("ptts_attention_noalloc", cabi.ptr(q), cabi.ptr(ck), cabi.ptr(cv), cabi.dtype_enum(dtype), B, HEADS, D,
              c_stride, 7, cabi.ptr(ctx_t), cabi.ptr(mask), 0 if mask is None else 30, cabi.ptr(got), cabi.stream_ptr())
    torch.cuda.synchronize()
    assert torch.isfinite(want.float()).all() and torch.isfinite(got.float()).all()
    # plain torch fp32 reference of the op, as a guard that both are right and not merely equal
    s = torch.einsum("bhd,bhjd->bhj", q.float().view(B, HEADS, D), ck.float()[:, :, :ctx])
    if mask is not None:
        s[:, :, :30] = torch.where(mask[:, None, :] != 0, s[:, :, :30], torch.finfo(dtype).min)
    ref = torch.einsum("bhj,bhjd->bhd", torch.softmax(s, -1), cv.float()[:, :, :ctx]).reshape(B, HEADS * D)
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-4
    assert float((want.float() - ref).abs().max()) < tol * float(ref.abs().max())
    return want, got


def test_noalloc_attention_equals_caller_scratch_attention_twice_in_a_row():
    lib = cabi.load()
    assert lib.ptts_attention_noalloc_max_floats() == 262144
    # small batches: the context is split over several CTAs per (utterance, head), i.e. the scratch IS used
    for B, c_stride, ctx, dtype, seed, masked in ((3, 700, 650, torch.bfloat16, 1, False),
                                                  (5, 400, 333, torch.float32, 2, True)):
        want, got = _run_both(B, c_stride, ctx, dtype, seed, masked)
        err = float((want.float() - got.float()).abs().max())
        assert err <= 1e-6, err  # same kernels, same order of operations


def test_noalloc_attention_rejects_sizes_beyond_the_static_array():
    B = 3364 // HEADS + 1  # B * heads > 3363
    t = torch.zeros(8, device=DEV, dtype=torch.bfloat16)
    rc = cabi.load().ptts_attention_noalloc(cabi.ptr(t), cabi.ptr(t), cabi.ptr(t), cabi.PTTS_BF16, B, HEADS, D, 64, 1,
                                            None, None, 0, cabi.ptr(t), cabi.stream_ptr())
    assert rc == 1  # PTTS_E_ARG, nothing launched
    assert b"static array" in cabi.load().ptts_last_error_string()
