"""GPU parity of the fast kernel family (weight-streaming bf16 GEMM with fused epilogues, flash-decoding attention,
LayerNorm with folded split-K reduction) against torch fp32 references of the same ops and against the CPU oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import decoder_oracle  # noqa: E402
from oracle.gen_golden import TINY  # noqa: E402

import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.engine import DecoderEngine, GenState  # noqa: E402

DEV = "cuda"


def _linear(x, w, bias=None, alpha=1.0, act=0, residual=None, row_mask=None, ydtype=torch.float32,
            impl=cabi.IMPL_FAST):
    M, K = x.shape
    N = w.shape[0]
    y = residual if (residual is not None and ydtype == torch.float32) else torch.empty(M, N, device=DEV, dtype=ydtype)
    cabi.call("ptts_linear", cabi.ptr(x), cabi.dtype_enum(x.dtype), cabi.ptr(w), cabi.dtype_enum(w.dtype),
              cabi.ptr(bias), M, N, K, float(alpha), act, cabi.ptr(residual), cabi.ptr(row_mask), cabi.ptr(y),
              cabi.dtype_enum(ydtype), impl, cabi.stream_ptr())
    return y


@pytest.mark.parametrize("M,N,K", [(64, 1024, 1024), (64, 3072, 1024), (1, 1024, 4096), (37, 9792, 1024),
                                   (200, 384, 256), (64, 4096, 1024), (300, 2048, 768), (5, 26, 128)])
def test_skinny_gemm_vs_torch_fp32(M, N, K):
    g = torch.Generator().manual_seed(M * 7 + N)
    x = (torch.randn(M, K, generator=g) * 0.5).to(DEV, torch.bfloat16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(DEV, torch.bfloat16)
    ref = x.float() @ w.float().t()
    got = _linear(x, w)
    tol = 2e-3 * float(ref.abs().max())
    assert float((got - ref).abs().max()) < tol
    # fused epilogue: alpha, bias, gelu, row mask, residual (in place), bf16 output
    bias = torch.randn(N, generator=g).to(DEV)
    rm = (torch.rand(M, generator=g) > 0.3).to(DEV, torch.int32)
    res = torch.randn(M, N, generator=g).to(DEV)
    want = torch.nn.functional.gelu(0.125 * ref + bias) * rm[:, None] + res
    got = _linear(x, w, bias=bias, alpha=0.125, act=cabi.ACT["gelu"], residual=res.clone(), row_mask=rm)
    assert float((got - want).abs().max()) < 2e-3 * float(want.abs().max()) + 1e-3
    got16 = _linear(x, w, bias=bias, act=cabi.ACT["gelu"], ydtype=torch.bfloat16)
    want16 = torch.nn.functional.gelu(ref + bias)
    assert float((got16.float() - want16).abs().max()) < 1e-2 * float(want16.abs().max())
    # the simple family on the same inputs (validates the dispatcher fallback too)
    simple = _linear(x, w, impl=cabi.IMPL_SIMPLE)
    assert float((simple - ref).abs().max()) < tol


@pytest.mark.parametrize("M,N,K", [(2112, 3072, 1024), (3072, 2304, 768), (3072, 768, 2048), (256, 130, 64),
                                   (1000, 1024, 4096), (257, 9792, 1024)])
def test_large_m_tile_gemm_vs_torch_fp32(M, N, K):
    """fast_gemm_big.cu (M >= 256: prompt prefill, cross K/V, T5 encoder shapes; ragged last tiles in M and N): persistent
    128 x 128 tcgen05 tiles with the fused STORE epilogue."""
    g = torch.Generator().manual_seed(M + 3 * N + K)
    x = (torch.randn(M, K, generator=g) * 0.5).to(DEV, torch.bfloat16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(DEV, torch.bfloat16)
    ref = x.float() @ w.float().t()
    got = _linear(x, w)
    assert float((got - ref).abs().max()) < 2e-3 * float(ref.abs().max())
    bias = torch.randn(N, generator=g).to(DEV)
    rm = (torch.rand(M, generator=g) > 0.3).to(DEV, torch.int32)
    res = torch.randn(M, N, generator=g).to(DEV)
    want = torch.nn.functional.gelu(0.125 * ref + bias) * rm[:, None] + res
    got = _linear(x, w, bias=bias, alpha=0.125, act=cabi.ACT["gelu"], residual=res.clone(), row_mask=rm)
    assert float((got - want).abs().max()) < 2e-3 * float(want.abs().max()) + 1e-3
    got16 = _linear(x, w, bias=bias, ydtype=torch.bfloat16)
    assert float((got16.float() - (ref + bias)).abs().max()) < 1e-2 * float((ref + bias).abs().max())
    # twice in a row (persistent CTAs, barrier phases start from scratch in every launch)
    again = _linear(x, w)
    assert float((again - ref).abs().max()) < 2e-3 * float(ref.abs().max())


@pytest.mark.parametrize("M,N,K", [(64, 4096, 1024), (64, 256, 512), (128, 1024, 768), (64, 64, 256), (32, 128, 1024),
                                   (96, 192, 256)])
def test_activation_gemm_full_k_vs_torch_fp32(M, N, K):
    """fc1 shape family of fast_gemm_act.cu (tcgen05 M = 64 weight rows x N = 32 batch rows per CTA over the whole K,
    activation in the epilogue): bf16 out = act(x . w^T), no bias / residual."""
    g = torch.Generator().manual_seed(M + N + K)
    x = (torch.randn(M, K, generator=g) * 0.5).to(DEV, torch.bfloat16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(DEV, torch.bfloat16)
    ref = x.float() @ w.float().t()
    for name, fn in (("gelu", torch.nn.functional.gelu), ("relu", torch.relu)):
        want = fn(ref)
        got = _linear(x, w, act=cabi.ACT[name], ydtype=torch.bfloat16).float()
        assert float((got - want).abs().max()) < 1e-2 * float(want.abs().max()), name
        simple = _linear(x, w, act=cabi.ACT[name], ydtype=torch.bfloat16, impl=cabi.IMPL_SIMPLE).float()
        assert float((got - simple).abs().max()) < 1e-2 * float(want.abs().max()), name
    # every element is written exactly once: a second call into a poisoned buffer gives the same result
    y = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    cabi.call("ptts_linear", cabi.ptr(x), cabi.dtype_enum(x.dtype), cabi.ptr(w), cabi.dtype_enum(w.dtype), None, M, N, K,
              1.0, cabi.ACT["gelu"], None, None, cabi.ptr(y), cabi.dtype_enum(torch.bfloat16), cabi.IMPL_FAST,
              cabi.stream_ptr())
    assert not bool(torch.isnan(y.float()).any())
    assert float((y.float() - torch.nn.functional.gelu(ref)).abs().max()) < 1e-2 * float(ref.abs().max())


@pytest.mark.parametrize("kw", [dict(do_sample=False), dict(do_sample=True, top_k=20, temperature=0.9, seed=7),
                                dict(do_sample=False, guidance_scale=2.0)])
def test_fused_step_heads_partials_equal_materialised_logits(kw):
    """Captured decode step at B=64: the 9 lm_heads run as split-K partial sums that the epilogue kernel folds while it
    reads them; the step-at-a-time host loop (stopping criteria present) materialises the logits matrix with the STORE
    GEMM and runs the same epilogue on it.  Same K slices, same summation order: the tokens must be identical."""
    kw2 = dict(vocab_size=1088, hidden_size=1024, num_hidden_layers=2, num_attention_heads=16, ffn_dim=4096,
               num_codebooks=9, max_position_embeddings=128, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
    dims = decoder_oracle.DecoderDims(**kw2)
    sd = decoder_oracle.random_decoder_state(dims, seed=5)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw2), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    g = torch.Generator().manual_seed(1)
    B = 64
    nb = 2 * B if "guidance_scale" in kw else B
    enc = (torch.randn(nb, 12, 1024, generator=g) * 0.3).to(DEV)
    prm = (torch.randn(B, 6, 1024, generator=g) * 0.02).to(DEV)
    procs = [ptts.SuppressTokensAtOrAbove(1024)]
    fused = m.generate(encoder_hidden_states=enc, prompt_hidden_states=prm, max_length=40, logits_processor=procs, **kw)
    never = [lambda ids, scores: False]
    stepwise = m.generate(encoder_hidden_states=enc, prompt_hidden_states=prm, max_length=40, logits_processor=procs,
                          stopping_criteria=never, **kw)
    assert fused.shape == stepwise.shape == (B, 9, 40 - 9)
    assert torch.equal(fused, stepwise)


def _attention(q, ck, cv, ctx, key_mask, impl):
    B, heads, c_stride, d = ck.shape
    out = torch.empty(B, heads * d, device=DEV, dtype=q.dtype)
    scratch = torch.zeros(1 << 20, device=DEV, dtype=torch.float32)
    ctx_t = torch.tensor([ctx - 3], dtype=torch.int32, device=DEV)
    cabi.call("ptts_attention", cabi.ptr(q), cabi.ptr(ck), cabi.ptr(cv), cabi.dtype_enum(q.dtype), B, heads, d, 1,
              c_stride, 3, cabi.ptr(ctx_t), 0, cabi.ptr(key_mask), 0 if key_mask is None else key_mask.shape[1],
              cabi.ptr(out), cabi.ptr(scratch), impl, cabi.stream_ptr())
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,heads,Tq,c_stride,ctx,causal,masked", [(64, 16, 33, 901, 33, 1, False), (3, 4, 33, 48, 48, 0, True),
                                                                  (2, 2, 7, 300, 257, 1, True), (1, 1, 2, 9, 2, 1, False)])
def test_prefill_attention_staged_vs_torch_fp32(dtype, B, heads, Tq, c_stride, ctx, causal, masked):
    """fast_attn_prefill.cu (Tq > 1: prompt prefill self-attention with the causal mask, cross-attention over the cached T5
    keys): K / V of a (head, utterance) pair staged in shared memory once for all its queries; masked keys take
    finfo(dtype).min like the reference's additive mask.  Also checked against the simple family on the same inputs."""
    g = torch.Generator().manual_seed(Tq * 131 + ctx)
    d = 64
    q = (torch.randn(B, Tq, heads * d, generator=g) * 0.3).to(DEV, dtype)
    ck = torch.randn(B, heads, c_stride, d, generator=g).to(DEV, dtype)
    cv = torch.randn(B, heads, c_stride, d, generator=g).to(DEV, dtype)
    km = None
    if masked:
        km = (torch.rand(B, ctx, generator=g) > 0.25).to(torch.int32)
        km[:, -1] = 1
        km = km.to(DEV)
    outs = {}
    for impl in (cabi.IMPL_FAST, cabi.IMPL_SIMPLE):
        out = torch.full((B, Tq, heads * d), float("nan"), device=DEV, dtype=dtype)
        cabi.call("ptts_attention", cabi.ptr(q), cabi.ptr(ck), cabi.ptr(cv), cabi.dtype_enum(dtype), B, heads, d, Tq,
                  c_stride, ctx, None, causal, cabi.ptr(km), 0 if km is None else ctx, cabi.ptr(out), None, impl,
                  cabi.stream_ptr())
        outs[impl] = out.float()
    qf = q.float().view(B, Tq, heads, d).transpose(1, 2)                      # (B, heads, Tq, d)
    sc = qf @ ck.float()[:, :, :ctx].transpose(-1, -2)                        # (B, heads, Tq, ctx)
    neg = torch.finfo(dtype).min
    keep = torch.ones(B, 1, Tq, ctx, dtype=torch.bool, device=DEV)
    if causal:
        t = torch.arange(Tq, device=DEV)[:, None]
        j = torch.arange(ctx, device=DEV)[None, :]
        keep = keep & (j < ctx - Tq + t + 1)[None, None]
    if km is not None:
        keep = keep & (km[:, None, None, :] != 0)
    sc = torch.where(keep, sc, torch.full_like(sc, neg))
    want = (torch.softmax(sc, dim=-1) @ cv.float()[:, :, :ctx]).transpose(1, 2).reshape(B, Tq, heads * d)
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    assert float((outs[cabi.IMPL_FAST] - want).abs().max()) < tol * max(1.0, float(want.abs().max()))
    assert float((outs[cabi.IMPL_FAST] - outs[cabi.IMPL_SIMPLE]).abs().max()) < tol * max(1.0, float(want.abs().max()))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,heads,c_stride,ctx", [(64, 16, 901, 466), (2, 16, 2700, 2609), (1, 2, 64, 1),
                                                  (3, 4, 130, 129), (16, 16, 48, 48)])
def test_flash_decode_attention_vs_torch_fp32(dtype, B, heads, c_stride, ctx):
    g = torch.Generator().manual_seed(B + ctx)
    d = 64
    q = (torch.randn(B, heads * d, generator=g) * 0.3).to(DEV, dtype)
    ck = torch.randn(B, heads, c_stride, d, generator=g).to(DEV, dtype)
    cv = torch.randn(B, heads, c_stride, d, generator=g).to(DEV, dtype)
    mlen = min(ctx, 7)
    km = (torch.rand(B, mlen, generator=g) > 0.4).to(DEV, torch.int32)
    km[0] = 0  # a row whose masked prefix is fully masked
    for mask in (None, km):
        s = torch.einsum("bhd,bhjd->bhj", q.float().view(B, heads, d), ck.float()[:, :, :ctx])
        if mask is not None:
            neg = torch.finfo(dtype).min
            s[:, :, :mlen] = torch.where(mask[:, None, :].bool(), s[:, :, :mlen], torch.full_like(s[:, :, :mlen], neg))
        p = torch.softmax(s, dim=-1)
        ref = torch.einsum("bhj,bhjd->bhd", p, cv.float()[:, :, :ctx]).reshape(B, heads * d)
        fast = _attention(q, ck, cv, ctx, mask, cabi.IMPL_FAST).float()
        simple = _attention(q, ck, cv, ctx, mask, cabi.IMPL_SIMPLE).float()
        tol = 2e-5 if dtype == torch.float32 else 1e-2
        assert float((fast - ref).abs().max()) < tol, (B, ctx, mask is not None)
        assert float((simple - ref).abs().max()) < tol


def test_layernorm_fast_vs_torch():
    g = torch.Generator().manual_seed(2)
    for rows, H in ((64, 1024), (3, 1536), (5, 128)):
        x = (torch.randn(rows, H, generator=g) * 2 + 0.3).to(DEV)
        gam, bet = torch.randn(H, generator=g).to(DEV), torch.randn(H, generator=g).to(DEV)
        ref = torch.nn.functional.layer_norm(x, (H,), gam, bet, 1e-5)
        for dt in (torch.float32, torch.bfloat16):
            y = torch.empty(rows, H, device=DEV, dtype=dt)
            cabi.call("ptts_layernorm", cabi.ptr(x), cabi.ptr(gam), cabi.ptr(bet), rows, H, 1e-5, cabi.ptr(y),
                      cabi.dtype_enum(dt), cabi.stream_ptr())
            tol = 2e-5 if dt == torch.float32 else 2e-2
            assert float((y.float() - ref).abs().max()) < tol * max(1.0, float(ref.abs().max()))


def _teacher_forced_decode(m, dims, ids, enc, prm, steps_from=1, enc_mask=None):
    """prefill on ids[:, :1] then cached decode steps that are forced to follow `ids`; returns logits per position."""
    w = m.packed_weights()
    K = dims.num_codebooks
    B, T = ids.shape[0] // K, ids.shape[1]
    st = GenState(w, B, enc.shape[1], prm.shape[1], 1, T + 1, impl=m.impl, enc_mask=enc_mask)
    st.mask.fill_(-1)
    st.ids.zero_()
    st.ids[:, :1] = ids[:, :1].to(DEV)
    st.ctrl.copy_(torch.tensor([1, 0, 0, B * K, 0, 0, 0, 0], dtype=torch.int32))
    eng = DecoderEngine(w)
    eng.cross_kv(st, enc.to(DEV))
    eng.prefill(st, prompt_hidden=prm.to(DEV, torch.float32).contiguous(), run_epilogue=False)
    out = []
    for t in range(1, T):
        forced = torch.full_like(st.logits, -1e9)
        forced.scatter_(1, ids[:, t : t + 1].to(DEV), 0.0)
        out.append(st.logits.clone())
        st.logits.copy_(forced)
        eng.epilogue(st)  # greedy argmax of the forced scores writes ids[:, t] and the next hidden state
        eng.step(st, run_epilogue=False)
    out.append(st.logits.clone())
    assert torch.equal(st.ids[:, :T].cpu(), ids)
    return torch.stack(out, dim=1)  # (B*K, T, V)


@pytest.mark.parametrize("impl", [cabi.IMPL_SIMPLE, cabi.IMPL_FAST])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_cached_decode_steps_teacher_forced_vs_oracle(impl, dtype):
    """Every cached decode step (static KV cache write in place + flash-decoding + fused epilogues) reproduces the
    oracle's full-sequence forward: 1e-3 absolute in fp32, 1e-2 of the logit scale in bf16 (BASELINE.json)."""
    kw = dict(TINY, hidden_size=256, num_attention_heads=4, ffn_dim=1024, num_hidden_layers=3, vocab_size=1088,
              num_codebooks=9, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=5, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV)
    if dtype != torch.float32:
        m = m.to(dtype)
    m.impl = impl
    g = torch.Generator().manual_seed(3)
    B, S, P, T = 5, 11, 6, 9
    enc, prm = torch.randn(B, S, 256, generator=g), torch.randn(B, P, 256, generator=g) * 0.5
    ids = torch.randint(0, 1024, (B * 9, T), generator=g)
    if dtype == torch.float32:
        want, _ = decoder_oracle.decoder_forward(sd, dims, ids, enc, None, prm, None, None)
    else:
        sd_bf = {k: (v.to(torch.bfloat16).float() if v.dim() == 2 and "embed" not in k else v) for k, v in sd.items()}
        want, _ = decoder_oracle.decoder_forward(sd_bf, dims, ids, enc.to(torch.bfloat16).float(), None, prm, None,
                                                 None)
    got = _teacher_forced_decode(m, dims, ids, enc.to(dtype), prm).cpu()
    want = want[:, P:, :]
    err = float((got - want).abs().max())
    if dtype == torch.float32:
        assert err < 1e-3, err
    else:
        assert err / float(want.abs().max()) < 1e-2, err


@pytest.mark.parametrize("B,H,heads,L,P,T", [(3, 256, 4, 2, 70, 340), (32, 1024, 16, 1, 40, 150)])
def test_fused_attention_prologue_long_context_vs_oracle(B, H, heads, L, P, T):
    """The fused flash-decode path (split-K q/k/v partials folded in the attention prologue, in-place cache append,
    pre-wait K staging) over a context of several hundred tokens: every step's bf16 logits stay within 1e-2 of the
    logit scale of the fp32 oracle's full-sequence forward."""
    # (3, 4 heads): split-KV + combine kernel; (32, 16 heads) = 512 (utterance, head) pairs: one CTA per pair
    kw = dict(TINY, hidden_size=H, num_attention_heads=heads, ffn_dim=1024, num_hidden_layers=L, vocab_size=1088,
              num_codebooks=9, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025, max_position_embeddings=512)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=8, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    g = torch.Generator().manual_seed(9)
    S = 11
    enc, prm = torch.randn(B, S, H, generator=g), torch.randn(B, P, H, generator=g) * 0.5
    ids = torch.randint(0, 1024, (B * 9, T), generator=g)
    sd_bf = {k: (v.to(torch.bfloat16).float() if v.dim() == 2 and "embed" not in k else v) for k, v in sd.items()}
    want, _ = decoder_oracle.decoder_forward(sd_bf, dims, ids, enc.to(torch.bfloat16).float(), None, prm, None, None)
    got = _teacher_forced_decode(m, dims, ids, enc.to(torch.bfloat16), prm).cpu()
    want = want[:, P:, :]
    err = (got - want).abs().amax(dim=(0, 2)) / float(want.abs().max())
    assert float(err.max()) < 1e-2, (int(err.argmax()), float(err.max()))


@pytest.mark.parametrize("H,heads,F,B", [(1536, 24, 6144, 32), (1024, 16, 4096, 64)])
def test_scaled_and_mini_layer_shapes_bf16_vs_oracle(H, heads, F, B):
    """BASELINE.json configs[3] (36 x 1536 decoder, 24 heads, ffn 6144, bs=32) and configs[1] (Mini, bs=64) layer
    shapes on the tcgen05 split-K path, two layers deep: bf16 logits within 1e-2 of the logit scale of the fp32
    oracle on bf16-rounded weights."""
    kw = dict(TINY, hidden_size=H, num_attention_heads=heads, ffn_dim=F, num_hidden_layers=2, vocab_size=1088,
              num_codebooks=9, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025, max_position_embeddings=128)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=6, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    g = torch.Generator().manual_seed(4)
    S, P, T = 48, 32, 4
    enc, prm = torch.randn(B, S, H, generator=g) * 0.3, torch.randn(B, P, H, generator=g) * 0.1
    ids = torch.randint(0, 1024, (B * 9, T), generator=g)
    sd_bf = {k: (v.to(torch.bfloat16).float() if v.dim() == 2 and "embed" not in k else v) for k, v in sd.items()}
    want, _ = decoder_oracle.decoder_forward(sd_bf, dims, ids, enc.to(torch.bfloat16).float(), None, prm, None, None)
    n0 = cabi.load().ptts_launch_count()
    got = _teacher_forced_decode(m, dims, ids, enc.to(torch.bfloat16), prm).cpu()
    assert cabi.load().ptts_launch_count() - n0 >= (T - 1) * (2 * 9 + 2)
    want = want[:, P:, :]
    assert float((got - want).abs().max()) / float(want.abs().max()) < 1e-2


@pytest.mark.parametrize("H,heads,B,S", [(1024, 16, 13, 37), (1536, 24, 8, 70), (1024, 16, 1, 3)])
def test_fused_cross_attention_block_masked_ragged_batch_vs_oracle(H, heads, B, S):
    """The one-launch cross-attention block of the decode step (fast_xblock.cu: fold + LayerNorm, per-head q
    projection, attention over the cached T5 K/V, per-head out_proj partial sums) with padded descriptions
    (encoder attention mask, one utterance with a single valid key) and a batch that is not a multiple of the 8
    utterances a CTA carries: bf16 logits within 1e-2 of the logit scale of the fp32 oracle."""
    kw = dict(TINY, hidden_size=H, num_attention_heads=heads, ffn_dim=1024, num_hidden_layers=2, vocab_size=1088,
              num_codebooks=9, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025, max_position_embeddings=128)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=12, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    g = torch.Generator().manual_seed(13)
    P, T = 5, 5
    lens = torch.randint(1, S + 1, (B,), generator=g)
    lens[0], lens[-1] = S, 1
    enc_mask = (torch.arange(S)[None, :] < lens[:, None]).to(torch.int64)
    enc = torch.randn(B, S, H, generator=g) * 0.5 * enc_mask[..., None]
    prm = torch.randn(B, P, H, generator=g) * 0.1
    ids = torch.randint(0, 1024, (B * 9, T), generator=g)
    sd_bf = {k: (v.to(torch.bfloat16).float() if v.dim() == 2 and "embed" not in k else v) for k, v in sd.items()}
    want, _ = decoder_oracle.decoder_forward(sd_bf, dims, ids, enc.to(torch.bfloat16).float(), enc_mask, prm, None,
                                             None)
    n0 = cabi.load().ptts_launch_count()
    got = _teacher_forced_decode(m, dims, ids, enc.to(torch.bfloat16), prm, enc_mask=enc_mask).cpu()
    launches = cabi.load().ptts_launch_count() - n0
    want = want[:, P:, :]
    assert float((got - want).abs().max()) / float(want.abs().max()) < 1e-2
    # 9 launches per layer and decode step (the unfused chain has 12): the fused block really ran
    per_step = 2 * 9 + 2 + 1
    prefill = launches - (T - 1) * per_step
    assert 0 < prefill < 2 * 20 + 8, (launches, prefill)


def test_fast_path_launches_its_own_kernels_and_graph_replays_match_eager():
    """bf16 fast path: a CUDA-graph replayed generation equals the eager launch sequence token for token, and the
    library's launch counter shows the fused layer (12 launches per layer at this width, where the one-launch
    cross-attention block does not apply, instead of the simple family's 15)."""
    kw = dict(TINY, hidden_size=256, num_attention_heads=4, ffn_dim=1024, num_hidden_layers=2)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=7, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    g = torch.Generator().manual_seed(5)
    enc, prm = torch.randn(3, 7, 256, generator=g).to(DEV), torch.randn(3, 4, 256, generator=g).to(DEV)
    outs = []
    for graph in (False, True):
        m.use_cuda_graph = graph
        outs.append(m.generate(encoder_hidden_states=enc, prompt_hidden_states=prm, max_length=30, do_sample=False,
                               logits_processor=[ptts.SuppressTokensAtOrAbove(32)]).cpu())
    assert torch.equal(outs[0], outs[1])
    st = m._state_cache[1]
    assert st.graph is not None and st.graph_launches == 2 * 12 + 3, st.graph_launches


def test_several_decode_steps_per_graph_replay_equal_single_step_replays():
    """engine.capture_multi: a graph of 4 consecutive steps (the position counter lives on the device) selects the same
    tokens as single-step replays, also when the step budget is not a multiple of 4 (the remainder replays single steps)."""
    from parler_tts_b200.engine import DecoderEngine

    kw = dict(TINY, hidden_size=256, num_attention_heads=4, ffn_dim=1024, num_hidden_layers=2)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=7, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    g = torch.Generator().manual_seed(5)
    enc, prm = torch.randn(3, 7, 256, generator=g).to(DEV), torch.randn(3, 4, 256, generator=g).to(DEV)
    kwargs = dict(encoder_hidden_states=enc, prompt_hidden_states=prm, max_length=31, do_sample=True, top_k=12, seed=4,
                  logits_processor=[ptts.SuppressTokensAtOrAbove(32)])
    want = m.generate(**kwargs).cpu()
    old = DecoderEngine.steps_per_graph
    try:
        DecoderEngine.steps_per_graph = 4
        m.reset_engine()
        got = m.generate(**kwargs).cpu()
        assert m._state_cache[1].graph_multi is not None and m._state_cache[1].steps_per_graph == 4
    finally:
        DecoderEngine.steps_per_graph = old
    assert torch.equal(got, want)


# ------------------------------------------------------------------------------------------------ DAC fast family
def _split(w):
    hi = w.to(torch.bfloat16)
    return hi.contiguous(), (w - hi.float()).to(torch.bfloat16).contiguous()


def _weights(w_tnc):
    """[taps][N][Cin] fp32 -> hi/mid bf16 planes in the pre-tiled stage image of the warp-specialised tcgen05 kernel"""
    hi, mid = _split(w_tnc)
    from parler_tts_b200.dac import DacDecoderEngine
    return DacDecoderEngine._tile(hi, mid)[0]


@pytest.mark.parametrize("Cin,Cout,k,dil,T", [(96, 96, 7, 9, 700), (192, 192, 1, 1, 300), (1024, 1536, 7, 1, 37),
                                              (384, 384, 7, 3, 129), (32, 26, 7, 1, 5), (192, 192, 7, 9, 1000),
                                              (96, 96, 1, 1, 257), (64, 40, 7, 3, 256)])
def test_dac_conv1d_nwc_vs_torch(Cin, Cout, k, dil, T):
    if Cout % 16:
        pytest.skip("pre-tiled kernel works on 16-column groups")
    g = torch.Generator().manual_seed(Cin + k + dil)
    B = 2
    x = torch.randn(B, Cin, T, generator=g).to(DEV)
    w = (torch.randn(Cout, Cin, k, generator=g) / (Cin * k) ** 0.5).to(DEV)
    bias = torch.randn(Cout, generator=g).to(DEV)
    alpha = (1.0 + 0.2 * torch.randn(Cin, generator=g)).abs().add(0.05).to(DEV)
    res = torch.randn(B, Cout, T, generator=g).to(DEV)
    xs = x + (1.0 / (alpha[None, :, None] + 1e-9)) * torch.sin(alpha[None, :, None] * x) ** 2
    want = torch.nn.functional.conv1d(xs.double(), w.double(), bias.double(), dilation=dil,
                                      padding=dil * (k - 1) // 2).float() + res
    wt = _weights(w.permute(2, 0, 1).contiguous())
    x_nwc, res_nwc = x.transpose(1, 2).contiguous(), res.transpose(1, 2).contiguous()
    y = torch.empty(B, T, Cout, device=DEV)
    cabi.call("ptts_dac_conv1d_px", cabi.ptr(x_nwc), None, B, T, Cin, cabi.ptr(wt), cabi.ptr(bias), Cout, k, dil,
              cabi.ptr(alpha), cabi.ptr(res_nwc), cabi.ptr(y), None, None, cabi.stream_ptr())
    err = float((y.transpose(1, 2) - want).abs().max())
    assert err < 2e-4 * max(1.0, float(want.abs().max())), err


@pytest.mark.parametrize("Cin,Cout,s,T", [(1536, 768, 8, 11), (192, 96, 2, 300), (64, 32, 4, 130), (384, 192, 4, 515)])
def test_dac_convtr1d_nwc_vs_torch(Cin, Cout, s, T):
    g = torch.Generator().manual_seed(Cin + s)
    B = 2
    x = torch.randn(B, Cin, T, generator=g).to(DEV)
    w = (torch.randn(Cin, Cout, 2 * s, generator=g) / (Cin * 2) ** 0.5).to(DEV)
    bias = torch.randn(Cout, generator=g).to(DEV)
    alpha = (1.0 + 0.2 * torch.randn(Cin, generator=g)).abs().add(0.05).to(DEV)
    xs = x + (1.0 / (alpha[None, :, None] + 1e-9)) * torch.sin(alpha[None, :, None] * x) ** 2
    want = torch.nn.functional.conv_transpose1d(xs.double(), w.double(), bias.double(), stride=s,
                                                padding=(s + 1) // 2).float()
    assert want.shape[-1] == T * s
    wt = _weights(w.permute(2, 1, 0).reshape(2, s * Cout, Cin).contiguous())
    x_nwc = x.transpose(1, 2).contiguous()
    y = torch.empty(B, T * s, Cout, device=DEV)
    cabi.call("ptts_dac_convtr1d_px", cabi.ptr(x_nwc), None, B, T, Cin, cabi.ptr(wt), cabi.ptr(bias), Cout, s,
              cabi.ptr(alpha), cabi.ptr(y), cabi.stream_ptr())
    err = float((y.transpose(1, 2) - want).abs().max())
    assert err < 2e-4 * max(1.0, float(want.abs().max())), err


def _to_planes(x_nct, alpha):
    """reference construction of activated hi/mid planes [B][2][C/8][8 + T + 264][8] from (B, C, T) fp32"""
    B, Cc, T = x_nct.shape
    xs = x_nct if alpha is None else x_nct + (1.0 / (alpha[None, :, None] + 1e-9)) * torch.sin(alpha[None, :, None] * x_nct) ** 2
    hi = xs.to(torch.bfloat16)
    mid = (xs - hi.float()).to(torch.bfloat16)
    p = torch.zeros(B, 2, Cc // 8, 8 + T + 264, 8, dtype=torch.bfloat16, device=x_nct.device)
    for i, pl in enumerate((hi, mid)):
        p[:, i, :, 8 : 8 + T] = pl.view(B, Cc // 8, 8, T).permute(0, 1, 3, 2)
    return p.contiguous(), xs


@pytest.mark.parametrize("Cin,Cout,k,dil,T", [(96, 96, 1, 1, 700), (192, 192, 1, 1, 513), (384, 384, 7, 1, 300)])
def test_dac_conv1d_planes_in_and_out(Cin, Cout, k, dil, T):
    """conv1d reading activated planes and writing planes for its consumer == torch fp64 on the same Snake'd input."""
    from parler_tts_b200.dac import DacDecoderEngine
    g = torch.Generator().manual_seed(Cin + k + T)
    B = 2
    x = torch.randn(B, Cin, T, generator=g).to(DEV)
    w = (torch.randn(Cout, Cin, k, generator=g) / (Cin * k) ** 0.5).to(DEV)
    bias = torch.randn(Cout, generator=g).to(DEV)
    a_in = (1.0 + 0.2 * torch.randn(Cin, generator=g)).abs().add(0.05).to(DEV)
    a_out = (1.0 + 0.2 * torch.randn(Cout, generator=g)).abs().add(0.05).to(DEV)
    res = torch.randn(B, Cout, T, generator=g).to(DEV)
    xp, xs = _to_planes(x, a_in)
    want = torch.nn.functional.conv1d(xs.double(), w.double(), bias.double(), dilation=dil,
                                      padding=dil * (k - 1) // 2).float() + res
    wt, _ = DacDecoderEngine._tile(*_split(w.permute(2, 0, 1).contiguous()))
    res_nwc = res.transpose(1, 2).contiguous()
    y = torch.empty(B, T, Cout, device=DEV)
    yp = torch.zeros(B, 2, Cout // 8, 8 + T + 264, 8, dtype=torch.bfloat16, device=DEV)
    cabi.call("ptts_dac_conv1d_px", None, cabi.ptr(xp), B, T, Cin, cabi.ptr(wt), cabi.ptr(bias), Cout, k, dil, None,
              cabi.ptr(res_nwc), cabi.ptr(y), cabi.ptr(yp), cabi.ptr(a_out), cabi.stream_ptr())
    tol = 2e-4 * max(1.0, float(want.abs().max()))
    assert float((y.transpose(1, 2) - want).abs().max()) < tol
    want_p, want_s = _to_planes(want, a_out)
    got = (yp[:, 0].float() + yp[:, 1].float())[:, :, 8 : 8 + T].permute(0, 1, 3, 2).reshape(B, Cout, T)
    assert float((got - want_s).abs().max()) < 2 * tol
    assert float(yp[:, :, :, :8].float().abs().max()) == 0 and float(yp[:, :, :, 8 + T :].float().abs().max()) == 0


@pytest.mark.parametrize("Cin,Cout,s,T", [(384, 192, 4, 515), (1536, 768, 8, 11)])
def test_dac_convtr1d_planes_in(Cin, Cout, s, T):
    from parler_tts_b200.dac import DacDecoderEngine
    g = torch.Generator().manual_seed(Cin + s + T)
    B = 2
    x = torch.randn(B, Cin, T, generator=g).to(DEV)
    w = (torch.randn(Cin, Cout, 2 * s, generator=g) / (Cin * 2) ** 0.5).to(DEV)
    bias = torch.randn(Cout, generator=g).to(DEV)
    alpha = (1.0 + 0.2 * torch.randn(Cin, generator=g)).abs().add(0.05).to(DEV)
    xp, xs = _to_planes(x, alpha)
    want = torch.nn.functional.conv_transpose1d(xs.double(), w.double(), bias.double(), stride=s,
                                                padding=(s + 1) // 2).float()
    wt, _ = DacDecoderEngine._tile(*_split(w.permute(2, 1, 0).reshape(2, s * Cout, Cin).contiguous()))
    y = torch.empty(B, T * s, Cout, device=DEV)
    cabi.call("ptts_dac_convtr1d_px", None, cabi.ptr(xp), B, T, Cin, cabi.ptr(wt), cabi.ptr(bias), Cout, s, None,
              cabi.ptr(y), cabi.stream_ptr())
    err = float((y.transpose(1, 2) - want).abs().max())
    assert err < 2e-4 * max(1.0, float(want.abs().max())), err


def test_dac_fast_equals_simple_on_ragged_batch():
    """Channels-last tensor-core DAC == fp32 CUDA-core DAC (1e-3) through the batched and the per-sample path."""
    from oracle import dac_oracle
    from parler_tts_b200.dac import DacDecoderEngine

    sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=4)
    fast, simple = DacDecoderEngine(sd, DEV, impl=cabi.IMPL_FAST), DacDecoderEngine(sd, DEV, impl=cabi.IMPL_SIMPLE)
    assert fast.nwc and not simple.nwc
    g = torch.Generator().manual_seed(1)
    codes = torch.randint(0, 1024, (3, 9, 40), generator=g).to(DEV)
    a, b = fast.decode(codes), simple.decode(codes)
    assert a.shape == b.shape == (3, 1, 512 * 40)
    assert float((a - b).abs().max()) < 1e-3
    codes[1, :, 30:] = 1024
    a, b = fast.decode_generated(codes, 1025, 1024, 1024), simple.decode_generated(codes, 1025, 1024, 1024)
    assert a.shape == b.shape and float((a - b).abs().max()) < 1e-3
