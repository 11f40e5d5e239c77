"""GPU: the generation-config surface of generate() (every field the reference's _get_logits_processor /
_get_logits_warper would honour, modeling_parler_tts.py:2561-2567, 2597), host-side processors together with guidance,
streamer / stopping criteria / num_return_sequences, shard independence, forward() with the static cache, and the fp8
KV-cache kernels - against the CPU oracle and plain torch references."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import decoder_oracle, generate_oracle  # noqa: E402
from oracle.gen_golden import TINY  # noqa: E402

import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402

DEV = "cuda"


def _tiny_model(sd=None, dtype=torch.float32, impl=cabi.IMPL_SIMPLE, seed=11):
    dims = decoder_oracle.DecoderDims(**TINY)
    if sd is None:
        sd = decoder_oracle.random_decoder_state(dims, seed=seed, ln_jitter=0.1)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**TINY), state_dict=sd).to(DEV)
    if dtype != torch.float32:
        m = m.to(dtype)
    m.impl = impl
    return dims, sd, m


def _eos_prone_state(dims, sd):
    """Biases every head towards EOS (id 32) so that rows would finish within a few steps."""
    sd2 = dict(sd)
    for k in range(dims.num_codebooks):
        w = sd2[f"lm_heads.{k}.weight"].clone()
        w[32] = w[32] * 0 + sd2["model.decoder.layer_norm.weight"] * 0.05
        sd2[f"lm_heads.{k}.weight"] = w
    sd2["model.decoder.layer_norm.bias"] = sd2["model.decoder.layer_norm.bias"] + 0.5
    return sd2


def _inputs(B=3, S=5, P=2, seed=9):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, S, 128, generator=g), torch.randn(B, P, 128, generator=g)


# ------------------------------------------------------------------------------------ generation-config fields
@pytest.mark.parametrize("kw", [
    dict(do_sample=True, top_k=0, top_p=0.8, seed=3),
    dict(do_sample=True, top_k=20, top_p=0.5, temperature=0.9, seed=4),
    dict(do_sample=True, top_p=0.05, seed=5),          # top_k falls back to the legacy default of 50
    dict(do_sample=False, repetition_penalty=1.3),
    dict(do_sample=True, top_k=8, repetition_penalty=0.8, seed=6),
    dict(do_sample=False, repetition_penalty=1.5, guidance_scale=2.0),
])
def test_fused_epilogue_fields_vs_oracle(kw):
    dims, sd, m = _tiny_model()
    enc, prm = _inputs()
    M = 18
    okw = dict(kw)
    if okw.get("do_sample") and "top_k" not in okw:
        okw["top_k"] = 50
    want = generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=M, **okw)
    for graph in (True, False):
        m.use_cuda_graph = graph
        codes = _generate(m, enc, prm, dict(kw, max_length=M))
        assert np.array_equal(codes.cpu().numpy(), want["codes"]), (kw, graph)


@pytest.mark.parametrize("kw", [dict(min_new_tokens=9), dict(min_length=12), dict(min_new_tokens=15, min_length=7)])
def test_min_length_suppresses_eos_vs_oracle(kw):
    dims, sd, _ = _tiny_model()
    sd2 = _eos_prone_state(dims, sd)
    _, _, m = _tiny_model(sd=sd2)
    enc, prm = _inputs(B=2, S=4, P=3, seed=4)
    base = generate_oracle.generate_codes(sd2, dims, enc, prompt_hidden_states=prm, max_length=40, do_sample=True,
                                          top_k=0, seed=3)
    want = generate_oracle.generate_codes(sd2, dims, enc, prompt_hidden_states=prm, max_length=40, do_sample=True,
                                          top_k=0, seed=3, **kw)
    assert base["sequences"].shape[1] < want["sequences"].shape[1] <= 40, "the constraint must change the outcome"
    out = m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=40, do_sample=True,
                     top_k=0, seed=3, **kw)
    assert np.array_equal(out.cpu().numpy(), want["codes"])


class _BanEven:
    """A user logits processor the fused epilogue knows nothing about."""

    def __call__(self, input_ids, scores):
        scores = scores.clone()
        scores[:, 0::2] -= 60.0  # finite: -inf on both guidance halves would turn the mix into NaN (as in the reference)
        return scores


def test_python_processors_with_and_without_guidance_vs_oracle():
    dims, sd, m = _tiny_model()
    enc, prm = _inputs(B=2, S=4, P=2, seed=21)
    for gs in (None, 2.5):
        want = generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=16, do_sample=False,
                                              guidance_scale=gs, logits_processors=[_BanEven()],
                                              repetition_penalty=1.2, suppress_from=32)
        codes = _generate(m, enc, prm, dict(max_length=16, do_sample=False, guidance_scale=gs, repetition_penalty=1.2),
                          logits_processor=[_BanEven(), ptts.SuppressTokensAtOrAbove(32)])
        got = codes.cpu().numpy()
        assert np.array_equal(got, want["codes"]), gs
        assert (got % 2 == 1).all()


def test_transformers_warpers_run_on_the_host_path():
    """typical_p / epsilon_cutoff / no_repeat_ngram_size are not in the fused epilogue: transformers' own classes run
    between the forward kernels and the epilogue, and the draw is reproducible."""
    dims, sd, m = _tiny_model()
    enc, prm = _inputs()
    kw = dict(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=16, do_sample=True,
              typical_p=0.6, epsilon_cutoff=0.01, no_repeat_ngram_size=2, top_k=30, seed=12,
              logits_processor=[ptts.SuppressTokensAtOrAbove(32)])
    a, b = m.generate(**kw), m.generate(**kw)
    assert a.shape == (3, 4, 12) and torch.equal(a, b) and int(a.max()) < 32
    seq = a.cpu().numpy()
    for r in seq.reshape(-1, seq.shape[-1]):  # the de-delayed rows are the raw rows shifted: no repeated bigram
        grams = [tuple(r[i:i + 2]) for i in range(len(r) - 1)]
        assert len(grams) == len(set(grams))
    with pytest.raises(NotImplementedError):
        m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=16,
                   encoder_repetition_penalty=1.5)
    with pytest.raises(ValueError):
        m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=16, num_beams=2)


# ------------------------------------------------------------------------------------ streamer / stopping / shards
class _Streamer:
    def __init__(self):
        self.chunks, self.ended = [], False

    def put(self, value):
        self.chunks.append(value.clone())

    def end(self):
        self.ended = True


def test_streamer_and_stopping_criteria():
    dims, sd, m = _tiny_model()
    enc, prm = _inputs()
    kw = dict(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=16, do_sample=False,
              logits_processor=[ptts.SuppressTokensAtOrAbove(32)])
    full = m.generate(**kw)
    st = _Streamer()
    codes = m.generate(streamer=st, **kw)
    assert torch.equal(codes, full) and st.ended
    assert st.chunks[0].shape == (12, 1) and int(st.chunks[0][0, 0]) == dims.bos_token_id  # the start column first
    cols = torch.stack([c.reshape(-1) for c in st.chunks[1:]], dim=1)  # then one raw token column per step
    assert cols.shape == (12, 15)
    seq = torch.cat([st.chunks[0], cols], dim=1).view(3, 4, 16)
    for k in range(4):  # de-delay: codes[b][k][f] = seq[b][k][f + k + 1]
        assert torch.equal(seq[:, k, k + 1 : k + 1 + 12], full[:, k].cpu())

    class StopAt:
        def __init__(self, n):
            self.n = n

        def __call__(self, input_ids, scores):
            return input_ids.shape[1] >= self.n

    short = m.generate(stopping_criteria=[StopAt(10)], **kw)  # Lout = 10 -> 10 - K = 6 frames, same tokens
    assert short.shape == (3, 4, 6)
    assert torch.equal(short, full[:, :, :6])


def test_sampling_does_not_depend_on_the_shard_layout():
    """row_offset keys the Philox counter by the GLOBAL row: a shard generates exactly the rows of the unsharded run."""
    dims, sd, m = _tiny_model()
    enc, prm = _inputs(B=4, S=5, P=2, seed=30)
    kw = dict(max_length=16, do_sample=True, top_k=10, seed=99)
    K = dims.num_codebooks
    whole, _, _ = _codes(m, enc, prm, 0, kw)
    for lo, hi in ((0, 1), (1, 4), (2, 4)):
        part, _, _ = _codes(m, enc[lo:hi], prm[lo:hi], lo, kw)
        assert torch.equal(part, whole[lo:hi]), (lo, hi)
    del K


def _codes(m, enc, prm, row_offset, kw, logits_processor=None):
    from parler_tts_b200 import modeling as M

    gc, _ = M._resolve_generation(m.generation_config, None, dict(kw), None)
    seed = kw.get("seed", 0)
    return M._generate_codes(m, enc.to(DEV), None, None, None, prm.to(DEV).float().contiguous(), None, None, gc, kw,
                             logits_processor, None, None, m.impl, m.use_cuda_graph, seed, row_offset)


def _generate(m, enc, prm, kw, logits_processor=None):
    """generate() of the decoder-only model; with guidance the conditioning is doubled the way the COMPOSITE model does
    it (null text, prompt repeated: modeling_parler_tts.py:2231-2237, 2115-2118), which is what the oracle restates
    (the decoder-only generate zeroes the prompt of the unconditional half instead, 1214-1229)."""
    if kw.get("guidance_scale") and kw["guidance_scale"] > 1:
        e2, p2 = torch.cat([enc, torch.zeros_like(enc)]), prm.repeat(2, 1, 1)
        return _codes(m, e2, p2, 0, kw, logits_processor)[0]
    return m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV),
                      logits_processor=logits_processor, **kw)


def test_num_return_sequences_repeats_inputs_like_the_reference():
    from test_gpu_generate import _composite  # tests/ is on sys.path (pytest rootdir import mode)

    model, dims, sd, dac_sd = _composite()
    g = torch.Generator().manual_seed(2)
    input_ids, prompt_ids = torch.randint(2, 100, (2, 7), generator=g), torch.randint(2, 100, (2, 4), generator=g)
    kw = dict(do_sample=True, top_k=12, seed=8, max_length=20, return_codes=True,
              logits_processor=[ptts.SuppressTokensAtOrAbove(32)])
    two = model.generate(input_ids=input_ids.to(DEV), prompt_input_ids=prompt_ids.to(DEV), num_return_sequences=2, **kw)
    rep = model.generate(input_ids=input_ids.repeat_interleave(2, 0).to(DEV),
                         prompt_input_ids=prompt_ids.repeat_interleave(2, 0).to(DEV), **kw)
    assert two.shape[0] == 4 and torch.equal(two, rep)
    assert not torch.equal(two[0], two[1])  # different Philox rows: the copies differ
    with pytest.raises(ValueError):
        model.generate(input_ids=input_ids.to(DEV), prompt_input_ids=prompt_ids.to(DEV), do_sample=False,
                       num_return_sequences=2)


def test_generation_config_ids_override_the_decoder_config():
    """generation_config.{decoder_start,pad,eos}_token_id drive mask and loop (modeling_parler_tts.py:2530-2535,
    2586-2587) even when they differ from the decoder config: eos -> 7 freezes rows on token 7 and pads with 5."""
    dims, sd, m = _tiny_model()
    enc, prm = _inputs(B=2, S=4, P=2, seed=17)
    got = m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=30,
                     do_sample=True, top_k=0, seed=2, eos_token_id=7, pad_token_id=5)
    import copy

    d2 = copy.copy(dims)
    d2.eos_token_id, d2.pad_token_id = 7, 5
    want = generate_oracle.generate_codes(sd, d2, enc, prompt_hidden_states=prm, max_length=30, do_sample=True,
                                          top_k=0, seed=2)
    assert np.array_equal(got.cpu().numpy(), want["codes"])


# ------------------------------------------------------------------------------------ forward() with the static cache
@pytest.mark.parametrize("impl", [cabi.IMPL_SIMPLE, cabi.IMPL_FAST])
def test_forward_with_static_cache_matches_reference_fixture(golden_dir, impl):
    """ParlerTTSForCausalLM.forward(past_key_values=...) (modeling_parler_tts.py:1096-1113): prefill on the first
    columns, then cached steps, equals the reference's teacher-forced logits of the whole sequence."""
    z = np.load(os.path.join(golden_dir, "decoder_tiny.npz"))
    dims, sd, m = _tiny_model(impl=impl)
    enc, prm = torch.from_numpy(z["enc"]).to(DEV), torch.from_numpy(z["prm"]).to(DEV)
    ids = torch.from_numpy(z["tf_ids"]).to(DEV)
    P = prm.shape[1]
    o = m(input_ids=ids[:, :2], encoder_hidden_states=enc, prompt_hidden_states=prm, use_cache=True)
    assert o.past_key_values.get_seq_length() == P + 2
    o2 = m(input_ids=ids[:, 2:5], past_key_values=o.past_key_values)
    o3 = m(input_ids=ids[:, 5:], past_key_values=o2.past_key_values)
    assert o3.past_key_values.get_seq_length() == P + 6
    got = torch.cat([o.logits, o2.logits, o3.logits], dim=1).cpu().numpy()
    np.testing.assert_allclose(got, z["tf_logits"], atol=1e-4, rtol=1e-3)
    with pytest.raises(TypeError):
        m(input_ids=ids[:, :1], past_key_values=((None,) * 4,))


@pytest.mark.parametrize("impl", [cabi.IMPL_SIMPLE, cabi.IMPL_FAST])
def test_cached_steps_match_the_reference_steps_masked_and_unmasked(golden_dir, impl):
    """The reference's own cached decode steps (ParlerTTSForCausalLM.forward with its tuple cache, modeling_parler_tts.py:
    292-304, 843-879; oracle/gen_golden.py gen_decoder) against prefill + three cached steps here, with and without
    ragged description / prompt masks: logits of every step, fp32."""
    z = np.load(os.path.join(golden_dir, "decoder_tiny.npz"))
    dims, sd, m = _tiny_model(impl=impl)
    enc, prm = torch.from_numpy(z["enc"]), torch.from_numpy(z["prm"])
    B = enc.shape[0]
    ids = torch.full((B * dims.num_codebooks, 1), dims.bos_token_id)
    for tag in ("nomask", "masked"):
        em = torch.from_numpy(z["enc_mask"]) if tag == "masked" else None
        pm = torch.from_numpy(z["prm_mask"]) if tag == "masked" else None
        e = enc if em is None else enc * em[..., None]
        o = m(input_ids=ids.to(DEV), encoder_hidden_states=e.to(DEV), encoder_attention_mask=None if em is None else em.to(DEV),
              prompt_hidden_states=prm.to(DEV), prompt_attention_mask=None if pm is None else pm.to(DEV), use_cache=True,
              max_cache_len=8)
        np.testing.assert_allclose(o.logits.cpu().numpy(), z[f"{tag}_prefill_logits"], atol=1e-4, rtol=1e-3)
        steps = torch.from_numpy(z[f"{tag}_step_ids"])  # (3, B*K, 1)
        past = o.past_key_values
        for t in range(3):
            o = m(input_ids=steps[t].to(DEV), past_key_values=past)
            past = o.past_key_values
            np.testing.assert_allclose(o.logits.cpu().numpy(), z[f"{tag}_step{t}_logits"], atol=1e-4, rtol=1e-3,
                                       err_msg=f"{tag} step {t}")


@pytest.mark.parametrize("impl", [cabi.IMPL_SIMPLE, cabi.IMPL_FAST])
def test_second_decoder_shape_matches_the_reference(golden_dir, impl):
    """Three heads, five codebooks, three layers, hidden 192 (not a power of two), ragged left- / right-padded description
    and prompt masks: prefill + five cached steps against the reference's own logits (oracle/gen_golden.py gen_decoder2)."""
    from oracle.gen_golden import TINY2

    z = np.load(os.path.join(golden_dir, "decoder_tiny2.npz"))
    dims = decoder_oracle.DecoderDims(**TINY2)
    sd = decoder_oracle.random_decoder_state(dims, seed=23, ln_jitter=0.2)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**TINY2), state_dict=sd).to(DEV)
    m.impl = impl
    B, S, P, n = (int(v) for v in z["meta"])
    enc, prm = torch.from_numpy(z["enc"]), torch.from_numpy(z["prm"])
    em, pm = torch.from_numpy(z["enc_mask"]), torch.from_numpy(z["prm_mask"])
    e = enc * em[..., None]
    ids = torch.full((B * dims.num_codebooks, 1), dims.bos_token_id)
    o = m(input_ids=ids.to(DEV), encoder_hidden_states=e.to(DEV), encoder_attention_mask=em.to(DEV),
          prompt_hidden_states=prm.to(DEV), prompt_attention_mask=pm.to(DEV), use_cache=True, max_cache_len=n + 2)
    np.testing.assert_allclose(o.logits.cpu().numpy(), z["prefill_logits"], atol=1e-4, rtol=1e-3)
    steps = torch.from_numpy(z["step_ids"])
    past = o.past_key_values
    for t in range(n):
        o = m(input_ids=steps[t].to(DEV), past_key_values=past)
        past = o.past_key_values
        np.testing.assert_allclose(o.logits.cpu().numpy(), z[f"step{t}_logits"], atol=1e-4, rtol=1e-3, err_msg=f"step {t}")
    assert past.get_seq_length() == P + 1 + n


# ------------------------------------------------------------------------------------ fp8 KV cache kernels
def _e4m3_to_float(u8):
    return u8.view(torch.float8_e4m3fn).float()


@pytest.mark.parametrize("B,ctx,masked", [(64, 466, False), (3, 700, False), (2, 300, True), (64, 33, False)])
def test_fp8_attention_kernel_vs_torch_on_the_dequantised_cache(B, ctx, masked):
    heads, d, c_max = 16, 64, ctx + 40
    g = torch.Generator(device=DEV).manual_seed(ctx)
    stage = (torch.randn(2, B, heads, ctx, d, device=DEV, generator=g) *
             torch.linspace(0.3, 3.0, heads, device=DEV).view(1, 1, heads, 1, 1)).to(torch.bfloat16)
    cache = torch.zeros(2, B, heads, c_max, d, device=DEV, dtype=torch.uint8)
    scale = torch.zeros(2, heads, device=DEV)
    ws = torch.zeros(2 * heads, device=DEV, dtype=torch.int32)
    cabi.call("ptts_kv_quantize_fp8", cabi.ptr(stage), 2, B, heads, d, ctx, c_max, cabi.ptr(ws), cabi.ptr(scale),
              cabi.ptr(cache), cabi.stream_ptr())
    # quantisation itself: scale = 4 * amax / 448 per (K|V, head); bytes = round-to-nearest-even e4m3 of x / scale
    amax = stage.float().abs().amax(dim=(1, 3, 4))
    torch.testing.assert_close(scale, amax * 4 / 448, rtol=1e-6, atol=0)
    want_bytes = (stage.float() * (1.0 / scale).view(2, 1, heads, 1, 1)).to(torch.float8_e4m3fn).view(torch.uint8)
    assert torch.equal(cache[:, :, :, :ctx], want_bytes)
    q = (torch.randn(B, heads * d, device=DEV, generator=g) * 0.4).to(torch.bfloat16)
    mask = None
    if masked:
        mask = (torch.rand(B, 50, device=DEV, generator=g) > 0.3).to(torch.int32)
        mask[:, -1] = 1
    out = torch.empty(B, heads * d, device=DEV, dtype=torch.bfloat16)
    scratch = torch.zeros(1 << 22, device=DEV)
    ctx_t = torch.tensor([ctx - 5], dtype=torch.int32, device=DEV)
    cabi.call("ptts_attention_fp8", cabi.ptr(q), cabi.ptr(cache[0]), cabi.ptr(cache[1]), cabi.ptr(scale[0]),
              cabi.ptr(scale[1]), B, heads, d, c_max, 5, cabi.ptr(ctx_t), cabi.ptr(mask), 0 if mask is None else 50,
              cabi.ptr(out), cabi.ptr(scratch), cabi.stream_ptr())
    k = _e4m3_to_float(cache[0, :, :, :ctx]) * scale[0].view(1, heads, 1, 1)
    v = _e4m3_to_float(cache[1, :, :, :ctx]) * scale[1].view(1, heads, 1, 1)
    s = torch.einsum("bhd,bhjd->bhj", q.float().view(B, heads, d), k)
    if mask is not None:
        s[:, :, :50] = torch.where(mask[:, None, :] != 0, s[:, :, :50], torch.finfo(torch.bfloat16).min)
    ref = torch.einsum("bhj,bhjd->bhd", torch.softmax(s, dim=-1), v).reshape(B, heads * d)
    err = float((out.float() - ref).abs().max() / ref.abs().max())
    assert err < 1e-2, err  # f16 q / f16 partial dot products / bf16 output rounding; the cache bytes are identical


def test_generate_with_fp8_kv_cache_tracks_the_bf16_cache():
    """kv_dtype='fp8' end to end on the fused tcgen05 path (graph replay and eager): prefill on the staging cache,
    quantisation, appends in the attention prologue.  Greedy tokens stay close to the bf16-cache run and the teacher-forced
    logits agree to a few percent of the logit scale (3 mantissa bits per cached element)."""
    kw = dict(TINY, num_hidden_layers=3)
    dims = decoder_oracle.DecoderDims(**kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=5, ln_jitter=0.05)
    m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV).to(torch.bfloat16)
    m.impl = cabi.IMPL_FAST
    enc, prm = _inputs(B=8, S=6, P=20, seed=2)
    ids = torch.randint(0, 32, (8 * 4, 12), generator=torch.Generator().manual_seed(1))
    logits = {}
    for kvd in (None, "fp8"):
        m.kv_dtype = kvd
        o = m(input_ids=ids[:, :1].to(DEV), encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV),
              use_cache=True, max_cache_len=16)
        logits[kvd] = m(input_ids=ids[:, 1:].to(DEV), past_key_values=o.past_key_values).logits.float().cpu()
    scale = float(logits[None].abs().max())
    err = float((logits["fp8"] - logits[None]).abs().max()) / scale
    assert 0 < err < 8e-2, err
    outs = {}
    for kvd in (None, "fp8"):
        m.kv_dtype = kvd
        for graph in (True, False):
            m.use_cuda_graph = graph
            outs[kvd, graph] = m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV),
                                          max_length=40, do_sample=False,
                                          logits_processor=[ptts.SuppressTokensAtOrAbove(32)]).cpu()
    m.kv_dtype = None
    assert torch.equal(outs["fp8", True], outs["fp8", False]), "graph replay == eager with the fp8 cache"
    assert torch.equal(outs[None, True], outs[None, False])
    with pytest.raises(cabi.ParlerB200Error):
        m32 = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw), state_dict=sd).to(DEV)
        m32.kv_dtype = "fp8"
        m32.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=12)
