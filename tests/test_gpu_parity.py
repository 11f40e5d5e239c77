"""GPU parity: the CUDA path (through the C ABI) against the committed reference fixtures and the CPU oracle."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dac_oracle, decoder_oracle, delay_pattern, generate_oracle  # noqa: E402
from oracle.gen_golden import TINY  # noqa: E402

import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.dac import DacDecoderEngine  # noqa: E402

DEV = "cuda"
IMPLS = [cabi.IMPL_SIMPLE, cabi.IMPL_FAST]


def _tiny_model(dtype=torch.float32, impl=cabi.IMPL_SIMPLE, dims_kw=TINY, seed=11, ln_jitter=0.1):
    dims = decoder_oracle.DecoderDims(**dims_kw)
    sd = decoder_oracle.random_decoder_state(dims, seed=seed, ln_jitter=ln_jitter)
    cfg = ptts.ParlerTTSDecoderConfig(**{k: v for k, v in dims_kw.items()})
    m = ptts.ParlerTTSForCausalLM(cfg, state_dict=sd).to(DEV)
    if dtype != torch.float32:
        m = m.to(dtype)
    m.impl = impl
    return dims, sd, m


# ------------------------------------------------------------------------------------------------ integer kernels
CASES = ["kat", "mini", "mini_long", "short", "edge", "prompt3", "prompt5"]


def test_delay_mask_bit_exact_vs_reference_fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "delay_patterns.npz"))
    for name in CASES:
        B, K, M, T, bos, pad = (int(v) for v in z[name + "_meta"])
        ids, mask = ptts.build_delay_pattern_mask(torch.from_numpy(z[name + "_in"]).to(DEV), bos, pad, M, K)
        assert np.array_equal(mask.cpu().numpy(), z[name + "_mask"]), name
        assert np.array_equal(ids.cpu().numpy(), z[name + "_ids"]), name
        applied = ptts.apply_delay_pattern_mask(torch.from_numpy(z[name + "_gen"]).to(DEV), mask)
        assert np.array_equal(applied.cpu().numpy(), z[name + "_applied"]), name


def test_delay_mask_bit_exact_vs_reference_randomised(golden_dir):
    """48 randomised reference cases (oracle/gen_golden.py delay_random): 1..9 codebooks, every admissible prompt length,
    prompt values colliding with BOS / PAD, the degenerate M < 2K - 1 mask."""
    z = np.load(os.path.join(golden_dir, "delay_patterns_random.npz"))
    for name in z["names"]:
        B, K, M, T, bos, pad = (int(v) for v in z[name + "_meta"])
        ids, mask = ptts.build_delay_pattern_mask(torch.from_numpy(z[name + "_in"]).to(DEV), bos, pad, M, K)
        assert np.array_equal(mask.cpu().numpy(), z[name + "_mask"]), (name, B, K, M, T)
        assert np.array_equal(ids.cpu().numpy(), z[name + "_ids"]), (name, B, K, M, T)
        applied = ptts.apply_delay_pattern_mask(torch.from_numpy(z[name + "_gen"]).to(DEV), mask)
        assert np.array_equal(applied.cpu().numpy(), z[name + "_applied"]), (name, B, K, M, T)


def test_delay_mask_full_size_properties():
    # BASELINE config 2/5 sizes: B=64, K=9, M=869 and M=2609 - structure checked without the oracle
    for B, M in ((64, 869), (16, 2609)):
        K, bos, pad = 9, 1025, 1024
        _, mask = ptts.build_delay_pattern_mask(torch.full((B * K, 1), bos, device=DEV), bos, pad, M, K)
        m = mask.view(B, K, M)
        for k in range(K):
            assert bool((m[:, k, : k + 1] == bos).all())
            assert bool((m[:, k, k + 1 : M - K + 1 + k] == -1).all())
            assert bool((m[:, k, M - K + 1 + k :] == pad).all())
        o_ids, o_mask = delay_pattern.build_delay_pattern_mask(np.full((B * K, 1), bos), bos, pad, M, K)
        assert np.array_equal(o_mask, mask.cpu().numpy())


def test_dedelay_and_frame_filter_vs_oracle():
    rng = np.random.default_rng(3)
    B, K, M, bos, pad = 3, 9, 40, 1025, 1024
    start = np.full((B * K, 1), bos, dtype=np.int64)
    _, mask = delay_pattern.build_delay_pattern_mask(start, bos, pad, M, K)
    for Lout in (M, 31, 17, 16, K):  # 16 and K are below 2K-1: the reference keeps every column there
        raw = rng.integers(0, 1088, size=(B * K, Lout)).astype(np.int64)
        want = delay_pattern.revert_delay_pattern(raw, start, mask, bos, pad, K)
        ids_d = torch.zeros(B * K, M, dtype=torch.int64, device=DEV)
        ids_d[:, :Lout] = torch.from_numpy(raw).to(DEV)
        mask_d = torch.from_numpy(mask).to(DEV)
        codes = torch.empty(B, K, want.shape[-1], dtype=torch.int64, device=DEV)
        cabi.call("ptts_dedelay", cabi.ptr(ids_d), M, cabi.ptr(mask_d), M, B, K, Lout, cabi.ptr(codes),
                  cabi.stream_ptr())
        assert np.array_equal(codes.cpu().numpy(), want)
    codes = torch.from_numpy(rng.integers(1000, 1040, size=(4, 9, 700))).to(DEV)
    codes[3] = 1024  # a sample with no valid frame
    packed = torch.empty_like(codes)
    counts = torch.zeros(4, dtype=torch.int32, device=DEV)
    cabi.call("ptts_filter_frames", cabi.ptr(codes), 4, 9, 700, 1024, cabi.ptr(packed), cabi.ptr(counts),
              cabi.stream_ptr())
    keep = delay_pattern.valid_frame_mask(codes.cpu().numpy(), 1024)
    for b in range(4):
        n = int(keep[b].sum())
        assert int(counts[b]) == n
        assert np.array_equal(packed[b, :, :n].cpu().numpy(), codes[b].cpu().numpy()[:, keep[b]])


# ------------------------------------------------------------------------------------------------ decoder forward
@pytest.mark.parametrize("impl", IMPLS)
def test_forward_logits_vs_reference_fixture_fp32(golden_dir, impl):
    z = np.load(os.path.join(golden_dir, "decoder_tiny.npz"))
    dims, sd, m = _tiny_model(impl=impl)
    enc, prm = torch.from_numpy(z["enc"]), torch.from_numpy(z["prm"])
    B = enc.shape[0]
    ids = torch.full((B * dims.num_codebooks, 1), dims.bos_token_id)
    for tag in ("nomask", "masked"):
        em = torch.from_numpy(z["enc_mask"]) if tag == "masked" else None
        pm = torch.from_numpy(z["prm_mask"]) if tag == "masked" else None
        e = enc if em is None else enc * em[..., None]
        out = m(input_ids=ids, encoder_hidden_states=e, encoder_attention_mask=em, prompt_hidden_states=prm,
                prompt_attention_mask=pm)
        np.testing.assert_allclose(out.logits.cpu().numpy(), z[f"{tag}_prefill_logits"], atol=1e-4, rtol=1e-3)
    out = m(input_ids=torch.from_numpy(z["tf_ids"]), encoder_hidden_states=enc, prompt_hidden_states=prm)
    np.testing.assert_allclose(out.logits.cpu().numpy(), z["tf_logits"], atol=1e-4, rtol=1e-3)


class _Recorder:
    def __init__(self):
        self.scores = []

    def __call__(self, input_ids, scores):
        self.scores.append(scores.detach().float().cpu().clone())
        return scores


@pytest.mark.parametrize("impl", IMPLS)
def test_greedy_loop_token_exact_vs_reference_fixture_fp32(golden_dir, impl):
    """prefill + cached decode steps + fused epilogue against the loop driven over the reference decoder."""
    z = np.load(os.path.join(golden_dir, "loop_tiny.npz"))
    dims, sd, m = _tiny_model(impl=impl)
    enc, prm = torch.from_numpy(z["enc"]).to(DEV), torch.from_numpy(z["prm"]).to(DEV)
    for tag in ("free", "noeos"):
        max_length, suppress, steps = (int(v) for v in z[f"{tag}_meta"])
        procs = [] if suppress < 0 else [ptts.SuppressTokensAtOrAbove(suppress)]
        for graph in (False, True):
            m.use_cuda_graph = graph
            rec = _Recorder()
            codes = m.generate(encoder_hidden_states=enc, prompt_hidden_states=prm, max_length=max_length,
                               do_sample=False, logits_processor=procs + ([] if graph else [rec]))
            assert np.array_equal(codes.cpu().numpy(), z[f"{tag}_codes"]), (tag, graph)
            if not graph:
                got = torch.stack(rec.scores).numpy()
                want = z[f"{tag}_scores"]
                fin = np.isfinite(want)
                np.testing.assert_allclose(got[fin], want[fin], atol=2e-4, rtol=1e-3)


@pytest.mark.parametrize("impl", IMPLS)
def test_bf16_logits_within_1e2_relative(impl):
    """BASELINE.json: bf16 logits within 1e-2 relative error (of the logit scale) of the fp32 oracle."""
    kw = dict(TINY, hidden_size=256, num_attention_heads=4, ffn_dim=1024, num_hidden_layers=4, vocab_size=1088,
              num_codebooks=9, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
    dims, sd, m = _tiny_model(torch.bfloat16, impl, kw, seed=5, ln_jitter=0.05)
    g = torch.Generator().manual_seed(1)
    B, S, P, T = 4, 12, 6, 5
    enc, prm = torch.randn(B, S, 256, generator=g), torch.randn(B, P, 256, generator=g) * 0.5
    ids = torch.randint(0, 1024, (B * 9, T), generator=g)
    sd_bf = {k: (v.to(torch.bfloat16).float() if v.dim() == 2 and "embed" not in k else v) for k, v in sd.items()}
    want, _ = decoder_oracle.decoder_forward(sd_bf, dims, ids, enc.to(torch.bfloat16).float(), None, prm, None, None)
    got = m(input_ids=ids, encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV)).logits.cpu()
    scale = want.abs().max()
    assert float((got - want).abs().max() / scale) < 1e-2


def test_mini_architecture_fp32_greedy_token_exact_margin_gated():
    """Config 1 shape (24 x 1024, 9 codebooks), B=2, short: greedy tokens equal the oracle's wherever the oracle's
    top-2 margin exceeds fp32 noise; logits agree to 1e-3."""
    kw = dict(vocab_size=1088, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, ffn_dim=4096,
              num_codebooks=9, max_position_embeddings=256, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
    dims, sd, m = _tiny_model(torch.float32, cabi.IMPL_FAST, kw, seed=0, ln_jitter=0.0)
    g = torch.Generator().manual_seed(1234)
    B, S, P, M = 2, 10, 6, 24
    enc, prm = torch.randn(B, S, 1024, generator=g) * 0.3, torch.randn(B, P, 1024, generator=g) * 0.02
    want = generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=M, do_sample=False,
                                          suppress_from=1024, return_trace=True)
    rec = _Recorder()
    codes = m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=M,
                       do_sample=False, logits_processor=[ptts.SuppressTokensAtOrAbove(1024), rec])
    got = torch.stack(rec.scores)[:, :, :1024]
    ref = want["scores"][:, :, :1024]
    top2 = ref.topk(2, dim=-1).values
    margins = top2[..., 0] - top2[..., 1]  # (steps, rows)
    err = float((got - ref).abs().max())
    assert err < 1e-3, err
    # every decision in front of the first one the oracle itself calls too close is compared UNCONDITIONALLY (greedy,
    # ids >= 1024 suppressed: the chosen token is the argmax of the recorded scores); the first step always is
    unsure = (margins <= 10 * err).any(dim=1)
    n_sure = int(unsure.float().argmax()) if bool(unsure.any()) else margins.shape[0]
    assert n_sure >= 1, (float(margins[0].min()), err)
    assert torch.equal(got[:n_sure].argmax(-1), ref[:n_sure].argmax(-1))
    if n_sure == margins.shape[0]:
        assert np.array_equal(codes.cpu().numpy(), want["codes"])
    # (the unconditional full-length statement at this depth is test_gpu_baseline_configs.py, whose inputs were chosen
    # for their margin against the reference decoder in float64)


# ------------------------------------------------------------------------------------------------ epilogue
def test_sampling_epilogue_vs_oracle_shared_uniforms():
    dims, sd, m = _tiny_model()
    g = torch.Generator().manual_seed(9)
    B, S, P, M = 3, 5, 2, 18
    enc, prm = torch.randn(B, S, 128, generator=g), torch.randn(B, P, 128, generator=g)
    for top_k, temp in ((5, 1.0), (50, 0.7), (0, 1.3)):
        want = generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=M, do_sample=True,
                                              top_k=top_k, temperature=temp, seed=77, return_trace=True)
        codes = m.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=M,
                           do_sample=True, top_k=top_k, temperature=temp, seed=77)
        assert np.array_equal(codes.cpu().numpy(), want["codes"]), (top_k, temp)


def test_eos_freeze_and_early_stop_match_oracle():
    """Rows that draw EOS are frozen to PAD and the loop stops when every row has finished (transformers loop)."""
    dims, sd, m = _tiny_model()
    sd2 = dict(sd)
    for k in range(dims.num_codebooks):  # bias every head towards EOS (id 32) so that rows finish quickly
        w = sd2[f"lm_heads.{k}.weight"].clone()
        w[32] = w[32] * 0 + sd2["model.decoder.layer_norm.weight"] * 0.05
        sd2[f"lm_heads.{k}.weight"] = w
    sd2["model.decoder.layer_norm.bias"] = sd2["model.decoder.layer_norm.bias"] + 0.5
    cfg = ptts.ParlerTTSDecoderConfig(**TINY)
    m2 = ptts.ParlerTTSForCausalLM(cfg, state_dict=sd2).to(DEV)
    m2.impl = cabi.IMPL_SIMPLE
    g = torch.Generator().manual_seed(4)
    enc, prm = torch.randn(2, 4, 128, generator=g), torch.randn(2, 3, 128, generator=g)
    want = generate_oracle.generate_codes(sd2, dims, enc, prompt_hidden_states=prm, max_length=40, do_sample=True,
                                          top_k=0, seed=3)
    out = m2.generate(encoder_hidden_states=enc.to(DEV), prompt_hidden_states=prm.to(DEV), max_length=40,
                      do_sample=True, top_k=0, seed=3)
    assert want["sequences"].shape[1] < 40, "fixture should stop early"
    assert np.array_equal(out.cpu().numpy(), want["codes"])
