"""CPU: the oracle restatement against fixtures produced by the reference's own code (oracle/gen_golden.py)."""
import os

import numpy as np
import torch

from oracle import dac_oracle, decoder_oracle, delay_pattern, generate_oracle
from oracle.gen_golden import TINY

CASES = ["kat", "mini", "mini_long", "short", "edge", "prompt3", "prompt5"]


def test_delay_mask_docstring_kat():
    # /root/reference/parler_tts/modeling_parler_tts.py:84-87 with B=9, P=8
    _, mask = delay_pattern.build_delay_pattern_mask(np.full((4, 1), 9), 9, 8, 8, 4)
    want = np.array([[9, -1, -1, -1, -1, 8, 8, 8], [9, 9, -1, -1, -1, -1, 8, 8],
                     [9, 9, 9, -1, -1, -1, -1, 8], [9, 9, 9, 9, -1, -1, -1, -1]])
    assert np.array_equal(mask, want)


def test_delay_mask_vs_reference_randomised(golden_dir):
    """48 randomised cases run through the reference (oracle/gen_golden.py delay_random): 1..9 codebooks, every prompt
    length up to M - K + 1, prompt values colliding with BOS / PAD, the degenerate M < 2K - 1 mask."""
    z = np.load(os.path.join(golden_dir, "delay_patterns_random.npz"))
    assert len(z["names"]) == 48
    for name in z["names"]:
        B, K, M, T, bos, pad = z[name + "_meta"]
        ids, mask = delay_pattern.build_delay_pattern_mask(z[name + "_in"], bos, pad, M, K)
        assert np.array_equal(mask, z[name + "_mask"]), (name, z[name + "_meta"])
        assert np.array_equal(ids, z[name + "_ids"]), (name, z[name + "_meta"])
        applied = delay_pattern.apply_delay_pattern_mask(z[name + "_gen"], mask)
        assert np.array_equal(applied, z[name + "_applied"]), (name, z[name + "_meta"])


def test_delay_then_dedelay_round_trip_property():
    """Size-independent property of a1 / a2 / a14 (modeling_parler_tts.py:68-139, 2633-2648): lay codes (B, K, F) out in
    the delay pattern (code f of codebook k in column f + k + 1, anything in the BOS / PAD triangles), and the de-delay
    of that matrix returns the codes - for every K, and for matrices cut short (the loop stopped early)."""
    rng = np.random.default_rng(5)
    for _ in range(40):
        K = int(rng.integers(1, 10))
        B = int(rng.integers(1, 4))
        F = int(rng.integers(K, 60))           # M = F + K >= 2K - 1
        M, bos, pad = F + K, 1025, 1024
        codes = rng.integers(0, 1024, size=(B, K, F)).astype(np.int64)
        start = np.full((B * K, 1), bos, dtype=np.int64)
        ids0, mask = delay_pattern.build_delay_pattern_mask(start, bos, pad, M, K)
        assert ids0.shape == (B * K, 1) and mask.shape == (B * K, M)
        raw = rng.integers(0, 1088, size=(B * K, M)).astype(np.int64)   # what the loop "selected" inside the triangles
        for k in range(K):
            raw[k::K, k + 1 : k + 1 + F] = codes[:, k]
        got = delay_pattern.revert_delay_pattern(raw, start, mask, bos, pad, K)
        assert got.shape == (B, K, F) and np.array_equal(got, codes)
        # stopped early at Lout columns: frames 0 .. Lout - K - 1 come back when Lout >= 2K - 1, otherwise the reference
        # keeps every column (its rebuilt mask is all -1)
        Lout = int(rng.integers(max(2 * K - 1, K + 1), M + 1))
        part = delay_pattern.revert_delay_pattern(raw[:, :Lout], start, mask, bos, pad, K)
        assert part.shape == (B, K, Lout - K) and np.array_equal(part, codes[:, :, : Lout - K])
        # frames holding an id outside the codebook are the ones the ragged path drops
        bad = codes.copy()
        bad[0, rng.integers(0, K), rng.integers(0, F)] = 1024
        keep = delay_pattern.valid_frame_mask(bad, 1024)
        assert keep.shape == (B, F) and int((~keep).sum()) == 1
        assert delay_pattern.needs_sequential_decode(bad, bos, pad, pad) and not delay_pattern.needs_sequential_decode(codes, bos, pad, pad)


def test_position_table_vs_reference(golden_dir):
    """ParlerTTSSinusoidalPositionalEmbedding.get_embedding (modeling_parler_tts.py:200-213): the table a random-init
    model is built with - the host's generator and the oracle's - equals the reference's bit for bit (even and odd
    widths; the 4096 x 1024 Mini table through sampled rows and a per-row fp64 sum)."""
    from parler_tts_b200.modeling import sinusoidal_table as host_table

    z = np.load(os.path.join(golden_dir, "positions.npz"))
    for fn in (host_table, decoder_oracle.sinusoidal_table):
        for n, d in ((64, 16), (37, 9), (5, 128)):
            got = fn(n, d).numpy()
            assert got.dtype == np.float32 and np.array_equal(got, z[f"tab_{n}_{d}"]), (fn.__module__, n, d)
        big = fn(4096, 1024)
        assert np.array_equal(big[torch.from_numpy(z["mini_rows"])].numpy(), z["mini_sample"])
        assert np.array_equal(big.double().sum(dim=1).numpy(), z["mini_row_sums"])


def test_delay_mask_vs_reference(golden_dir):
    z = np.load(os.path.join(golden_dir, "delay_patterns.npz"))
    for name in CASES:
        B, K, M, T, bos, pad = z[name + "_meta"]
        ids, mask = delay_pattern.build_delay_pattern_mask(z[name + "_in"], bos, pad, M, K)
        assert np.array_equal(mask, z[name + "_mask"]), name
        assert np.array_equal(ids, z[name + "_ids"]), name
        applied = delay_pattern.apply_delay_pattern_mask(z[name + "_gen"], mask)
        assert np.array_equal(applied, z[name + "_applied"]), name


def _tiny():
    dims = decoder_oracle.DecoderDims(**TINY)
    return dims, decoder_oracle.random_decoder_state(dims, seed=11, ln_jitter=0.1)


def test_decoder_forward_vs_reference(golden_dir):
    z = np.load(os.path.join(golden_dir, "decoder_tiny.npz"))
    dims, sd = _tiny()
    enc, prm = torch.from_numpy(z["enc"]), torch.from_numpy(z["prm"])
    B = enc.shape[0]
    for tag in ("nomask", "masked"):
        em = torch.from_numpy(z["enc_mask"]) if tag == "masked" else None
        pm = torch.from_numpy(z["prm_mask"]) if tag == "masked" else None
        e = enc if em is None else enc * em[..., None]
        ids = torch.full((B * dims.num_codebooks, 1), dims.bos_token_id)
        logits, past = decoder_oracle.decoder_forward(sd, dims, ids, e, em, prm, pm, None)
        np.testing.assert_allclose(logits.numpy(), z[f"{tag}_prefill_logits"], atol=2e-5, rtol=1e-4)
        steps = torch.from_numpy(z[f"{tag}_step_ids"])
        for t in range(3):
            logits, past = decoder_oracle.decoder_forward(sd, dims, steps[t], e, em, None, pm, past)
            np.testing.assert_allclose(logits.numpy(), z[f"{tag}_step{t}_logits"], atol=2e-5, rtol=1e-4)
        np.testing.assert_allclose(past[1][0].numpy(), z[f"{tag}_k_self_l1"], atol=1e-5)
        np.testing.assert_allclose(past[0][3].numpy(), z[f"{tag}_v_cross_l0"], atol=1e-5)
    logits, _ = decoder_oracle.decoder_forward(sd, dims, torch.from_numpy(z["tf_ids"]), enc, None, prm, None, None)
    np.testing.assert_allclose(logits.numpy(), z["tf_logits"], atol=2e-5, rtol=1e-4)


def test_decoder_forward_vs_reference_second_shape(golden_dir):
    """Three heads, five codebooks, three layers, hidden 192, ragged left- / right-padded masks, five cached steps
    (oracle/gen_golden.py gen_decoder2): logits of every step and the final self-attention cache."""
    from oracle.gen_golden import TINY2

    z = np.load(os.path.join(golden_dir, "decoder_tiny2.npz"))
    dims = decoder_oracle.DecoderDims(**TINY2)
    sd = decoder_oracle.random_decoder_state(dims, seed=23, ln_jitter=0.2)
    B, S, P, n = (int(v) for v in z["meta"])
    enc, prm = torch.from_numpy(z["enc"]), torch.from_numpy(z["prm"])
    em, pm = torch.from_numpy(z["enc_mask"]), torch.from_numpy(z["prm_mask"])
    e = enc * em[..., None]
    ids = torch.full((B * dims.num_codebooks, 1), dims.bos_token_id)
    logits, past = decoder_oracle.decoder_forward(sd, dims, ids, e, em, prm, pm, None)
    np.testing.assert_allclose(logits.numpy(), z["prefill_logits"], atol=2e-5, rtol=1e-4)
    steps = torch.from_numpy(z["step_ids"])
    for t in range(n):
        logits, past = decoder_oracle.decoder_forward(sd, dims, steps[t], e, em, None, pm, past)
        np.testing.assert_allclose(logits.numpy(), z[f"step{t}_logits"], atol=2e-5, rtol=1e-4, err_msg=f"step {t}")
    np.testing.assert_allclose(past[-1][0].numpy(), z["k_self_last"], atol=1e-5)
    np.testing.assert_allclose(past[0][1].numpy(), z["v_self_l0"], atol=1e-5)
    assert past[0][0].shape[2] == P + 1 + n


def test_loop_vs_reference_driven_probe(golden_dir):
    z = np.load(os.path.join(golden_dir, "loop_tiny.npz"))
    dims, sd = _tiny()
    enc, prm = torch.from_numpy(z["enc"]), torch.from_numpy(z["prm"])
    for tag in ("free", "noeos"):
        max_length, suppress, steps = z[f"{tag}_meta"]
        res = generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=int(max_length),
                                             do_sample=False, suppress_from=None if suppress < 0 else int(suppress),
                                             return_trace=True)
        # margin-gated: the two forwards differ by fp32 rounding; decisions with a clear top-2 gap must agree
        ref_scores = z[f"{tag}_scores"]
        top2 = np.sort(ref_scores, axis=-1)[..., -2:]
        assert (top2[..., 1] - top2[..., 0]).min() > 1e-4
        assert np.array_equal(res["sequences"], z[f"{tag}_sequences"]), tag
        assert np.array_equal(res["codes"], z[f"{tag}_codes"]), tag
        assert res["steps"] == steps


def test_dac_vs_transformers_equivalent(golden_dir):
    z = np.load(os.path.join(golden_dir, "dac_tiny.npz"))
    c = z["cfg"]
    cfg = dict(n_codebooks=int(c[0]), codebook_size=int(c[1]), codebook_dim=int(c[2]), latent_dim=int(c[3]),
               decoder_dim=int(c[4]), rates=tuple(int(x) for x in c[5:]))
    sd = dac_oracle.random_dac_state(cfg, seed=31)
    codes = torch.from_numpy(z["codes"])
    zlat = dac_oracle.from_codes(sd, codes, cfg)
    np.testing.assert_allclose(zlat.numpy(), z["z"], atol=1e-5, rtol=1e-5)
    wav = dac_oracle.decode_codes(sd, codes, cfg).squeeze(1)
    assert wav.shape == z["wav"].shape and wav.shape[-1] == 512 * codes.shape[-1]
    np.testing.assert_allclose(wav.numpy(), z["wav"], atol=2e-6)


def test_philox_known_answer():
    # Random123 known-answer test for philox4x32-10: counter = key = 0 and the all-ones vector
    out = generate_oracle.philox4x32(np.zeros(4, np.uint32), np.zeros(2, np.uint32))
    assert [hex(int(x)) for x in out] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    out = generate_oracle.philox4x32(np.full(4, 0xFFFFFFFF, np.uint32), np.full(2, 0xFFFFFFFF, np.uint32))
    assert [hex(int(x)) for x in out] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]


def test_sampling_candidates_and_distribution():
    rng = np.random.default_rng(0)
    s = rng.normal(size=200).astype(np.float32)
    keep = generate_oracle.topk_candidates(s, 50)
    assert keep.sum() == 50 and s[keep].min() >= np.sort(s)[-50]
    draws = [generate_oracle.sample_row(s, generate_oracle.uniform_for(5, 0, i), 1.0, 50) for i in range(400)]
    assert all(keep[d] for d in draws)
    assert len(set(draws)) > 20


def test_t5_encoder_vs_transformers(golden_dir):
    """oracle/t5_oracle.py against transformers' T5EncoderModel outputs (tests/golden/t5_tiny.npz: gated-gelu and
    relu stacks, 150 tokens > max_distance, padded and almost-empty rows)."""
    from oracle import t5_oracle
    from oracle.gen_golden import T5_CASES

    z = np.load(os.path.join(golden_dir, "t5_tiny.npz"))
    for name, cfg in T5_CASES.items():
        sd = t5_oracle.random_t5_state(cfg, seed=5)
        got = t5_oracle.t5_encode(sd, cfg, torch.from_numpy(z[name + "_ids"]), torch.from_numpy(z[name + "_mask"]))
        assert float((got - torch.from_numpy(z[name + "_out"])).abs().max()) < 1e-5, name


def test_t5_bucket_table_matches_transformers():
    from transformers.models.t5.modeling_t5 import T5Attention

    from parler_tts_b200.text_encoder import MAX_REL, relative_position_buckets

    rel = torch.arange(-MAX_REL, MAX_REL + 1)
    for nb, md in ((32, 128), (16, 64)):
        want = T5Attention._relative_position_bucket(rel, bidirectional=True, num_buckets=nb, max_distance=md)
        assert torch.equal(relative_position_buckets(nb, md).long(), want)


def test_dac_encode_vs_transformers(golden_dir):
    """oracle/dac_oracle.py encode path (encoder conv stack + residual VQ) against transformers' DacModel.encode
    (tests/golden/dac_enc_tiny.npz); also the right-padding of a ragged length."""
    z = np.load(os.path.join(golden_dir, "dac_enc_tiny.npz"))
    cfg = dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=128, decoder_dim=64, rates=(8, 8, 4, 2))
    sd = dict(dac_oracle.random_dac_state(cfg, seed=31))
    sd.update(dac_oracle.random_dac_encoder_state(cfg, seed=31, encoder_dim=8))
    audio = torch.from_numpy(z["audio"])
    lat = dac_oracle.encode_latents(sd, audio, cfg)
    assert float((lat - torch.from_numpy(z["z"])).abs().max()) < 1e-5
    assert np.array_equal(dac_oracle.rvq_encode(sd, lat, cfg).numpy(), z["codes"])
    assert dac_oracle.encode_audio(sd, audio[..., : 512 * 8 + 100], cfg).shape == (2, 3, 9)


def test_dac_decoder_receptive_field_is_ten_frames():
    """The chunked (streaming) DAC decode of parler_tts_b200/dac.py relies on this: an output frame of the 44.1 kHz
    decoder sees exactly 10 frames of context on either side (first conv 3 + block 1: 7), so a chunk decoded with a
    halo >= 10 equals the same frames of the one-shot decode; 9 is not enough."""
    import torch

    from oracle import dac_oracle

    sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=2)
    F, f0, f1 = 36, 12, 24
    codes = torch.randint(0, 1024, (1, 9, F), generator=torch.Generator().manual_seed(6))
    full = dac_oracle.decode_codes(sd, codes, dtype=torch.float64)[:, :, f0 * 512 : f1 * 512]
    err = {}
    for halo in (9, 10):
        lo, hi = f0 - halo, f1 + halo
        part = dac_oracle.decode_codes(sd, codes[:, :, lo:hi], dtype=torch.float64)
        err[halo] = float((part[:, :, (f0 - lo) * 512 : (f1 - lo) * 512] - full).abs().max())
    assert err[10] == 0.0 and err[9] > 1e-7, err


def test_dac_oracle_vs_transformers_live_on_other_shapes():
    """The codec oracle against the installed transformers' DacModel (the arithmetic-equivalent of the un-vendored
    descript-audio-codec), live, on shapes the committed fixture does not hold: other numbers of blocks and codebooks,
    other widths, and descript's 24 kHz rate set with its ODD stride 5 (5 T - 1 samples from that block)."""
    from oracle.gen_golden import hf_dac_with_oracle_weights

    shapes = [
        dict(n_codebooks=2, codebook_size=16, codebook_dim=4, latent_dim=64, decoder_dim=96, rates=(8, 5, 4, 2)),
        dict(n_codebooks=4, codebook_size=64, codebook_dim=8, latent_dim=32, decoder_dim=32, rates=(2, 2)),
        dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=64, decoder_dim=64, rates=(4, 2, 8, 2)),
        dict(n_codebooks=1, codebook_size=8, codebook_dim=2, latent_dim=16, decoder_dim=48, rates=(3, 4, 4)),
    ]
    for i, cfg in enumerate(shapes):
        sd = dac_oracle.random_dac_state(cfg, seed=40 + i)
        hf = hf_dac_with_oracle_weights(cfg, sd)
        codes = torch.randint(0, cfg["codebook_size"], (2, cfg["n_codebooks"], 6),
                              generator=torch.Generator().manual_seed(i))
        with torch.no_grad():
            want = hf.decode(audio_codes=codes).audio_values
            zq = hf.quantizer.from_codes(codes)[0]
        got = dac_oracle.decode_codes(sd, codes, cfg).squeeze(1)
        assert got.shape == want.shape, (cfg["rates"], got.shape, want.shape)
        if all(r % 2 == 0 for r in cfg["rates"]):  # an odd stride s gives s * T - 1 samples in that block
            assert got.shape == (2, 6 * int(np.prod(cfg["rates"])))
        np.testing.assert_allclose(dac_oracle.from_codes(sd, codes, cfg).numpy(), zq.numpy(), atol=1e-5, rtol=1e-5)
        np.testing.assert_allclose(got.numpy(), want.numpy(), atol=2e-5, rtol=1e-4, err_msg=str(cfg["rates"]))


def test_dac_encode_oracle_vs_transformers_live_on_other_shapes():
    """The codec ENCODE oracle (audio-prompt path, modeling_parler_tts.py:2247-2292 -> dac_wrapper/modeling_dac.py:25-96)
    against the installed transformers' DacModel.encode, live, on shapes the committed fixture does not hold: latents to
    1e-5, codes index-exact wherever the nearest-codebook margin is clear (and the margins are required to be clear)."""
    from oracle.gen_golden import hf_dac_encoder_with_oracle_weights

    shapes = [
        (dict(n_codebooks=2, codebook_size=16, codebook_dim=4, latent_dim=64, decoder_dim=32, rates=(4, 2, 8, 2)), 4),
        (dict(n_codebooks=4, codebook_size=64, codebook_dim=3, latent_dim=32, decoder_dim=32, rates=(2, 4)), 8),
        (dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=48, decoder_dim=48, rates=(8, 5, 4, 2)), 3),
    ]
    for i, (cfg, enc_dim) in enumerate(shapes):
        assert enc_dim * 2 ** len(cfg["rates"]) == cfg["latent_dim"]
        sd = dict(dac_oracle.random_dac_state(cfg, seed=50 + i))
        sd.update(dac_oracle.random_dac_encoder_state(cfg, seed=50 + i, encoder_dim=enc_dim))
        hf = hf_dac_encoder_with_oracle_weights(cfg, sd)
        hop = int(np.prod(cfg["rates"]))
        audio = torch.randn(2, 1, hop * 7, generator=torch.Generator().manual_seed(i)) * 0.5
        with torch.no_grad():
            want_z = hf.encoder(audio)
            want = hf.encode(audio).audio_codes
        z = dac_oracle.encode_latents(sd, audio, cfg)
        np.testing.assert_allclose(z.numpy(), want_z.numpy(), atol=1e-5, rtol=1e-4, err_msg=str(cfg["rates"]))
        got, margins = dac_oracle.rvq_encode(sd, z, cfg, return_margins=True)
        assert got.shape == want.shape == (2, cfg["n_codebooks"], 7)
        sure = margins > 1e-5
        assert float(sure.float().mean()) > 0.98 and torch.equal(got[sure], want[sure]), str(cfg["rates"])


def test_t5_oracle_vs_transformers_live_on_other_shapes():
    """The T5 encoder oracle against the installed transformers' T5EncoderModel, live, on shapes the committed fixture
    does not hold: flan-t5-base's proportions (12 heads of d_kv = d_model / 12, gated-GELU, 3 layers here), a stack whose
    heads do not tile d_model (d_kv * heads != d_model), few buckets with a short max_distance, and no mask at all."""
    from oracle import t5_oracle
    from oracle.gen_golden import hf_t5

    base = dict(vocab_size=90, relative_attention_num_buckets=32, relative_attention_max_distance=128,
                layer_norm_epsilon=1e-6)
    shapes = [
        dict(base, d_model=96, d_kv=8, num_heads=12, d_ff=256, num_layers=3, gated=True, dense_act_fn="gelu_new",
             feed_forward_proj="gated-gelu"),
        dict(base, d_model=48, d_kv=20, num_heads=3, d_ff=64, num_layers=2, gated=False, dense_act_fn="relu",
             feed_forward_proj="relu", relative_attention_num_buckets=8, relative_attention_max_distance=16),
    ]
    for i, cfg in enumerate(shapes):
        sd = t5_oracle.random_t5_state(cfg, seed=60 + i)
        g = torch.Generator().manual_seed(i)
        ids = torch.randint(0, cfg["vocab_size"], (3, 41), generator=g)
        mask = torch.ones(3, 41, dtype=torch.long)
        mask[0, 30:] = 0
        mask[2, 1:] = 0
        hf = hf_t5(cfg, sd)
        for m in (mask, None):
            with torch.no_grad():
                want = hf(input_ids=ids, attention_mask=m).last_hidden_state
            got = t5_oracle.t5_encode(sd, cfg, ids, m)
            np.testing.assert_allclose(got.numpy(), want.numpy(), atol=2e-5, rtol=1e-4, err_msg=f"shape {i}, mask {m is not None}")


def test_oracle_processors_match_transformers_classes():
    """The oracle's top-p / repetition-penalty restatements against the installed transformers' own classes."""
    from transformers.generation import logits_process as lp

    g = torch.Generator().manual_seed(0)
    scores = torch.randn(6, 96, generator=g) * 2
    ids = torch.randint(0, 96, (6, 11), generator=g)
    want = lp.RepetitionPenaltyLogitsProcessor(1.7)(ids, scores.clone())
    got = generate_oracle.repetition_penalty_(scores, ids.numpy(), 1.7)
    assert torch.equal(want, got)
    for top_p in (0.3, 0.9):
        want = lp.TopPLogitsWarper(top_p)(ids, scores.clone())
        for r in range(6):
            keep = generate_oracle.topp_candidates(scores[r].numpy(), np.ones(96, dtype=bool), top_p)
            assert np.array_equal(keep, torch.isfinite(want[r]).numpy()), (top_p, r)


def test_c1_fixture_reproduced_by_the_oracle_at_full_depth(golden_dir):
    """BASELINE.json configs[0] (Mini v0.1, bs=1, greedy, 172 frames): the oracle's own decoder + loop reproduce the
    tokens the reference decoder produced (oracle/gen_golden.py c1), i.e. the restatement is pinned at 24 layers."""
    from oracle.gen_golden import C1, MINI

    z = np.load(os.path.join(golden_dir, "c1_mini.npz"))
    seed, max_length, weight_seed = (int(v) for v in z["meta"])
    assert max_length == C1["max_length"]
    dims = decoder_oracle.DecoderDims(**MINI)
    sd = decoder_oracle.random_decoder_state(dims, seed=weight_seed)
    res = generate_oracle.generate_codes(sd, dims, torch.from_numpy(z["enc"]), prompt_hidden_states=torch.from_numpy(z["prm"]),
                                         max_length=max_length, do_sample=False, suppress_from=1024, return_trace=True)
    assert np.array_equal(res["sequences"], z["sequences"])
    assert np.array_equal(res["codes"], z["codes"]) and z["codes"].shape == (1, 9, 172)
    err = float((res["scores"][-1][:, :1024] - torch.from_numpy(z["last_scores"][:, :1024])).abs().max())
    margin = float(z["margin"][0])
    assert err < 1e-3 and margin > 10 * err, (margin, err)
