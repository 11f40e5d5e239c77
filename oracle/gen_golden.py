"""Generates the committed fixtures under tests/golden/ by running the UNMODIFIED reference code
(/root/reference, imported through oracle/ref_shim.py) and, for the un-vendored DAC dependency, the
arithmetic-equivalent implementation in the installed `transformers`.

Run in the build container only (the reference does not travel to the GPU box):
    python -m oracle.gen_golden
Weights are never stored: they are regenerated from seeds by oracle.decoder_oracle.random_decoder_state /
oracle.dac_oracle.random_dac_state, so a fixture = (seeds, inputs, reference outputs).
"""
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")

from oracle import dac_oracle, decoder_oracle, generate_oracle  # noqa: E402
from oracle.ref_shim import import_reference  # noqa: E402

warnings.filterwarnings("ignore")

TINY = dict(
    vocab_size=96, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, ffn_dim=256, num_codebooks=4,
    max_position_embeddings=64, pad_token_id=32, eos_token_id=32, bos_token_id=33,
)


def gen_delay(ref):
    cases = {}
    # known-answer pattern of the reference docstring (modeling_parler_tts.py:84-87): K=4, M=8, B=9, P=8
    specs = [
        ("kat", 1, 4, 8, 1, 9, 8),
        ("mini", 2, 9, 20, 1, 1025, 1024),
        ("mini_long", 3, 9, 181, 1, 1025, 1024),
        ("short", 2, 9, 16, 1, 1025, 1024),  # M < 2K-1: all -1
        ("edge", 1, 9, 17, 1, 1025, 1024),  # M == 2K-1
        ("prompt3", 2, 4, 12, 3, 33, 32),  # audio-prompt continuation
        ("prompt5", 1, 9, 30, 5, 1025, 1024),
    ]
    rng = np.random.default_rng(7)
    for name, B, K, M, T, bos, pad in specs:
        if T == 1:
            ids = np.full((B * K, 1), bos, dtype=np.int64)
        else:
            ids = rng.integers(0, pad, size=(B * K, T)).astype(np.int64)
            ids[:, 0] = bos
        out_ids, mask = ref.build_delay_pattern_mask(torch.from_numpy(ids), bos, pad, M, K)
        cases[name + "_in"] = ids
        cases[name + "_meta"] = np.array([B, K, M, T, bos, pad], dtype=np.int64)
        cases[name + "_ids"] = out_ids.numpy()
        cases[name + "_mask"] = mask.numpy()
        gen = rng.integers(0, pad + 2, size=(B * K, min(M, 11))).astype(np.int64)
        cases[name + "_gen"] = gen
        cases[name + "_applied"] = ref.apply_delay_pattern_mask(torch.from_numpy(gen), mask).numpy()
    np.savez_compressed(os.path.join(GOLD, "delay_patterns.npz"), **cases)
    print("delay_patterns.npz:", len(specs), "cases")


def gen_delay_random(ref, n=48, seed=20):
    """Randomised cases of the same two reference functions (modeling_parler_tts.py:68-139): 1..9 codebooks, prompts of
    every admissible length 1 <= T <= M - K + 1 (T = M - K + 1 is the "nothing left to fill" branch, :131-133), prompt
    values that collide with BOS / PAD, M on both sides of 2K - 1 (the degenerate all -1 mask, :105-106)."""
    rng = np.random.default_rng(seed)
    cases, names = {}, []
    for i in range(n):
        K = int(rng.integers(1, 10))
        B = int(rng.integers(1, 4))
        degenerate = i % 8 == 7 and K >= 2
        M = int(rng.integers(1, 2 * K - 1)) if degenerate else int(rng.integers(2 * K - 1, 2 * K + 40))
        t_max = M if degenerate else M - K + 1
        T = t_max if i % 6 == 5 else int(rng.integers(1, t_max + 1))
        pad = int(rng.integers(8, 2000))
        bos = pad + 1 if i % 3 else int(rng.integers(0, pad))
        ids = rng.integers(0, pad + 2, size=(B * K, T)).astype(np.int64)
        if i % 2 == 0:
            ids[:, 0] = bos
        out_ids, mask = ref.build_delay_pattern_mask(torch.from_numpy(ids), bos, pad, M, K)
        name = f"r{i:02d}"
        names.append(name)
        cases[name + "_in"] = ids
        cases[name + "_meta"] = np.array([B, K, M, T, bos, pad], dtype=np.int64)
        cases[name + "_ids"] = out_ids.numpy()
        cases[name + "_mask"] = mask.numpy()
        gen = rng.integers(0, pad + 2, size=(B * K, int(rng.integers(1, M + 1)))).astype(np.int64)
        cases[name + "_gen"] = gen
        cases[name + "_applied"] = ref.apply_delay_pattern_mask(torch.from_numpy(gen), mask).numpy()
    cases["names"] = np.array(names)
    np.savez_compressed(os.path.join(GOLD, "delay_patterns_random.npz"), **cases)
    print("delay_patterns_random.npz:", n, "cases;",
          sum(1 for nm in names if (cases[nm + "_mask"] == -1).all()), "degenerate,",
          sum(1 for nm in names if not (cases[nm + "_mask"][:: int(cases[nm + "_meta"][1])] == -1).any()),
          "with nothing left to fill in the first codebook")


def _ref_decoder(ref, dims, sd):
    cfg = ref.ParlerTTSDecoderConfig(
        vocab_size=dims.vocab_size, max_position_embeddings=dims.max_position_embeddings,
        num_hidden_layers=dims.num_hidden_layers, ffn_dim=dims.ffn_dim, num_attention_heads=dims.num_attention_heads,
        hidden_size=dims.hidden_size, num_codebooks=dims.num_codebooks, pad_token_id=dims.pad_token_id,
        eos_token_id=dims.eos_token_id, bos_token_id=dims.bos_token_id, dropout=0.1,
    )
    m = ref.ParlerTTSForCausalLM(cfg).eval()
    missing = m.load_state_dict(sd, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return m


def gen_decoder(ref):
    dims = decoder_oracle.DecoderDims(**TINY)
    sd = decoder_oracle.random_decoder_state(dims, seed=11, ln_jitter=0.1)
    model = _ref_decoder(ref, dims, sd)
    g = torch.Generator().manual_seed(12)
    B, S, P, K = 3, 5, 4, dims.num_codebooks
    enc = torch.randn(B, S, dims.hidden_size, generator=g)
    prm = torch.randn(B, P, dims.hidden_size, generator=g) * 0.5
    enc_mask = torch.tensor([[1, 1, 1, 1, 1], [1, 1, 1, 0, 0], [1, 1, 1, 1, 0]])
    prm_mask = torch.tensor([[1, 1, 1, 1], [0, 1, 1, 1], [1, 1, 0, 0]])
    out = {"meta": np.array([B, S, P], dtype=np.int64)}
    for tag, em, pm in (("nomask", None, None), ("masked", enc_mask, prm_mask)):
        e = enc if em is None else enc * em[..., None]
        ids = torch.full((B * K, 1), dims.bos_token_id)
        steps = torch.randint(0, dims.vocab_size, (3, B * K, 1), generator=g)
        with torch.no_grad():
            o = model(input_ids=ids, encoder_hidden_states=e, encoder_attention_mask=em, prompt_hidden_states=prm,
                      prompt_attention_mask=pm, use_cache=True)
            out[f"{tag}_prefill_logits"] = o.logits.numpy()
            past = o.past_key_values
            for t in range(3):
                o = model(input_ids=steps[t], encoder_hidden_states=e, encoder_attention_mask=em,
                          prompt_attention_mask=pm, past_key_values=past, use_cache=True)
                past = o.past_key_values
                out[f"{tag}_step{t}_logits"] = o.logits.numpy()
        out[f"{tag}_step_ids"] = steps.numpy()
        out[f"{tag}_k_self_l1"] = past[1][0].numpy()
        out[f"{tag}_v_cross_l0"] = past[0][3].numpy()
    out["enc"], out["prm"], out["enc_mask"], out["prm_mask"] = enc.numpy(), prm.numpy(), enc_mask.numpy(), prm_mask.numpy()
    # teacher-forced multi-token forward without cache (ParlerTTSForCausalLM.forward on a whole sequence)
    seq = torch.randint(0, dims.vocab_size, (B * K, 6), generator=g)
    with torch.no_grad():
        o = model(input_ids=seq, encoder_hidden_states=enc, prompt_hidden_states=prm, use_cache=False)
    out["tf_ids"], out["tf_logits"] = seq.numpy(), o.logits.numpy()
    np.savez_compressed(os.path.join(GOLD, "decoder_tiny.npz"), **out)
    print("decoder_tiny.npz written")
    return dims, sd, model


TINY2 = dict(
    vocab_size=70, hidden_size=192, num_hidden_layers=3, num_attention_heads=3, ffn_dim=320, num_codebooks=5,
    max_position_embeddings=48, pad_token_id=64, eos_token_id=64, bos_token_id=65,
)


def gen_decoder2(ref):
    """A second decoder shape through the reference's forward (three heads, five codebooks, three layers, a hidden size
    that is not a power of two, LayerNorm gains / biases away from 1 / 0): prefill and five cached steps with ragged
    description and prompt masks (left- and right-padded), logits of every step and the whole final cache."""
    dims = decoder_oracle.DecoderDims(**TINY2)
    sd = decoder_oracle.random_decoder_state(dims, seed=23, ln_jitter=0.2)
    model = _ref_decoder(ref, dims, sd)
    g = torch.Generator().manual_seed(24)
    B, S, P, K, n = 4, 7, 5, dims.num_codebooks, 5
    enc = torch.randn(B, S, dims.hidden_size, generator=g)
    prm = torch.randn(B, P, dims.hidden_size, generator=g) * 0.5
    enc_mask = torch.tensor([[1, 1, 1, 1, 1, 1, 1], [1, 1, 1, 1, 0, 0, 0], [1, 0, 0, 0, 0, 0, 0], [1, 1, 1, 1, 1, 1, 0]])
    prm_mask = torch.tensor([[1, 1, 1, 1, 1], [0, 0, 1, 1, 1], [1, 1, 1, 0, 0], [0, 1, 1, 1, 1]])
    e = enc * enc_mask[..., None]
    ids = torch.full((B * K, 1), dims.bos_token_id)
    steps = torch.randint(0, dims.vocab_size, (n, B * K, 1), generator=g)
    out = {"meta": np.array([B, S, P, n], dtype=np.int64), "enc": enc.numpy(), "prm": prm.numpy(),
           "enc_mask": enc_mask.numpy(), "prm_mask": prm_mask.numpy(), "step_ids": steps.numpy()}
    with torch.no_grad():
        o = model(input_ids=ids, encoder_hidden_states=e, encoder_attention_mask=enc_mask, prompt_hidden_states=prm,
                  prompt_attention_mask=prm_mask, use_cache=True)
        out["prefill_logits"] = o.logits.numpy()
        past = o.past_key_values
        for t in range(n):
            o = model(input_ids=steps[t], encoder_hidden_states=e, encoder_attention_mask=enc_mask,
                      prompt_attention_mask=prm_mask, past_key_values=past, use_cache=True)
            past = o.past_key_values
            out[f"step{t}_logits"] = o.logits.numpy()
    out["k_self_last"] = past[-1][0].numpy()
    out["v_self_l0"] = past[0][1].numpy()
    np.savez_compressed(os.path.join(GOLD, "decoder_tiny2.npz"), **out)
    print("decoder_tiny2.npz written:", {k: v.shape for k, v in out.items() if k.endswith("logits")})


def gen_loop(ref, dims, sd, model):
    """Drives the reference decoder + the reference mask functions through the restated loop (greedy), so that the
    fixture pins prepare_inputs/apply-mask/cache semantics end to end on the reference's own forward."""
    g = torch.Generator().manual_seed(21)
    B, S, P = 2, 6, 3
    enc = torch.randn(B, S, dims.hidden_size, generator=g)
    prm = torch.randn(B, P, dims.hidden_size, generator=g) * 0.5
    out = {"enc": enc.numpy(), "prm": prm.numpy()}
    for tag, max_length, suppress in (("free", 24, None), ("noeos", 20, dims.pad_token_id)):
        def fwd(inp, prompt, past):
            with torch.no_grad():
                o = model(input_ids=inp, encoder_hidden_states=enc, prompt_hidden_states=prompt,
                          past_key_values=past, use_cache=True)
            return o.logits, o.past_key_values

        # the loop with reference functions substituted for the oracle's
        res = generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=max_length,
                                             do_sample=False, suppress_from=suppress, forward_fn=fwd, return_trace=True)
        # cross-check mask/apply with the reference implementation
        start = torch.full((B * dims.num_codebooks, 1), dims.bos_token_id)
        _, rmask = ref.build_delay_pattern_mask(start, dims.bos_token_id, dims.pad_token_id, max_length, dims.num_codebooks)
        assert np.array_equal(rmask.numpy(), res["pattern_mask"])
        applied = ref.apply_delay_pattern_mask(torch.from_numpy(res["sequences"]), rmask)
        _, m2 = ref.build_delay_pattern_mask(start, dims.bos_token_id, dims.pad_token_id, applied.shape[1], dims.num_codebooks)
        keep = (m2 != dims.bos_token_id) & (m2 != dims.pad_token_id)
        codes = applied[keep].reshape(B, dims.num_codebooks, -1).numpy()
        assert np.array_equal(codes, res["codes"])
        out[f"{tag}_sequences"], out[f"{tag}_codes"] = res["sequences"], codes
        out[f"{tag}_scores"] = res["scores"].numpy()
        out[f"{tag}_meta"] = np.array([max_length, -1 if suppress is None else suppress, res["steps"]], dtype=np.int64)
        print(f"loop {tag}: Lout={res['sequences'].shape[1]} codes={codes.shape}")
    np.savez_compressed(os.path.join(GOLD, "loop_tiny.npz"), **out)


def hf_dac_with_oracle_weights(cfg, sd):
    """transformers' DacModel (arithmetic-equivalent to descript DAC) carrying the oracle's random decoder + quantizer
    weights `sd` of the descript layout `cfg` (dac_oracle.random_dac_state); the encoder side keeps its random init."""
    from transformers import DacConfig, DacModel

    rates, n_q = tuple(cfg["rates"]), cfg["n_codebooks"]
    nb = len(rates)
    assert cfg["latent_dim"] % 2 ** nb == 0
    hf_cfg = DacConfig(encoder_hidden_size=cfg["latent_dim"] // 2 ** nb, downsampling_ratios=list(reversed(rates)),
                       decoder_hidden_size=cfg["decoder_dim"], n_codebooks=n_q, codebook_size=cfg["codebook_size"],
                       codebook_dim=cfg["codebook_dim"], sampling_rate=44100)
    assert hf_cfg.hidden_size == cfg["latent_dim"] and list(hf_cfg.upsampling_ratios) == list(rates)
    hf = DacModel(hf_cfg).eval()
    new = {}
    for i in range(n_q):
        new[f"quantizer.quantizers.{i}.codebook.weight"] = sd[f"quantizer.quantizers.{i}.codebook.weight"]
        w, b = dac_oracle.wn(sd, f"quantizer.quantizers.{i}.out_proj")
        new[f"quantizer.quantizers.{i}.out_proj.weight"], new[f"quantizer.quantizers.{i}.out_proj.bias"] = w, b
    w, b = dac_oracle.wn(sd, "decoder.model.0")
    new["decoder.conv1.weight"], new["decoder.conv1.bias"] = w, b
    for bi in range(nb):
        p, h = f"decoder.model.{bi + 1}.block", f"decoder.block.{bi}"
        new[h + ".snake1.alpha"] = sd[p + ".0.alpha"]
        w, b = dac_oracle.wn(sd, p + ".1")
        new[h + ".conv_t1.weight"], new[h + ".conv_t1.bias"] = w, b
        for ri in range(3):
            r, hr = f"{p}.{ri + 2}.block", f"{h}.res_unit{ri + 1}"
            new[hr + ".snake1.alpha"], new[hr + ".snake2.alpha"] = sd[r + ".0.alpha"], sd[r + ".2.alpha"]
            w, b = dac_oracle.wn(sd, r + ".1")
            new[hr + ".conv1.weight"], new[hr + ".conv1.bias"] = w, b
            w, b = dac_oracle.wn(sd, r + ".3")
            new[hr + ".conv2.weight"], new[hr + ".conv2.bias"] = w, b
    new["decoder.snake1.alpha"] = sd[f"decoder.model.{nb + 1}.alpha"]
    w, b = dac_oracle.wn(sd, f"decoder.model.{nb + 2}")
    new["decoder.conv2.weight"], new["decoder.conv2.bias"] = w, b
    res = hf.load_state_dict(new, strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    assert all(k.startswith("encoder.") or ".in_proj" in k for k in res.missing_keys), res.missing_keys
    return hf


def gen_dac():
    """transformers' DacModel (arithmetic-equivalent to descript DAC) on the oracle's random weights."""
    cfg = dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=128, decoder_dim=64, rates=(8, 8, 4, 2))
    sd = dac_oracle.random_dac_state(cfg, seed=31)
    hf = hf_dac_with_oracle_weights(cfg, sd)
    g = torch.Generator().manual_seed(32)
    codes = torch.randint(0, 32, (2, 3, 7), generator=g)
    with torch.no_grad():
        wav = hf.decode(audio_codes=codes).audio_values
        z = hf.quantizer.from_codes(codes)[0]
    np.savez_compressed(os.path.join(GOLD, "dac_tiny.npz"), codes=codes.numpy(), wav=wav.numpy(), z=z.numpy(),
                        cfg=np.array([3, 32, 4, 128, 64, 8, 8, 4, 2], dtype=np.int64))
    print("dac_tiny.npz:", wav.shape)


def hf_dac_encoder_with_oracle_weights(cfg, sd):
    """transformers' DacModel carrying the oracle's random ENCODER + quantizer weights (descript layout:
    dac_oracle.random_dac_state + random_dac_encoder_state); the decoder keeps its random init."""
    from transformers import DacConfig, DacModel

    rates, n_q = tuple(cfg["rates"]), cfg["n_codebooks"]
    nb = len(rates)
    hf_cfg = DacConfig(encoder_hidden_size=cfg["latent_dim"] // 2 ** nb, downsampling_ratios=list(reversed(rates)),
                       decoder_hidden_size=cfg["decoder_dim"], n_codebooks=n_q, codebook_size=cfg["codebook_size"],
                       codebook_dim=cfg["codebook_dim"], sampling_rate=44100)
    hf = DacModel(hf_cfg).eval()
    new = {}

    def put(dst, src):
        new[dst + ".weight"], new[dst + ".bias"] = dac_oracle.wn(sd, src)

    for i in range(n_q):
        q = f"quantizer.quantizers.{i}"
        new[q + ".codebook.weight"] = sd[q + ".codebook.weight"]
        put(q + ".out_proj", q + ".out_proj")
        put(q + ".in_proj", q + ".in_proj")
    put("encoder.conv1", "encoder.block.0")
    for bi in range(nb):
        p, h = f"encoder.block.{bi + 1}.block", f"encoder.block.{bi}"
        for ri in range(3):
            r, hr = f"{p}.{ri}.block", f"{h}.res_unit{ri + 1}"
            new[hr + ".snake1.alpha"], new[hr + ".snake2.alpha"] = sd[r + ".0.alpha"], sd[r + ".2.alpha"]
            put(hr + ".conv1", r + ".1")
            put(hr + ".conv2", r + ".3")
        new[h + ".snake1.alpha"] = sd[p + ".3.alpha"]
        put(h + ".conv1", p + ".4")
    new["encoder.snake1.alpha"] = sd[f"encoder.block.{nb + 1}.alpha"]
    put("encoder.conv2", f"encoder.block.{nb + 2}")
    res = hf.load_state_dict(new, strict=False)
    assert not res.unexpected_keys and all(k.startswith("decoder.") for k in res.missing_keys), res
    return hf


def gen_dac_encode():
    """transformers' DacModel.encode (encoder + residual VQ) on the oracle's random weights -> tests/golden/dac_enc_tiny.npz."""
    cfg = dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=128, decoder_dim=64, rates=(8, 8, 4, 2))
    sd = dict(dac_oracle.random_dac_state(cfg, seed=31))
    sd.update(dac_oracle.random_dac_encoder_state(cfg, seed=31, encoder_dim=8))
    hf = hf_dac_encoder_with_oracle_weights(cfg, sd)
    g = torch.Generator().manual_seed(33)
    audio = torch.randn(2, 1, 512 * 9, generator=g) * 0.5
    with torch.no_grad():
        out = hf.encode(audio)
        z = hf.encoder(audio)
    codes = out.audio_codes
    got_z = dac_oracle.encode_latents(sd, audio, cfg)
    got, margins = dac_oracle.rvq_encode(sd, got_z, cfg, return_margins=True)
    print(f"dac encode: latents max abs {float((got_z - z).abs().max()):.2e}, codes equal "
          f"{bool(torch.equal(got, codes))}, min margin {float(margins.min()):.2e}")
    np.savez_compressed(os.path.join(GOLD, "dac_enc_tiny.npz"), audio=audio.numpy(), z=z.numpy(), codes=codes.numpy(),
                        margins=margins.numpy())


T5_TINY = dict(vocab_size=120, d_model=64, d_kv=16, d_ff=96, num_layers=2, num_heads=4,
               relative_attention_num_buckets=32, relative_attention_max_distance=128, layer_norm_epsilon=1e-6)
T5_CASES = {"gated": dict(T5_TINY, gated=True, dense_act_fn="gelu_new", feed_forward_proj="gated-gelu"),
            "relu": dict(T5_TINY, gated=False, dense_act_fn="relu", feed_forward_proj="relu", d_kv=32, num_heads=2)}


def hf_t5(cfg, sd):
    """transformers' own T5EncoderModel (what the reference's AutoModelForTextEncoding builds) with our seeded state."""
    from transformers import T5Config, T5EncoderModel

    keys = ("vocab_size", "d_model", "d_kv", "d_ff", "num_layers", "num_heads", "relative_attention_num_buckets",
            "relative_attention_max_distance", "layer_norm_epsilon", "feed_forward_proj")
    m = T5EncoderModel(T5Config(**{k: cfg[k] for k in keys}, tie_word_embeddings=False)).eval()
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("embed_tokens" in k for k in missing), (missing, unexpected)
    return m


def gen_t5():
    from oracle import t5_oracle

    out = {}
    for name, cfg in T5_CASES.items():
        sd = t5_oracle.random_t5_state(cfg, seed=5)
        g = torch.Generator().manual_seed(17)
        ids = torch.randint(0, cfg["vocab_size"], (3, 150), generator=g)  # 150 > max_distance: clamped buckets too
        mask = torch.ones(3, 150, dtype=torch.long)
        mask[1, 97:] = 0
        mask[2, 5:] = 0
        with torch.no_grad():
            want = hf_t5(cfg, sd)(input_ids=ids, attention_mask=mask).last_hidden_state
        out[name + "_ids"], out[name + "_mask"], out[name + "_out"] = ids.numpy(), mask.numpy(), want.numpy()
        got = t5_oracle.t5_encode(sd, cfg, ids, mask)
        print(f"t5 {name}: oracle vs transformers max abs {float((got - want).abs().max()):.2e}")
    np.savez_compressed(os.path.join(GOLD, "t5_tiny.npz"), **out)


MINI = dict(vocab_size=1088, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, ffn_dim=4096,
            num_codebooks=9, max_position_embeddings=4096, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
# BASELINE.json configs[0] (SURVEY.md section 8d "C1"): Mini v0.1 architecture, random-init, bs=1, greedy, ~2 s of audio
C1 = dict(B=1, S=48, P=32, max_length=181, weight_seed=0)


def c1_inputs(seed):
    g = torch.Generator().manual_seed(seed)
    enc = torch.randn(C1["B"], C1["S"], MINI["hidden_size"], generator=g) * 0.3
    prm = torch.randn(C1["B"], C1["P"], MINI["hidden_size"], generator=g) * 0.02
    return enc, prm


def gen_c1(ref, seeds=(1234, 1235, 1236, 1237, 1238, 1239)):
    """C1 at full depth on the REFERENCE decoder: the restated greedy loop drives the unmodified
    ParlerTTSForCausalLM.forward (24 x 1024, fp32, KV-cache tuples) for max_length - 1 = 180 forward passes; ids >= 1024
    are suppressed (a user logits_processor) so that random weights produce all 172 frames.  The input seed with the
    largest minimum top-2 logit margin is kept, and that margin is re-measured with the reference decoder in float64:
    the GPU test asserts the tokens unconditionally and asserts the margin it relies on."""
    dims = decoder_oracle.DecoderDims(**MINI)
    sd = decoder_oracle.random_decoder_state(dims, seed=C1["weight_seed"])
    model = _ref_decoder(ref, dims, sd)

    def run(m, enc, prm, dtype):
        enc, prm = enc.to(dtype), prm.to(dtype)

        def fwd(inp, prompt, past):
            with torch.no_grad():
                o = m(input_ids=inp, encoder_hidden_states=enc, prompt_hidden_states=prompt, past_key_values=past,
                      use_cache=True)
            return o.logits, o.past_key_values

        return generate_oracle.generate_codes(sd, dims, enc, prompt_hidden_states=prm, max_length=C1["max_length"],
                                              do_sample=False, suppress_from=1024, forward_fn=fwd, return_trace=True)

    def min_margin(res):
        sc = res["scores"][:, :, :1024].double()
        top2 = sc.topk(2, dim=-1).values
        gap = (top2[..., 0] - top2[..., 1]).numpy()  # (steps, rows)
        # only decisions whose token is used matter: row k's column j is overridden by BOS for j <= k and by PAD
        # for j >= M - K + 1 + k (the delay mask), everything else feeds the next step
        mask = res["pattern_mask"][:, 1:] == -1  # (rows, steps)
        return float(gap[mask.T[: gap.shape[0]]].min())

    best = None
    for seed in seeds:
        enc, prm = c1_inputs(seed)
        res = run(model, enc, prm, torch.float32)
        mm = min_margin(res)
        print(f"c1 seed {seed}: min top-2 margin (fp32) {mm:.3e}")
        if best is None or mm > best[1]:
            best = (seed, mm, res)
    seed, mm32, res32 = best
    enc, prm = c1_inputs(seed)
    torch.set_default_dtype(torch.float64)  # transformers' mask helper builds finfo(dtype).min with the default dtype
    try:
        res64 = run(model.double(), enc, prm, torch.float64)
    finally:
        torch.set_default_dtype(torch.float32)
    mm64 = min_margin(res64)
    same = np.array_equal(res32["sequences"], res64["sequences"])
    print(f"c1: seed {seed}, fp64 margin {mm64:.3e}, fp32 == fp64 tokens: {same}")
    assert same, "fp32 and fp64 reference runs disagree: pick another seed"
    err = float((res32["scores"][:, :, :1024].double() - res64["scores"][:, :, :1024]).abs().max())
    np.savez_compressed(os.path.join(GOLD, "c1_mini.npz"), enc=enc.numpy(), prm=prm.numpy(),
                        sequences=res64["sequences"], codes=res64["codes"],
                        meta=np.array([seed, C1["max_length"], C1["weight_seed"]], dtype=np.int64),
                        margin=np.array([mm64, err], dtype=np.float64),
                        last_scores=res64["scores"][-1].float().numpy())
    print(f"c1_mini.npz written: codes {res64['codes'].shape}, fp32-vs-fp64 reference logit error {err:.2e}")


def gen_positions(ref):
    """Known answers of the reference's position-table generator (modeling_parler_tts.py:200-213,
    ParlerTTSSinusoidalPositionalEmbedding.get_embedding): small tables whole, the Mini table (4096 x 1024) as a
    sample of rows plus an fp64 sum per row."""
    from parler_tts.modeling_parler_tts import ParlerTTSSinusoidalPositionalEmbedding as Pos

    out = {}
    for n, d in ((64, 16), (37, 9), (5, 128)):
        out[f"tab_{n}_{d}"] = Pos.get_embedding(n, d).numpy()
    big = Pos.get_embedding(4096, 1024)
    rows = np.array([0, 1, 2, 33, 466, 901, 2609, 4095])
    out["mini_rows"] = rows
    out["mini_sample"] = big[rows].numpy()
    out["mini_row_sums"] = big.double().sum(dim=1).numpy()
    np.savez_compressed(os.path.join(GOLD, "positions.npz"), **out)
    print("positions.npz:", {k: v.shape for k, v in out.items()})


def gen_layout(ref):
    """Checkpoint / config layout of the reference: state-dict key names and shapes of the reference's
    ParlerTTSForCausalLM for two decoder shapes, and the default `to_dict()` of its three config classes."""
    import json

    from parler_tts.configuration_parler_tts import ParlerTTSDecoderConfig as RefDecCfg
    from parler_tts.dac_wrapper.configuration_dac import DACConfig as RefDacCfg
    from parler_tts.modeling_parler_tts import ParlerTTSForCausalLM as RefLM

    out = {"decoder_state_dict": {}, "config_defaults": {}}
    for name, kw in (("tiny", TINY), ("tiny2", TINY2)):
        sd = RefLM(RefDecCfg(**kw)).state_dict()
        out["decoder_state_dict"][name] = {k: list(v.shape) for k, v in sd.items()}
    drop = ("transformers_version",)
    out["config_defaults"]["decoder"] = {k: v for k, v in RefDecCfg().to_dict().items() if k not in drop}
    out["config_defaults"]["dac"] = {k: v for k, v in RefDacCfg().to_dict().items() if k not in drop}
    with open(os.path.join(GOLD, "layout.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True, default=str)
    print("layout.json:", {k: len(v) for k, v in out["decoder_state_dict"].items()}, "keys;",
          len(out["config_defaults"]["decoder"]), "+", len(out["config_defaults"]["dac"]), "config fields")


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = import_reference()
    gen_delay(ref)
    gen_delay_random(ref)
    gen_positions(ref)
    gen_layout(ref)
    dims, sd, model = gen_decoder(ref)
    gen_decoder2(ref)
    gen_loop(ref, dims, sd, model)
    gen_dac()
    gen_dac_encode()
    gen_t5()
    gen_c1(ref)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "layout":
    gen_layout(import_reference())
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "decoder2":
    gen_decoder2(import_reference())
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "delay_random":
    gen_delay_random(import_reference())
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "positions":
    gen_positions(import_reference())
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "c1":
    gen_c1(import_reference())
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "t5":
    gen_t5()  # transformers only: does not need /root/reference
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "dac_encode":
    gen_dac_encode()
elif __name__ == "__main__":
    main()
