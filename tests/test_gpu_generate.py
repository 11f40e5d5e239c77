"""GPU parity of the public generate() API and of the DAC decoder against the oracle / reference fixtures."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dac_oracle, decoder_oracle, generate_oracle  # noqa: E402
from oracle.gen_golden import TINY  # noqa: E402

import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402
from parler_tts_b200.dac import DacDecoderEngine  # noqa: E402

DEV = "cuda"
DAC_TINY = dict(n_codebooks=4, codebook_size=32, codebook_dim=4, latent_dim=128, decoder_dim=64, rates=(8, 8, 4, 2))


def _composite(seed=0):
    from transformers import T5Config

    dims = decoder_oracle.DecoderDims(**TINY)
    sd = decoder_oracle.random_decoder_state(dims, seed=11, ln_jitter=0.1)
    dac_sd = dac_oracle.random_dac_state(DAC_TINY, seed=31)
    dec_cfg = ptts.ParlerTTSDecoderConfig(**TINY)
    t5 = T5Config(d_model=64, num_layers=2, d_ff=128, num_heads=2, d_kv=32, vocab_size=100)
    dac_cfg = ptts.DACConfig(num_codebooks=4, codebook_size=32, latent_dim=128)
    cfg = ptts.ParlerTTSConfig.from_sub_models_config(t5, dac_cfg, dec_cfg, vocab_size=100)
    torch.manual_seed(seed)
    model = ptts.ParlerTTSForConditionalGeneration(
        cfg, decoder=ptts.ParlerTTSForCausalLM(dec_cfg, state_dict=sd),
        audio_encoder=ptts.DACModel(dac_cfg, state_dict=dac_sd, decoder_dim=64, codebook_dim=4))
    model.generation_config.max_length = 24
    return model.to(DEV), dims, sd, dac_sd


def _oracle_generate(model, dims, sd, dac_sd, input_ids, prompt_ids, attention_mask=None, prompt_mask=None,
                     decode=True, **kw):
    """The oracle's view of generate(): T5 (third-party, same module on the CPU) -> projection -> loop -> DAC."""
    te = model.text_encoder.to("cpu").float()
    am = torch.ones_like(input_ids) if attention_mask is None else attention_mask
    with torch.no_grad():
        hs = te(input_ids=input_ids, attention_mask=am).last_hidden_state
    model.text_encoder.to(DEV)
    w, b = model.enc_to_dec_proj.weight.detach().cpu().float(), model.enc_to_dec_proj.bias.detach().cpu().float()
    guidance = kw.get("guidance_scale")
    enc = decoder_oracle.project_encoder(hs, am, w, b)
    if guidance is not None and guidance > 1:  # null conditioning is appended BEFORE the projection in the reference
        hs2 = torch.cat([hs, torch.zeros_like(hs)])
        am2 = torch.cat([am, torch.zeros_like(am)])
        enc_full = decoder_oracle.project_encoder(hs2, am2, w, b)
    prm = model.embed_prompts.weight.detach().cpu().float()[prompt_ids]
    if guidance is not None and guidance > 1:
        # generate_codes doubles enc itself with zeros; the projected null half equals bias*0-mask = 0 too
        assert float(enc_full[enc.shape[0]:].abs().max()) == 0.0
    res = generate_oracle.generate_codes(sd, dims, enc, encoder_attention_mask=am, prompt_hidden_states=prm,
                                         prompt_attention_mask=prompt_mask, **kw)
    if not decode:
        return res, None
    codes = torch.from_numpy(res["codes"])
    audio = dac_oracle.decode_generated(dac_sd, codes, dims.bos_token_id, dims.pad_token_id, dims.eos_token_id, 32,
                                        DAC_TINY)
    return res, audio


def test_generate_end_to_end_greedy_vs_oracle():
    model, dims, sd, dac_sd = _composite()
    g = torch.Generator().manual_seed(2)
    input_ids, prompt_ids = torch.randint(2, 100, (3, 7), generator=g), torch.randint(2, 100, (3, 4), generator=g)
    # ids >= 32 are emittable but are not codebook entries (the reference's DAC would raise on them): cut them
    res, want = _oracle_generate(model, dims, sd, dac_sd, input_ids, prompt_ids, max_length=24, do_sample=False,
                                 suppress_from=32)
    procs = [ptts.SuppressTokensAtOrAbove(32)]
    audio = model.generate(input_ids=input_ids.to(DEV), prompt_input_ids=prompt_ids.to(DEV), do_sample=False,
                           logits_processor=procs)
    codes = model.generate(input_ids=input_ids.to(DEV), prompt_input_ids=prompt_ids.to(DEV), do_sample=False,
                           logits_processor=procs, return_codes=True)
    assert np.array_equal(codes.cpu().numpy(), res["codes"])
    assert audio.shape == want.shape
    np.testing.assert_allclose(audio.cpu().numpy(), want.numpy(), atol=1e-3)


def test_generate_masks_sampling_and_guidance_vs_oracle():
    model, dims, sd, dac_sd = _composite()
    g = torch.Generator().manual_seed(3)
    input_ids, prompt_ids = torch.randint(2, 100, (2, 6), generator=g), torch.randint(2, 100, (2, 5), generator=g)
    am = torch.tensor([[1, 1, 1, 1, 1, 1], [1, 1, 1, 1, 0, 0]])
    pm = torch.tensor([[1, 1, 1, 1, 1], [0, 0, 1, 1, 1]])
    for kw in (dict(do_sample=False), dict(do_sample=True, top_k=10, temperature=0.8, seed=5),
               dict(do_sample=False, guidance_scale=2.5)):
        res, want = _oracle_generate(model, dims, sd, dac_sd, input_ids, prompt_ids, am, pm, decode=False,
                                     max_length=20, **kw)
        codes = model.generate(input_ids=input_ids.to(DEV), attention_mask=am.to(DEV),
                               prompt_input_ids=prompt_ids.to(DEV), prompt_attention_mask=pm.to(DEV), max_length=20,
                               return_codes=True, **kw)
        assert np.array_equal(codes.cpu().numpy(), res["codes"]), kw


def test_generate_masks_without_a_zero_equal_no_mask():
    """An attention mask of ones masks nothing (the reference multiplies by it and adds a zero bias,
    modeling_parler_tts.py:2010-2017): generate() hands the decoder no mask then, the mask-free kernels run, and the
    codes equal those of a call without masks - also for a caller's explicit masks of ones, with and without guidance."""
    model, dims, sd, dac_sd = _composite()
    g = torch.Generator().manual_seed(6)
    input_ids, prompt_ids = torch.randint(2, 100, (3, 6), generator=g).to(DEV), torch.randint(2, 100, (3, 5), generator=g).to(DEV)
    for kw in (dict(do_sample=False), dict(do_sample=True, top_k=10, seed=4), dict(do_sample=False, guidance_scale=2.0)):
        want = model.generate(input_ids=input_ids, prompt_input_ids=prompt_ids, max_length=20, return_codes=True, **kw)
        got = model.generate(input_ids=input_ids, attention_mask=torch.ones_like(input_ids), prompt_input_ids=prompt_ids,
                             prompt_attention_mask=torch.ones_like(prompt_ids).cpu(), max_length=20, return_codes=True, **kw)
        assert torch.equal(got, want), kw
        if not kw.get("guidance_scale"):
            st = model.decoder._state_cache[1]
            assert st.enc_mask is None and st.prompt_mask is None  # the loop ran without the mask loads
    # the oracle (which applies the mask of ones) agrees
    res, _ = _oracle_generate(model, dims, sd, dac_sd, input_ids.cpu(), prompt_ids.cpu(), torch.ones(3, 6, dtype=torch.int64),
                              None, decode=False, max_length=20, do_sample=False)
    codes = model.generate(input_ids=input_ids, attention_mask=torch.ones_like(input_ids), prompt_input_ids=prompt_ids,
                           max_length=20, return_codes=True, do_sample=False)
    assert np.array_equal(codes.cpu().numpy(), res["codes"])


def test_generate_ragged_audio_and_out_of_range_codes():
    """Unsuppressed random-init generation emits EOS/unused ids: special ids route to the per-sample path with zero
    right-padding (modeling_parler_tts.py:2664-2680); an id >= codebook_size without any special id raises like the
    reference's codebook lookup would."""
    model, dims, sd, dac_sd = _composite()
    g = torch.Generator().manual_seed(12)
    input_ids, prompt_ids = torch.randint(2, 100, (3, 7), generator=g), torch.randint(2, 100, (3, 4), generator=g)
    res, _ = _oracle_generate_codes_only(model, dims, sd, input_ids, prompt_ids, max_length=24, do_sample=True,
                                         top_k=0, seed=1)
    codes = torch.from_numpy(res["codes"])
    special = bool(((codes == 33) | (codes == 32)).any())
    if special:
        want = dac_oracle.decode_generated(dac_sd, codes, 33, 32, 32, 32, DAC_TINY)
        audio = model.generate(input_ids=input_ids.to(DEV), prompt_input_ids=prompt_ids.to(DEV), do_sample=True,
                               top_k=0, seed=1, max_length=24)
        assert audio.shape == want.shape
        np.testing.assert_allclose(audio.cpu().numpy(), want.numpy(), atol=1e-3)
    eng = model.audio_encoder.engine()
    bad = torch.full((1, 4, 5), 40, dtype=torch.int64, device=DEV)
    with pytest.raises(IndexError):
        eng.decode_generated(bad, 33, 32, 32)


def _oracle_generate_codes_only(model, dims, sd, input_ids, prompt_ids, **kw):
    te = model.text_encoder.to("cpu").float()
    am = torch.ones_like(input_ids)
    with torch.no_grad():
        hs = te(input_ids=input_ids, attention_mask=am).last_hidden_state
    model.text_encoder.to(DEV)
    w, b = model.enc_to_dec_proj.weight.detach().cpu().float(), model.enc_to_dec_proj.bias.detach().cpu().float()
    enc = decoder_oracle.project_encoder(hs, am, w, b)
    prm = model.embed_prompts.weight.detach().cpu().float()[prompt_ids]
    return generate_oracle.generate_codes(sd, dims, enc, encoder_attention_mask=am, prompt_hidden_states=prm, **kw), None


def test_generate_argument_errors_match_reference():
    model, *_ = _composite()
    ids = torch.randint(2, 100, (1, 4)).to(DEV)
    with pytest.raises(ValueError):
        model.generate(input_ids=ids, prompt_input_ids=ids, num_beams=2)
    with pytest.raises(ValueError):
        model.generate(input_ids=ids, prompt_input_ids=ids, do_sample=False, num_return_sequences=2)
    with pytest.raises(ValueError):
        model.generate(input_ids=ids, prompt_input_ids=ids, max_length=30, min_length=40)


# ------------------------------------------------------------------------------------------------ DAC
def test_dac_decode_vs_transformers_fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "dac_tiny.npz"))
    c = z["cfg"]
    cfg = dict(n_codebooks=int(c[0]), codebook_size=int(c[1]), codebook_dim=int(c[2]), latent_dim=int(c[3]),
               decoder_dim=int(c[4]), rates=tuple(int(x) for x in c[5:]))
    sd = dac_oracle.random_dac_state(cfg, seed=31)
    for impl in (cabi.IMPL_SIMPLE, cabi.IMPL_FAST):
        eng = DacDecoderEngine(sd, DEV, cfg["rates"], impl)
        codes = torch.from_numpy(z["codes"]).to(DEV)
        zlat = eng.dequant(codes)
        np.testing.assert_allclose(zlat.cpu().numpy(), z["z"], atol=1e-4, rtol=1e-4)
        wav = eng.decode(codes).squeeze(1)
        assert wav.shape == z["wav"].shape
        np.testing.assert_allclose(wav.cpu().numpy(), z["wav"], atol=1e-3)


@pytest.mark.parametrize("impl", [cabi.IMPL_SIMPLE, cabi.IMPL_FAST])
def test_dac_44k_architecture_vs_oracle_waveform_1e3(impl):
    """Full 44.1 kHz decoder shape (1024 -> 1536 ... 96 -> 1, 9 codebooks): waveform within 1e-3 max-abs."""
    sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=2)
    eng = DacDecoderEngine(sd, DEV, impl=impl)
    g = torch.Generator().manual_seed(6)
    codes = torch.randint(0, 1024, (2, 9, 11), generator=g)
    want = dac_oracle.decode_codes(sd, codes).squeeze(1)
    got = eng.decode(codes.to(DEV)).squeeze(1)
    assert got.shape == (2, 512 * 11)
    assert float((got.cpu() - want).abs().max()) < 1e-3


@pytest.mark.parametrize("rates", [(4, 2, 8, 2), (2, 2, 2, 2), (8, 4)])
def test_dac_rates_read_from_the_checkpoint_vs_oracle(rates):
    """A codec whose up-sampling rates are not the 44.1 kHz model's: DACModel takes them from the checkpoint's transposed
    convs (modeling.py _dac_rates_of) and the decode equals the oracle's waveform."""
    cfg = dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=64, decoder_dim=64, rates=rates)
    sd = dac_oracle.random_dac_state(cfg, seed=5)
    codes = torch.randint(0, 32, (2, 3, 9), generator=torch.Generator().manual_seed(1))
    want = dac_oracle.decode_codes(sd, codes, cfg).squeeze(1)
    m = ptts.DACModel(ptts.DACConfig(num_codebooks=3, codebook_size=32, latent_dim=64), state_dict=sd).to(DEV)
    assert m.rates == rates
    for impl in (cabi.IMPL_SIMPLE, cabi.IMPL_FAST):
        got = m.engine(impl).decode(codes.to(DEV)).squeeze(1).cpu()
        assert got.shape == want.shape == (2, 9 * int(np.prod(rates)))
        assert float((got - want).abs().max()) < 1e-4


def test_dac_chunked_decode_equals_one_shot_decode():
    """Streaming form of the DAC decode: chunks decoded with 16 frames of real context on
    both sides concatenate to the one-shot decode (the decoder's receptive field is 10 frames, see the oracle test)."""
    sd = dac_oracle.random_dac_state(dac_oracle.DAC_44K, seed=3)
    eng = DacDecoderEngine(sd, DEV, impl=cabi.IMPL_FAST)
    codes = torch.randint(0, 1024, (3, 9, 150), generator=torch.Generator().manual_seed(9)).to(DEV)
    full = eng.decode(codes)
    chunks = list(eng.decode_chunks(codes, chunk_frames=64))
    assert [f0 for f0, _ in chunks] == [0, 64, 128]
    got = torch.cat([w for _, w in chunks], dim=-1)
    assert got.shape == full.shape
    assert float((got - full).abs().max()) < 2e-5  # same arithmetic per sample; tile boundaries may differ


def test_dac_ragged_path_vs_oracle():
    sd = dac_oracle.random_dac_state(DAC_TINY, seed=31)
    eng = DacDecoderEngine(sd, DEV, DAC_TINY["rates"], cabi.IMPL_SIMPLE)
    g = torch.Generator().manual_seed(8)
    codes = torch.randint(0, 32, (3, 4, 9), generator=g)
    codes[0, 2, 3] = 33  # bos inside -> frame dropped
    codes[1, :, 5:] = 32  # trailing pad frames
    codes[2] = 32  # nothing valid
    want = dac_oracle.decode_generated(sd, codes, 33, 32, 32, 32, DAC_TINY)
    got = eng.decode_generated(codes.to(DEV), 33, 32, 32)
    assert got.shape == want.shape
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), atol=1e-3)


# ------------------------------------------------------------------------------------ audio-prompt continuation
def _enc_state(cfg, encoder_dim):
    sd = dict(dac_oracle.random_dac_state(cfg, seed=31))
    sd.update(dac_oracle.random_dac_encoder_state(cfg, seed=31, encoder_dim=encoder_dim))
    return sd


def test_dac_encode_vs_transformers_fixture(golden_dir):
    """ptts_dac_conv1d_strided + ptts_dac_vq_encode against transformers' DacModel.encode: latents within 1e-4,
    codes identical wherever the reference's own best-vs-second margin exceeds 1e-4 (all of them here)."""
    from parler_tts_b200.dac import DacEncoderEngine

    z = np.load(os.path.join(golden_dir, "dac_enc_tiny.npz"))
    cfg = dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=128, decoder_dim=64, rates=(8, 8, 4, 2))
    for impl in (cabi.IMPL_SIMPLE, cabi.IMPL_FAST):
        eng = DacEncoderEngine(_enc_state(cfg, 8), DEV, cfg["rates"], impl)
        audio = torch.from_numpy(z["audio"]).to(DEV)
        lat = eng.encode_latents(audio)
        assert float((lat.cpu() - torch.from_numpy(z["z"])).abs().max()) < 1e-4
        codes = eng.quantize(lat).cpu().numpy()
        sure = z["margins"] > 1e-4
        assert sure.all() and np.array_equal(codes[sure], z["codes"][sure])
        # ragged length: zero right-padding to a hop multiple (DAC.preprocess)
        want = dac_oracle.encode_audio(_enc_state(cfg, 8), torch.from_numpy(z["audio"])[..., :4200], cfg)
        assert np.array_equal(eng.encode(audio[..., :4200]).cpu().numpy(), want.numpy())


def test_dac_encode_44k_architecture_vs_oracle():
    """Full 44.1 kHz encoder shape (1 -> 64 -> ... -> 1024, strides 2,4,8,8, 9 codebooks x 1024 x 8): latents within
    1e-3 of the fp32 oracle, codes identical wherever the oracle's margin exceeds 1e-3."""
    from parler_tts_b200.dac import DacEncoderEngine

    cfg = dac_oracle.DAC_44K
    sd = _enc_state(cfg, 64)
    audio = torch.randn(2, 1, 512 * 12, generator=torch.Generator().manual_seed(5)) * 0.3
    want_z = dac_oracle.encode_latents(sd, audio, cfg)
    want, margins = dac_oracle.rvq_encode(sd, want_z, cfg, return_margins=True)
    eng = DacEncoderEngine(sd, DEV, cfg["rates"])
    lat = eng.encode_latents(audio.to(DEV))
    assert float((lat.cpu() - want_z).abs().max()) < 1e-3 * max(1.0, float(want_z.abs().max()))
    # quantise the ORACLE's latents so that every codebook sees the same residual: index-exact up to near ties
    codes = eng.quantize(want_z.to(DEV)).cpu()
    sure = margins > 1e-3
    assert float(sure.float().mean()) > 0.95 and torch.equal(codes[sure], want[sure])


def test_generate_with_audio_prompt_vs_oracle():
    """generate(input_values=...) (modeling_parler_tts.py:2247-2292): the prompt is DAC-encoded on the CUDA path, its
    codes seed the delay pattern, and the continuation equals the oracle's greedy loop on the same codes."""
    model, dims, sd, dac_sd = _composite()
    enc_sd = dac_oracle.random_dac_encoder_state(DAC_TINY, seed=31, encoder_dim=8)
    full = dict(dac_sd)
    full.update(enc_sd)
    model.audio_encoder = ptts.DACModel(model.config.audio_encoder, state_dict=full, decoder_dim=64,
                                        codebook_dim=4).to(DEV)
    g = torch.Generator().manual_seed(21)
    input_ids, prompt_ids = torch.randint(2, 100, (2, 6), generator=g), torch.randint(2, 100, (2, 4), generator=g)
    audio = torch.randn(2, 1, 512 * 3, generator=g) * 0.4
    prompt_codes = dac_oracle.encode_audio(full, audio, DAC_TINY)  # (2, 4, 3)
    got_codes = model.audio_encoder.encode(audio.to(DEV)).audio_codes[0].cpu()
    assert torch.equal(got_codes, prompt_codes)
    kw = dict(max_length=24, do_sample=False, suppress_from=32)
    res, _ = _oracle_generate(model, dims, sd, dac_sd, input_ids, prompt_ids, decode=False,
                              decoder_input_ids=prompt_codes.reshape(2 * 4, 3).numpy(), **kw)
    codes = model.generate(input_ids=input_ids.to(DEV), prompt_input_ids=prompt_ids.to(DEV),
                           input_values=audio.to(DEV), max_length=24, do_sample=False, return_codes=True,
                           logits_processor=[ptts.SuppressTokensAtOrAbove(32)])
    assert np.array_equal(codes.cpu().numpy(), res["codes"])
