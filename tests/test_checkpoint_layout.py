"""CPU: the checkpoint / config layout of the drop-in (SURVEY.md section 8b "Checkpoint / config layout (unchanged)")
against the reference's own classes (tests/golden/layout.json, written by oracle/gen_golden.py layout from the
unmodified /root/reference), and the save_pretrained -> from_pretrained round trip over that layout."""
import json
import os

import torch

import parler_tts_b200 as ptts
from oracle import dac_oracle
from oracle.gen_golden import TINY, TINY2


def _layout(golden_dir):
    with open(os.path.join(golden_dir, "layout.json")) as f:
        return json.load(f)


def test_decoder_state_dict_keys_and_shapes_are_the_references(golden_dir):
    """configuration_parler_tts.py / modeling_parler_tts.py:755-1094: `model.decoder.embed_tokens.{k}.weight`,
    `model.decoder.embed_positions.weights`, `model.decoder.layers.{i}.{self_attn,encoder_attn}.{q,k,v,out}_proj.weight`,
    the three LayerNorms per layer, `fc1` / `fc2`, `model.decoder.layer_norm`, `lm_heads.{k}.weight` - names and shapes."""
    want = _layout(golden_dir)["decoder_state_dict"]
    for name, kw in (("tiny", TINY), ("tiny2", TINY2)):
        m = ptts.ParlerTTSForCausalLM(ptts.ParlerTTSDecoderConfig(**kw))
        got = {k: list(v.shape) for k, v in m.state_dict().items()}
        assert got == want[name], (name, sorted(set(got) ^ set(want[name]))[:6])


def test_config_defaults_are_the_references(golden_dir):
    """ParlerTTSDecoderConfig() / DACConfig() (configuration_parler_tts.py:82-127, dac_wrapper/configuration_dac.py:8-23):
    every field of the reference's `to_dict()` exists here with the same default."""
    want = _layout(golden_dir)["config_defaults"]
    for key, cfg in (("decoder", ptts.ParlerTTSDecoderConfig()), ("dac", ptts.DACConfig())):
        got = json.loads(json.dumps(cfg.to_dict(), default=str))
        ref = want[key]
        missing = sorted(set(ref) - set(got))
        assert not missing, (key, missing)
        diff = {k: (ref[k], got[k]) for k in ref if ref[k] != got[k]}
        assert not diff, (key, diff)
    assert ptts.ParlerTTSDecoderConfig.model_type == "parler_tts_decoder" and ptts.DACConfig.model_type == "dac"
    assert ptts.ParlerTTSConfig.model_type == "parler_tts"


def test_save_pretrained_from_pretrained_round_trip(tmp_path):
    """config.json with the three sub-configs under `text_encoder` / `audio_encoder` / `decoder`, generation_config.json,
    one safetensors file with the reference's key prefixes; loading it back gives the same tensors, and the codec's
    up-sampling rates come from the checkpoint."""
    from transformers import T5Config

    dec_cfg = ptts.ParlerTTSDecoderConfig(**TINY)
    t5 = T5Config(d_model=64, num_layers=2, d_ff=128, num_heads=2, d_kv=32, vocab_size=100)
    dac_cfg = ptts.DACConfig(num_codebooks=4, codebook_size=32, latent_dim=128)
    cfg = ptts.ParlerTTSConfig.from_sub_models_config(t5, dac_cfg, dec_cfg, vocab_size=100)
    rates = (4, 2, 8, 2)
    dac_sd = dac_oracle.random_dac_state(dict(n_codebooks=4, codebook_size=32, codebook_dim=4, latent_dim=128,
                                              decoder_dim=64, rates=rates), seed=3)
    torch.manual_seed(0)
    m = ptts.ParlerTTSForConditionalGeneration(cfg, decoder=ptts.ParlerTTSForCausalLM(dec_cfg),
                                               audio_encoder=ptts.DACModel(dac_cfg, state_dict=dac_sd))
    m.generation_config.max_length = 77
    sd = m.state_dict()
    assert {k.split(".")[0] for k in sd} == {"text_encoder", "audio_encoder", "decoder", "enc_to_dec_proj", "embed_prompts"}
    assert sd["enc_to_dec_proj.weight"].shape == (128, 64) and "enc_to_dec_proj.bias" in sd       # modeling_parler_tts.py:1639
    assert sd["embed_prompts.weight"].shape == (100, 128)                                         # :1642
    assert "decoder.model.decoder.layers.1.encoder_attn.out_proj.weight" in sd and "decoder.lm_heads.3.weight" in sd
    assert "audio_encoder.model.decoder.model.1.block.1.weight_v" in sd                           # descript keys, verbatim
    m.save_pretrained(str(tmp_path))
    assert {"config.json", "generation_config.json", "model.safetensors"} <= set(os.listdir(tmp_path))
    c = json.load(open(tmp_path / "config.json"))
    assert c["model_type"] == "parler_tts" and c["is_encoder_decoder"] is True and c["vocab_size"] == 100
    assert c["text_encoder"]["model_type"] == "t5" and c["audio_encoder"]["model_type"] == "dac"
    assert c["decoder"]["model_type"] == "parler_tts_decoder" and c["decoder"]["num_codebooks"] == TINY["num_codebooks"]
    assert json.load(open(tmp_path / "generation_config.json"))["max_length"] == 77
    m2 = ptts.ParlerTTSForConditionalGeneration.from_pretrained(str(tmp_path))
    sd2 = m2.state_dict()
    assert set(sd) == set(sd2) and all(torch.equal(sd[k], sd2[k]) for k in sd)
    assert m2.audio_encoder.rates == rates and m2.generation_config.max_length == 77
    assert m2.config.decoder.to_dict() == m.config.decoder.to_dict()


def test_model_init_script_flow_from_sub_models_pretrained(tmp_path):
    """The way the reference builds its random-init checkpoints (helpers/model_init_scripts/init_model_600M.py:24-70 -
    the source of BASELINE.json's "random-init" configurations): a decoder from its config, saved alone; the composite
    assembled from three checkpoints by `from_sub_models_pretrained(..., vocab_size=...)`; generation defaults set on
    it; saved; loaded."""
    import pytest
    from transformers import T5Config, T5EncoderModel

    t5_dir, dac_dir, dec_dir, out_dir = (str(tmp_path / n) for n in ("t5", "dac", "decoder", "parler"))
    torch.manual_seed(1)
    T5EncoderModel(T5Config(d_model=64, num_layers=2, d_ff=128, num_heads=2, d_kv=32, vocab_size=100)).save_pretrained(t5_dir)
    dac_cfg = ptts.DACConfig(num_codebooks=4, codebook_size=32, latent_dim=128)
    dac_sd = dac_oracle.random_dac_state(dict(n_codebooks=4, codebook_size=32, codebook_dim=4, latent_dim=128,
                                              decoder_dim=64, rates=(8, 8, 4, 2)), seed=3)
    ptts.DACModel(dac_cfg, state_dict=dac_sd).save_pretrained(dac_dir)
    assert all(k.startswith("model.") for k in ptts.DACModel.from_pretrained(dac_dir).state_dict())

    decoder_config = ptts.ParlerTTSDecoderConfig(**dict(TINY, vocab_size=32 + 64, pad_token_id=32, eos_token_id=32,
                                                        bos_token_id=33))
    decoder = ptts.ParlerTTSForCausalLM(decoder_config)
    decoder.save_pretrained(dec_dir)
    back = ptts.ParlerTTSForCausalLM.from_pretrained(dec_dir)
    assert back.config.to_dict() == decoder.config.to_dict()
    assert all(torch.equal(v, back.state_dict()[k]) for k, v in decoder.state_dict().items())

    model = ptts.ParlerTTSForConditionalGeneration.from_sub_models_pretrained(
        text_encoder_pretrained_model_name_or_path=t5_dir, audio_encoder_pretrained_model_name_or_path=dac_dir,
        decoder_pretrained_model_name_or_path=dec_dir, vocab_size=100)
    assert model.config.vocab_size == 100 and model.config.decoder.num_codebooks == TINY["num_codebooks"]
    assert model.embed_prompts.weight.shape == (100, TINY["hidden_size"])
    assert model.enc_to_dec_proj.weight.shape == (TINY["hidden_size"], 64)
    assert all(torch.equal(v, model.decoder.state_dict()[k]) for k, v in decoder.state_dict().items())
    model.generation_config.decoder_start_token_id = 33
    model.generation_config.pad_token_id = 32
    model.generation_config.eos_token_id = 32
    model.generation_config.max_length = int(30 * model.audio_encoder.config.frame_rate)
    model.generation_config.do_sample = True
    model.generation_config.guidance_scale = 1
    model.config.pad_token_id = 32
    model.config.decoder_start_token_id = 33
    model.save_pretrained(out_dir)
    again = ptts.ParlerTTSForConditionalGeneration.from_pretrained(out_dir)
    assert again.generation_config.max_length == 2580 and again.generation_config.do_sample is True
    assert again.generation_config.decoder_start_token_id == 33
    assert again.config.pad_token_id == 32 and again.config.decoder_start_token_id == 33
    sd, sd2 = model.state_dict(), again.state_dict()
    assert set(sd) == set(sd2) and all(torch.equal(sd[k], sd2[k]) for k in sd)
    # a composite checkpoint also serves as the decoder's source (modeling_parler_tts.py:1893-1894)
    dec2 = ptts.ParlerTTSForCausalLM.from_pretrained(out_dir)
    assert all(torch.equal(v, dec2.state_dict()[k]) for k, v in decoder.state_dict().items())
    # models handed over directly, and the reference's errors for a missing source
    m3 = ptts.ParlerTTSForConditionalGeneration.from_sub_models_pretrained(
        text_encoder_model=model.text_encoder, audio_encoder_model=model.audio_encoder, decoder_model=decoder, vocab_size=100)
    assert m3.decoder is decoder
    with pytest.raises(ValueError, match="text_encoder_pretrained_model_name_or_path"):
        ptts.ParlerTTSForConditionalGeneration.from_sub_models_pretrained(decoder_pretrained_model_name_or_path=dec_dir)
    with pytest.raises(ValueError, match="decoder_pretrained_model_name_or_path"):
        ptts.ParlerTTSForConditionalGeneration.from_sub_models_pretrained(
            text_encoder_pretrained_model_name_or_path=t5_dir, audio_encoder_pretrained_model_name_or_path=dac_dir)
