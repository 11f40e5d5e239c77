"""Configuration objects with the reference's JSON layout (so existing `config.json` files load unchanged).

Field names and defaults follow /root/reference/parler_tts/configuration_parler_tts.py:82-127 (decoder),
:193-238 (composite) and /root/reference/parler_tts/dac_wrapper/configuration_dac.py:8-23 (DAC).
"""
from transformers import AutoConfig
from transformers.configuration_utils import PretrainedConfig

_DECODER_DEFAULTS = dict(
    vocab_size=2049,
    max_position_embeddings=2048,
    num_hidden_layers=24,
    ffn_dim=4096,
    num_attention_heads=16,
    layerdrop=0.0,
    use_cache=True,
    activation_function="gelu",
    hidden_size=1024,
    dropout=0.1,
    attention_dropout=0.0,
    activation_dropout=0.0,
    initializer_factor=0.02,
    scale_embedding=False,
    num_codebooks=4,
)

_DAC_DEFAULTS = dict(num_codebooks=9, model_bitrate=8, codebook_size=1024, latent_dim=1024, frame_rate=86,
                     sampling_rate=44100)


class DACConfig(PretrainedConfig):
    """`audio_encoder` section of a Parler-TTS checkpoint (model_type "dac")."""

    model_type = "dac"

    def __init__(self, **kwargs):
        for name, default in _DAC_DEFAULTS.items():
            setattr(self, name, kwargs.pop(name, default))
        super().__init__(**kwargs)


class ParlerTTSDecoderConfig(PretrainedConfig):
    """`decoder` section (model_type "parler_tts_decoder")."""

    model_type = "parler_tts_decoder"
    keys_to_ignore_at_inference = ["past_key_values"]

    def __init__(self, pad_token_id=2048, bos_token_id=2049, eos_token_id=2048, tie_word_embeddings=False, **kwargs):
        for name, default in _DECODER_DEFAULTS.items():
            setattr(self, name, kwargs.pop(name, default))
        super().__init__(pad_token_id=pad_token_id, bos_token_id=bos_token_id, eos_token_id=eos_token_id,
                         tie_word_embeddings=tie_word_embeddings, **kwargs)
        # transformers 5.x no longer stores these on every config; the kernels need them
        self.pad_token_id, self.bos_token_id, self.eos_token_id = pad_token_id, bos_token_id, eos_token_id


def _sub_config(section):
    section = dict(section)
    model_type = section.pop("model_type")
    if model_type == "dac" and ("latent_dim" in section or "model_bitrate" in section or not section):
        return DACConfig(**section)
    if model_type == "parler_tts_decoder":
        return ParlerTTSDecoderConfig(**section)
    return AutoConfig.for_model(model_type, **section)


class ParlerTTSConfig(PretrainedConfig):
    """Composite config: {vocab_size, text_encoder{...}, audio_encoder{...}, decoder{...}} (model_type "parler_tts")."""

    model_type = "parler_tts"
    is_composition = True
    has_no_defaults_at_init = True

    def __init__(self, vocab_size=1024, **kwargs):
        sections = {}
        for name in ("text_encoder", "audio_encoder", "decoder"):
            if name not in kwargs:
                raise ValueError("Config has to be initialized with text_encoder, audio_encoder and decoder config")
            sections[name] = kwargs.pop(name)
        # sub-configs first: PretrainedConfig.__init__ of transformers 5.x validates through to_dict()
        self.vocab_size = vocab_size
        self.text_encoder = _sub_config(sections["text_encoder"])
        self.audio_encoder = _sub_config(sections["audio_encoder"])
        dec = dict(sections["decoder"])
        dec.pop("model_type", None)
        self.decoder = ParlerTTSDecoderConfig(**dec)
        super().__init__(**kwargs)
        self.is_encoder_decoder = True

    @classmethod
    def from_sub_models_config(cls, text_encoder_config, audio_encoder_config, decoder_config, **kwargs):
        return cls(text_encoder=text_encoder_config.to_dict(), audio_encoder=audio_encoder_config.to_dict(),
                   decoder=decoder_config.to_dict(), **kwargs)

    def to_dict(self):
        out = super().to_dict()
        for name in ("text_encoder", "audio_encoder", "decoder"):
            sub = self.__dict__.get(name)
            if sub is None:
                continue
            d = sub.to_dict()
            d["model_type"] = sub.model_type
            out[name] = d
        out["model_type"] = self.model_type
        return out

    @property
    def sampling_rate(self):
        return self.audio_encoder.sampling_rate
