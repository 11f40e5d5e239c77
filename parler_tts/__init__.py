"""Drop-in import name of the reference package: `from parler_tts import ParlerTTSForConditionalGeneration` resolves to
the B200-native implementation in `parler_tts_b200` (same exported names as /root/reference/parler_tts/__init__.py:4-16).

The reference also registers its DAC wrapper with transformers' Auto classes (`AutoConfig.register("dac", DACConfig)`,
`AutoModel.register(DACConfig, DACModel)`, __init__.py:15-16).  Newer transformers releases ship their own "dac" model
type, where that registration raises; it is attempted here and skipped when the name is already taken.
"""
from parler_tts_b200 import (  # noqa: F401
    DACConfig,
    DACModel,
    ParlerTTSConfig,
    ParlerTTSDecoderConfig,
    ParlerTTSForCausalLM,
    ParlerTTSForConditionalGeneration,
    SuppressTokensAtOrAbove,
    __version__,
    apply_delay_pattern_mask,
    build_delay_pattern_mask,
)

try:  # pragma: no cover - depends on the installed transformers
    from transformers import AutoConfig, AutoModel

    AutoConfig.register("dac", DACConfig)
    AutoModel.register(DACConfig, DACModel)
except Exception:  # "dac" already belongs to transformers' own DacConfig
    pass

__all__ = [
    "DACConfig", "DACModel", "ParlerTTSConfig", "ParlerTTSDecoderConfig", "ParlerTTSForCausalLM",
    "ParlerTTSForConditionalGeneration", "apply_delay_pattern_mask", "build_delay_pattern_mask",
]
