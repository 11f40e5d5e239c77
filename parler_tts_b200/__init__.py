"""parler_tts_b200: B200-native (sm_100a) implementation of Parler-TTS's codec-token generation path.

Exports the same names as the reference package (/root/reference/parler_tts/__init__.py:4-12)."""
__version__ = "0.1"

from .configuration import DACConfig, ParlerTTSConfig, ParlerTTSDecoderConfig
from .modeling import (
    DACModel,
    ParlerTTSForCausalLM,
    ParlerTTSForConditionalGeneration,
    SuppressTokensAtOrAbove,
    apply_delay_pattern_mask,
    build_delay_pattern_mask,
)
from .streaming import AudioChunkStreamer

__all__ = [
    "AudioChunkStreamer", "DACConfig", "DACModel", "ParlerTTSConfig", "ParlerTTSDecoderConfig", "ParlerTTSForCausalLM",
    "ParlerTTSForConditionalGeneration", "SuppressTokensAtOrAbove", "apply_delay_pattern_mask",
    "build_delay_pattern_mask",
]
