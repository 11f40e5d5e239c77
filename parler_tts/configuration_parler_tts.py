"""Import path of the reference module (/root/reference/parler_tts/configuration_parler_tts.py)."""
from parler_tts_b200.configuration import ParlerTTSConfig, ParlerTTSDecoderConfig  # noqa: F401

__all__ = ["ParlerTTSConfig", "ParlerTTSDecoderConfig"]
