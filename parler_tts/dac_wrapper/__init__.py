"""Import path of the reference sub-package (/root/reference/parler_tts/dac_wrapper/__init__.py:1-2)."""
from parler_tts_b200.configuration import DACConfig  # noqa: F401
from parler_tts_b200.modeling import DACModel  # noqa: F401

__all__ = ["DACConfig", "DACModel"]
