"""Import path of /root/reference/parler_tts/dac_wrapper/configuration_dac.py."""
from parler_tts_b200.configuration import DACConfig  # noqa: F401
