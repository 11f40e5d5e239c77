"""Import path of /root/reference/parler_tts/dac_wrapper/modeling_dac.py."""
from parler_tts_b200.modeling import DACModel  # noqa: F401
