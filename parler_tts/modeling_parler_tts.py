"""Import path of the reference module (/root/reference/parler_tts/modeling_parler_tts.py): the names of it that are on
the generation path, served by parler_tts_b200.  The module-internal building blocks of the reference (attention /
decoder-layer / positional-embedding nn.Modules) have no counterpart: that work lives in the CUDA library."""
from parler_tts_b200.modeling import (  # noqa: F401
    ParlerTTSForCausalLM,
    ParlerTTSForConditionalGeneration,
    apply_delay_pattern_mask,
    build_delay_pattern_mask,
)

__all__ = ["ParlerTTSForCausalLM", "ParlerTTSForConditionalGeneration", "apply_delay_pattern_mask",
           "build_delay_pattern_mask"]
