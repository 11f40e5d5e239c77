"""Import shim for the UNMODIFIED reference package at /root/reference (container only).

TEST INFRASTRUCTURE.  Used solely by oracle/gen_golden.py to produce the committed fixtures under
tests/golden/.  Nothing on the GPU box imports this (the reference does not travel).

Two things stop `import parler_tts` in this image:
  1. `from dac.model import DAC` (parler_tts/dac_wrapper/modeling_dac.py:7) - descript-audio-codec is absent;
  2. `AutoConfig.register("dac", ...)` (parler_tts/modeling_parler_tts.py:51, parler_tts/__init__.py:15) clashes
     with the "dac" model type that transformers 5.x ships.
The shim injects a stub `dac.model.DAC` and makes the two Auto registrations tolerant.  No reference
source is modified or copied.
"""
import sys
import types

REFERENCE_ROOT = "/root/reference"


def import_reference():
    import torch.nn as nn
    import transformers

    if "dac" not in sys.modules:
        dac = types.ModuleType("dac")
        dac_model = types.ModuleType("dac.model")

        class DAC(nn.Module):  # placeholder: the codec arithmetic is restated in oracle/dac_oracle.py
            def __init__(self, *a, **k):
                super().__init__()

        dac_model.DAC = DAC
        dac.model = dac_model
        sys.modules["dac"] = dac
        sys.modules["dac.model"] = dac_model

    def _tolerant(fn):
        def wrapped(*a, **k):
            try:
                return fn(*a, **k)
            except ValueError:
                return None
        return wrapped

    transformers.AutoConfig.register = staticmethod(_tolerant(transformers.AutoConfig.register))
    transformers.AutoModel.register = classmethod(
        lambda cls, *a, **k: None
    )
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import parler_tts  # noqa: the reference package

    assert parler_tts.__file__.startswith(REFERENCE_ROOT), parler_tts.__file__
    return parler_tts
