"""CPU: the C-ABI shared library loads and exports every symbol that include/parler_b200.h declares."""
import os
import re

import pytest

from parler_tts_b200 import cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "parler_b200.h")).read()
    return sorted(set(re.findall(r"PTTS_API\s+[\w\s\*]+?\b(ptts_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    names = _declared()
    assert len(names) >= 18
    assert sorted(cabi.PROTOTYPES) == names


def test_library_exports_every_declared_symbol():
    if not os.path.exists(cabi.LIB_PATH):
        pytest.fail(f"{cabi.LIB_PATH} missing: run __graft_entry__.build()")
    lib = cabi.load()
    for name in _declared():
        assert hasattr(lib, name), name
    assert lib.ptts_abi_version() == cabi.ABI_VERSION
    assert lib.ptts_last_error_string() is not None


def test_struct_layout_matches_header():
    import ctypes as C

    assert C.sizeof(cabi.PttsCtrl) == 32
    assert C.sizeof(cabi.PttsLayer) == 13 * 8
    assert cabi.PttsModel.embed.offset == 56 and C.sizeof(cabi.PttsModel) == 56 + 6 * 8
    assert cabi.PttsState.seed.offset == 48 and cabi.PttsState.ctrl.offset == 64
    assert cabi.PttsState.kv_dtype.offset == 64 + 15 * 8 and cabi.PttsState.kv_scale.offset == 64 + 15 * 8 + 24
    assert C.sizeof(cabi.PttsState) == 64 + 15 * 8 + 24 + 4 * 8


def test_product_library_has_no_knobs_and_no_diag_exports():
    """The shipped library reads no PTTS_* environment variable and exports none of include/parler_b200_diag.h."""
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", cabi.LIB_PATH], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert not [n for n in exported if n.startswith("ptts_diag")], exported
    assert exported == set(cabi.PROTOTYPES), exported ^ set(cabi.PROTOTYPES)
    blob = open(cabi.LIB_PATH, "rb").read()
    assert b"getenv" not in blob or b"PTTS_SKIP" not in blob
    for knob in (b"PTTS_SKIP", b"PTTS_GEMM", b"PTTS_XB_", b"PTTS_L2PF", b"PTTS_TIMELINE", b"PTTS_PDL"):
        assert knob not in blob, knob


def test_reference_import_name_resolves():
    import parler_tts
    import parler_tts_b200

    for name in ("ParlerTTSForConditionalGeneration", "ParlerTTSForCausalLM", "ParlerTTSConfig", "ParlerTTSDecoderConfig",
                 "DACConfig", "DACModel", "build_delay_pattern_mask", "apply_delay_pattern_mask"):
        assert getattr(parler_tts, name) is getattr(parler_tts_b200, name), name
    # the reference's module paths resolve too (parler_tts/modeling_parler_tts.py, configuration_parler_tts.py, dac_wrapper/)
    from parler_tts.configuration_parler_tts import ParlerTTSConfig, ParlerTTSDecoderConfig
    from parler_tts.dac_wrapper import DACConfig, DACModel
    from parler_tts.dac_wrapper.configuration_dac import DACConfig as C2
    from parler_tts.dac_wrapper.modeling_dac import DACModel as M2
    from parler_tts.modeling_parler_tts import (ParlerTTSForCausalLM, ParlerTTSForConditionalGeneration,
                                                apply_delay_pattern_mask, build_delay_pattern_mask)

    assert ParlerTTSConfig is parler_tts_b200.ParlerTTSConfig and ParlerTTSDecoderConfig is parler_tts_b200.ParlerTTSDecoderConfig
    assert DACConfig is C2 is parler_tts_b200.DACConfig and DACModel is M2 is parler_tts_b200.DACModel
    assert ParlerTTSForCausalLM is parler_tts_b200.ParlerTTSForCausalLM
    assert ParlerTTSForConditionalGeneration is parler_tts_b200.ParlerTTSForConditionalGeneration
    assert build_delay_pattern_mask is parler_tts_b200.build_delay_pattern_mask
    assert apply_delay_pattern_mask is parler_tts_b200.apply_delay_pattern_mask


def test_argument_errors_are_reported_without_a_gpu():
    lib = cabi.load()
    rc = lib.ptts_build_delay_mask(None, 1, 9, 1, 20, 1025, 1024, None, None, None)
    assert rc == 1
    assert b"null" in lib.ptts_last_error_string()


def test_transposed_conv_rejects_strides_it_cannot_launch():
    """descript's 16 / 24 kHz codecs use a stride of 5, and 16 would exceed the default dynamic shared memory: both are
    argument errors with a message, not launch failures (checked before anything touches the device)."""
    import ctypes

    lib = cabi.load()
    dummy = ctypes.c_void_p(256)  # never dereferenced: the stride check comes first
    for stride in (5, 16, 3, 1, 0):
        rc = lib.ptts_dac_convtr1d(dummy, 1, 32, 4, dummy, None, 16, stride, None, dummy, cabi.IMPL_SIMPLE, None)
        assert rc == 1, stride
        assert b"stride must be 2, 4 or 8" in lib.ptts_last_error_string()


def test_no_cpu_fallback_in_public_api():
    import torch

    import parler_tts_b200 as p

    if torch.cuda.is_available():
        pytest.skip("only meaningful on a CPU-only box")
    with pytest.raises(cabi.ParlerB200Error):
        p.build_delay_pattern_mask(torch.full((9, 1), 1025), 1025, 1024, 20, 9)
