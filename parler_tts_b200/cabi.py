"""ctypes binding of libparler_b200.so (the C ABI declared in include/parler_b200.h).

There is NO CPU fallback: if the shared library is missing or cannot be loaded, importing the compute path raises.
Struct layouts mirror include/parler_b200.h field by field.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PTTS_LIBRARY: another build of the same library (the diagnostic build of profiles/scripts, A/B runs of compile-time
# variants); default = the in-tree product build.  bench.py refuses to run with it set.
LIB_PATH = os.environ.get("PTTS_LIBRARY") or os.path.join(_HERE, "libparler_b200.so")

PTTS_F32, PTTS_BF16, PTTS_FP8 = 0, 1, 2
ACT = {"none": 0, "gelu": 1, "relu": 2, "silu": 3, "gelu_new": 4}
IMPL_SIMPLE, IMPL_FAST = 0, 1
MAX_CODEBOOKS = 16
ABI_VERSION = 2

c_i32, c_i64, c_f32, c_u64, c_vp = C.c_int32, C.c_int64, C.c_float, C.c_uint64, C.c_void_p


class PttsCtrl(C.Structure):
    _fields_ = [("len", c_i32), ("done", c_i32), ("stop_len", c_i32), ("n_unfinished", c_i32), ("ticket", c_i32),
                ("steps", c_i32), ("reserved", c_i32 * 2)]


class PttsLayer(C.Structure):
    _fields_ = [(n, c_vp) for n in ("w_qkv", "w_o", "w_qc", "w_kvc", "w_oc", "w_fc1", "w_fc2", "ln1_g", "ln1_b",
                                    "ln2_g", "ln2_b", "ln3_g", "ln3_b")]


class PttsModel(C.Structure):
    _fields_ = [(n, c_i32) for n in ("n_layers", "hidden", "n_heads", "head_dim", "ffn", "n_codebooks", "vocab",
                                     "max_pos", "wdtype", "act_fn", "pad_id", "eos_id", "bos_id", "reserved0")] + [
        ("embed", c_vp), ("pos", c_vp), ("lnf_g", c_vp), ("lnf_b", c_vp), ("w_heads", c_vp),
        ("layers", C.POINTER(PttsLayer))]


class PttsState(C.Structure):
    _fields_ = [("B", c_i32), ("S", c_i32), ("P", c_i32), ("T0", c_i32), ("c_max", c_i32), ("max_len", c_i32),
                ("impl", c_i32), ("do_sample", c_i32), ("top_k", c_i32), ("temperature", c_f32),
                ("suppress_from", c_i32), ("guidance_scale", c_f32), ("seed", c_u64), ("row_offset", c_i64),
                ("ctrl", c_vp), ("kv_self", c_vp), ("kv_cross", c_vp), ("enc_mask", c_vp), ("prompt_mask", c_vp),
                ("h", c_vp), ("xn", c_vp), ("q", c_vp), ("attn", c_vp), ("ffn", c_vp), ("logits", c_vp),
                ("scratch", c_vp), ("ids", c_vp), ("mask", c_vp), ("unfinished", c_vp),
                ("kv_dtype", c_i32), ("min_length", c_i32), ("top_p", c_f32), ("repetition_penalty", c_f32),
                ("logits_premixed", c_i32), ("self_attn_kernel", c_i32), ("kv_scale", c_vp), ("kv_stage", c_vp), ("pad_id", c_i64), ("eos_id", c_i64)]


class PttsT5Layer(C.Structure):
    _fields_ = [(n, c_vp) for n in ("w_qkv", "w_o", "w_wi", "w_wo", "ln0", "ln1")]


class PttsT5(C.Structure):
    _fields_ = [(n, c_i32) for n in ("n_layers", "d_model", "n_heads", "d_kv", "d_ff", "gated", "act_fn", "wdtype",
                                     "vocab", "n_buckets", "max_rel", "reserved0")] + [
        ("eps", c_f32), ("reserved1", c_f32), ("embed", c_vp), ("rel_bias", c_vp), ("bucket_of", c_vp),
        ("ln_f", c_vp), ("layers", C.POINTER(PttsT5Layer))]


_PM, _PS, _PT5 = C.POINTER(PttsModel), C.POINTER(PttsState), C.POINTER(PttsT5)

# name -> (restype, argtypes); must list every PTTS_API symbol of include/parler_b200.h
PROTOTYPES = {
    "ptts_abi_version": (c_i32, []),
    "ptts_last_error_string": (C.c_char_p, []),
    "ptts_scratch_floats": (c_i64, [_PM, c_i32]),
    "ptts_launch_count": (c_i64, []),
    "ptts_build_delay_mask": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_i32, c_i64, c_i64, c_vp, c_vp, c_vp]),
    "ptts_apply_delay_mask": (C.c_int, [c_vp, c_i64, c_vp, c_i64, c_i32, c_i32, c_vp, c_i64, c_vp]),
    "ptts_dedelay": (C.c_int, [c_vp, c_i64, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "ptts_filter_frames": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_i64, c_vp, c_vp, c_vp]),
    "ptts_embed_prefill": (C.c_int, [_PM, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "ptts_embed_tokens": (C.c_int, [_PM, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "ptts_layernorm": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_i32, c_f32, c_vp, c_i32, c_vp]),
    "ptts_linear": (C.c_int, [c_vp, c_i32, c_vp, c_i32, c_vp, c_i32, c_i32, c_i32, c_f32, c_i32, c_vp, c_vp, c_vp,
                              c_i32, c_i32, c_vp]),
    "ptts_attention": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_i32,
                                 c_vp, c_i32, c_vp, c_vp, c_i32, c_vp]),
    "ptts_attention_noalloc_max_floats": (c_i64, []),
    "ptts_attention_noalloc": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32,
                                         c_vp, c_vp]),
    "ptts_attention_fp8": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32,
                                     c_vp, c_vp, c_vp]),
    "ptts_attention_ws": (C.c_int, [c_vp, c_i32, c_i64, c_i32, c_f32, c_i32, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_i32,
                                    c_i32, c_i32, c_i32, c_vp, c_i32, c_vp, c_vp]),
    "ptts_kv_quantize_fp8": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "ptts_cross_kv": (C.c_int, [_PM, _PS, c_vp, c_vp]),
    "ptts_prefill": (C.c_int, [_PM, _PS, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp]),
    "ptts_decode_step": (C.c_int, [_PM, _PS, c_i32, c_vp]),
    "ptts_step_epilogue": (C.c_int, [_PM, _PS, c_vp]),
    "ptts_dac_dequant": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_i64, c_i64, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp]),
    "ptts_dac_conv1d": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32,
                                  c_vp, c_i32, c_vp]),
    "ptts_dac_convtr1d": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_i32, c_vp]),
    "ptts_dac_dequant_nwc": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_i64, c_i64, c_vp, c_i32, c_i32, c_vp, c_vp]),
    "ptts_dac_conv_out_nwc": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "ptts_dac_tile_n": (C.c_int, [c_i32]),
    "ptts_dac_set_cta_limit": (C.c_int, [c_i32]),
    "ptts_dac_conv1d_px": (C.c_int, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp,
                                     c_vp, c_vp, c_vp]),
    "ptts_dac_convtr1d_px": (C.c_int, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "ptts_dac_conv1d_strided": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp,
                                          c_vp]),
    "ptts_dac_vq_encode": (C.c_int, [c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                     c_vp, c_vp]),
    "ptts_t5_workspace_bytes": (c_i64, [_PT5, c_i32]),
    "ptts_t5_encode": (C.c_int, [_PT5, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_i32, c_vp]),
}

# include/parler_b200_diag.h: present only in libparler_b200_diag.so (make DIAG=1; profiles/scripts select it with
# PTTS_LIBRARY).  Bound when the loaded library exports them, never required.
DIAG_PROTOTYPES = {
    "ptts_diag_chain": (C.c_int, [c_i32, c_i32, c_i32, c_vp, c_vp]),
    "ptts_diag_gemm_stamps": (C.c_int, [c_vp, c_i32]),
    "ptts_diag_timeline": (C.c_int, [c_i32, c_vp, c_vp, c_i32]),
    "ptts_diag_xblock": (C.c_int, [c_vp, c_i32]),
}

_lib = None


class ParlerB200Error(RuntimeError):
    pass


def load():
    """Loads the shared library (once).  Raises if it is missing: the product path never falls back to the CPU."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ParlerB200Error(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C parler_tts_b200/csrc).  There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    for name, (res, args) in DIAG_PROTOTYPES.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.restype, fn.argtypes = res, args
    v = lib.ptts_abi_version()
    if v != ABI_VERSION:
        raise ParlerB200Error(f"ABI mismatch: library {v}, binding {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().ptts_last_error_string().decode("utf-8", "replace")
        raise ParlerB200Error(f"{what} failed (code {rc}): {msg}")


def call(name, *args):
    check(getattr(load(), name)(*args), name)


def ptr(t):
    """Device (or host) pointer of a torch tensor, or None."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(stream=None):
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def dtype_enum(dt):
    import torch

    if dt == torch.float32:
        return PTTS_F32
    if dt == torch.bfloat16:
        return PTTS_BF16
    raise ParlerB200Error(f"unsupported dtype {dt}: the CUDA path stores fp32 or bf16")
