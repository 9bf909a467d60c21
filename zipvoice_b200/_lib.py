"""ctypes binding of libzipvoice_b200.so (C ABI in include/zipvoice_b200.h).

The library is the product: if it is missing or fails to load we raise -- there is no
Python / PyTorch fallback for any of its entry points.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libzipvoice_b200.so")
ZV_MAX_STACKS = 8


class ZvConfig(C.Structure):
    _fields_ = [
        ("feat_dim", C.c_int),
        ("fm_num_stacks", C.c_int),
        ("fm_downsampling_factor", C.c_int * ZV_MAX_STACKS),
        ("fm_num_layers", C.c_int * ZV_MAX_STACKS),
        ("fm_cnn_module_kernel", C.c_int * ZV_MAX_STACKS),
        ("fm_feedforward_dim", C.c_int), ("fm_num_heads", C.c_int), ("fm_dim", C.c_int),
        ("text_num_layers", C.c_int), ("text_feedforward_dim", C.c_int), ("text_cnn_module_kernel", C.c_int),
        ("text_num_heads", C.c_int), ("text_dim", C.c_int),
        ("query_head_dim", C.c_int), ("value_head_dim", C.c_int), ("pos_head_dim", C.c_int), ("pos_dim", C.c_int),
        ("time_embed_dim", C.c_int), ("text_embed_dim", C.c_int),
        ("vocab_size", C.c_int), ("pad_id", C.c_int),
        ("use_guidance_scale_embed", C.c_int), ("dialog", C.c_int), ("spk_a_id", C.c_int), ("spk_b_id", C.c_int),
        ("stereo", C.c_int),
    ]


_P, _I, _F, _D, _LL = C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_longlong
_SIGS = {
    "zv_create": (_I, [C.POINTER(ZvConfig), C.POINTER(_P)]),
    "zv_destroy": (None, [_P]),
    "zv_last_error": (C.c_char_p, [_P]),
    "zv_set_weight": (_I, [_P, C.c_char_p, _P, C.POINTER(C.c_int64), _I, _I]),
    "zv_finalize_weights": (_I, [_P]),
    "zv_set_option": (_I, [_P, C.c_char_p, _I]),
    "zv_time_steps": (_I, [_D, _D, _I, _D, C.POINTER(_F)]),
    "zv_velocity": (_I, [_P, _I, _F, _P, _P, _P, _P, _I, _I, C.POINTER(_F), _I, _P, _P]),
    "zv_forward_fm_decoder": (_I, [_P, C.POINTER(_F), _I, _P, _P, _P, _P, _I, _I, C.POINTER(_F), _I, _P, _P]),
    "zv_sample": (_I, [_P, _I, _P, _P, _P, _P, _I, _I, _I, C.POINTER(_F), _I, _D, _D, _D, _P, _P]),
    "zv_text_embed": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "zv_text_condition": (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P, _P]),
    "zv_fbank": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "zv_fbank_bigvgan": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "zv_vocos_decode": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "zv_resample": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _I, _P]),
    "zv_resample_static": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _I, _P]),
    "zv_prof_enable": (_I, [_P, _I]),
    "zv_prof_num_classes": (_I, []),
    "zv_prof_class_name": (C.c_char_p, [_I]),
    "zv_prof_class_is_tensor": (_I, [_I]),
    "zv_prof_read": (_I, [_P, C.POINTER(_D), C.POINTER(_D), C.POINTER(_LL)]),
    "zv_launch_count": (_LL, []),
    "zv_launch_count_reset": (None, []),
}
# private test / bench hooks (zipvoice_b200/csrc/zv_debug.h), not part of the reference-facing surface
_DEBUG_SIGS = {
    "zv_debug_gemm": (_I, [_I, _I, _I, _I, _I, _I, _I] + [_P] * 12),
    "zv_debug_attention": (_I, [_I, _I, _I, _I, _I, _P, _I, _P, _P, _P, _I, _P, _P, _P]),
    "zv_debug_set_attn_impl": (_I, [_I]),
    "zv_debug_attention_ksplit": (_I, [_I]),
    "zv_debug_attention_bench": (_I, [_I] * 8 + [C.POINTER(_F)]),
    "zv_debug_gemm_bench": (_I, [_I] * 9 + [C.POINTER(_F)]),
    "zv_debug_pos_table": (_I, [_P, _I, _I, _I, _I, _P, _P]),
}
EXPORTS = tuple(_SIGS)
DEBUG_EXPORTS = tuple(_DEBUG_SIGS)

_lib = None


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (and, in a source checkout with nvcc, build) the shared library."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise RuntimeError(f"{LIB_PATH} is missing; run `python -m zipvoice_b200.build`")
        from . import build as _build
        _build.build()
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in list(_SIGS.items()) + list(_DEBUG_SIGS.items()):
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def loaded_path() -> str:
    return LIB_PATH
