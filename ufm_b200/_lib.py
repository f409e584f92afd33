"""ctypes binding of ``libufm_b200.so`` (C ABI in ``include/ufm_b200.h``).

The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no fallback:
if it is missing, or no sm_100 GPU is present when an engine is created, the caller gets an exception.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libufm_b200.so")

UFM_F32, UFM_BF16, UFM_U8 = 0, 1, 2
UFM_NCHW, UFM_NHWC = 0, 1
GEMM_GELU, GEMM_RELU_BF16, GEMM_TAIL, GEMM_TAIL_SIGMOID, GEMM_TAIL_COORD = 1, 2, 4, 8, 16
ADAPT_MASK, ADAPT_COV_EXP_TANH, ADAPT_CONF_EXP, ADAPT_CONF_SIGMOID, ADAPT_CONF_SOFTMAX = 1, 2, 3, 4, 5
PROFILE_CATS = ("gemm", "conv3x3", "attention", "layernorm", "resize", "move", "refine")


class UfmConfig(C.Structure):
    _fields_ = [
        ("img_h", C.c_int), ("img_w", C.c_int), ("patch", C.c_int),
        ("enc_dim", C.c_int), ("enc_depth", C.c_int), ("enc_heads", C.c_int),
        ("enc_take_first", C.c_int), ("enc_take_last", C.c_int),
        ("info_dim", C.c_int), ("info_depth", C.c_int), ("info_heads", C.c_int), ("info_hidden", C.c_int),
        ("info_take0", C.c_int), ("info_take1", C.c_int),
        ("info_alternating", C.c_int),
        ("info_qk_norm", C.c_int), ("info_layerscale", C.c_int), ("info_has_proj", C.c_int),
        ("dpt_layer_dims", C.c_int * 4),
        ("dpt_feature_dim", C.c_int), ("dpt_hidden", C.c_int),
        ("flow_out_ch", C.c_int), ("unc_out_ch", C.c_int),
        ("flow_scale", C.c_float * 2), ("flow_shift", C.c_float * 2),
        ("refine", C.c_int),
        ("cls_in", C.c_int), ("cls_hidden", C.c_int), ("cls_out", C.c_int), ("cls_patch", C.c_int),
        ("refine_P", C.c_int),
        ("temperature", C.c_float),
        ("unc_n_adaptors", C.c_int), ("unc_kind", C.c_int * 4),
        ("unc_vmin", C.c_float * 4), ("unc_vmax", C.c_float * 4),
        ("unet", C.c_int),
        ("info_rope", C.c_int),
        ("flow_coord", C.c_int),
        ("enc_norm_intermediate", C.c_int), ("info_norm_intermediate", C.c_int),
        ("img_mean", C.c_float * 3), ("img_std", C.c_float * 3),
        ("hooks_set", C.c_int), ("dpt_hooks", C.c_int * 4), ("unc_hooks", C.c_int * 4),
        ("enc_registers", C.c_int),
    ]


# every symbol include/ufm_b200.h declares: name -> (restype, argtypes)
_P, _I, _L, _F, _U = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint32
SYMBOLS = {
    "ufm_create": (_I, [C.POINTER(UfmConfig), _I, C.POINTER(_P)]),
    "ufm_destroy": (None, [_P]),
    "ufm_last_error": (C.c_char_p, [_P]),
    "ufm_set_weight": (_I, [_P, C.c_char_p, _P, _I, C.POINTER(_L), _I]),
    "ufm_prepare": (_I, [_P, _I]),
    "ufm_forward": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P]),
    "ufm_forward_host": (_I, [_P, _P, _P, _I, _I, _I, _P, _P]),
    "ufm_set_keep_aux": (_I, [_P, _I]),
    "ufm_set_view_index": (_I, [_P, _I]),
    "ufm_stage_desc": (_I, [_P, C.c_char_p, C.POINTER(_I), C.POINTER(_L), C.POINTER(_I)]),
    "ufm_stage_copy": (_I, [_P, C.c_char_p, _P, _L]),
    "ufm_last_launch_count": (_I, [_P]),
    "ufm_set_profile": (_I, [_P, _I]),
    "ufm_profile_read": (_I, [_P, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(_L)]),
    "ufm_debug_flag": (_I, [_I, _I]),
    "ufm_op_gemm": (_I, [_P, _I, _I, _I, _I, _L, _L, _L, _I, _I, _P, _I, _I, _I, _P, _P, _P, _P, _L, _L, _L, _P, _P,
                         _L, _L, _L, _L, _U, _P, _P, _P, _P, _I, _P, _P, C.c_char_p, _I]),
    "ufm_op_attention": (_I, [_P, _I, _I, _I, _F, _P, _P, C.c_char_p, _I]),
    "ufm_op_layernorm": (_I, [_P, _I, _I, _I, _I, _I, _P, _P, _F, _I, _P, _L, _L, _P, _L, _L, _P]),
    "ufm_op_qk_norm": (_I, [_P, _L, _I, _P, _P, _P, _P, _F, _P]),
    "ufm_op_rope": (_I, [_P, _L, _I, _I, _I, _P]),
    "ufm_op_resize_ac": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "ufm_op_shuffle": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "ufm_op_im2col_s2": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "ufm_op_refine": (_I, [_P, _I, _I, _I, _P, _P, _F, _P, _P, _P, _P]),
    "ufm_op_prep_resize": (_I, [_P, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P]),
    "ufm_op_unmap": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P]),
    "ufm_op_conv3x3_s2_static": (_I, [_P, _I, _I, _I, _I, _P, _I, _P, _P, _P, C.c_char_p, _I]),
}
CONV3X3_S2_STATIC_MAX_BYTES = 16 << 20  # UFM_CONV3X3_S2_STATIC_MAX_BYTES of include/ufm_b200.h

_lib = None


def load():
    """Load the shared library (raises if it was not built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "ufm_b200 has no CPU or PyTorch fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
