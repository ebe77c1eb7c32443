"""ctypes binding of libb200net.so (the C-ABI declared in include/b200net.h).

There is no fallback: if the library is missing the import fails loudly, and every non-zero status becomes a
RuntimeError carrying b200_last_error() -- the same way torch surfaces kernel errors to the reference's scripts.
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int32, c_int64, c_uint64, c_uint8, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200net.so")

F32, BF16 = 0, 1
CONV_FPROP, CONV_DGRAD, CONV_WGRAD = 0, 1, 2

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "b200net: %s not found. Build it first: python -c 'import __graft_entry__ as g; g.build()' "
        "(nvcc, sm_100a). There is no CPU or cuDNN fallback." % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)


class ConvDesc(Structure):
    _fields_ = [("N", c_int32), ("H", c_int32), ("W", c_int32), ("C", c_int32), ("K", c_int32), ("R", c_int32),
                ("S", c_int32), ("stride", c_int32), ("pad", c_int32), ("dtype", c_int32)]


P = c_void_p
I32, I64, F, U64 = c_int32, c_int64, c_float, c_uint64
_PP = POINTER(c_void_p)
_PI64 = POINTER(c_int64)

# name -> (restype, argtypes); every symbol declared in include/b200net.h appears here (tests check the two agree)
SIGNATURES = {
    "b200_version": (I32, []),
    "b200_last_error": (c_char_p, []),
    "b200_launch_count": (I64, []),
    "b200_gemm_timing_enable": (I32, [I32]),
    "b200_gemm_timing_read": (I32, [POINTER(ctypes.c_double), POINTER(c_int64), POINTER(ctypes.c_double)]),
    "b200_nchw_to_nhwc": (I32, [P, P, I32, I32, I32, I32, I32, I32, P]),
    "b200_nhwc_to_nchw": (I32, [P, P, I32, I32, I32, I32, I32, P]),
    "b200_conv_weight_ld": (I32, [I32, I32, I32]),
    "b200_pack_weight": (I32, [P, P, I32, I32, I32, I32, I32, I32, I32, P]),
    "b200_pack_weight_multi": (I32, [I32, _PP, _PP, POINTER(c_int32), POINTER(c_int32), POINTER(c_int32),
                                     POINTER(c_int32), POINTER(c_int32), I32, P]),
    "b200_im2col": (I32, [P, P, I32, I32, I32, I32, I32, I32, I32, I32, I32, I32, P]),
    "b200_stem_stage": (I32, [P, P, I32, I32, I32, I32, I32, I32, P]),
    "b200_stem_stage_nchw": (I32, [P, P, I32, I32, I32, I32, I32, I32, I32, P]),
    "b200_stem_pack_weight": (I32, [P, P, I32, I32, I32, I32, P]),
    "b200_preprocess_u8": (I32, [P, P, P, P, I32, I32, I32, P, I32, I32, P]),
    "b200_cast": (I32, [P, P, I64, I32, I32, P]),
    "b200_conv_workspace_bytes": (I64, [POINTER(ConvDesc), I32]),
    "b200_conv_fprop": (I32, [POINTER(ConvDesc), P, P, P, P, I64, P]),
    "b200_conv_fprop_bnstats_rows": (I64, [POINTER(ConvDesc)]),
    "b200_conv_fprop_bnstats": (I32, [POINTER(ConvDesc), P, P, P, P, I64, P]),
    "b200_conv_fprop_affine": (I32, [POINTER(ConvDesc), P, P, P, P, P, P, I32, P]),
    "b200_conv_dgrad": (I32, [POINTER(ConvDesc), P, P, P, I32, P, I64, P]),
    "b200_conv_wgrad": (I32, [POINTER(ConvDesc), P, P, P, P, I64, P]),
    "b200_gemm_f32": (I32, [P, P, P, I32, I32, I32, I64, I64, I64, I32, I32, P, I32, I32, P]),
    "b200_gemm_bf16": (I32, [P, P, P, I32, I32, I32, I64, I64, I64, I32, I32, I32, I32, I32, P, I64, P]),
    "b200_bn_workspace_bytes": (I64, [I64, I32]),
    "b200_bn_train_stats": (I32, [P, I64, I32, I32, P, P, P, P, F, F, P, P, P, P, P, I64, P]),
    "b200_bn_finalize_stats": (I32, [P, I64, I64, I32, P, P, P, P, F, F, P, P, P, P, P]),
    "b200_bn_train_stats_noalloc": (I32, [P, I64, I32, I32, P, P, P, P, F, F, P, P, P, P, P]),
    "b200_bn_eval_coeffs": (I32, [P, P, P, P, F, P, P, I32, P]),
    "b200_bn_mask_bytes": (I64, [I64, I32, I32]),
    "b200_bn_apply": (I32, [P, P, P, P, I64, I32, I32, P, P, I32, I32, P]),
    "b200_bn_backward": (I32, [P, P, P, P, P, P, I64, I32, I32, P, P, P, I32, P, P, P, I64, I32, P]),
    "b200_add": (I32, [P, P, P, I64, I32, P]),
    "b200_relu_bwd": (I32, [P, P, P, I64, I32, P]),
    "b200_maxpool3x3s2_fwd": (I32, [P, P, P, I32, I32, I32, I32, I32, P]),
    "b200_maxpool3x3s2_bwd": (I32, [P, P, P, I32, I32, I32, I32, I32, P]),
    "b200_bn_relu_maxpool_fwd": (I32, [P, P, P, P, P, I32, I32, I32, I32, I32, P]),
    "b200_maxpool_bn_bwd": (I32, [P, P, P, P, P, I32, I32, I32, I32, I32, P, P, P, P, P, P, I64, P]),
    "b200_maxpool3x3s2_bwd_rows": (I32, [P, P, P, I32, I32, I32, I32, I32, P]),
    "b200_wgrad_defer": (I32, [I32]),
    "b200_wgrad_flush": (I32, [P]),
    "b200_avgpool_fwd": (I32, [P, P, I32, I32, I32, I32, P]),
    "b200_avgpool_bwd": (I32, [P, P, I32, I32, I32, I32, P]),
    "b200_relu_fwd_f32": (I32, [P, P, I64, P]),
    "b200_dropout_fwd": (I32, [P, P, P, I64, F, U64, U64, P]),
    "b200_dropout_bwd": (I32, [P, P, P, I64, F, P]),
    "b200_log_softmax_fwd": (I32, [P, P, I32, I32, P]),
    "b200_log_softmax_bwd": (I32, [P, P, P, I32, I32, P]),
    "b200_nll_loss_fwd": (I32, [P, P, P, I32, I32, P]),
    "b200_nll_loss_bwd": (I32, [P, P, P, I32, I32, P]),
    "b200_colsum_f32": (I32, [P, P, I32, I32, P]),
    "b200_multi_tensor_sgd": (I32, [I32, _PP, _PP, _PP, _PI64, F, F, F, F, I32, I32, F, P]),
    "b200_multi_tensor_adam": (I32, [I32, _PP, _PP, _PP, _PP, _PI64, F, F, F, F, F, I32, I64, F, P]),
    "b200_l2norm_workspace_bytes": (I64, []),
    "b200_multi_tensor_l2norm_sq": (I32, [I32, _PP, _PI64, P, P, I64, P]),
    "b200_multi_tensor_clip_scale": (I32, [I32, _PP, _PI64, P, F, P]),
    "b200_multi_tensor_zero": (I32, [I32, _PP, _PI64, P]),
    "b200_bucket_pack": (I32, [I32, _PP, _PI64, P, F, P]),
    "b200_bucket_unpack": (I32, [I32, _PP, _PI64, P, F, P]),
    "b200_multi_tensor_copy": (I32, [I32, _PP, _PP, _PI64, F, P]),
    "b200_dropout_counter": (I32, [P]),
    "b200_embed_ln_fwd": (I32, [P, P, P, P, P, P, P, F, I32, I32, I32, I32, I32, F, U64, U64, P, P, P, P, P, I32, P]),
    "b200_embed_bwd": (I32, [P, P, P, I32, I32, I32, I32, I32, P, P, P, P, I64, I32, P]),
    "b200_ln_fwd": (I32, [P, P, P, P, P, P, P, P, F, I64, I32, F, U64, U64, I32, P]),
    "b200_ln_bwd_workspace_bytes": (I64, [I64, I32]),
    "b200_ln_bwd": (I32, [P, P, P, P, P, P, I64, I32, F, U64, U64, I32, P, P, P, P, P, P, I64, I32, P]),
    "b200_gelu_fwd": (I32, [P, P, I64, I32, P]),
    "b200_gelu_bwd": (I32, [P, P, P, I64, I32, P]),
    "b200_gelu_bwd_bias": (I32, [P, P, P, P, I64, I32, I32, P, I64, P]),
    "b200_colsum_workspace_bytes": (I64, [I64, I32]),
    "b200_colsum": (I32, [P, P, I64, I32, I32, P, I64, P]),
    "b200_attention_fwd": (I32, [P, P, P, P, I32, I32, I32, I32, F, U64, U64, I32, P]),
    "b200_attention_bwd": (I32, [P, P, P, P, P, I32, I32, I32, I32, F, U64, U64, I32, P]),
    "b200_attention_flash_supported": (I32, [I32, I32]),
    "b200_attention_flash_fwd": (I32, [P, P, P, P, P, I32, I32, I32, I32, F, U64, U64, P]),
    "b200_attention_flash_bwd": (I32, [P, P, P, P, P, P, P, P, I32, I32, I32, I32, F, P]),
    "b200_maxpool2x2_fwd": (I32, [P, P, P, I32, I32, I32, I32, I32, P]),
    "b200_maxpool2x2_bwd": (I32, [P, P, P, I32, I32, I32, I32, I32, P]),
    "b200_adaptive_avgpool_fwd": (I32, [P, P, I32, I32, I32, I32, I32, I32, I32, P]),
    "b200_adaptive_avgpool_bwd": (I32, [P, P, I32, I32, I32, I32, I32, I32, I32, P]),
    "b200_dropout_act": (I32, [P, P, I64, F, U64, U64, I32, P]),
    "b200_tanh_fwd_f32": (I32, [P, P, I64, P]),
    "b200_tanh_bwd_f32": (I32, [P, P, P, I64, P]),
    "b200_gather_rows": (I32, [P, P, I32, I64, I32, I32, P]),
    "b200_scatter_rows": (I32, [P, P, I32, I64, I32, I32, P]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header/library mismatch: fail at import
    _fn.restype = _res
    _fn.argtypes = _args


def last_error():
    msg = lib.b200_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc, what=""):
    if rc != 0:
        raise RuntimeError("b200net %s failed (status %d): %s" % (what, rc, last_error()))


def ptr_array(ptrs):
    return (c_void_p * len(ptrs))(*ptrs)


def i32_array(vals):
    return (c_int32 * len(vals))(*vals)


def i64_array(vals):
    return (c_int64 * len(vals))(*vals)
