"""ctypes binding of include/sav_abi.h (libsav_b200.so).

This is the only door from Python to the device: every call enqueues hand-written sm_100a
kernels on the current torch CUDA stream.  There is NO fallback: if the shared library is missing
or the device is not a B200, importing/using these functions raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libsav_b200.so")

_lib = None


class SavError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("D", C.c_void_p),
        ("M", C.c_int64), ("N", C.c_int64), ("K", C.c_int64),
        ("lda", C.c_int64), ("ldb", C.c_int64), ("ldd", C.c_int64),
        ("a_mn", C.c_int32), ("b_mn", C.c_int32), ("d_fp32", C.c_int32), ("split_k", C.c_int32),
        ("bias", C.c_void_p), ("act", C.c_int32), ("accumulate", C.c_int32),
        ("aux_out", C.c_void_p), ("aux_in", C.c_void_p),
        ("col_scale", C.c_void_p), ("row_scale", C.c_void_p), ("rows_per_scale", C.c_int64),
        ("residual", C.c_void_p), ("ldr", C.c_int64),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_int64),
        ("dgelu_colsum", C.c_void_p),
        ("ln_out", C.c_void_p), ("ldy", C.c_int64), ("ln_gamma", C.c_void_p), ("ln_beta", C.c_void_p),
        ("ln_mean", C.c_void_p), ("ln_rstd", C.c_void_p), ("ln_eps", C.c_float), ("reserved", C.c_int32),
    ]


class AttnFwdArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("o", C.c_void_p),
        ("o_lo", C.c_void_p), ("lse", C.c_void_p),
        ("B", C.c_int64), ("H", C.c_int64), ("Nq", C.c_int64), ("Nk", C.c_int64), ("D", C.c_int64),
        ("q_batch_stride", C.c_int64), ("q_row_stride", C.c_int64),
        ("k_batch_stride", C.c_int64), ("k_row_stride", C.c_int64),
        ("v_batch_stride", C.c_int64), ("v_row_stride", C.c_int64),
        ("o_batch_stride", C.c_int64), ("o_row_stride", C.c_int64),
        ("scale", C.c_float),
    ]


class AttnBwdArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("o", C.c_void_p),
        ("o_lo", C.c_void_p), ("d_o", C.c_void_p), ("lse", C.c_void_p),
        ("dq", C.c_void_p), ("dk", C.c_void_p), ("dv", C.c_void_p),
        ("delta", C.c_void_p), ("dq_accum", C.c_void_p),
        ("B", C.c_int64), ("H", C.c_int64), ("Nq", C.c_int64), ("Nk", C.c_int64), ("D", C.c_int64),
        ("q_batch_stride", C.c_int64), ("q_row_stride", C.c_int64),
        ("k_batch_stride", C.c_int64), ("k_row_stride", C.c_int64),
        ("v_batch_stride", C.c_int64), ("v_row_stride", C.c_int64),
        ("o_batch_stride", C.c_int64), ("o_row_stride", C.c_int64),
        ("scale", C.c_float),
    ]


class BGemmArgs(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("D", C.c_void_p),
        ("M", C.c_int64), ("N", C.c_int64), ("K", C.c_int64),
        ("batch_outer", C.c_int64), ("batch_inner", C.c_int64),
        ("lda", C.c_int64), ("a_stride_inner", C.c_int64), ("a_stride_outer", C.c_int64),
        ("ldb", C.c_int64), ("b_stride_inner", C.c_int64), ("b_stride_outer", C.c_int64),
        ("ldd", C.c_int64), ("d_stride_inner", C.c_int64), ("d_stride_outer", C.c_int64),
        ("a_mn", C.c_int32), ("b_mn", C.c_int32), ("d_fp32", C.c_int32), ("reserved", C.c_int32),
    ]


class Conv2dGeom(C.Structure):
    _fields_ = [
        ("B", C.c_int64), ("H", C.c_int64), ("W", C.c_int64), ("C", C.c_int64),
        ("Ho", C.c_int64), ("Wo", C.c_int64),
        ("kh", C.c_int32), ("kw", C.c_int32), ("sh", C.c_int32), ("sw", C.c_int32),
        ("pad_top", C.c_int32), ("pad_left", C.c_int32),
        ("xs_b", C.c_int64), ("xs_h", C.c_int64), ("xs_w", C.c_int64), ("xs_c", C.c_int64),
        ("ld", C.c_int64),
    ]


class ThSoftmaxArgs(C.Structure):
    _fields_ = [
        ("S", C.c_void_p), ("P", C.c_void_p), ("dP", C.c_void_p), ("dS", C.c_void_p),
        ("mix_pre", C.c_void_p), ("mix_post", C.c_void_p),
        ("dmix_pre", C.c_void_p), ("dmix_post", C.c_void_p),
        ("rel", C.c_void_p), ("drel", C.c_void_p),
        ("B", C.c_int64), ("H", C.c_int64), ("Nq", C.c_int64), ("Nk", C.c_int64),
        ("ld", C.c_int64), ("ldr", C.c_int64),
        ("rel_w", C.c_int32), ("rel_h", C.c_int32), ("scale", C.c_float), ("reserved", C.c_int32),
    ]


_P, _I64, _I32, _F, _D = C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_double

# name -> (restype, argtypes); also the list the CPU test checks against include/sav_abi.h
SIGNATURES = {
    "sav_abi_version": (C.c_int, []),
    "sav_last_error": (C.c_char_p, []),
    "sav_device_check": (C.c_int, []),
    "sav_launch_count": (C.c_int64, []),
    "sav_gemm_ln_tiles": (C.c_int32, [_I64]),
    "sav_gemm_workspace_bytes": (C.c_int64, [C.POINTER(GemmArgs)]),
    "sav_gemm": (C.c_int, [C.POINTER(GemmArgs), _P]),
    "sav_bgemm": (C.c_int, [C.POINTER(BGemmArgs), _P]),
    "sav_thsoftmax_fwd": (C.c_int, [C.POINTER(ThSoftmaxArgs), _P]),
    "sav_thsoftmax_bwd": (C.c_int, [C.POINTER(ThSoftmaxArgs), _P]),
    "sav_scale_rows_cols": (C.c_int, [_P, _I64, _P, _P, _I64, _P, _I64, _I64, _I64, _P]),
    "sav_layerscale_grad": (C.c_int, [_P, _I64, _P, _I64, _I64, _I64, _P, _P, _P, _P, _P]),
    "sav_copy2d_f32": (C.c_int, [_P, _I64, _P, _I64, _I64, _I64, _I32, _P]),
    "sav_dwconv3x3_out_hw": (C.c_int, [_I64, _I64, _I32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "sav_dwconv3x3_fwd": (C.c_int, [_P, _P, _P, _P, _P, _I64, _I64, _I64, _I64, _I32, _P]),
    "sav_dwconv3x3_bwd": (C.c_int, [_P, _P, _P, _P, _P, _I64, _I64, _I64, _I64, _I32, _I32, _P]),
    "sav_batchnorm_fwd": (C.c_int, [_P, _P, _P, _P, _P, _F, _P, _P, _P, _P, _I64, _I64, _P]),
    "sav_batchnorm_bwd": (C.c_int, [_P] * 9 + [_I64, _I64, _P]),
    "sav_colsum_bf16": (C.c_int, [_P, _I64, _I64, _I64, _P, _I32, _P]),
    "sav_colsum_workspace_bytes": (C.c_int64, [_I64, _I64]),
    "sav_colsum_bf16_ws": (C.c_int, [_P, _I64, _I64, _I64, _P, _I32, _P, _I64, _P]),
    "sav_colsum_static_max_bytes": (C.c_int64, []),
    "sav_colsum_bf16_static": (C.c_int, [_P, _I64, _I64, _I64, _P, _I32, _P]),
    "sav_layernorm_fwd": (C.c_int, [_P, _I64, _P, _P, _P, _I64, _P, _P, _I64, _I64, _F, _P]),
    "sav_layernorm_bwd": (C.c_int, [_P, _I64, _P, _I64] + [_P] * 8 + [_I64, _I64, _P]),
    "sav_layernorm_bwd_colsum": (C.c_int, [_P, _I64, _P, _I64] + [_P] * 9 + [_I64, _I64, _P]),
    "sav_patchify": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _I64, _I64, _I32, _P]),
    "sav_embed_fwd": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _I32, _P]),
    "sav_embed_bwd": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _I32, _I32, _P]),
    "sav_attn_fwd": (C.c_int, [C.POINTER(AttnFwdArgs), _P]),
    "sav_attn_bwd": (C.c_int, [C.POINTER(AttnBwdArgs), _P]),
    "sav_softmax_xent": (C.c_int, [_P, _I64, _P, _I64, _I64, _F, _F, _P, _P, _P]),
    "sav_softmax_xent_mix": (C.c_int, [_P, _I64, _P, _P, _P, _I64, _I64, _F, _F, _P, _P, _P]),
    "sav_schedule_step": (C.c_int, [_P, _P, _I32, _F, _F, _I64, _I64, _F, _D, _D, _P]),
    "sav_drop_path_factors": (C.c_int, [C.c_uint64, _P, _I64, _I64, _I64, _F, _P, _P]),
    "sav_layernorm_fwd_f32": (C.c_int, [_P, _I64, _P, _P, _P, _I64, _P, _P, _I64, _I64, _F, _P]),
    "sav_layernorm_bwd_f32": (C.c_int, [_P, _I64, _P, _I64] + [_P] * 8 + [_I64, _I64, _P]),
    "sav_cast2d_bf16_f32": (C.c_int, [_P, _I64, _P, _I64, _I64, _I64, _I32, _P]),
    "sav_dp_allreduce_sum_f32": (C.c_int, [C.c_uint64, _P, _P, _I64, _I64, _I64, _I32, _I32, _I32, _P]),
    "sav_maxpool3x3s2_valid_fwd": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _P]),
    "sav_maxpool3x3s2_valid_bwd": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _P]),
    "sav_patchify_act": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _I64, _I64, _I32, _P]),
    "sav_pixel_patchify": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _I64, _I64, _I64, _I64, _I32, _P]),
    "sav_attn_small_fwd": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _F, _P]),
    "sav_attn_small_bwd": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _F, _P]),
    "sav_transpose_pad": (C.c_int, [_P, _I32, _I64, _P, _I32, _I64, _P, _I64, _I64, _I64, _I64, _I64, _P]),
    "sav_conv_same_geom": (C.c_int, [_I64, _I32, _I32, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "sav_im2col": (C.c_int, [_P, _I32, _P, C.POINTER(Conv2dGeom), _P]),
    "sav_col2im": (C.c_int, [_P, _P, _I32, _I32, C.POINTER(Conv2dGeom), _P]),
    "sav_bn_stats": (C.c_int, [_P, _I64, _I64, _P, _P, _P]),
    "sav_bn_finalize": (C.c_int, [_P, _P, _I64, _F, _P, _P, _P, _I64, _P]),
    "sav_bn_update_running": (C.c_int, [_P, _P, _P, _P, _F, _I64, _P]),
    "sav_batchnorm_eval": (C.c_int, [_P, _P, _P, _P, _P, _F, _P, _I64, _I64, _P]),
    "sav_act_fwd": (C.c_int, [_P, _P, _I64, _I32, _P]),
    "sav_act_bwd": (C.c_int, [_P, _P, _P, _I64, _I32, _P]),
    "sav_bn_act_fwd": (C.c_int, [_P, _P, _P, _P, _P, _P, _I64, _I64, _I32, _P]),
    "sav_bn_act_bwd": (C.c_int, [_P] * 10 + [_I64, _I64, _I32, _P]),
    "sav_gate_res_act_fwd": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _I32, _P]),
    "sav_gate_res_act_bwd": (C.c_int, [_P] * 7 + [_I64, _I64, _I64, _I32, _P]),
    "sav_spatial_mean_fwd": (C.c_int, [_P, _P, _I64, _I64, _I64, _P]),
    "sav_spatial_mean_bwd": (C.c_int, [_P, _P, _I64, _I64, _I64, _P]),
    "sav_maxpool3x3s2_fwd": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _P]),
    "sav_maxpool3x3s2_bwd": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _P]),
    "sav_rotary": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _I32, _I32, _P]),
    "sav_rotary_table": (C.c_int, [_P, _I64, _I64, _P]),
    "sav_topk_correct": (C.c_int, [_P, _I64, _P, _I64, _I64, _I32, _I32, _I32, _I32, _I32, _P, _P]),
    "sav_cast_f32_bf16": (C.c_int, [_P, _P, _I64, _P]),
    "sav_sumsq": (C.c_int, [_P, _I64, _P, _P]),
    "sav_adam_step": (C.c_int, [_P, _P, _P, _P, _P, _I64, _P, _F, _F, _D, _D, _F, _F, _I64, _P, _P]),
    "sav_add_bf16": (C.c_int, [_P, _P, _P, _I64, _P]),
    "sav_fill_f32": (C.c_int, [_P, _F, _I64, _P]),
    "sav_copy2d_bf16": (C.c_int, [_P, _I64, _P, _I64, _I64, _I64, _P]),
}


def lib():
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SavError(
                f"{LIB_PATH} not found: run `python __graft_entry__.py` (build()) first. "
                "There is no CPU or PyTorch fallback for the hot path.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        if l.sav_abi_version() != 3:
            raise SavError("libsav_b200.so ABI version mismatch")
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        raise SavError(f"sav error {rc}: {lib().sav_last_error().decode()}")


def launch_count():
    return int(lib().sav_launch_count())
