"""ctypes binding of libb200glue.so (include/b200glue.h).  No CPU fallback: a missing library raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200GLUE_LIB points at an alternative build of the same library (kernel experiments); never a fallback
LIB_PATH = os.environ.get("B200GLUE_LIB") or os.path.join(_HERE, "libb200glue.so")

EPI_BIAS_BF16, EPI_QKV_ROTARY, EPI_LN_GELU, EPI_RESID, EPI_F32 = 0, 1, 2, 3, 4
EPI_CONV_RELU_BN, EPI_CONV_BN_L2, EPI_CONV_BN_SOFTMAX_D2S = 5, 6, 7
ABI_VERSION = 9


class GemmDesc(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("a_rows", C.c_longlong), ("lda", C.c_longlong),
        ("b", C.c_void_p), ("b_rows", C.c_longlong), ("ldb", C.c_longlong),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int), ("batch", C.c_int),
        ("a_batch_rows", C.c_int), ("b_batch_rows", C.c_int),
        ("block_n", C.c_int), ("epilogue", C.c_int),
        ("conv", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Cin", C.c_int),
        ("bias", C.c_void_p), ("scale", C.c_void_p), ("shift", C.c_void_p),
        ("relu", C.c_int), ("pool", C.c_int), ("alpha", C.c_float),
        ("resid", C.c_void_p), ("cossin", C.c_void_p), ("rot_cols", C.c_int),
        ("ln_gamma", C.c_void_p), ("ln_beta", C.c_void_p), ("ln_eps", C.c_float),
        ("out16", C.c_void_p), ("ld16", C.c_longlong),
        ("out32", C.c_void_p), ("ld32", C.c_longlong), ("out32_batch_stride", C.c_longlong),
        ("cells_h", C.c_int), ("cells_w", C.c_int),
        ("tile_live", C.c_void_p),
    ]


class AttnSeq(C.Structure):
    _fields_ = [
        ("q_row0", C.c_int), ("q_len", C.c_int), ("kv_row0", C.c_int), ("kv_len", C.c_int),
        ("q_len_slot", C.c_int), ("kv_len_slot", C.c_int),
    ]


class AttnDesc(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("q_rows", C.c_longlong), ("ldq", C.c_longlong), ("q_col0", C.c_int),
        ("k", C.c_void_p), ("k_rows", C.c_longlong), ("ldk", C.c_longlong), ("k_col0", C.c_int),
        ("v", C.c_void_p), ("v_rows", C.c_longlong), ("ldv", C.c_longlong), ("v_col0", C.c_int),
        ("seqs", C.c_void_p), ("num_seqs", C.c_int), ("max_q_len", C.c_int), ("heads", C.c_int),
        ("len_override", C.c_void_p), ("scale", C.c_float),
        ("out", C.c_void_p), ("ld_out", C.c_longlong), ("out_col0", C.c_int),
    ]


class ExportField(C.Structure):
    _fields_ = [("src", C.c_void_p), ("scales", C.c_void_p), ("elems_per_pair", C.c_longlong),
                ("dst_offset", C.c_longlong), ("kind", C.c_int), ("reserved", C.c_int)]


EXPORT_F32, EXPORT_F32_KEYPOINTS, EXPORT_I64 = 0, 1, 2
EXPORT_MAX_FIELDS = 16

_SIGNATURES = {
    "b200_abi_version": (C.c_int, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_sm_count": (C.c_int, [C.POINTER(C.c_int)]),
    "b200_gemm": (C.c_int, [C.POINTER(GemmDesc), C.c_void_p]),
    "b200_attention": (C.c_int, [C.POINTER(AttnDesc), C.c_void_p]),
    "b200_sp_conv1a": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_sp_conv_stem": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int, C.c_void_p, C.c_void_p]),
    "b200_sp_conv1a_u8": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_sp_conv_stem_u8": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_void_p, C.c_void_p]),
    "b200_sp_nms_border": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "b200_sp_nms_border_list": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "b200_sp_sample_descriptors_mode": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                                  C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "b200_sg_keypoint_encoder_head": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_sg_split_hilo": (C.c_int, [C.c_void_p, C.c_longlong, C.c_longlong, C.c_void_p, C.c_void_p]),
    "b200_split_hilo": (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p]),
    "b200_nn_match_workspace": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "b200_nn_match": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_sg_sinkhorn_workspace": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "b200_sg_sinkhorn_assignment": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_void_p]),
    "b200_sp_select_topk_list": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p,
                                           C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p]),
    "b200_sp_count_candidates": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "b200_sp_select_topk": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_sp_sample_descriptors": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                             C.c_int, C.c_void_p, C.c_void_p]),
    "b200_lg_ingest": (C.c_int, [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]),
    "b200_lg_posenc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p]),
    "b200_lg_rowdot": (C.c_int, [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200_lg_assignment_workspace": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "b200_lg_assignment": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p]),
    "b200_lg_assignment_static_capacity": (C.c_size_t, []),
    "b200_lg_assignment_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_lg_adaptive_decide": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                          C.c_void_p]),
    "b200_lg_adaptive_prune": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float] +
                               [C.c_void_p] * 7 + [C.c_longlong] + [C.c_void_p] * 7),
    "b200_lg_adaptive_commit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
    "b200_lg_adaptive_snapshot": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 10),
    "b200_lg_assignment_live": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                          C.c_float] + [C.c_void_p] * 7),
    "b200_lg_adaptive_scatter": (C.c_int, [C.c_void_p] * 6 + [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5),
    "b200_prep_resize": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                   C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                   C.c_void_p]),
    "b200_export_pack": (C.c_int, [C.POINTER(ExportField), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_longlong,
                                   C.c_void_p]),
    "b200_eval_matches_homography": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                               C.c_void_p, C.c_void_p]),
    "b200_eval_matches_epipolar": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_eval_homography_dlt": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_lg_count_below": (C.c_int, [C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "b200_lg_prune": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_lg_scatter_matches": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


class B200Error(RuntimeError):
    pass


def declared_symbols():
    return sorted(_SIGNATURES)


def lib():
    """Load the shared library once; raise if it is absent (there is no fallback path)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(
                f"{LIB_PATH} is missing: build it with `python -m glue_factory_b200.build` "
                "(the B200 path has no CPU / PyTorch fallback)"
            )
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        if handle.b200_abi_version() != ABI_VERSION:
            raise B200Error("libb200glue.so ABI version mismatch; rebuild the library")
        _lib = handle
    return _lib


def check(code: int):
    if code != 0:
        raise B200Error(lib().b200_last_error().decode() or f"b200glue call failed with code {code}")
