"""ctypes binding of libowl_dit.so (the C ABI declared in include/owl_dit.h).

There is no CPU fallback: if the shared library is missing the import of any op fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libowl_dit.so")

i32, i64, f32, vp = C.c_int32, C.c_int64, C.c_float, C.c_void_p

OWL_GEMM_NT, OWL_GEMM_NN, OWL_GEMM_TN = 0, 1, 2
OWL_EPI_STORE, OWL_EPI_BIAS_SILU, OWL_EPI_GATE_RESIDUAL, OWL_EPI_DSILU, OWL_EPI_F32 = 0, 1, 2, 3, 4


class GemmEpilogue(C.Structure):
    """struct owl_gemm_epilogue (include/owl_dit.h)."""

    _fields_ = [
        ("mode", i32), ("accumulate", i32), ("rows_per_frame", i32), ("reserved", i32),
        ("bias", vp), ("out0", vp), ("ld0", i64), ("out1", vp), ("ld1", i64),
        ("outf", vp), ("ldf", i64), ("aux0", vp), ("lda0", i64), ("aux1", vp), ("lda1", i64),
    ]


class AdaLNPrologue(C.Structure):
    """struct owl_adaln_prologue (include/owl_dit.h)."""

    _fields_ = [("a", vp), ("b", vp), ("ld_ab", i64), ("rows_per_frame", i32), ("eps", f32)]


class AttnMask(C.Structure):
    """struct owl_attn_mask (include/owl_dit.h)."""

    _fields_ = [
        ("tokens_per_frame", i32), ("window_frames", i32), ("q_offset", i32), ("causal", i32),
        ("doc_id", vp), ("doc_ld", i64), ("doc_first", vp), ("doc_last", vp), ("doc_breaks", vp),
    ]


_ATTN_FWD = [vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, i64, i64, i64,
             C.POINTER(AttnMask), vp]
_ATTN_BWD = [vp, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, vp, vp, vp, i64, i64,
             vp, i64, i64, vp, i64, i64, i64, i64, i64, i64, i64, C.POINTER(AttnMask), vp]

# name -> argtypes; every symbol include/owl_dit.h declares (tests check the list against the header)
SIGNATURES = {
    "owl_abi_version": ([], C.c_int),
    "owl_error_string": ([C.c_int], C.c_char_p),
    "owl_launch_count": ([], C.c_uint64),
    "owl_set_reserved_sms": ([C.c_int], C.c_int),
    "owl_set_coresident_reserve": ([C.c_int], C.c_int),
    "owl_cast_f32_bf16_multi": ([vp, vp, vp, C.c_int, vp], C.c_int),
    "owl_debug_timestamp": ([vp, vp], C.c_int),
    "owl_gemm_bf16": ([C.c_int, i64, i64, i64, vp, i64, vp, i64, C.POINTER(GemmEpilogue), vp], C.c_int),
    "owl_gemm_smallm_workspace_bytes": ([i64], i64),
    "owl_gemm_smallm": ([i64, i64, i64, vp, i64, vp, i64, C.POINTER(GemmEpilogue), C.POINTER(AdaLNPrologue), vp, i64, vp,
                         i64, vp], C.c_int),
    "owl_gemm_smallm_static_max_n": ([], i64),
    "owl_gemm_smallm_static": ([i64, i64, i64, vp, i64, vp, i64, C.POINTER(GemmEpilogue), vp], C.c_int),
    "owl_gemm_bf16_check": ([C.c_int, i64, i64, i64, vp, i64, vp, i64, vp, i64, vp], C.c_int),
    "owl_cast_f32_bf16": ([vp, vp, i64, vp], C.c_int),
    "owl_silu_bf16": ([vp, vp, i64, vp], C.c_int),
    "owl_sincos_embed": ([vp, vp, vp, i64, C.c_int, f32, vp], C.c_int),
    "owl_dsilu_mul_bf16": ([vp, vp, vp, i64, vp], C.c_int),
    "owl_adaln_modulate_fwd": ([vp, vp, vp, i64, vp, vp, i64, i64, i64, f32, vp], C.c_int),
    "owl_residual_adaln_fwd": ([vp, vp, vp, i64, vp, vp, i64, vp, vp, vp, i64, i64, i64, f32, vp], C.c_int),
    "owl_adaln_modulate_bwd": ([vp, vp, vp, i64, vp, vp, vp, vp, vp, i64, i64, i64, i64, vp], C.c_int),
    "owl_gate_fwd": ([vp, vp, i64, vp, vp, i64, i64, i64, vp], C.c_int),
    "owl_gate_bwd": ([vp, vp, vp, i64, vp, vp, i64, vp, C.c_int, i64, i64, i64, vp], C.c_int),
    "owl_adaln_gate_bwd": ([vp, vp, vp, i64, vp, vp, vp, vp, vp, i64, vp, vp, i64, vp, vp, i64, vp, C.c_int,
                            i64, i64, i64, vp], C.c_int),
    "owl_colsum_bf16_multi": ([vp, vp, vp, vp, vp, C.c_int, vp], C.c_int),
    "owl_colsum_bf16": ([vp, i64, vp, C.c_int, i64, i64, vp], C.c_int),
    "owl_qk_norm_rope_fwd": ([vp, vp, vp, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, vp,
                              i64, i64, i64, i64, f32, vp], C.c_int),
    "owl_qk_norm_rope_bwd": ([vp, vp, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, vp, vp, vp, i64, vp,
                              i64, i64, i64, i64, vp], C.c_int),
    "owl_rope_apply": ([vp, vp, vp, i64, vp, i64, i64, i64, C.c_int, vp], C.c_int),
    "owl_kv_append": ([vp, i64, i64, vp, i64, i64, i64, i64, i64, i64, i64, vp], C.c_int),
    "owl_layer_norm_fwd": ([vp, vp, i64, i64, f32, vp], C.c_int),
    "owl_layer_norm_bwd": ([vp, vp, vp, i64, i64, f32, vp], C.c_int),
    "owl_attn_doc_extents": ([vp, i64, i64, i64, vp, vp, vp, vp], C.c_int),
    "owl_attn_decode_workspace_bytes": ([i64, i64, i64, i64], i64),
    "owl_attn_decode": ([vp, i64, i64, vp, vp, i64, i64, i64, vp, i64, vp, i64, i64, i64, i64, i64, i64, vp, i64, vp],
                        C.c_int),
    "owl_attn_decode_fused": ([vp, vp, vp, i64, vp, vp, i64, i64, i64, vp, i64, vp, i64, i64, i64, i64, i64, i64, f32, vp, i64,
                               vp], C.c_int),
    "owl_qk_norm_rope_decode": ([vp, vp, vp, i64, vp, vp, i64, i64, vp, vp, i64, i64, i64, i64, i64, i64, i64, f32, vp],
                                C.c_int),
    "owl_decode_state_advance": ([vp, i64, i64, i64, vp], C.c_int),
    "owl_shard_reduce_peers": ([vp, vp, C.c_int, i64, f32, C.c_int, vp], C.c_int),
    "owl_mc_allreduce": ([vp, i64, f32, C.c_int, vp], C.c_int),
    "owl_shard_reduce": ([vp, vp, i64, C.c_int, i64, f32, C.c_int, vp], C.c_int),
    "owl_rf_noise": ([vp, vp, vp, vp, vp, i64, i64, vp], C.c_int),
    "owl_mse_fwd": ([vp, vp, vp, i64, vp], C.c_int),
    "owl_mse_bwd": ([vp, vp, vp, vp, i64, vp], C.c_int),
    "owl_attn_fwd": (_ATTN_FWD, C.c_int),
    "owl_attn_bwd": (_ATTN_BWD, C.c_int),
    "owl_attn_fwd_check": (_ATTN_FWD, C.c_int),
    "owl_attn_bwd_check": (_ATTN_BWD, C.c_int),
}

_lib = None


def load() -> C.CDLL:
    """Load libowl_dit.so and bind every entry point.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m owl_audio_exps_b200.build` "
            "(the sm_100a kernels are the only implementation; there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    missing = []
    for name, (argtypes, restype) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            missing.append(name)
            continue
        fn.argtypes = argtypes
        fn.restype = restype
    if missing:
        raise RuntimeError(f"{LIB_PATH} does not export {missing}: stale build, rebuild it")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().owl_error_string(rc).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {rc})")
