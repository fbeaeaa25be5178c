"""ctypes binding of libdedloc_b200.so (C ABI in include/dedloc_b200.h).

The library is built in-tree by `dedloc_b200._build.build()`. On a machine with a GPU a missing or
unloadable library is a hard error (no silent PyTorch fallback on the CUDA path); on a CPU-only
machine `available()` is False and only the CPU reference paths may be used.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

MAX_PEERS = 16
IPC_HANDLE_BYTES = 64

_LIB: Optional[C.CDLL] = None
_LIB_ERR: Optional[str] = None
LIB_PATH = Path(__file__).resolve().parent / "libdedloc_b200.so"


class NativeError(RuntimeError):
    pass


class ARParams(C.Structure):
    """Mirror of DedlocARParams."""

    _fields_ = [
        ("P", C.c_int), ("me", C.c_int), ("mode", C.c_int), ("codec", C.c_int),
        ("phases", C.c_int), ("xgpu_barriers", C.c_int), ("tag", C.c_uint32), ("T", C.c_int),
        ("timeout_ns", C.c_ulonglong),
        ("replicate_state", C.c_int), ("max_blocks", C.c_int),
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("bias_correction", C.c_float), ("clamp_value", C.c_float), ("momentum", C.c_float),
        ("trust", C.c_float),
        ("w", C.c_float * MAX_PEERS),
        ("slot", C.c_int * MAX_PEERS),
        ("cbeg", C.c_longlong * MAX_PEERS),
        ("cend", C.c_longlong * MAX_PEERS),
        ("send", C.c_void_p * MAX_PEERS),
        ("aux", C.c_void_p * MAX_PEERS),
        ("out32", C.c_void_p * MAX_PEERS),
        ("out16", C.c_void_p * MAX_PEERS),
        ("m", C.c_void_p * MAX_PEERS),
        ("v", C.c_void_p * MAX_PEERS),
        ("table", C.c_void_p * MAX_PEERS),
        ("flags", C.c_void_p * MAX_PEERS),
        ("gavg", C.c_void_p), ("chunk_stats", C.c_void_p), ("chunk_tid", C.c_void_p),
        ("tensor_cbeg", C.c_void_p), ("t_wd", C.c_void_p), ("t_lr_scale", C.c_void_p),
        ("t_flags", C.c_void_p), ("status", C.c_void_p),
    ]


AR2_MAX_STAGES = 64


class AR2Params(C.Structure):
    """Mirror of DedlocAR2Params (push-based, stage-pipelined all-reduce kernel)."""

    _fields_ = [
        ("P", C.c_int), ("me", C.c_int), ("mode", C.c_int), ("codec", C.c_int), ("T", C.c_int), ("K", C.c_int),
        ("tag", C.c_uint32), ("src_f32", C.c_int), ("timeout_ns", C.c_ulonglong),
        ("nP", C.c_int), ("nR", C.c_int), ("nU", C.c_int), ("nA", C.c_int),
        ("NS", C.c_int), ("replicate_state", C.c_int),
        ("Ls", C.c_longlong), ("C", C.c_longlong), ("slot_bytes", C.c_longlong), ("aux_slot_floats", C.c_longlong),
        ("scale", C.c_float),
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("bias_correction", C.c_float), ("clamp_value", C.c_float), ("momentum", C.c_float), ("trust", C.c_float),
        ("w", C.c_float * MAX_PEERS),
        ("owner_of_k", C.c_int * MAX_PEERS),
        ("kidx", C.c_int * MAX_PEERS),
        ("acc", C.c_void_p), ("send", C.c_void_p), ("aux", C.c_void_p),
        ("p32", C.c_void_p), ("p16", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
        ("gavg", C.c_void_p), ("chunk_stats", C.c_void_p), ("chunk_tid", C.c_void_p),
        ("tensor_cbeg", C.c_void_p), ("t_wd", C.c_void_p), ("t_lr_scale", C.c_void_p), ("t_flags", C.c_void_p),
        ("status", C.c_void_p), ("counters", C.c_void_p),
        ("stage_in", C.c_void_p * MAX_PEERS),
        ("aux_in", C.c_void_p * MAX_PEERS),
        ("delta", C.c_void_p * MAX_PEERS),
        ("inflag", C.c_void_p * MAX_PEERS),
        ("outflag", C.c_void_p * MAX_PEERS),
        ("table", C.c_void_p * MAX_PEERS),
        ("abort_word", C.c_void_p * MAX_PEERS),
        ("out32", C.c_void_p * MAX_PEERS),
        ("out16", C.c_void_p * MAX_PEERS),
        ("m_all", C.c_void_p * MAX_PEERS),
        ("v_all", C.c_void_p * MAX_PEERS),
        ("trace", C.c_void_p), ("debug_flags", C.c_int), ("sequential", C.c_int),
    ]


class NvlsParams(C.Structure):
    """Mirror of DedlocNvlsParams."""

    _fields_ = [
        ("mc_base", C.c_void_p), ("uc_base", C.c_void_p), ("n", C.c_longlong),
        ("off_flags", C.c_longlong), ("off_payload", C.c_longlong), ("off_result", C.c_longlong),
        ("P", C.c_int), ("me", C.c_int), ("round", C.c_uint32), ("ctas", C.c_int),
        ("payload_is_bf16", C.c_int), ("reserved", C.c_int), ("timeout_ns", C.c_ulonglong),
        ("status", C.c_void_p), ("out16", C.c_void_p), ("out32", C.c_void_p),
    ]


class GemmArgs(C.Structure):
    """Mirror of DedlocGemmArgs."""

    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("C", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("lda", C.c_longlong), ("ldb", C.c_longlong), ("ldc", C.c_longlong),
        ("a_mn_major", C.c_int), ("b_mn_major", C.c_int), ("is_fp16", C.c_int),
        ("epilogue", C.c_int), ("out_kind", C.c_int), ("alpha", C.c_float),
        ("bias", C.c_void_p), ("residual", C.c_void_p), ("ld_residual", C.c_longlong),
        ("aux_in", C.c_void_p), ("ld_aux", C.c_longlong),
        ("pre_out", C.c_void_p), ("ld_pre", C.c_longlong),
        ("split_k", C.c_int), ("block_n", C.c_int), ("max_ctas", C.c_int), ("cta_pair", C.c_int),
    ]


# epilogue / mode constants (include/dedloc_b200.h)
EPI_BIAS, EPI_BIAS_F32, EPI_GELU, EPI_DGELU, EPI_RESIDUAL, EPI_SAVE_PRE = 1, 2, 4, 8, 16, 32
EPI_SAVE_DGELU, EPI_MUL_AUX = 64, 128
OUT_BF16, OUT_F32, OUT_F32_ATOMIC = 0, 1, 2
AR_MEAN, AR_LAMB, AR_LARS = 0, 1, 2
TF_FROZEN, TF_NO_LARS, TF_ROWNORM128 = 1, 2, 4
AR_STATUS = {0: "ok", 1: "abort_before_reduce", 2: "abort_before_update", 3: "abort_after_push",
             4: "skipped_nonfinite"}


def _declare(lib: C.CDLL) -> None:
    vp, ll, i, f = C.c_void_p, C.c_longlong, C.c_int, C.c_float
    lib.dedloc_last_error.restype = C.c_char_p
    lib.dedloc_launch_count.restype = ll
    lib.dedloc_add_launch_count.argtypes = [ll]
    lib.dedloc_add_launch_count.restype = None
    lib.dedloc_device_info.argtypes = [C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(ll)]
    lib.dedloc_symm_alloc.argtypes = [ll, C.POINTER(vp), C.c_char_p]
    lib.dedloc_symm_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.dedloc_symm_close.argtypes = [vp]
    lib.dedloc_symm_free.argtypes = [vp]
    lib.dedloc_enable_peer_access.argtypes = [i]
    lib.dedloc_memset_async.argtypes = [vp, i, ll, vp]
    lib.dedloc_memcpy_async.argtypes = [vp, vp, ll, vp]
    lib.dedloc_memcpy_peer_async.argtypes = [vp, i, vp, i, ll, vp]
    lib.dedloc_memcpy2d_async.argtypes = [vp, ll, vp, ll, ll, ll, vp]
    lib.dedloc_memcpy2d_async.restype = i
    lib.dedloc_sumsq.argtypes = [vp, i, ll, vp, i, vp, vp]
    lib.dedloc_accumulate_clip.argtypes = [vp, vp, i, ll, vp, f, i, vp]
    lib.dedloc_encode.argtypes = [i, vp, ll, f, vp, vp, vp]
    lib.dedloc_codec_aux_floats.argtypes = [i, ll]
    lib.dedloc_codec_aux_floats.restype = ll
    lib.dedloc_codec_payload_bytes_per_elem.argtypes = [i]
    lib.dedloc_fused_allreduce.argtypes = [C.POINTER(ARParams), vp]
    lib.dedloc_fused_allreduce_v2.argtypes = [C.POINTER(AR2Params), i, vp, vp]
    lib.dedloc_fused_allreduce_v2.restype = i
    lib.dedloc_ar2_params_size.restype = i
    lib.dedloc_ar2_max_ctas.argtypes = [i, i]
    lib.dedloc_ar2_max_ctas.restype = i
    lib.dedloc_nvls_allreduce_mean.argtypes = [C.POINTER(NvlsParams), vp]
    lib.dedloc_nvls_allreduce_mean.restype = i
    lib.dedloc_nvls_params_size.restype = i
    lib.dedloc_gemm.argtypes = [C.POINTER(GemmArgs), vp]
    lib.dedloc_layernorm_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i, i, f, vp]
    lib.dedloc_layernorm_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i, i, vp]
    lib.dedloc_colsum_accumulate.argtypes = [vp, vp, i, i, ll, vp]
    lib.dedloc_sinkhorn_scratch_floats.argtypes = [i, i]
    lib.dedloc_sinkhorn_scratch_floats.restype = ll
    lib.dedloc_sinkhorn.argtypes = [vp, i, ll, i, i, i, f, i, f, vp, ll, vp, vp]
    lib.dedloc_sinkhorn.restype = i
    lib.dedloc_attention_fwd.argtypes = [vp, vp, vp, vp, i, i, i, i, f, vp]
    lib.dedloc_attention_fwd.restype = i
    lib.dedloc_attention_bwd_scratch_floats.argtypes = [i]
    lib.dedloc_attention_bwd_scratch_floats.restype = ll
    lib.dedloc_attention_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, i, i, i, f, vp]
    lib.dedloc_attention_bwd.restype = i
    newer = {
        "dedloc_embed_ln_fwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, f, vp],
        "dedloc_embed_ln_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, ll, i, vp],
        "dedloc_ln128_fwd": [vp, vp, vp, vp, vp, vp, i, f, vp],
        "dedloc_ln128_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i, vp],
        "dedloc_l2norm128_fwd": [vp, vp, i, f, vp],
        "dedloc_l2norm128_bwd": [vp, vp, vp, i, f, vp],
        "dedloc_swav_ce_fwd": [vp, i, ll, vp, i, vp, i, i, i, f, vp, vp, vp],
        "dedloc_swav_ce_bwd": [vp, i, ll, vp, i, vp, i, i, i, f, vp, vp, vp, vp],
    }
    for name, argtypes in newer.items():
        if not hasattr(lib, name):
            raise NativeError(f"{LIB_PATH} is older than its Python binding (no {name}): run `python -m dedloc_b200._build`")
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = argtypes, i
    lib.dedloc_ce_fwd.argtypes = [vp, ll, vp, i, i, vp, vp, vp]
    lib.dedloc_ce_fwd.restype = i
    lib.dedloc_ce_bwd.argtypes = [vp, ll, vp, vp, vp, vp, ll, i, i, vp]
    lib.dedloc_ce_bwd.restype = i
    for name in ("dedloc_layernorm_fwd", "dedloc_layernorm_bwd", "dedloc_colsum_accumulate"):
        getattr(lib, name).restype = i
    for name in ("dedloc_symm_alloc", "dedloc_symm_open", "dedloc_symm_close", "dedloc_symm_free",
                 "dedloc_enable_peer_access", "dedloc_memset_async", "dedloc_memcpy_async",
                 "dedloc_memcpy_peer_async", "dedloc_sumsq", "dedloc_accumulate_clip", "dedloc_encode",
                 "dedloc_fused_allreduce", "dedloc_gemm", "dedloc_device_info", "dedloc_ar_params_size",
                 "dedloc_chunk_elems", "dedloc_codec_payload_bytes_per_elem"):
        getattr(lib, name).restype = i
    # optional entry points (declared when present so that older builds keep loading)
    for name, argtypes in _OPTIONAL.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = i


_OPTIONAL: dict = {}


def register_optional(name: str, argtypes: list) -> None:
    """Modules that own an optional entry point declare its signature here before first use."""
    _OPTIONAL[name] = argtypes
    if _LIB is not None and hasattr(_LIB, name):
        fn = getattr(_LIB, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int


def load(required: bool = False) -> Optional[C.CDLL]:
    """Load (once) and return the native library; None when it is not built and not required."""
    global _LIB, _LIB_ERR
    if _LIB is not None:
        return _LIB
    path = Path(os.environ.get("DEDLOC_B200_LIB", LIB_PATH))
    if not path.exists():
        _LIB_ERR = f"{path} does not exist; run `python -m dedloc_b200._build`"
    else:
        try:
            lib = C.CDLL(str(path))
            _declare(lib)
            if lib.dedloc_ar_params_size() != C.sizeof(ARParams):
                raise NativeError(
                    f"ABI mismatch: DedlocARParams is {lib.dedloc_ar_params_size()} bytes in the library, "
                    f"{C.sizeof(ARParams)} in the binding")
            if lib.dedloc_ar2_params_size() != C.sizeof(AR2Params):
                raise NativeError(
                    f"ABI mismatch: DedlocAR2Params is {lib.dedloc_ar2_params_size()} bytes in the library, "
                    f"{C.sizeof(AR2Params)} in the binding")
            _LIB = lib
            return lib
        except OSError as e:  # e.g. libcudart missing
            _LIB_ERR = f"cannot load {path}: {e}"
    if required:
        raise NativeError(_LIB_ERR)
    return None


def available() -> bool:
    return load() is not None


def lib() -> C.CDLL:
    return load(required=True)


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().dedloc_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def launch_count() -> int:
    return int(lib().dedloc_launch_count()) if available() else 0


def require_native_on_gpu() -> None:
    """Fail loudly if CUDA is visible but the extension is missing (no silent eager fallback)."""
    import torch

    if torch.cuda.is_available() and not available():
        raise NativeError(f"CUDA device present but native library unavailable: {_LIB_ERR}")
