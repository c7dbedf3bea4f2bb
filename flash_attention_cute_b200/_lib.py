"""ctypes binding of the C ABI in include/fa_sm100.h (libfa_sm100.so).

This is the FFI stub that stands where the reference's pybind module
``flash_attention_extension`` stood (csrc/flash_attention_api.cpp:137-141).  There is no CPU or
PyTorch fallback behind it: if the library cannot be built or loaded the import fails loudly.
"""
from __future__ import annotations

import ctypes
import threading
from ctypes import POINTER, Structure, c_char_p, c_float, c_int32, c_int64, c_uint32, c_uint64, c_void_p

from . import _build

FA_SM100_FP16 = 0
FA_SM100_BF16 = 1
FLAG_NO_PACK_GQA = 1
FLAG_FORCE_PACK_GQA = 2
FLAG_NO_DECODE = 8


def tune_flags(emu: int | None = None, spec: bool | None = None, fence: bool | None = None) -> int:
    """Tuning bits of fa_sm100_fwd_params.flags (honoured by FA_SM100_TUNE=1 builds only): `emu` =
    exp2 evaluations per 8 element pairs moved to the FMA pipe, `spec` = speculative stale-max
    softmax, `fence` = tcgen05.fence::after_thread_sync in front of P·V."""
    f = 0
    if emu is not None:
        f |= (int(emu) + 1) << 8
    if spec is not None:
        f |= (2 if spec else 1) << 12
    if fence is not None:
        f |= (2 if fence else 1) << 14
    return f

STATUS_NAMES = {
    0: "OK",
    1: "INVALID_ARGUMENT",
    2: "UNSUPPORTED_DEVICE",
    3: "CUDA",
    4: "KERNEL_ABORT",
}

# Every symbol include/fa_sm100.h declares (tests check the library exports all of them).
EXPORTED_SYMBOLS = (
    "fa_sm100_fwd",
    "fa_sm100_fwd_decode",
    "fa_sm100_decode_plan",
    "fa_sm100_fwd_decode_static",
    "fa_sm100_static_scratch_bytes",
    "fa_sm100_paged_decode_plan",
    "fa_sm100_fwd_decode_paged",
    "fa_sm100_varlen_workspace_bytes",
    "fa_sm100_fwd_varlen",
    "fa_sm100_rope",
    "fa_sm100_last_error",
    "fa_sm100_check_device",
    "fa_sm100_set_watchdog_ms",
    "fa_sm100_launch_count",
    "fa_sm100_build_info",
    "fa_sm100_plan",
    "fa_sm100_kernel_name",
    "fa_sm100_selftest_umma",
)


class FwdParams(Structure):
    """Mirror of ``struct fa_sm100_fwd_params``."""

    _fields_ = [
        ("q", c_void_p),
        ("k", c_void_p),
        ("v", c_void_p),
        ("o", c_void_p),
        ("batch", c_int64),
        ("heads_q", c_int64),
        ("heads_kv", c_int64),
        ("seqlen_q", c_int64),
        ("seqlen_kv", c_int64),
        ("headdim", c_int64),
        ("q_strides", c_int64 * 3),
        ("k_strides", c_int64 * 3),
        ("v_strides", c_int64 * 3),
        ("o_strides", c_int64 * 3),
        ("softmax_scale", c_float),
        ("causal", c_int32),
        ("dtype", c_int32),
        ("device", c_int32),
        ("flags", c_uint32),
    ]


class VarlenParams(Structure):
    """Mirror of ``struct fa_sm100_varlen_params``."""

    _fields_ = [
        ("q", c_void_p),
        ("k", c_void_p),
        ("v", c_void_p),
        ("o", c_void_p),
        ("cu_seqlens_q", c_void_p),   # device int32 [num_seqs + 1]
        ("cu_seqlens_kv", c_void_p),  # device int32 [num_seqs + 1]
        ("num_seqs", c_int64),
        ("total_q", c_int64),
        ("total_kv", c_int64),
        ("max_seqlen_q", c_int64),
        ("heads_q", c_int64),
        ("heads_kv", c_int64),
        ("headdim", c_int64),
        ("q_strides", c_int64 * 2),
        ("k_strides", c_int64 * 2),
        ("v_strides", c_int64 * 2),
        ("o_strides", c_int64 * 2),
        ("softmax_scale", c_float),
        ("causal", c_int32),
        ("dtype", c_int32),
        ("device", c_int32),
        ("flags", c_uint32),
    ]


class PagedDecodeParams(Structure):
    """Mirror of ``struct fa_sm100_paged_decode_params``."""

    _fields_ = [
        ("q", c_void_p),
        ("k_cache", c_void_p),
        ("v_cache", c_void_p),
        ("o", c_void_p),
        ("block_table", c_void_p),  # device int32 [B, block_table_stride]
        ("seqlens_k", c_void_p),    # device int32 [B]
        ("batch", c_int64),
        ("heads_q", c_int64),
        ("heads_kv", c_int64),
        ("headdim", c_int64),
        ("page_size", c_int64),
        ("max_seqlen_k", c_int64),
        ("block_table_stride", c_int64),
        ("q_strides", c_int64 * 2),
        ("o_strides", c_int64 * 2),
        ("k_strides", c_int64 * 3),
        ("v_strides", c_int64 * 3),
        ("softmax_scale", c_float),
        ("dtype", c_int32),
        ("device", c_int32),
        ("flags", c_uint32),
    ]


class RopeParams(Structure):
    """Mirror of fa_sm100_rope_params (include/fa_sm100.h)."""

    _fields_ = [
        ("q", c_void_p),
        ("k", c_void_p),
        ("q_out", c_void_p),
        ("k_out", c_void_p),
        ("cos", c_void_p),
        ("sin", c_void_p),
        ("batch", c_int64),
        ("heads_q", c_int64),
        ("heads_kv", c_int64),
        ("seqlen", c_int64),
        ("headdim", c_int64),
        ("q_strides", c_int64 * 3),
        ("k_strides", c_int64 * 3),
        ("q_out_strides", c_int64 * 3),
        ("k_out_strides", c_int64 * 3),
        ("cos_sin_strides", c_int64 * 2),
        ("dtype", c_int32),
        ("device", c_int32),
    ]


_lock = threading.Lock()
_lib = None


def load() -> ctypes.CDLL:
    """Build (if stale) and dlopen libfa_sm100.so; set up prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = _build.LIB
        import os as _os

        override = _os.environ.get("FA_SM100_LIB_OVERRIDE")  # development aid: A/B against an older build
        if override:
            path = override
        elif not _build.is_current():
            try:
                path = _build.build()
            except Exception as exc:
                # Never load a library whose sources (kernels, ABI structs) have moved on: that is a
                # silent mismatch.  FA_SM100_ALLOW_STALE=1 overrides, for machines without nvcc.
                import os
                import warnings

                if not _build.LIB.exists() or os.environ.get("FA_SM100_ALLOW_STALE", "") in ("", "0"):
                    raise RuntimeError(
                        "flash_attention (sm_100a): libfa_sm100.so is missing or older than its sources "
                        "and could not be rebuilt; there is no fallback implementation "
                        "(set FA_SM100_ALLOW_STALE=1 to load the stale library anyway)"
                    ) from exc
                warnings.warn("flash_attention (sm_100a): loading a STALE libfa_sm100.so "
                              f"(rebuild failed: {exc})", RuntimeWarning)
                path = _build.LIB
        lib = ctypes.CDLL(str(path))
        lib.fa_sm100_fwd.argtypes = [POINTER(FwdParams), c_void_p]
        lib.fa_sm100_fwd.restype = c_int32
        lib.fa_sm100_last_error.argtypes = []
        lib.fa_sm100_last_error.restype = c_char_p
        lib.fa_sm100_check_device.argtypes = [c_int32]
        lib.fa_sm100_check_device.restype = c_int32
        lib.fa_sm100_launch_count.argtypes = []
        lib.fa_sm100_launch_count.restype = c_uint64
        lib.fa_sm100_build_info.argtypes = []
        lib.fa_sm100_build_info.restype = c_char_p
        lib.fa_sm100_plan.argtypes = [POINTER(FwdParams)] + [POINTER(c_int32)] * 4
        lib.fa_sm100_plan.restype = c_int32
        lib.fa_sm100_decode_plan.argtypes = [POINTER(FwdParams), POINTER(c_int32), POINTER(c_int32), POINTER(c_int64)]
        lib.fa_sm100_decode_plan.restype = c_int32
        lib.fa_sm100_fwd_decode.argtypes = [POINTER(FwdParams), c_void_p, c_int64, c_void_p]
        lib.fa_sm100_fwd_decode.restype = c_int32
        lib.fa_sm100_fwd_decode_static.argtypes = [POINTER(FwdParams), c_void_p]
        lib.fa_sm100_fwd_decode_static.restype = c_int32
        lib.fa_sm100_static_scratch_bytes.argtypes = []
        lib.fa_sm100_static_scratch_bytes.restype = c_int64
        lib.fa_sm100_paged_decode_plan.argtypes = [POINTER(PagedDecodeParams), POINTER(c_int32), POINTER(c_int32),
                                                   POINTER(c_int64)]
        lib.fa_sm100_paged_decode_plan.restype = c_int32
        lib.fa_sm100_fwd_decode_paged.argtypes = [POINTER(PagedDecodeParams), c_void_p, c_int64, c_void_p]
        lib.fa_sm100_fwd_decode_paged.restype = c_int32
        lib.fa_sm100_varlen_workspace_bytes.argtypes = [c_int64]
        lib.fa_sm100_varlen_workspace_bytes.restype = c_int64
        lib.fa_sm100_fwd_varlen.argtypes = [POINTER(VarlenParams), c_void_p, c_int64, c_void_p]
        lib.fa_sm100_fwd_varlen.restype = c_int32
        lib.fa_sm100_set_watchdog_ms.argtypes = [c_int64]
        lib.fa_sm100_set_watchdog_ms.restype = None
        lib.fa_sm100_kernel_name.argtypes = [POINTER(FwdParams)]
        lib.fa_sm100_kernel_name.restype = c_char_p
        lib.fa_sm100_rope.argtypes = [POINTER(RopeParams), c_void_p]
        lib.fa_sm100_rope.restype = c_int32
        lib.fa_sm100_selftest_umma.argtypes = [
            c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
            c_int32, c_int32, c_float, c_uint32, c_uint32, c_int32, c_void_p,
        ]
        lib.fa_sm100_selftest_umma.restype = c_int32
        _lib = lib
        return lib


def last_error() -> str:
    return load().fa_sm100_last_error().decode("utf-8", "replace")


def raise_for_status(rc: int) -> None:
    if rc == 0:
        return
    msg = f"fa_sm100 [{STATUS_NAMES.get(rc, rc)}]: {last_error()}"
    if rc == 1:
        raise RuntimeError(msg)  # the reference raises RuntimeError from TORCH_CHECK
    raise RuntimeError(msg)


def launch_count() -> int:
    return int(load().fa_sm100_launch_count())


def check_device(device: int) -> None:
    """Synchronise and raise if an in-kernel watchdog fired (tests / smoke)."""
    raise_for_status(load().fa_sm100_check_device(int(device)))
