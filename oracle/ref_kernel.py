"""TEST INFRASTRUCTURE ONLY — launcher for the reference's own sm80 CuTe kernel on B200.

oracle/_ref/libref_fa.so is the reference's kernel TU compiled for sm_100a from the sources where
they lie (oracle/build_ref.sh).  This wrapper redoes the torch-side part of the reference host
function: validation is assumed done by the caller, the Sq == 1 q-head packing reshape follows
csrc/flash_attention_api.cpp:72-83,128-133, ``o = empty_like(q)`` follows :85.
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_int64, c_void_p, POINTER
from pathlib import Path

import torch

LIB_PATH = Path(__file__).resolve().parent / "_ref" / "libref_fa.so"
_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(str(LIB_PATH))
        _lib.ref_fa_fwd.argtypes = [c_void_p] * 4 + [c_int64] * 7 + [POINTER(c_int64)] * 4 + [
            c_float, c_int, c_int, c_void_p]
        _lib.ref_fa_fwd.restype = c_int
    return _lib


def ref_flash_attn_func(q, k, v, softmax_scale=None, causal=False):
    lib = _load()
    assert q.is_cuda and q.dtype in (torch.float16, torch.bfloat16) and q.size(3) % 8 == 0
    softmax_scale = q.size(-1) ** -0.5 if softmax_scale is None else softmax_scale
    bs, hq, sq, d = q.shape
    hkv, skv = k.size(1), k.size(2)
    g = hq // hkv
    pack = sq == 1
    q_in = q
    if pack:
        hq_k, sq_k, causal = hkv, sq * g, False
        q_in = q.reshape(bs, hq_k, sq_k, d)
    else:
        hq_k, sq_k = hq, sq
    o = torch.empty_like(q_in)

    def st(t):
        return (c_int64 * 3)(t.stride(0), t.stride(1), t.stride(2))

    stream = torch.cuda.current_stream(q.device).cuda_stream
    rc = lib.ref_fa_fwd(q_in.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(),
                        bs, hq_k, hkv, sq_k, skv, d, 1 if pack else g,
                        st(q_in), st(k), st(v), st(o),
                        float(softmax_scale), int(bool(causal)), int(q.dtype == torch.bfloat16),
                        c_void_p(stream))
    if rc != 0:
        raise RuntimeError(f"reference kernel launch failed with cudaError {rc}")
    return o.reshape(bs, hq, sq, d) if pack else o
