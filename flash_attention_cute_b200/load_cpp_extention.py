"""Counterpart of the reference's flash_attention/load_cpp_extention.py:11-54.

``load_extension()`` returns an object exposing ``flash_attention_fwd(q, k, v, softmax_scale,
causal) -> Tensor`` — the one function the reference's pybind module ``flash_attention_extension``
exports (csrc/flash_attention_api.cpp:137-141).  Here it is a thin host shim: the reference's
argument checks (csrc/flash_attention_api.cpp:22-59, same messages), ``o = empty_like(q)``
(:85, so O inherits q's strides), then one call through the C ABI ``fa_sm100_fwd`` on the
current torch stream (:115).  PyTorch only provides device memory and the stream.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_DTYPES = {torch.float16: _lib.FA_SM100_FP16, torch.bfloat16: _lib.FA_SM100_BF16}


FLAG_NO_DECODE = 8       # include/fa_sm100.h: FA_SM100_FLAG_NO_DECODE
DECODE_MAX_GROUP = 16    # csrc/fa_sm100_decode.cuh: kRows


def _check(cond: bool, msg: str) -> None:
    if not cond:
        raise RuntimeError(msg)


def _tma_compatible(t: torch.Tensor) -> bool:
    if t.data_ptr() % 16:
        return False
    for size, stride in zip(t.shape[:3], t.stride()[:3]):
        if size > 1 and (stride <= 0 or stride % 8):
            return False
    return True


def _compatible_copy(t: torch.Tensor) -> torch.Tensor:
    """One dense copy with an aligned base (`.contiguous()` alone returns a contiguous tensor as it is,
    misaligned storage offset included)."""
    c = t.contiguous()
    return c.clone() if c.data_ptr() % 16 else c


def fill_params(q, k, v, o, softmax_scale: float, causal: bool, flags: int = 0) -> _lib.FwdParams:
    p = _lib.FwdParams()
    p.q, p.k, p.v, p.o = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr()
    p.batch, p.heads_q, p.seqlen_q, p.headdim = q.shape
    p.heads_kv, p.seqlen_kv = k.shape[1], k.shape[2]
    for dst, t in ((p.q_strides, q), (p.k_strides, k), (p.v_strides, v), (p.o_strides, o)):
        dst[0], dst[1], dst[2] = t.stride(0), t.stride(1), t.stride(2)
    p.softmax_scale = float(softmax_scale)
    p.causal = 1 if causal else 0
    p.dtype = _DTYPES[q.dtype]
    if q.is_cuda:
        p.device = q.device.index if q.device.index is not None else torch.cuda.current_device()
    else:
        p.device = 0
    p.flags = int(flags)
    return p


class FlashAttentionExtension:
    """Stands in for the pybind module ``flash_attention_extension``."""

    name = "flash_attention_extension"

    def __init__(self) -> None:
        self._lib = _lib.load()
        self._workspaces = {}
        self._plans = {}  # problem key -> (FwdParams bytes, use_decode, workspace bytes, o strides)
        self._rope_plans = {}  # problem key -> RopeParams bytes

    def flash_attention_fwd(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                            softmax_scale: float, causal: bool, flags: int = 0) -> torch.Tensor:
        # Fast path: this exact problem (shapes, strides, dtypes, devices, scale, flags) has been
        # validated and planned before — only the four pointers change.  The checks, the params
        # fill and the plan query below cost 25-35 us of Python per call, more than a decode kernel.
        key = (q.shape, q.stride(), k.shape, k.stride(), v.shape, v.stride(), q.dtype, k.dtype, v.dtype,
               q.device, k.device, v.device, float(softmax_scale), bool(causal), int(flags))
        ent = self._plans.get(key)
        if ent is not None and not ((q.data_ptr() | k.data_ptr() | v.data_ptr()) & 15):
            o = torch.empty_like(q)
            if o.stride() == ent[3] and not (o.data_ptr() & 15):
                params = _lib.FwdParams.from_buffer_copy(ent[0])
                params.q, params.k, params.v, params.o = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr()
                stream = torch.cuda.current_stream(q.device).cuda_stream
                if ent[1]:
                    ws = self._workspace_for(ent[2], q.device, stream)
                    rc = self._lib.fa_sm100_fwd_decode(ctypes.byref(params),
                                                       ctypes.c_void_p(ws.data_ptr() if ws is not None else 0),
                                                       ctypes.c_int64(ws.numel() if ws is not None else 0),
                                                       ctypes.c_void_p(stream))
                else:
                    rc = self._lib.fa_sm100_fwd(ctypes.byref(params), ctypes.c_void_p(stream))
                _lib.raise_for_status(rc)
                return o

        # --- the reference's checks, csrc/flash_attention_api.cpp:22-59 ---
        _check(q.dim() == 4 and k.dim() == 4 and v.dim() == 4, "q, k, v must be 4-D [batch, heads, seq, dim]")
        _check(q.size(0) == k.size(0) and q.size(0) == v.size(0), "q, k, v must have the same batch size")
        _check(k.size(1) == v.size(1), "k, v must have the same number of heads")
        _check(k.size(2) == v.size(2), "k, v must have the same sequence length")
        _check(q.size(3) == k.size(3) and q.size(3) == v.size(3), "q, k, v must have the same hidden dimension")
        _check(q.size(0) > 0 and q.size(1) > 0 and q.size(2) > 0 and q.size(3) > 0,
               "q, k, v must have at least one element")
        _check(k.size(1) > 0 and k.size(2) > 0, "q, k, v must have at least one element")
        _check(q.size(1) >= k.size(1),
               "number of heads in q must be greater or equal to number of heads in k and v")
        _check(q.size(1) % k.size(1) == 0,
               "number of heads in q must be multiple of number of heads in k and v")
        _check(q.dtype == k.dtype and q.dtype == v.dtype, "q, k, v must have the same data type")
        _check(q.dtype in _DTYPES, "q, k, v only support fp16 or bf16 data type")
        _check(q.stride(3) == 1, "q must be contiguous in the last dimension")
        _check(k.stride(3) == 1, "k must be contiguous in the last dimension")
        _check(v.stride(3) == 1, "v must be contiguous in the last dimension")
        _check(q.size(3) % 8 == 0, "hidden dimension must be multiple of 8")
        _check(q.size(3) <= 128, "only support hidden dimension <= 128")
        _check(q.is_cuda and k.is_cuda and v.is_cuda, "q, k, v must be on CUDA device")
        _check(q.device == k.device and q.device == v.device, "q, k, v must be on the same CUDA device")

        # TMA needs 16-byte aligned bases and strides; anything else gets one contiguous copy.
        cacheable = True
        if not _tma_compatible(q):
            q, cacheable = _compatible_copy(q), False
        if not _tma_compatible(k):
            k, cacheable = _compatible_copy(k), False
        if not _tma_compatible(v):
            v, cacheable = _compatible_copy(v), False
        o = torch.empty_like(q)  # csrc/flash_attention_api.cpp:85 — O keeps q's strides
        if not _tma_compatible(o) or o.stride(3) != 1:
            o, cacheable = torch.empty(q.shape, dtype=q.dtype, device=q.device), False

        params = fill_params(q, k, v, o, softmax_scale, causal, flags)
        stream = torch.cuda.current_stream(q.device).cuda_stream
        # Sq == 1 (the reference's packing branch, csrc/flash_attention_api.cpp:72-83) goes to the
        # split-KV decode kernel whenever it applies: G <= 16 q heads per kv head, positive scale.
        use_decode = (q.size(2) == 1 and q.size(1) // k.size(1) <= DECODE_MAX_GROUP and softmax_scale > 0
                      and not (flags & FLAG_NO_DECODE))
        ws_bytes = 0
        if use_decode:
            nbytes = ctypes.c_int64()
            _lib.raise_for_status(self._lib.fa_sm100_decode_plan(ctypes.byref(params), None, None,
                                                                ctypes.byref(nbytes)))
            ws_bytes = nbytes.value
            ws = self._workspace_for(ws_bytes, q.device, stream)
            rc = self._lib.fa_sm100_fwd_decode(ctypes.byref(params),
                                               ctypes.c_void_p(ws.data_ptr() if ws is not None else 0),
                                               ctypes.c_int64(ws.numel() if ws is not None else 0),
                                               ctypes.c_void_p(stream))
        else:
            rc = self._lib.fa_sm100_fwd(ctypes.byref(params), ctypes.c_void_p(stream))
        _lib.raise_for_status(rc)
        if cacheable:
            if len(self._plans) >= 4096:
                self._plans.clear()
            self._plans[key] = (bytes(params), use_decode, ws_bytes, o.stride())
        return o

    def rope(self, q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor):
        """(q·cos + rotate_half(q)·sin, k·cos + rotate_half(k)·sin) in one kernel, bit-equal to the
        reference expression (models/rope_attn_fwd.py:14-38).  q [B,Hq,S,D], k [B,Hkv,S,D] (strided
        views welcome), cos / sin [B or 1, S, D].  Outputs keep the inputs' strides."""
        # Fast path, as in flash_attention_fwd: this exact problem has been validated before, only
        # the pointers change (the checks and the params fill are ~20 us of Python, more than the
        # kernel on decode-sized inputs).
        key = (q.shape, q.stride(), k.shape, k.stride(), cos.shape, cos.stride(), sin.stride(), q.dtype, k.dtype,
               cos.dtype, sin.dtype, q.device, k.device, cos.device, sin.device)
        ent = self._rope_plans.get(key)
        if ent is not None and not ((q.data_ptr() | k.data_ptr() | cos.data_ptr() | sin.data_ptr()) & 15):
            q_out, k_out = torch.empty_like(q), torch.empty_like(k)
            if q_out.stride() == q.stride() and k_out.stride() == k.stride() and \
                    not ((q_out.data_ptr() | k_out.data_ptr()) & 15):
                p = _lib.RopeParams.from_buffer_copy(ent)
                p.q, p.k, p.q_out, p.k_out = q.data_ptr(), k.data_ptr(), q_out.data_ptr(), k_out.data_ptr()
                p.cos, p.sin = cos.data_ptr(), sin.data_ptr()
                stream = torch.cuda.current_stream(q.device).cuda_stream
                _lib.raise_for_status(self._lib.fa_sm100_rope(ctypes.byref(p), ctypes.c_void_p(stream)))
                return q_out, k_out
        q_in, k_in, cos_in, sin_in = q, k, cos, sin
        _check(q.dim() == 4 and k.dim() == 4, "q, k must be 4-D [batch, heads, seq, dim]")
        _check(cos.dim() == 3 and sin.dim() == 3 and cos.shape == sin.shape, "cos, sin must be 3-D [batch, seq, dim]")
        _check(q.size(0) == k.size(0) and q.size(2) == k.size(2) and q.size(3) == k.size(3),
               "q, k must agree in batch, sequence length and hidden dimension")
        _check(cos.size(0) in (1, q.size(0)) and cos.size(1) == q.size(2) and cos.size(2) == q.size(3),
               "cos, sin must be [batch or 1, seq, dim]")
        _check(q.dtype == k.dtype == cos.dtype == sin.dtype and q.dtype in _DTYPES,
               "q, k, cos, sin must share one of fp16 / bf16")
        _check(q.size(3) % 16 == 0, "rope: hidden dimension must be a multiple of 16")
        _check(q.is_cuda and k.is_cuda and cos.is_cuda and sin.is_cuda, "q, k, cos, sin must be on CUDA device")
        if not _tma_compatible(q) or q.stride(3) != 1:
            q = _compatible_copy(q)
        if not _tma_compatible(k) or k.stride(3) != 1:
            k = _compatible_copy(k)
        if cos.stride(2) != 1 or cos.stride(1) % 8 or cos.data_ptr() % 16 or (cos.size(0) > 1 and cos.stride(0) % 8):
            cos = cos.contiguous()
        if sin.stride() != cos.stride() or sin.data_ptr() % 16:
            cos, sin = cos.contiguous(), sin.contiguous()
        q_out, k_out = torch.empty_like(q), torch.empty_like(k)
        if q_out.stride() != q.stride():
            q_out = torch.empty(q.shape, dtype=q.dtype, device=q.device)
        if k_out.stride() != k.stride():
            k_out = torch.empty(k.shape, dtype=k.dtype, device=k.device)
        p = _lib.RopeParams()
        p.q, p.k, p.q_out, p.k_out = q.data_ptr(), k.data_ptr(), q_out.data_ptr(), k_out.data_ptr()
        p.cos, p.sin = cos.data_ptr(), sin.data_ptr()
        p.batch, p.heads_q, p.seqlen, p.headdim = q.shape
        p.heads_kv = k.size(1)
        for dst, t in ((p.q_strides, q), (p.k_strides, k), (p.q_out_strides, q_out), (p.k_out_strides, k_out)):
            dst[0], dst[1], dst[2] = t.stride(0), t.stride(1), t.stride(2)
        p.cos_sin_strides[0] = cos.stride(0) if cos.size(0) > 1 else 0
        p.cos_sin_strides[1] = cos.stride(1)
        p.dtype = _DTYPES[q.dtype]
        p.device = q.device.index if q.device.index is not None else torch.cuda.current_device()
        stream = torch.cuda.current_stream(q.device).cuda_stream
        _lib.raise_for_status(self._lib.fa_sm100_rope(ctypes.byref(p), ctypes.c_void_p(stream)))
        if q is q_in and k is k_in and cos is cos_in and sin is sin_in and q_out.stride() == q.stride() \
                and k_out.stride() == k.stride():   # nothing was copied: the plan holds for this key
            if len(self._rope_plans) >= 1024:
                self._rope_plans.clear()
            self._rope_plans[key] = bytes(p)
        return q_out, k_out

    def decode_plan(self, q, k, v) -> dict:
        """Split the decode kernel would use for these shapes (no launch; works without a GPU)."""
        params = fill_params(q, k, v, q, 1.0, False, 0)
        nsplit, chunk, nbytes = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int64()
        rc = self._lib.fa_sm100_decode_plan(ctypes.byref(params), ctypes.byref(nsplit), ctypes.byref(chunk),
                                            ctypes.byref(nbytes))
        _lib.raise_for_status(rc)
        return {"nsplit": nsplit.value, "chunk": chunk.value, "workspace_bytes": nbytes.value}

    def _workspace_for(self, nbytes: int, device, stream):
        """Grow-only decode workspace per (device, stream): launches on one stream are ordered, so
        they can share it.  It only holds the partial results of one launch — no state survives a
        launch, whatever the shapes of consecutive calls (include/fa_sm100.h: fa_sm100_fwd_decode;
        the arrival counters live in a library-owned ring)."""
        if nbytes == 0:
            return None
        key = (device.index, stream)
        ws = self._workspaces.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = torch.empty(max(nbytes, 1 << 22), dtype=torch.uint8, device=device)
            self._workspaces[key] = ws
        return ws

    def plan(self, q, k, v, causal: bool = False, flags: int = 0) -> dict:
        """Describe the launch fa_sm100_fwd would do (no launch; works without a GPU)."""
        params = fill_params(q, k, v, q, 1.0, causal, flags)
        out = [ctypes.c_int32() for _ in range(4)]
        rc = self._lib.fa_sm100_plan(ctypes.byref(params), *[ctypes.byref(x) for x in out])
        _lib.raise_for_status(rc)
        return dict(zip(("grid", "smem_bytes", "pack_g", "rows_per_head"), (x.value for x in out)))


_ext = None


def load_extension() -> FlashAttentionExtension:
    global _ext
    if _ext is None:
        _ext = FlashAttentionExtension()
    return _ext
