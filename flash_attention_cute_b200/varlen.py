"""Variable-length (packed) batches: ``flash_attn_varlen_func``.

The reference lists varlen as a TODO (README.md:18) and has no code for it; the call shape follows
the flash-attn comparator the reference benchmarks against (scripts/benchmark_kernel.py:10,49-55):
packed ``[total_tokens, heads, dim]`` tensors and int32 ``cu_seqlens``.  Each sequence is computed
exactly like one batch entry of ``flash_attn_func`` (same GQA rule, bottom-right causal alignment,
csrc/mask.cuh:37-39; bit-equal results).  The work is done by the C ABI ``fa_sm100_fwd_varlen``
(include/fa_sm100.h): ONE persistent launch of the sm_100a forward kernel whose tile scheduler
walks the DEVICE ``cu_seqlens`` (plus a one-CTA plan kernel in front) — no host read of the
offsets, no synchronisation, CUDA-graph capturable.  CUDA only — there is no CPU or PyTorch
fallback.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_workspaces: dict = {}


def _device_offsets(cu, name, device):
    """int32 offsets on `device`.  A device tensor is used as it is; a host list / tensor costs one
    small async H2D copy (no sync).  Returns (tensor, host list or None)."""
    if isinstance(cu, torch.Tensor):
        if cu.dtype != torch.int32 or cu.dim() != 1:
            raise RuntimeError(f"{name} must be a 1-D int32 tensor")
        if cu.numel() < 1:
            raise RuntimeError(f"{name} must hold num_seqs + 1 offsets")
        if cu.is_cuda:
            if cu.device != device:
                raise RuntimeError(f"{name} must be on the same CUDA device as q")
            return (cu if cu.is_contiguous() else cu.contiguous()), None
        host = cu.tolist()
    else:
        host = [int(x) for x in cu]
        if len(host) < 1:
            raise RuntimeError(f"{name} must hold num_seqs + 1 offsets")
    if host[0] != 0:
        raise RuntimeError("cu_seqlens must start at 0")
    if any(b < a for a, b in zip(host, host[1:])):
        raise RuntimeError("cu_seqlens must be non-decreasing")
    return torch.tensor(host, dtype=torch.int32).to(device, non_blocking=True), host


def _workspace(nbytes: int, device, stream) -> torch.Tensor:
    key = (device.index, stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q=None, max_seqlen_k=None,
                           softmax_scale=None, causal=False):
    """q: [total_q, Hq, D]; k, v: [total_k, Hkv, D]; cu_seqlens_*: int32 [num_seqs + 1] (device
    tensor — used in place, nothing is read back — or host list / host tensor).  Returns
    o: [total_q, Hq, D] with q's strides.  ``max_seqlen_q`` only steers the GQA packing choice
    (queries shorter than 256 rows pack the heads of a kv group into one tile); ``max_seqlen_k``
    is accepted for call compatibility.  Device offsets are trusted (they must start at 0, be
    non-decreasing and end at the number of packed tokens); host offsets are checked."""
    if not (q.is_cuda and k.is_cuda and v.is_cuda):
        raise NotImplementedError(
            "flash_attn_varlen_func is implemented for CUDA (sm_100a / B200) tensors only; "
            f"got device {q.device}. There is no CPU or PyTorch fallback.")
    if q.dim() != 3 or k.dim() != 3 or v.dim() != 3:
        raise RuntimeError("q, k, v must be [total_tokens, heads, dim]")
    if q.dtype not in (torch.float16, torch.bfloat16) or k.dtype != q.dtype or v.dtype != q.dtype:
        raise RuntimeError("q, k, v only support fp16 or bf16 data type")
    if k.shape != v.shape or q.size(2) != k.size(2):
        raise RuntimeError("k and v must have the same shape and q's head dimension")
    if q.device != k.device or q.device != v.device:
        raise RuntimeError("q, k, v must be on the same CUDA device")
    cu_q, lq = _device_offsets(cu_seqlens_q, "cu_seqlens_q", q.device)
    cu_k, lk = _device_offsets(cu_seqlens_k, "cu_seqlens_k", q.device)
    if cu_q.numel() != cu_k.numel():
        raise RuntimeError("cu_seqlens_q and cu_seqlens_k must have the same length")
    n = cu_q.numel() - 1
    if lq is not None:
        if lq[-1] != q.size(0):
            raise RuntimeError("cu_seqlens must end at the number of packed tokens")
        if max_seqlen_q is not None and n and max(b - a for a, b in zip(lq, lq[1:])) > max_seqlen_q:
            raise RuntimeError("max_seqlen_q is smaller than the longest query sequence")
        if max_seqlen_q is None and n:
            max_seqlen_q = max(b - a for a, b in zip(lq, lq[1:]))
    if lk is not None:
        if lk[-1] != k.size(0):
            raise RuntimeError("cu_seqlens must end at the number of packed tokens")
        if max_seqlen_k is not None and n and max(b - a for a, b in zip(lk, lk[1:])) > max_seqlen_k:
            raise RuntimeError("max_seqlen_k is smaller than the longest key sequence")

    def _ok(t):  # 16-byte aligned base, token / head strides multiples of 8 elements, unit D stride
        return t.stride(2) == 1 and t.data_ptr() % 16 == 0 and t.stride(0) % 8 == 0 and t.stride(1) % 8 == 0 \
            and t.stride(0) > 0 and t.stride(1) > 0

    def _fix(t):
        if t.numel() == 0 or _ok(t):
            return t
        c = t.contiguous()
        return c.clone() if c.data_ptr() % 16 else c

    q, k, v = _fix(q), _fix(k), _fix(v)
    o = torch.empty_like(q)
    if q.numel() and not _ok(o):
        o = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    if q.numel() == 0 or n == 0:
        return o
    lib = _lib.load()
    p = _lib.VarlenParams()
    p.q, p.k, p.v, p.o = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr()
    p.cu_seqlens_q, p.cu_seqlens_kv = cu_q.data_ptr(), cu_k.data_ptr()
    p.num_seqs = n
    p.total_q, p.total_kv = q.size(0), k.size(0)
    p.max_seqlen_q = int(max_seqlen_q) if max_seqlen_q else 0
    p.heads_q, p.heads_kv, p.headdim = q.size(1), k.size(1), q.size(2)
    for name, t in (("q", q), ("k", k), ("v", v), ("o", o)):
        getattr(p, name + "_strides")[0] = t.stride(0) if t.size(0) > 1 else t.size(1) * t.size(2)
        getattr(p, name + "_strides")[1] = t.stride(1) if t.size(1) > 1 else t.size(2)
    p.softmax_scale = float(q.size(2) ** -0.5 if softmax_scale is None else softmax_scale)
    p.causal = int(bool(causal))
    p.dtype = _lib.FA_SM100_BF16 if q.dtype == torch.bfloat16 else _lib.FA_SM100_FP16
    p.device = q.device.index if q.device.index is not None else torch.cuda.current_device()
    stream = torch.cuda.current_stream(q.device).cuda_stream
    ws = _workspace(int(lib.fa_sm100_varlen_workspace_bytes(n)), q.device, stream)
    _lib.raise_for_status(lib.fa_sm100_fwd_varlen(ctypes.byref(p), ctypes.c_void_p(ws.data_ptr()),
                                                 ctypes.c_int64(ws.numel()), ctypes.c_void_p(stream)))
    return o
