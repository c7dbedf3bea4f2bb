"""Variable-length (packed) batches: ``flash_attn_varlen_func``.

The reference lists varlen as a TODO (README.md:18) and has no code for it; the call shape follows
the flash-attn comparator the reference benchmarks against (scripts/benchmark_kernel.py:10,49-55):
packed ``[total_tokens, heads, dim]`` tensors and int32 ``cu_seqlens``.  Each sequence is computed
exactly like one batch entry of ``flash_attn_func`` (same GQA rule, bottom-right causal alignment,
csrc/mask.cuh:37-39).  The work is done by the C ABI ``fa_sm100_fwd_varlen``
(include/fa_sm100.h): one launch of the sm_100a forward kernel per non-empty sequence on the
current stream.  CUDA only — there is no CPU or PyTorch fallback.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def _host_offsets(cu, name):
    if isinstance(cu, torch.Tensor):
        if cu.dtype != torch.int32 or cu.dim() != 1:
            raise RuntimeError(f"{name} must be a 1-D int32 tensor")
        cu = cu.tolist()  # a device tensor costs one small D2H read; pass a host tensor/list to avoid it
    cu = [int(x) for x in cu]
    if len(cu) < 1:
        raise RuntimeError(f"{name} must hold num_seqs + 1 offsets")
    return (ctypes.c_int32 * len(cu))(*cu), cu


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q=None, max_seqlen_k=None,
                           softmax_scale=None, causal=False):
    """q: [total_q, Hq, D]; k, v: [total_k, Hkv, D]; cu_seqlens_*: int32 [num_seqs + 1] (host list,
    host tensor or device tensor).  Returns o: [total_q, Hq, D] with q's strides.
    ``max_seqlen_*`` are accepted for call compatibility and only checked."""
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
    cu_q, lq = _host_offsets(cu_seqlens_q, "cu_seqlens_q")
    cu_k, lk = _host_offsets(cu_seqlens_k, "cu_seqlens_k")
    if len(lq) != len(lk):
        raise RuntimeError("cu_seqlens_q and cu_seqlens_k must have the same length")
    if lq[-1] != q.size(0) or lk[-1] != k.size(0):
        raise RuntimeError("cu_seqlens must end at the number of packed tokens")
    n = len(lq) - 1
    if max_seqlen_q is not None and n and max(b - a for a, b in zip(lq, lq[1:])) > max_seqlen_q:
        raise RuntimeError("max_seqlen_q is smaller than the longest query sequence")
    if max_seqlen_k is not None and n and max(b - a for a, b in zip(lk, lk[1:])) > max_seqlen_k:
        raise RuntimeError("max_seqlen_k is smaller than the longest key sequence")
    q = q if q.stride(2) == 1 else q.contiguous()
    k = k if k.stride(2) == 1 else k.contiguous()
    v = v if v.stride(2) == 1 else v.contiguous()
    o = torch.empty_like(q)
    if q.numel() == 0:
        return o
    p = _lib.VarlenParams()
    p.q, p.k, p.v, p.o = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr()
    p.cu_seqlens_q, p.cu_seqlens_kv = cu_q, cu_k
    p.num_seqs = n
    p.heads_q, p.heads_kv, p.headdim = q.size(1), k.size(1), q.size(2)
    for name, t in (("q", q), ("k", k), ("v", v), ("o", o)):
        getattr(p, name + "_strides")[0] = t.stride(0)
        getattr(p, name + "_strides")[1] = t.stride(1)
    p.softmax_scale = float(q.size(2) ** -0.5 if softmax_scale is None else softmax_scale)
    p.causal = int(bool(causal))
    p.dtype = _lib.FA_SM100_BF16 if q.dtype == torch.bfloat16 else _lib.FA_SM100_FP16
    p.device = q.device.index if q.device.index is not None else torch.cuda.current_device()
    stream = torch.cuda.current_stream(q.device).cuda_stream
    _lib.raise_for_status(_lib.load().fa_sm100_fwd_varlen(ctypes.byref(p), ctypes.c_void_p(stream)))
    return o
