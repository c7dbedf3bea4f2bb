"""Decode against a paged KV cache: ``flash_attn_decode_paged``.

"Paged KV cache" is a TODO of the reference (README.md:21); the contract is its single-query path
(csrc/flash_attention_api.cpp:72-83: one query row per sequence sees every key of that sequence)
applied to the keys a block table names.  Layouts follow the flash-attn comparator
(``flash_attn_with_kvcache``): ``k_cache`` / ``v_cache`` ``[num_pages, page_size, Hkv, D]`` (any
strides with unit D stride, so ``[num_pages, Hkv, page_size, D]`` storage works through a
``transpose(1, 2)`` view), ``block_table`` int32 ``[B, max_pages_per_seq]``, ``seqlens_k`` int32
``[B]``, both on the device.  One launch of the split-KV kernel through the C ABI
``fa_sm100_fwd_decode_paged``; CUDA only, no fallback.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_workspaces: dict = {}


FLAG_PAGED_ONE_WAVE = 16  # include/fa_sm100.h: FA_SM100_FLAG_PAGED_ONE_WAVE
_FLAT_BLOCK = 64          # keys per block of the flattened key space (fa_sm100_decode.cuh: kFlatBlock)


def prefer_one_wave(seqlens_host, heads_kv: int, num_sms: int = 148) -> bool:
    """Host-side choice between the two ways fa_sm100_fwd_decode_paged cuts the work, for callers
    that have the lengths on the host (serving engines do).  The flattened key space gives every
    CTA total / (2 SMs) blocks whatever the batch looks like; the one-wave split gives the CTAs of
    the longest sequence max_len / nsplit.  The latter has fewer partial results, which is worth
    some imbalance: it is chosen while its longest CTA stays within 1.3x the fair share."""
    lens = [int(n) for n in seqlens_host]
    if not lens or len(lens) > 2048:
        return True
    slots = 2 * num_sms
    units = len(lens) * heads_kv
    total = heads_kv * sum(max(1, -(-n // _FLAT_BLOCK)) for n in lens)
    fair = max(4.0, total / slots)
    max_len = max(max(lens), 1)
    nsplit = max(1, min(slots // units if units <= slots else 1, max_len // 256))
    longest = -(-max_len // nsplit) / _FLAT_BLOCK
    return longest <= 1.3 * fair


def flash_attn_decode_paged(q, k_cache, v_cache, block_table, seqlens_k, softmax_scale=None,
                            max_seqlen_k=None, seqlens_k_host=None, one_wave=None):
    """q: [B, Hq, D] or [B, Hq, 1, D]; returns o of the same shape.  ``max_seqlen_k`` (an upper
    bound of seqlens_k known on the host) sizes the split; default: the table's capacity.
    ``one_wave``: True = split every sequence alike (FA_SM100_FLAG_PAGED_ONE_WAVE), False = the
    CTAs share the flattened key space (robust against ragged batches), None = decide from
    ``seqlens_k_host`` (the same lengths as a host list / CPU tensor) when given, else False."""
    if not (q.is_cuda and k_cache.is_cuda and v_cache.is_cuda):
        raise NotImplementedError(
            "flash_attn_decode_paged is implemented for CUDA (sm_100a / B200) tensors only; "
            f"got device {q.device}. There is no CPU or PyTorch fallback.")
    squeeze = q.dim() == 4
    if squeeze:
        if q.size(2) != 1:
            raise RuntimeError("paged decode needs exactly one query position per sequence")
        q3 = q[:, :, 0, :]
    else:
        q3 = q
    if q3.dim() != 3 or k_cache.dim() != 4 or v_cache.dim() != 4:
        raise RuntimeError("q must be [B, Hq, D]; k_cache, v_cache must be [pages, page_size, Hkv, D]")
    if q3.dtype not in (torch.float16, torch.bfloat16) or k_cache.dtype != q3.dtype or v_cache.dtype != q3.dtype:
        raise RuntimeError("q, k, v only support fp16 or bf16 data type")
    if k_cache.shape != v_cache.shape or k_cache.size(3) != q3.size(2):
        raise RuntimeError("k_cache and v_cache must have the same shape and q's head dimension")
    if block_table.dtype != torch.int32 or seqlens_k.dtype != torch.int32 or not block_table.is_cuda \
            or not seqlens_k.is_cuda or block_table.dim() != 2 or seqlens_k.dim() != 1:
        raise RuntimeError("block_table [B, pages] and seqlens_k [B] must be int32 CUDA tensors")
    B, Hq, D = q3.shape
    if block_table.size(0) != B or seqlens_k.size(0) != B:
        raise RuntimeError("block_table and seqlens_k must have one row per sequence")
    if q3.stride(2) != 1 or q3.data_ptr() % 16 or any(s % 8 for s in q3.stride()[:2]):
        q3 = q3.contiguous()
    if k_cache.stride(3) != 1 or v_cache.stride(3) != 1:
        raise RuntimeError("k_cache, v_cache must be contiguous in the last dimension")
    block_table = block_table if block_table.stride(1) == 1 else block_table.contiguous()
    seqlens_k = seqlens_k.contiguous()
    o = torch.empty((B, Hq, D), dtype=q3.dtype, device=q3.device)
    page_size = k_cache.size(1)
    lib = _lib.load()
    p = _lib.PagedDecodeParams()
    p.q, p.k_cache, p.v_cache, p.o = q3.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), o.data_ptr()
    p.block_table, p.seqlens_k = block_table.data_ptr(), seqlens_k.data_ptr()
    p.batch, p.heads_q, p.heads_kv, p.headdim = B, Hq, k_cache.size(2), D
    p.page_size = page_size
    p.max_seqlen_k = int(max_seqlen_k) if max_seqlen_k else block_table.size(1) * page_size
    p.block_table_stride = block_table.stride(0)
    p.q_strides[0], p.q_strides[1] = q3.stride(0), q3.stride(1)
    p.o_strides[0], p.o_strides[1] = o.stride(0), o.stride(1)
    for dst, t in ((p.k_strides, k_cache), (p.v_strides, v_cache)):
        dst[0], dst[1], dst[2] = t.stride(0), t.stride(2), t.stride(1)  # page, head, in-page token
    p.softmax_scale = float(D ** -0.5 if softmax_scale is None else softmax_scale)
    p.dtype = _lib.FA_SM100_BF16 if q3.dtype == torch.bfloat16 else _lib.FA_SM100_FP16
    p.device = q3.device.index if q3.device.index is not None else torch.cuda.current_device()
    if one_wave is None:
        one_wave = seqlens_k_host is not None and prefer_one_wave(
            seqlens_k_host.tolist() if hasattr(seqlens_k_host, "tolist") else seqlens_k_host, k_cache.size(2))
    p.flags = FLAG_PAGED_ONE_WAVE if one_wave else 0
    stream = torch.cuda.current_stream(q3.device).cuda_stream
    nbytes = ctypes.c_int64()
    _lib.raise_for_status(lib.fa_sm100_paged_decode_plan(ctypes.byref(p), None, None, ctypes.byref(nbytes)))
    ws_ptr, ws_len = 0, 0
    if nbytes.value:
        key = (q3.device.index, stream)
        ws = _workspaces.get(key)
        if ws is None or ws.numel() < nbytes.value:
            ws = torch.empty(max(nbytes.value, 1 << 22), dtype=torch.uint8, device=q3.device)
            _workspaces[key] = ws
        ws_ptr, ws_len = ws.data_ptr(), ws.numel()
    _lib.raise_for_status(lib.fa_sm100_fwd_decode_paged(ctypes.byref(p), ctypes.c_void_p(ws_ptr),
                                                       ctypes.c_int64(ws_len), ctypes.c_void_p(stream)))
    return o.unsqueeze(2) if squeeze else o
