"""TEST INFRASTRUCTURE ONLY — the oracle.

Restatement of the reference's fp32 eager attention, ``eager_attention`` in
scripts/benchmark_kernel.py:15-44, which is the only correctness reference the upstream repo has
(it is compared against the kernel at scripts/benchmark_kernel.py:114-123).

Parity status: PINNED.  oracle/make_golden.py lifts the reference function out of its script by
AST (importing the script would JIT-build the CUDA extension) and checks this restatement against
it bit-for-bit on CPU for every golden case; the resulting vectors are committed under
tests/golden/ and re-checked by tests/test_oracle.py on every run.

Arithmetic, line by line (scripts/benchmark_kernel.py):
  :22      scale = 1/sqrt(D) unless given
  :26-28   q, k, v upcast to fp32
  :30-35   additive bias, -inf above the diagonal
  :37-38   GQA: k, v repeat_interleave(Hq/Hkv) along the head axis (q head h -> kv head h // G)
  :40-43   softmax(q·kᵀ·scale + bias)·v, all fp32
  :44      cast back to the input dtype
Two documented deviations:
  (i)  the mask is built on q's device (the reference builds it on CPU, :33);
  (ii) causal alignment is bottom-right, ``tril(diagonal=Sk-Sq)``, which is what the reference
       *kernel* implements (csrc/mask.cuh:37-39); identical to the reference's ``tril(0)`` whenever
       Sq == Sk, i.e. in every BASELINE config.
Rows with no visible key (causal, Sq > Sk) are defined as zeros (the kernel's documented
behaviour) instead of the NaN softmax(-inf) gives.
"""
from __future__ import annotations

import torch


def eager_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                    softmax_scale: float | None = None, is_causal: bool = False,
                    out_dtype: torch.dtype | None = None) -> torch.Tensor:
    scale_factor = 1 / (query.size(-1) ** 0.5) if softmax_scale is None else softmax_scale
    L, S = query.size(-2), key.size(-2)
    origin_dtype = query.dtype if out_dtype is None else out_dtype
    query = query.float()
    key = key.float()
    value = value.float()

    attn_bias = torch.zeros(L, S, dtype=query.dtype, device=query.device)
    if is_causal:
        keep = torch.ones(L, S, dtype=torch.bool, device=query.device).tril(diagonal=S - L)
        attn_bias.masked_fill_(keep.logical_not(), float("-inf"))

    group = query.size(-3) // key.size(-3)
    key = key.repeat_interleave(group, -3)
    value = value.repeat_interleave(group, -3)

    attn_weights = torch.matmul(query, key.transpose(-2, -1)) * scale_factor
    attn_weights += attn_bias
    attn_weights = torch.softmax(attn_weights, dim=-1)
    if is_causal and L > S:
        attn_weights = torch.nan_to_num(attn_weights, nan=0.0)  # fully masked rows -> zeros
    attn_output = torch.matmul(attn_weights, value)
    return attn_output.to(origin_dtype)


def eager_attention_chunked(query, key, value, softmax_scale=None, is_causal=False,
                            out_dtype=None, max_score_elems: int = 1 << 28) -> torch.Tensor:
    """Same arithmetic, evaluated per (batch, kv-head group, query-row chunk) so the fp32 score
    matrix stays below ``max_score_elems`` elements (for S >= 8K)."""
    B, Hq, L, D = query.shape
    Hkv, S = key.size(1), key.size(2)
    G = Hq // Hkv
    origin_dtype = query.dtype if out_dtype is None else out_dtype
    scale = 1 / (D ** 0.5) if softmax_scale is None else softmax_scale
    out = torch.empty(B, Hq, L, value.size(-1), dtype=origin_dtype, device=query.device)
    rows = max(1, min(L, max_score_elems // max(1, G * S)))
    for b in range(B):
        for h in range(Hkv):
            kf = key[b, h].float()
            vf = value[b, h].float()
            for r0 in range(0, L, rows):
                r1 = min(L, r0 + rows)
                qf = query[b, h * G:(h + 1) * G, r0:r1].float()  # [G, r, D]
                w = torch.matmul(qf, kf.t()) * scale
                if is_causal:
                    ri = torch.arange(r0, r1, device=query.device).view(-1, 1)
                    ci = torch.arange(S, device=query.device).view(1, -1)
                    w = w.masked_fill(ci > ri + (S - L), float("-inf"))
                w = torch.softmax(w, dim=-1)
                if is_causal and L > S:
                    w = torch.nan_to_num(w, nan=0.0)
                out[b, h * G:(h + 1) * G, r0:r1] = torch.matmul(w, vf).to(origin_dtype)
    return out


def attention_flops(batch: int, heads_q: int, seqlen_q: int, seqlen_k: int, headdim: int,
                    causal: bool) -> float:
    """Algorithmic FLOPs: 4·B·Hq·Sq·Sk·D, halved for causal (BASELINE.json north_star)."""
    f = 4.0 * batch * heads_q * seqlen_q * seqlen_k * headdim
    return f / 2 if causal else f
