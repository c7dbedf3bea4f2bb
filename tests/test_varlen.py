"""Variable-length (packed) batches — SURVEY §8(f) rank 4 (reference TODO at README.md:18).

CPU: argument checks of the C ABI (host logic, no launch).  GPU: every sequence of a packed batch
must equal (i) the fp32 eager oracle of that sequence alone and (ii) bit for bit the fixed-length
operator called on that sequence alone — and tokens of neighbouring sequences must never leak."""
import ctypes

import pytest
import torch

from common import assert_close
from flash_attention_cute_b200 import _lib
from oracle.eager_attention import eager_attention


def _params(cu_q, cu_k, stride=1024, hq=8, hkv=8, d=128):
    p = _lib.VarlenParams()
    p.q = p.k = p.v = p.o = 256
    a = (ctypes.c_int32 * len(cu_q))(*cu_q)
    b = (ctypes.c_int32 * len(cu_k))(*cu_k)
    p.cu_seqlens_q, p.cu_seqlens_kv = a, b
    p.num_seqs = len(cu_q) - 1
    p.heads_q, p.heads_kv, p.headdim = hq, hkv, d
    for name in ("q", "k", "v", "o"):
        getattr(p, name + "_strides")[0] = stride
        getattr(p, name + "_strides")[1] = d
    p.softmax_scale = d ** -0.5
    p._keep = (a, b)
    return p


# ------------------------------------------------------------------ CPU: host logic
def test_struct_layout():
    # 4 ptr + 2 ptr + 4 i64 + 8 i64 + float + 3 i32 + u32 = 48 + 32 + 64 + 20 = 164 -> 168
    assert ctypes.sizeof(_lib.VarlenParams) == 168
    assert _lib.VarlenParams.q_strides.offset == 80


@pytest.mark.parametrize("cu_q, cu_k, stride, msg", [
    ([1, 4], [0, 4], 1024, "start at 0"),
    ([0, 4, 2], [0, 4, 8], 1024, "non-decreasing"),
    ([0, 4], [0, 4], 1028, "multiples of 8"),
])
def test_validation_errors(cu_q, cu_k, stride, msg):
    lib = _lib.load()
    p = _params(cu_q, cu_k, stride)
    assert lib.fa_sm100_fwd_varlen(ctypes.byref(p), None) == 1
    assert msg in _lib.last_error()


def test_all_empty_is_a_no_op():
    lib = _lib.load()
    n0 = _lib.launch_count()
    p = _params([0, 0, 0], [0, 5, 9])
    assert lib.fa_sm100_fwd_varlen(ctypes.byref(p), None) == 0
    assert _lib.launch_count() == n0


def test_no_cpu_fallback():
    from flash_attention import flash_attn_varlen_func

    q = torch.zeros(4, 2, 64, dtype=torch.float16)
    with pytest.raises(NotImplementedError, match="no CPU"):
        flash_attn_varlen_func(q, q, q, [0, 4], [0, 4])


# ------------------------------------------------------------------ GPU: parity
def _packed(lens_q, lens_k, hq, hkv, d, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, hq, d, generator=g, device="cuda").to(dtype)
    k = torch.randn(tk, hkv, d, generator=g, device="cuda").to(dtype)
    v = torch.randn(tk, hkv, d, generator=g, device="cuda").to(dtype)
    cu_q = [0]
    for n in lens_q:
        cu_q.append(cu_q[-1] + n)
    cu_k = [0]
    for n in lens_k:
        cu_k.append(cu_k[-1] + n)
    return q, k, v, cu_q, cu_k


CASES = [
    # lens_q, lens_k, Hq, Hkv, D
    ([128, 1000, 37, 512], [128, 1000, 37, 512], 8, 8, 128),
    ([300, 1, 2049], [300, 777, 2049], 32, 8, 128),      # a decode row inside a prefill batch
    ([64, 0, 129, 5], [200, 33, 129, 700], 14, 2, 64),   # an empty sequence, Sq != Sk
    ([4096, 17], [4096, 17], 14, 1, 64),
]


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("case", range(len(CASES)))
def test_varlen_parity(case, causal, dtype):
    from flash_attention import flash_attn_func, flash_attn_varlen_func

    lens_q, lens_k, hq, hkv, d = CASES[case]
    q, k, v, cu_q, cu_k = _packed(lens_q, lens_k, hq, hkv, d, dtype, 4000 + case)
    n0 = _lib.launch_count()
    o = flash_attn_varlen_func(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), causal=causal)
    assert _lib.launch_count() - n0 == sum(1 for n in lens_q if n > 0)
    _lib.check_device(0)
    assert o.shape == q.shape and o.stride() == q.stride()
    for i, (sq, sk) in enumerate(zip(lens_q, lens_k)):
        if sq == 0:
            continue
        qi = q[cu_q[i]:cu_q[i + 1]].transpose(0, 1).unsqueeze(0)   # [1, H, S, D] view
        ki = k[cu_k[i]:cu_k[i + 1]].transpose(0, 1).unsqueeze(0)
        vi = v[cu_k[i]:cu_k[i + 1]].transpose(0, 1).unsqueeze(0)
        oi = o[cu_q[i]:cu_q[i + 1]].transpose(0, 1).unsqueeze(0)
        if causal and sq > sk:
            continue  # rows without a visible key: documented difference (include/fa_sm100.h)
        ref = eager_attention(qi.cpu(), ki.cpu(), vi.cpu(), None, causal)
        assert_close(oi.cpu(), ref.float(), dtype, f"case {case} seq {i}")
        if sq == 1:
            continue  # flash_attn_func routes a single query row to the split-KV decode kernel
        same = flash_attn_func(qi, ki, vi, causal=causal)
        assert torch.equal(same, oi), f"case {case} seq {i}: differs from the fixed-length operator"


@pytest.mark.gpu
def test_varlen_device_offsets_strided_views_and_no_leak():
    from flash_attention import flash_attn_varlen_func

    lens = [257, 64, 900]
    q, k, v, cu_q, cu_k = _packed(lens, lens, 8, 2, 128, torch.bfloat16, 4100)
    # q, k, v as column slices of one fused projection output [T, (Hq + 2 Hkv) * D]
    fused = torch.cat([q.flatten(1), k.flatten(1), v.flatten(1)], dim=1).contiguous()
    qv = fused[:, : 8 * 128].view(-1, 8, 128)
    kv = fused[:, 8 * 128: 10 * 128].view(-1, 2, 128)
    vv = fused[:, 10 * 128:].view(-1, 2, 128)
    cu = torch.tensor(cu_q, dtype=torch.int32, device="cuda")
    o1 = flash_attn_varlen_func(qv, kv, vv, cu, cu, causal=True)
    o0 = flash_attn_varlen_func(q, k, v, cu_q, cu_k, causal=True)
    assert torch.equal(o0, o1)
    # poisoning the neighbours of sequence 1 must not change sequence 1
    k2, v2 = k.clone(), v.clone()
    k2[: cu_k[1]] = float("nan"); v2[: cu_k[1]] = float("nan")
    k2[cu_k[2]:] = float("nan"); v2[cu_k[2]:] = float("nan")
    o2 = flash_attn_varlen_func(q, k2, v2, cu_q, cu_k, causal=False)
    o3 = flash_attn_varlen_func(q, k, v, cu_q, cu_k, causal=False)
    assert torch.equal(o2[cu_q[1]:cu_q[2]], o3[cu_q[1]:cu_q[2]])
    _lib.check_device(0)
