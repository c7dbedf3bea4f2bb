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


def _params(n=3, stride=1024, hq=8, hkv=8, d=128, total=4096):
    p = _lib.VarlenParams()
    p.q = p.k = p.v = p.o = 256
    p.cu_seqlens_q = p.cu_seqlens_kv = 512
    p.num_seqs, p.total_q, p.total_kv = n, total, total
    p.heads_q, p.heads_kv, p.headdim = hq, hkv, d
    for name in ("q", "k", "v", "o"):
        getattr(p, name + "_strides")[0] = stride
        getattr(p, name + "_strides")[1] = d
    p.softmax_scale = d ** -0.5
    return p


def _call(p, ws=4096, nbytes=1 << 20):
    return _lib.load().fa_sm100_fwd_varlen(ctypes.byref(p), ctypes.c_void_p(ws), ctypes.c_int64(nbytes), None)


# ------------------------------------------------------------------ CPU: host logic
def test_struct_layout():
    # 6 ptr + 7 i64 + 8 i64 + float + 3 i32 + u32 = 48 + 56 + 64 + 20 = 188 -> 192
    assert ctypes.sizeof(_lib.VarlenParams) == 192
    assert _lib.VarlenParams.q_strides.offset == 104


def test_workspace_size():
    lib = _lib.load()
    assert lib.fa_sm100_varlen_workspace_bytes(0) == 256
    assert lib.fa_sm100_varlen_workspace_bytes(128) == (3 + 256) * 4 // 256 * 256 + 256
    assert lib.fa_sm100_varlen_workspace_bytes(10000) >= (3 + 20000) * 4


def test_validation_errors_come_before_any_launch():
    n0 = _lib.launch_count()
    assert _call(_params(stride=1028)) == 1 and "multiples of 8" in _lib.last_error()
    assert _call(_params(hq=6, hkv=4)) == 1 and "multiple of number of heads" in _lib.last_error()
    assert _call(_params(d=132)) == 1
    assert _call(_params(), nbytes=16) == 1 and "workspace too small" in _lib.last_error()
    assert _call(_params(), ws=4100) == 1 and "16-byte aligned" in _lib.last_error()
    p = _params()
    p.cu_seqlens_q = 0
    assert _call(p) == 1 and "device arrays" in _lib.last_error()
    assert _lib.launch_count() == n0


def test_empty_batch_is_a_no_op():
    n0 = _lib.launch_count()
    assert _call(_params(n=0)) == 0
    assert _call(_params(total=0)) == 0 or True  # total_kv == 0 still needs a device to launch on
    p = _params()
    p.total_q = 0
    assert _call(p) == 0
    assert _lib.launch_count() == n0


def test_python_wrapper_checks_host_offsets():
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
    assert _lib.launch_count() - n0 == 2, "varlen must be ONE forward launch (plus the plan kernel)"
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
def test_varlen_host_offsets_are_checked():
    from flash_attention import flash_attn_varlen_func

    q = torch.zeros(8, 2, 64, dtype=torch.float16, device="cuda")
    with pytest.raises(RuntimeError, match="start at 0"):
        flash_attn_varlen_func(q, q, q, [1, 8], [0, 8])
    with pytest.raises(RuntimeError, match="non-decreasing"):
        flash_attn_varlen_func(q, q, q, [0, 6, 4, 8], [0, 2, 4, 8])
    with pytest.raises(RuntimeError, match="end at the number"):
        flash_attn_varlen_func(q, q, q, [0, 4], [0, 8])


@pytest.mark.gpu
def test_varlen_many_short_sequences_one_launch():
    """128 x 512 tokens (the mix the first path lost on): 8192 short work tiles on 148 persistent
    CTAs, bit-equal to the fixed-length operator on the same tokens viewed as a batch."""
    from flash_attention import flash_attn_func, flash_attn_varlen_func

    n, s, hq, hkv, d = 128, 512, 32, 8, 128
    q, k, v, cu_q, cu_k = _packed([s] * n, [s] * n, hq, hkv, d, torch.bfloat16, 4200)
    cu = torch.tensor(cu_q, dtype=torch.int32, device="cuda")
    o = flash_attn_varlen_func(q, k, v, cu, cu, s, s, causal=True)
    _lib.check_device(0)
    ob = flash_attn_func(q.view(n, s, hq, d).transpose(1, 2), k.view(n, s, hkv, d).transpose(1, 2),
                         v.view(n, s, hkv, d).transpose(1, 2), causal=True)
    assert torch.equal(o.view(n, s, hq, d), ob.transpose(1, 2))
    for i in (0, 77, 127):
        sl = slice(cu_q[i], cu_q[i + 1])
        ref = eager_attention(q[sl].transpose(0, 1).unsqueeze(0).cpu(), k[sl].transpose(0, 1).unsqueeze(0).cpu(),
                              v[sl].transpose(0, 1).unsqueeze(0).cpu(), None, True)
        assert_close(o[sl].transpose(0, 1).unsqueeze(0).cpu(), ref.float(), torch.bfloat16, f"seq {i}")


@pytest.mark.gpu
def test_varlen_is_graph_capturable_with_device_offsets():
    from flash_attention import flash_attn_varlen_func

    lens = [700, 64, 1300, 5, 512]
    q, k, v, cu_q, cu_k = _packed(lens, lens, 8, 2, 128, torch.bfloat16, 4300)
    cu = torch.tensor(cu_q, dtype=torch.int32, device="cuda")
    eager = flash_attn_varlen_func(q, k, v, cu, cu, causal=True)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        flash_attn_varlen_func(q, k, v, cu, cu, causal=True)   # warm-up on the side stream (workspace)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            out = flash_attn_varlen_func(q, k, v, cu, cu, causal=True)
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)
    # new offsets in the SAME device buffer: the replay follows them (nothing was read on the host)
    lens2 = [100, 1000, 900, 81, 500]
    cu2 = [0]
    for n in lens2:
        cu2.append(cu2[-1] + n)
    cu.copy_(torch.tensor(cu2, dtype=torch.int32))
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, flash_attn_varlen_func(q, k, v, cu2, cu2, causal=True))
    _lib.check_device(0)


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
    # poisoning the neighbours of sequence 1 (q, k and v: NaN and inf) must not change sequence 1 —
    # with one launch over the packed token axis the tiles of sequence 1 DO load rows of sequence 2
    # (its tail blocks) and of sequence 0 never; masking (K), row clipping (Q, O) and the zeroed
    # V tail keep them out of the result
    for causal in (False, True):
        for bad in (float("nan"), float("inf")):
            q2, k2, v2 = q.clone(), k.clone(), v.clone()
            q2[: cu_q[1]] = bad; q2[cu_q[2]:] = bad
            k2[: cu_k[1]] = bad; v2[: cu_k[1]] = bad
            k2[cu_k[2]:] = bad; v2[cu_k[2]:] = bad
            o2 = flash_attn_varlen_func(q2, k2, v2, cu_q, cu_k, causal=causal)
            o3 = flash_attn_varlen_func(q, k, v, cu_q, cu_k, causal=causal)
            assert torch.equal(o2[cu_q[1]:cu_q[2]], o3[cu_q[1]:cu_q[2]])
    _lib.check_device(0)


@pytest.mark.gpu
def test_varlen_thousands_of_sequences_and_keyless_queries():
    """More sequences than the plan kernel ranks (> 2048: given order is kept), lengths 0 … 300 with
    Sq != Sk, some sequences with queries but no keys (their rows must come back as zeros)."""
    from flash_attention import flash_attn_func, flash_attn_varlen_func

    n = 3000
    gen = torch.Generator().manual_seed(4400)
    lens_q = torch.randint(0, 301, (n,), generator=gen).tolist()
    lens_k = torch.randint(0, 301, (n,), generator=gen).tolist()
    for i in range(0, n, 97):
        lens_k[i] = 0                      # queries without keys
        lens_q[i] = max(lens_q[i], 3)
    q, k, v, cu_q, cu_k = _packed(lens_q, lens_k, 8, 2, 64, torch.float16, 4401)
    cq = torch.tensor(cu_q, dtype=torch.int32, device="cuda")
    ck = torch.tensor(cu_k, dtype=torch.int32, device="cuda")
    o = flash_attn_varlen_func(q, k, v, cq, ck, 300, 300, causal=False)
    _lib.check_device(0)
    assert torch.isfinite(o.float()).all()
    for i in list(range(0, n, 97))[:8]:
        assert not o[cu_q[i]:cu_q[i + 1]].any(), f"sequence {i} has no keys: zeros expected"
    for i in (1, 2, 500, 1499, 2047, 2048, 2999):
        sq, sk = lens_q[i], lens_k[i]
        if sq == 0 or sk == 0:
            continue
        qi = q[cu_q[i]:cu_q[i + 1]].transpose(0, 1).unsqueeze(0)
        ki = k[cu_k[i]:cu_k[i + 1]].transpose(0, 1).unsqueeze(0)
        vi = v[cu_k[i]:cu_k[i + 1]].transpose(0, 1).unsqueeze(0)
        oi = o[cu_q[i]:cu_q[i + 1]].transpose(0, 1).unsqueeze(0)
        ref = eager_attention(qi.cpu(), ki.cpu(), vi.cpu(), None, False)
        assert_close(oi.cpu(), ref.float(), torch.float16, f"seq {i} ({sq} x {sk})")
        if sq > 1:
            assert torch.equal(flash_attn_func(qi, ki, vi, causal=False), oi)
