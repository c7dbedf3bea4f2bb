"""Split-KV decode kernel (seqlen_q == 1; reference anchor: the packing
branch csrc/flash_attention_api.cpp:72-83 and the decode loop scripts/benchmark_llm.py:76-88).

CPU: the split plan (host logic, no launch).  GPU: parity against the fp32 eager oracle through the
public op, against the prefill kernel on the same inputs, KV-cache style strided views, ragged and
tiny sequence lengths, every group size up to 16, repeated launches on one workspace."""
import pytest
import torch

from common import assert_close, make_qkv_fast
from oracle.eager_attention import eager_attention


def _ext():
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    return load_extension()


def _plan(B, Hq, Hkv, Sk, D=128):
    q = torch.empty(B, Hq, 1, D, dtype=torch.float16)
    k = torch.empty(B, Hkv, Sk, D, dtype=torch.float16)
    return _ext().decode_plan(q, k, k)


# ------------------------------------------------------------------ CPU: host logic
def test_plan_reference_decode_case():
    # scripts/benchmark_kernel.py usage example: B=16, 64/8 heads, Sk=1024 -> 128 units, 2 splits
    p = _plan(16, 64, 8, 1024)
    assert p["nsplit"] == 2 and p["chunk"] == 512
    assert p["workspace_bytes"] == 128 * 2 * 16 * (128 + 2) * 4  # partials only: no counter region


def test_plan_one_wave_and_chunk_granularity():
    for B, Hkv, Sk in [(1, 8, 65536), (1, 1, 300), (4, 2, 8192), (2, 8, 1000)]:
        p = _plan(B, Hkv * 4, Hkv, Sk)
        assert p["chunk"] % 64 == 0
        assert (p["nsplit"] - 1) * p["chunk"] < Sk <= p["nsplit"] * p["chunk"]
        assert B * Hkv * p["nsplit"] <= 2 * 148 or p["nsplit"] == 1
        assert p["nsplit"] == 1 or p["chunk"] >= 256


def test_plan_many_units_need_no_workspace():
    p = _plan(64, 32, 8, 4096)  # 512 units > one wave
    assert p["nsplit"] == 1 and p["workspace_bytes"] == 0


def test_plan_rejects_what_belongs_to_prefill():
    q = torch.empty(1, 8, 2, 128, dtype=torch.float16)
    k = torch.empty(1, 8, 64, 128, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="seqlen_q == 1"):
        _ext().decode_plan(q, k, k)
    q = torch.empty(1, 64, 1, 128, dtype=torch.float16)
    k = torch.empty(1, 2, 64, 128, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="at most 16"):
        _ext().decode_plan(q, k, k)


# ------------------------------------------------------------------ GPU: parity
def _run(q, k, v, flags=0, scale=None):
    from flash_attention_cute_b200 import _lib

    scale = q.size(-1) ** -0.5 if scale is None else scale
    before = _lib.launch_count()
    out = _ext().flash_attention_fwd(q, k, v, scale, False, flags)
    _lib.check_device(q.device.index or 0)
    assert _lib.launch_count() == before + 1, "decode must be a single launch"
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
@pytest.mark.parametrize("B,Hq,Hkv,Sk,D", [
    (16, 64, 8, 1024, 128),   # the reference's decode example
    (2, 32, 8, 1000, 128),    # ragged
    (1, 8, 8, 4096, 128),     # MHA (G = 1), many splits
    (3, 14, 2, 777, 64),      # G = 7, D = 64
    (2, 14, 1, 2048, 64),     # G = 14
    (1, 16, 1, 65, 128),      # G = 16, barely two tiles
    (2, 4, 4, 1, 128),        # a single key: O == V
    (1, 8, 2, 15, 96),        # shorter than a tile, D not a power of two
    (1, 32, 8, 32768, 128),   # long context, one wave of splits
])
def test_decode_matches_oracle(cuda_device, dtype, B, Hq, Hkv, Sk, D):
    q, k, v = make_qkv_fast(B, Hq, Hkv, 1, Sk, D, dtype, 11, cuda_device)
    out = _run(q, k, v)
    assert out.shape == q.shape and out.dtype == dtype
    assert_close(out, eager_attention(q, k, v, None, False, out_dtype=torch.float32), dtype, "decode")


@pytest.mark.gpu
def test_decode_equals_prefill_kernel(cuda_device):
    from flash_attention_cute_b200.load_cpp_extention import FLAG_NO_DECODE

    q, k, v = make_qkv_fast(4, 32, 8, 1, 3000, 128, torch.bfloat16, 3, cuda_device)
    a = _run(q, k, v)
    b = _run(q, k, v, flags=FLAG_NO_DECODE)
    ref = eager_attention(q, k, v, None, False, out_dtype=torch.float32)
    assert_close(a, ref, torch.bfloat16, "split-KV")
    assert_close(b, ref, torch.bfloat16, "prefill kernel, Sq=1")
    assert (a.float() - b.float()).abs().max().item() <= 2e-2


@pytest.mark.gpu
def test_decode_kv_cache_views_and_custom_scale(cuda_device):
    # a [B, S_max, H, D] cache of which the first Sk positions are valid, viewed as [B, H, S, D]
    B, Hq, Hkv, Smax, Sk, D = 2, 16, 4, 640, 517, 128
    g = torch.Generator(device=cuda_device).manual_seed(5)
    cache_k = torch.randn(B, Smax, Hkv, D, generator=g, device=cuda_device).half()
    cache_v = torch.randn(B, Smax, Hkv, D, generator=g, device=cuda_device).half()
    q = torch.randn(B, 1, Hq, D, generator=g, device=cuda_device).half().transpose(1, 2)
    k = cache_k[:, :Sk].transpose(1, 2)
    v = cache_v[:, :Sk].transpose(1, 2)
    out = _run(q, k, v, scale=0.05)
    assert out.stride() == q.stride()
    assert_close(out, eager_attention(q, k, v, 0.05, False, out_dtype=torch.float32), torch.float16, "cache view")


@pytest.mark.gpu
def test_decode_workspace_is_reusable_and_large_logits(cuda_device):
    q, k, v = make_qkv_fast(1, 8, 2, 1, 6000, 128, torch.float16, 9, cuda_device)
    q = q * 6.0  # logits of a few tens: the running max moves, partial maxima differ between splits
    ref = eager_attention(q, k, v, None, False, out_dtype=torch.float32)
    outs = [_run(q, k, v) for _ in range(5)]  # same workspace, counters must re-arm
    for o in outs:
        assert_close(o, ref, torch.float16, "repeat")
    assert all(torch.equal(outs[0], o) for o in outs[1:])


@pytest.mark.gpu
def test_cached_plan_path_matches_first_call(cuda_device):
    """The operator caches the validated launch plan per problem key; the second call with the same
    shapes (different data, different buffers) goes through the cache and must agree with the oracle,
    for the prefill and the decode route alike; a misaligned base pointer falls back to the full path."""
    ext = _ext()
    for Sq, Sk in [(1, 900), (200, 200)]:
        for seed in (1, 2, 3):
            q, k, v = make_qkv_fast(2, 8, 2, Sq, Sk, 128, torch.float16, seed, cuda_device)
            out = ext.flash_attention_fwd(q, k, v, 128 ** -0.5, True)
            assert_close(out, eager_attention(q, k, v, None, True, out_dtype=torch.float32), torch.float16,
                         f"cached call Sq={Sq} seed={seed}")
    assert len(ext._plans) >= 2
    flat = torch.randn(2 * 8 * 200 * 128 + 1, device=cuda_device).half()
    q_odd = flat[1:].view(2, 8, 200, 128)  # 2-byte offset: not TMA compatible
    q, k, v = make_qkv_fast(2, 8, 2, 200, 200, 128, torch.float16, 4, cuda_device)
    out = ext.flash_attention_fwd(q_odd, k, v, 128 ** -0.5, True)
    assert_close(out, eager_attention(q_odd, k, v, None, True, out_dtype=torch.float32), torch.float16, "misaligned q")


@pytest.mark.gpu
def test_decode_shape_changes_on_one_stream(cuda_device):
    """one grow-only workspace per stream serves every shape.  A call with few
    units and many splits followed by one with more units must not find its arrival counters
    clobbered by the first call's partials (the counters now live in a library-owned ring, the
    workspace holds partial results only)."""
    shapes = [(1, 32, 8, 6000), (16, 64, 8, 1024), (2, 32, 8, 3000), (16, 64, 8, 1024), (1, 8, 8, 9000),
              (9, 64, 8, 2048), (16, 64, 8, 1024)]
    for i, (B, Hq, Hkv, Sk) in enumerate(shapes):
        q, k, v = make_qkv_fast(B, Hq, Hkv, 1, Sk, 128, torch.bfloat16, 50 + i, cuda_device)
        plan = _ext().decode_plan(q, k, k)
        assert plan["nsplit"] > 1
        out = _run(q, k, v)
        assert_close(out, eager_attention(q, k, v, None, False, out_dtype=torch.float32), torch.bfloat16,
                     f"call {i}: B={B} units={B * Hkv} nsplit={plan['nsplit']}")


@pytest.mark.gpu
def test_decode_is_graph_capturable(cuda_device):
    """Cluster launch (2 <= nsplit <= 8, merge through DSMEM), the L2 merge (nsplit > 8) and the
    flattened paged kernel all record into a CUDA graph and replay to the eager bits."""
    from flash_attention import flash_attn_decode_paged

    ext = _ext()
    cases = []
    for B, Hq, Hkv, Sk in [(16, 64, 8, 1024), (1, 32, 8, 20000), (3, 16, 4, 2000)]:
        q, k, v = make_qkv_fast(B, Hq, Hkv, 1, Sk, 128, torch.bfloat16, 70 + B, cuda_device)
        assert (2 <= ext.decode_plan(q, k, k)["nsplit"] <= 8) == (B != 1)
        cases.append(lambda q=q, k=k, v=v: ext.flash_attention_fwd(q, k, v, 128 ** -0.5, False))
    pq, kc, vc, bt, sl, _, _ = _paged_case(6, 32, 8, 128, 64, [3000, 10, 0, 777, 64, 1500], torch.bfloat16,
                                           cuda_device, 9)
    cases.append(lambda: flash_attn_decode_paged(pq, kc, vc, bt, sl, max_seqlen_k=3000))
    cases.append(lambda: flash_attn_decode_paged(pq, kc, vc, bt, sl, max_seqlen_k=3000, one_wave=True))
    eager = [fn() for fn in cases]   # also creates the per-stream workspaces outside the capture
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for fn in cases:
            fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            outs = [fn() for fn in cases]
    for _ in range(3):
        for o in outs:
            o.zero_()
        g.replay()
        torch.cuda.synchronize()
        for i, (o, e) in enumerate(zip(outs, eager)):
            assert torch.equal(o, e), f"case {i}"
    from flash_attention_cute_b200 import _lib
    _lib.check_device(0)


# ------------------------------------------------------------------ paged KV cache
def _paged_case(B, Hq, Hkv, D, page, lens, dtype, device, seed, layout="pshd", shared=False):
    """Random pages in random order; returns (q, k_cache, v_cache, block_table, seqlens, dense k, dense v)."""
    g = torch.Generator(device=device).manual_seed(seed)
    max_pages = max(1, (max(lens) + page - 1) // page)
    n_pages = B * max_pages + 3
    if layout == "pshd":
        kc = torch.randn(n_pages, page, Hkv, D, generator=g, device=device).to(dtype)
        vc = torch.randn(n_pages, page, Hkv, D, generator=g, device=device).to(dtype)
    else:  # storage [pages, Hkv, page, D], handed over as a [pages, page, Hkv, D] view
        kc = torch.randn(n_pages, Hkv, page, D, generator=g, device=device).to(dtype).transpose(1, 2)
        vc = torch.randn(n_pages, Hkv, page, D, generator=g, device=device).to(dtype).transpose(1, 2)
    perm = torch.randperm(n_pages, generator=g, device=device).to(torch.int32)
    bt = perm[: B * max_pages].view(B, max_pages).contiguous()
    if shared:  # every sequence starts with the same (prefix) page
        bt[:, 0] = bt[0, 0]
    q = torch.randn(B, Hq, D, generator=g, device=device).to(dtype)
    sl = torch.tensor(lens, dtype=torch.int32, device=device)
    dense_k = kc[bt.long()].reshape(B, max_pages * page, Hkv, D).transpose(1, 2)  # [B, Hkv, S, D]
    dense_v = vc[bt.long()].reshape(B, max_pages * page, Hkv, D).transpose(1, 2)
    return q, kc, vc, bt, sl, dense_k, dense_v


def test_prefer_one_wave_from_host_lengths():
    from flash_attention_cute_b200.paged import prefer_one_wave

    assert prefer_one_wave([8192] * 16, 8)                    # uniform, chunks align with units
    assert prefer_one_wave([4096] * 64, 8)                    # 512 equal units, one CTA each
    assert not prefer_one_wave([16384] + [512] * 31, 8)       # one long sequence among short ones
    assert not prefer_one_wave([1024 + 478 * i for i in range(16)], 8)   # 1 K ... 8 K
    assert prefer_one_wave([64 + 4 * i for i in range(256)], 8)          # 2048 short units
    assert not prefer_one_wave([131072] + [100] * 255, 8)     # ... unless one of them is huge
    assert prefer_one_wave([100] * 3000, 8)                   # beyond what the flat kernel takes


def test_paged_plan_validation():
    import ctypes

    from flash_attention_cute_b200 import _lib

    lib = _lib.load()
    p = _lib.PagedDecodeParams()
    p.q = p.k_cache = p.v_cache = p.o = p.block_table = p.seqlens_k = 256
    p.batch, p.heads_q, p.heads_kv, p.headdim = 16, 64, 8, 128
    p.page_size, p.max_seqlen_k, p.block_table_stride = 128, 8192, 64
    p.q_strides[0], p.q_strides[1] = 64 * 128, 128
    p.o_strides[0], p.o_strides[1] = 64 * 128, 128
    for st in (p.k_strides, p.v_strides):
        st[0], st[1], st[2] = 128 * 8 * 128, 128, 8 * 128
    p.softmax_scale = 0.088
    ns, ch, nb = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int64()
    assert lib.fa_sm100_paged_decode_plan(ctypes.byref(p), ctypes.byref(ns), ctypes.byref(ch), ctypes.byref(nb)) == 0
    # per-sequence lengths are device data: one wave of CTAs (2 per SM) shares the flattened key
    # space, two partial slots per CTA, at least 256 keys per CTA; nsplit 0 says so
    assert ns.value == 0 and ch.value == 256 and nb.value == 2 * 296 * 16 * (128 + 2) * 4
    p.batch = 4096   # more sequences than a CTA keeps prefix sums for: one CTA per (unit, split)
    assert lib.fa_sm100_paged_decode_plan(ctypes.byref(p), ctypes.byref(ns), ctypes.byref(ch), ctypes.byref(nb)) == 0
    assert ns.value == 1 and nb.value == 0
    p.batch = 16
    p.flags = 16     # FA_SM100_FLAG_PAGED_ONE_WAVE: the split of the contiguous decode kernel
    assert lib.fa_sm100_paged_decode_plan(ctypes.byref(p), ctypes.byref(ns), ctypes.byref(ch), ctypes.byref(nb)) == 0
    assert ns.value == 2 and ch.value == 4096 and nb.value == 128 * 2 * 16 * (128 + 2) * 4
    p.flags = 0
    p.page_size = 24
    assert lib.fa_sm100_paged_decode_plan(ctypes.byref(p), None, None, None) == 1
    assert "multiple of 16" in _lib.last_error()
    p.page_size, p.block_table_stride = 128, 8
    assert lib.fa_sm100_paged_decode_plan(ctypes.byref(p), None, None, None) == 1
    assert "smaller than max_seqlen_k" in _lib.last_error()


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
@pytest.mark.parametrize("B,Hq,Hkv,D,page,lens,layout,shared", [
    (4, 32, 8, 128, 128, [1000, 517, 128, 1], "pshd", False),      # ragged last pages, a single key
    (2, 14, 2, 64, 16, [333, 40], "pshd", False),                   # smallest page, D = 64
    (3, 16, 1, 128, 256, [4096, 0, 2500], "pshd", False),           # an empty sequence -> zeros
    (16, 64, 8, 128, 128, [8192] * 16, "phsd", False),              # the measured shape, head-major pages
    (5, 8, 8, 128, 64, [700, 64, 65, 63, 1024], "pshd", True),      # shared prefix page
    (1, 32, 8, 128, 128, [32768], "pshd", False),                   # one long sequence, many splits
])
@pytest.mark.parametrize("one_wave", [False, True], ids=["flat", "one_wave"])
def test_paged_decode_matches_oracle(cuda_device, dtype, B, Hq, Hkv, D, page, lens, layout, shared, one_wave):
    from flash_attention import flash_attn_decode_paged
    from flash_attention_cute_b200 import _lib

    q, kc, vc, bt, sl, dk, dv = _paged_case(B, Hq, Hkv, D, page, lens, dtype, cuda_device, 77, layout, shared)
    n0 = _lib.launch_count()
    out = flash_attn_decode_paged(q, kc, vc, bt, sl, max_seqlen_k=max(max(lens), 1), one_wave=one_wave)
    _lib.check_device(0)
    assert _lib.launch_count() == n0 + 1 and out.shape == q.shape
    for b, n in enumerate(lens):
        if n == 0:
            assert not out[b].any()
            continue
        ref = eager_attention(q[b:b + 1].unsqueeze(2), dk[b:b + 1, :, :n], dv[b:b + 1, :, :n], None, False,
                              out_dtype=torch.float32)
        assert_close(out[b:b + 1].unsqueeze(2), ref, dtype, f"paged seq {b} len {n}")
    # same keys through the contiguous decode kernel: bit-equal when the split is the same
    if len(set(lens)) == 1 and lens[0] % page == 0:
        dense = _ext().flash_attention_fwd(q.unsqueeze(2), dk.contiguous(), dv.contiguous(), D ** -0.5, False)
        if one_wave:
            assert torch.equal(dense[:, :, 0], out)
        else:
            assert_close(out, dense[:, :, 0].float(), dtype, "paged (flattened) vs contiguous decode")


def _check_paged(cuda_device, B, Hq, Hkv, D, page, lens, dtype, seed, sample=None):
    from flash_attention import flash_attn_decode_paged
    from flash_attention_cute_b200 import _lib

    q, kc, vc, bt, sl, dk, dv = _paged_case(B, Hq, Hkv, D, page, lens, dtype, cuda_device, seed)
    n0 = _lib.launch_count()
    out = flash_attn_decode_paged(q, kc, vc, bt, sl, max_seqlen_k=max(max(lens), 1))
    _lib.check_device(0)
    assert _lib.launch_count() == n0 + 1
    again = flash_attn_decode_paged(q, kc, vc, bt, sl, max_seqlen_k=max(max(lens), 1))
    assert torch.equal(out, again), "two launches on the same inputs differ (stale counters or partials)"
    for b in (range(B) if sample is None else sample):
        n = lens[b]
        if n == 0:
            assert not out[b].any()
            continue
        ref = eager_attention(q[b:b + 1].unsqueeze(2), dk[b:b + 1, :, :n], dv[b:b + 1, :, :n], None, False,
                              out_dtype=torch.float32)
        assert_close(out[b:b + 1].unsqueeze(2), ref, dtype, f"paged seq {b} len {n}")
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["many_short", "skewed", "tiny_total", "all_empty", "exact_blocks", "d64_mqa"])
def test_paged_decode_flattened_key_space(cuda_device, name):
    """The CTAs share the flattened key space (fa_decode_flat_sm100): ranges that hold many whole
    units, units cut into 2 .. 40 pieces, a unit that starts exactly at a range boundary, batches
    whose total is below one range, empty sequences between non-empty ones."""
    import random

    rng = random.Random(11)
    if name == "many_short":      # 300 x 2 units, 0 .. 600 keys: a CTA walks several units
        lens = [rng.choice([0, 1, 63, 64, 65, rng.randrange(1, 600)]) for _ in range(300)]
        _check_paged(cuda_device, 300, 8, 2, 128, 64, lens, torch.bfloat16, 1, sample=range(0, 300, 7))
    elif name == "skewed":        # one 16 K sequence cut into many pieces, 31 short ones
        lens = [16384] + [rng.randrange(1, 700) for _ in range(31)]
        _check_paged(cuda_device, 32, 64, 8, 128, 128, lens, torch.bfloat16, 2)
    elif name == "tiny_total":    # fewer blocks than CTAs: most CTAs have nothing to do
        _check_paged(cuda_device, 2, 8, 1, 128, 16, [100, 30], torch.float16, 3)
    elif name == "all_empty":
        out = _check_paged(cuda_device, 5, 16, 4, 128, 16, [0] * 5, torch.float16, 4)
        assert not out.any()
    elif name == "exact_blocks":  # 37 units of exactly 8 blocks: W = 4 -> every unit starts a range
        _check_paged(cuda_device, 37, 4, 1, 128, 64, [512] * 37, torch.bfloat16, 5)
    else:                          # D = 64, MQA with 14 q heads, ragged
        lens = [rng.randrange(0, 3000) for _ in range(24)]
        _check_paged(cuda_device, 24, 14, 1, 64, 32, lens, torch.float16, 6)


@pytest.mark.gpu
@pytest.mark.parametrize("one_wave", [False, True], ids=["flat", "one_wave"])
def test_paged_decode_page_size_not_a_power_of_two(cuda_device, one_wave):
    """48- and 80-key pages: the kernels fall back from shifts to divisions (page_shift = -1)."""
    from flash_attention import flash_attn_decode_paged

    for page, lens in [(48, [1000, 47, 48, 49, 0, 2500]), (80, [81, 4000, 160])]:
        q, kc, vc, bt, sl, dk, dv = _paged_case(len(lens), 16, 4, 128, page, lens, torch.bfloat16, cuda_device, 21)
        out = flash_attn_decode_paged(q, kc, vc, bt, sl, max_seqlen_k=max(lens), one_wave=one_wave)
        for b, n in enumerate(lens):
            if n == 0:
                assert not out[b].any()
                continue
            ref = eager_attention(q[b:b + 1].unsqueeze(2), dk[b:b + 1, :, :n], dv[b:b + 1, :, :n], None, False,
                                  out_dtype=torch.float32)
            assert_close(out[b:b + 1].unsqueeze(2), ref, torch.bfloat16, f"page {page} seq {b} len {n}")


@pytest.mark.gpu
@pytest.mark.parametrize("one_wave", [False, True], ids=["flat", "one_wave"])
def test_paged_decode_clamps_lengths_outside_the_contract(cuda_device, one_wave):
    """seqlens_k[b] > max_seqlen_k or < 0 breaks the contract; the kernels clamp instead of
    walking out of the block table: the result is that of the clamped lengths."""
    from flash_attention import flash_attn_decode_paged

    lens = [1000, 700, 333, 1000]
    q, kc, vc, bt, sl, dk, dv = _paged_case(4, 16, 4, 128, 64, lens, torch.bfloat16, cuda_device, 33)
    good = flash_attn_decode_paged(q, kc, vc, bt, sl, max_seqlen_k=1000, one_wave=one_wave)
    bad = torch.tensor([50000, 700, -7, 1000], dtype=torch.int32, device=cuda_device)
    out = flash_attn_decode_paged(q, kc, vc, bt, bad, max_seqlen_k=1000, one_wave=one_wave)
    for b in (0, 1, 3):   # (the flattened split points move with the total: close, not bit-equal)
        assert_close(out[b:b + 1], good[b:b + 1].float(), torch.bfloat16, f"seq {b}")
    assert not out[2].any()


@pytest.mark.gpu
def test_paged_decode_exactly_as_many_sequences_as_the_flat_kernel_takes(cuda_device):
    """B = 2048: the prefix sums fill the shared-memory array to its last entry."""
    import random

    rng = random.Random(6)
    lens = [rng.randrange(0, 130) for _ in range(2048)]
    lens[-1] = 129
    _check_paged(cuda_device, 2048, 4, 1, 128, 16, lens, torch.bfloat16, 12, sample=[0, 1, 1023, 2046, 2047])


@pytest.mark.gpu
def test_paged_decode_more_sequences_than_the_flat_kernel_takes(cuda_device):
    """B > 2048: one CTA per unit (fa_decode_sm100<paged>), lengths still per sequence."""
    import random

    rng = random.Random(5)
    lens = [rng.randrange(0, 80) for _ in range(2100)]
    _check_paged(cuda_device, 2100, 4, 1, 128, 16, lens, torch.bfloat16, 8, sample=range(0, 2100, 97))


@pytest.mark.gpu
def test_paged_decode_ignores_pages_past_the_sequence(cuda_device):
    from flash_attention import flash_attn_decode_paged

    q, kc, vc, bt, sl, dk, dv = _paged_case(3, 32, 8, 128, 128, [300, 1024, 129], torch.bfloat16, cuda_device, 5)
    ref = flash_attn_decode_paged(q, kc, vc, bt, sl)
    used = torch.zeros(kc.size(0), dtype=torch.bool, device=cuda_device)
    for b, n in enumerate([300, 1024, 129]):
        used[bt[b, : (n + 127) // 128].long()] = True
    kc2, vc2 = kc.clone(), vc.clone()
    kc2[~used] = float("nan")
    vc2[~used] = float("nan")
    bt2 = bt.clone()
    for b, n in enumerate([300, 1024, 129]):   # table entries past the sequence may hold anything valid
        bt2[b, (n + 127) // 128:] = 0
    out = flash_attn_decode_paged(q, kc2, vc2, bt2, sl)
    assert torch.equal(out, ref)
