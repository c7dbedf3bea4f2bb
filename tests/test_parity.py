"""GPU parity: the CUDA path (through the public op -> C ABI) against the fp32 eager oracle, the
committed golden vectors generated from the reference's own eager function, and the reference's
own sm80 CuTe kernel rebuilt for sm_100a (oracle/_ref/libref_fa.so) on identical seeded inputs.
Tolerances: BASELINE.json north_star (see tests/common.py)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from common import TOL, assert_close, errors, make_qkv, make_qkv_fast
from oracle import ref_kernel
from oracle.eager_attention import eager_attention, eager_attention_chunked

pytestmark = pytest.mark.gpu

GOLDEN = sorted((Path(__file__).parent / "golden").glob("eager_*.npz"))


def fa(q, k, v, causal=False, scale=None, flags=0):
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    if flags:
        scale = q.size(-1) ** -0.5 if scale is None else scale
        out = load_extension().flash_attention_fwd(q, k, v, scale, causal, flags)
    else:
        from flash_attention import flash_attn_func

        out = flash_attn_func(q, k, v, softmax_scale=scale, causal=causal)
    _lib.check_device(q.device.index or 0)
    return out


def oracle32(q, k, v, causal, scale=None):
    if q.size(2) * k.size(2) * q.size(1) * q.size(0) > (1 << 27):
        return eager_attention_chunked(q, k, v, scale, causal, out_dtype=torch.float32)
    return eager_attention(q, k, v, scale, causal, out_dtype=torch.float32)


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
def test_against_reference_golden(cuda_device, path):
    g = np.load(path)
    B, Hq, Hkv, Sq, Sk, D = (int(x) for x in g["shape"])
    dtype = getattr(torch, str(g["dtype"]))
    run_dtype = torch.float16 if dtype == torch.float32 else dtype  # the op is 16-bit only
    q, k, v = make_qkv(B, Hq, Hkv, Sq, Sk, D, dtype, int(g["seed"]))
    if dtype == torch.float32:
        pytest.skip("fp32 vectors pin the oracle; the operator accepts fp16/bf16 only")
    out = fa(q.to(cuda_device, run_dtype), k.to(cuda_device, run_dtype), v.to(cuda_device, run_dtype),
             bool(g["causal"]))
    rows = out.float()[:, :, :: int(g["row_step"]), :].cpu()
    tol = TOL[run_dtype]
    # golden rows were rounded to `dtype` by the reference; allow one more rounding
    gold = torch.from_numpy(g["rows"])
    assert ((rows - gold).abs() / gold.abs().clamp_min(1.0)).max().item() <= tol["max_abs"]


SHAPES = [
    # B, Hq, Hkv, Sq, Sk, D
    (1, 8, 8, 128, 128, 128),
    (2, 8, 8, 512, 512, 128),
    (1, 32, 8, 1024, 1024, 128),
    (1, 14, 2, 512, 512, 64),
    (1, 14, 1, 384, 384, 64),
    (1, 8, 8, 1000, 1000, 128),   # ragged
    (2, 4, 2, 333, 777, 64),      # ragged, Sq != Sk
    (1, 8, 2, 128, 1024, 128),    # Sq < Sk
    (1, 4, 4, 300, 200, 64),      # Sq > Sk (fully masked rows when causal)
    (1, 8, 8, 256, 256, 96),      # D not in {64,128}
    (1, 8, 8, 256, 256, 40),
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True], ids=["full", "causal"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_vs_eager_oracle(cuda_device, shape, causal, dtype):
    B, Hq, Hkv, Sq, Sk, D = shape
    q, k, v = make_qkv_fast(B, Hq, Hkv, Sq, Sk, D, dtype, 1234, cuda_device)
    out = fa(q, k, v, causal)
    assert out.shape == q.shape and out.dtype == dtype and out.stride() == q.stride()
    assert_close(out, oracle32(q, k, v, causal), dtype, f"{shape} causal={causal}")


@pytest.mark.parametrize("causal", [False, True], ids=["full", "causal"])
@pytest.mark.parametrize("shape", [(1, 32, 8, 1024, 1024, 128), (1, 14, 2, 640, 640, 64), (1, 14, 1, 200, 333, 64),
                                   (2, 8, 2, 100, 100, 128), (1, 6, 2, 300, 300, 64)],
                         ids=lambda s: "x".join(map(str, s)))
def test_gqa_packed_equals_unpacked(cuda_device, shape, causal):
    from flash_attention_cute_b200 import _lib

    B, Hq, Hkv, Sq, Sk, D = shape
    q, k, v = make_qkv_fast(B, Hq, Hkv, Sq, Sk, D, torch.bfloat16, 99, cuda_device)
    ref = oracle32(q, k, v, causal)
    a = fa(q, k, v, causal, flags=_lib.FLAG_FORCE_PACK_GQA)
    b = fa(q, k, v, causal, flags=_lib.FLAG_NO_PACK_GQA)
    assert_close(a, ref, torch.bfloat16, "packed")
    assert_close(b, ref, torch.bfloat16, "unpacked")


@pytest.mark.parametrize("Hq,Hkv", [(32, 8), (14, 2), (14, 1), (8, 8)])
def test_decode_sq1(cuda_device, Hq, Hkv):
    q, k, v = make_qkv_fast(2, Hq, Hkv, 1, 1024, 64 if Hq == 14 else 128, torch.float16, 5, cuda_device)
    out = fa(q, k, v, True)
    assert_close(out, oracle32(q, k, v, False), torch.float16, "decode")


def test_bshd_strided_views(cuda_device):
    # the HF model path: [B,S,H,D] memory viewed as [B,H,S,D] (models/rope_attn_fwd.py:81-85)
    B, S, Hq, Hkv, D = 2, 512, 32, 8, 128
    g = torch.Generator(device=cuda_device).manual_seed(3)
    q = torch.randn(B, S, Hq, D, generator=g, device=cuda_device).bfloat16().transpose(1, 2)
    k = torch.randn(B, S, Hkv, D, generator=g, device=cuda_device).bfloat16().transpose(1, 2)
    v = torch.randn(B, S, Hkv, D, generator=g, device=cuda_device).bfloat16().transpose(1, 2)
    out = fa(q, k, v, True)
    assert out.stride() == q.stride()  # O inherits q's strides (csrc/flash_attention_api.cpp:85)
    assert out.transpose(1, 2).reshape(B, S, -1).is_contiguous()
    assert_close(out, oracle32(q, k, v, True), torch.bfloat16, "bshd")


def test_softmax_scale_and_inputs_untouched(cuda_device):
    q, k, v = make_qkv_fast(1, 4, 4, 256, 256, 64, torch.float16, 8, cuda_device)
    q0, k0, v0 = q.clone(), k.clone(), v.clone()
    out = fa(q, k, v, False, scale=0.3)
    assert_close(out, oracle32(q, k, v, False, 0.3), torch.float16, "scale")
    assert torch.equal(q, q0) and torch.equal(k, k0) and torch.equal(v, v0)


def test_large_logits_lazy_rescale(cuda_device):
    # growing row maxima force the rescale path; large scale stresses the exp2 range
    q, k, v = make_qkv_fast(1, 2, 2, 512, 2048, 128, torch.bfloat16, 11, cuda_device)
    k = (k.float() * torch.linspace(0.2, 6.0, 2048, device=cuda_device).view(1, 1, -1, 1)).bfloat16()
    out = fa(q, k, v, False)
    assert_close(out, oracle32(q, k, v, False), torch.bfloat16, "rescale")


def test_headdim_padding_path(cuda_device):
    # flash_attention/flash_attention.py:26-37: D % 8 != 0 is padded then sliced
    q, k, v = make_qkv_fast(1, 2, 2, 128, 128, 44, torch.float16, 2, cuda_device)
    out = fa(q, k, v, True, scale=44 ** -0.5)
    assert out.shape == q.shape
    assert_close(out, oracle32(q, k, v, True, 44 ** -0.5), torch.float16, "pad")


@pytest.mark.skipif(not ref_kernel.available(), reason="oracle/_ref/libref_fa.so not built")
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True], ids=["full", "causal"])
@pytest.mark.parametrize("shape", [(2, 8, 8, 512, 512, 128), (1, 32, 8, 1024, 1024, 128), (1, 14, 2, 777, 777, 64),
                                   (1, 8, 2, 128, 1024, 128), (2, 14, 2, 1, 512, 64)],
                         ids=lambda s: "x".join(map(str, s)))
def test_vs_reference_kernel(cuda_device, shape, causal, dtype):
    """Identical inputs through the reference's own kernel (compiled unmodified for sm_100a)."""
    B, Hq, Hkv, Sq, Sk, D = shape
    q, k, v = make_qkv_fast(B, Hq, Hkv, Sq, Sk, D, dtype, 4321, cuda_device)
    ours = fa(q, k, v, causal)
    theirs = ref_kernel.ref_flash_attn_func(q, k, v, None, causal)
    torch.cuda.synchronize()
    tol = TOL[dtype]
    max_abs, _ = errors(ours, theirs)
    assert max_abs <= tol["max_abs"], f"vs reference kernel: max_abs {max_abs:.3e}"
    # and the reference kernel itself sits within the same band of the oracle
    assert_close(theirs, oracle32(q, k, v, causal), dtype, "reference kernel vs oracle")


def test_full_size_c2_properties(cuda_device):
    """BASELINE config 2 at full size: rows of softmax sum to one => with V == 1 the output is 1;
    plus a sampled-row check against the oracle."""
    B, H, S, D = 4, 32, 4096, 128
    q, k, v = make_qkv_fast(B, H, H, S, S, D, torch.bfloat16, 1002, cuda_device)
    ones = torch.ones_like(v)
    out1 = fa(q, k, ones, False)
    assert (out1.float() - 1).abs().max().item() <= 1e-2
    out = fa(q, k, v, False)
    ref = oracle32(q[:1, :2], k[:1, :2], v[:1, :2], False)
    assert_close(out[:1, :2], ref, torch.bfloat16, "C2 sample")


def test_full_size_c3_causal_gqa(cuda_device):
    """BASELINE config 3: Llama-3-8B shape, S=8192, causal GQA; first-row identity (row 0 attends
    only to key 0 => O[0] == V[0]) and a sampled head against the oracle, packed and unpacked."""
    from flash_attention_cute_b200 import _lib

    q, k, v = make_qkv_fast(1, 32, 8, 8192, 8192, 128, torch.bfloat16, 1003, cuda_device)
    for flags in (_lib.FLAG_NO_PACK_GQA, _lib.FLAG_FORCE_PACK_GQA):
        out = fa(q, k, v, True, flags=flags)
        assert torch.equal(out[0, :, 0], v[0].repeat_interleave(4, 0)[:, 0])
        ref = oracle32(q[:, 4:8], k[:, 1:2], v[:, 1:2], True)
        assert_close(out[:, 4:8], ref, torch.bfloat16, f"C3 flags={flags}")


# ---------------------------------------------------------------------------------------------------
# Many short work tiles per persistent CTA.  A one-block M-tile can finish its
# softmax before the epilogue has read the previous tile's row sums; without back-pressure the
# RowSum barrier ran two phases ahead and the CTA dead-locked (then limped after the watchdog).
# ---------------------------------------------------------------------------------------------------
def _sampled_units_close(out, q, k, v, causal, dtype, units, what):
    G = q.size(1) // k.size(1)
    for b, hkv in units:
        ref = oracle32(q[b:b + 1, hkv * G:(hkv + 1) * G], k[b:b + 1, hkv:hkv + 1], v[b:b + 1, hkv:hkv + 1], causal)
        assert_close(out[b:b + 1, hkv * G:(hkv + 1) * G], ref, dtype, f"{what} unit (b={b}, kv head={hkv})")


def _timed_ms(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


SHORT_TILE_CASES = [
    # B, Hq, Hkv, S, D, dtype, causal, time bound (ms)
    (128, 32, 8, 512, 128, torch.bfloat16, True, 1.0),
    (64, 14, 2, 512, 64, torch.float16, True, 0.5),
    (64, 14, 2, 512, 64, torch.float16, False, 0.5),
    (128, 32, 8, 256, 128, torch.bfloat16, True, 0.6),
    (64, 32, 8, 1024, 128, torch.bfloat16, True, 1.0),
    (256, 8, 8, 128, 128, torch.bfloat16, False, 0.5),   # one-block tiles only, non-causal
    (512, 16, 2, 64, 64, torch.float16, True, 0.5),      # seqlen < one M-tile, GQA-packed
]


@pytest.mark.parametrize("case", SHORT_TILE_CASES,
                         ids=lambda c: f"B{c[0]}-H{c[1]}/{c[2]}-S{c[3]}-D{c[4]}-{'causal' if c[6] else 'full'}")
def test_many_short_tiles_per_cta(cuda_device, case):
    from flash_attention_cute_b200 import _lib

    B, Hq, Hkv, S, D, dtype, causal, bound_ms = case
    q, k, v = make_qkv_fast(B, Hq, Hkv, S, S, D, dtype, 777, cuda_device)
    for _ in range(3):  # the dead-lock was a race: give it several launches
        out = fa(q, k, v, causal)
    units = [(0, 0), (B // 2, Hkv - 1), (B - 1, Hkv // 2)]
    _sampled_units_close(out, q, k, v, causal, dtype, units, f"short tiles {case[:5]}")
    ms = _timed_ms(lambda: fa(q, k, v, causal), 10)
    _lib.check_device(0)
    assert ms <= bound_ms, f"{case[:5]}: {ms:.3f} ms per call (bound {bound_ms} ms)"
    # zeros in, zeros out (the r01 slow leg used zero inputs)
    z = fa(torch.zeros_like(q), torch.zeros_like(k), torch.zeros_like(v), causal)
    assert not z.any()


# ---------------------------------------------------------------------------------------------------
# BASELINE config 4 at sweep sizes: fp16, D = 64, Qwen2-0.5B heads (14 q / 2 kv) and MQA (14 / 1),
# B = 32768 / S, causal and non-causal; parity on sampled (batch, kv-head) units (oracle in chunks).
# The reference runs D = 64 inside its D = 128 kernel (csrc/kernel_dispatcher.h:45-52).
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("causal", [False, True], ids=["full", "causal"])
@pytest.mark.parametrize("Hkv", [2, 1])
@pytest.mark.parametrize("S", [4096, 32768])
def test_c4_sweep_sizes(cuda_device, S, Hkv, causal):
    B = 32768 // S
    q, k, v = make_qkv_fast(B, 14, Hkv, S, S, 64, torch.float16, 4004, cuda_device)
    out = fa(q, k, v, causal)
    if S == 32768:   # sample query rows: the oracle over 32K x 32K x 14 heads is too slow for a test
        rows = torch.cat([torch.arange(0, 256), torch.arange(16000, 16128), torch.arange(S - 256, S)]).to(cuda_device)
        hs = slice(0, 14 // Hkv)
        qs = q[:, hs].index_select(2, rows)
        kk, vv = k[:, :1], v[:, :1]
        sc = 64 ** -0.5
        s = torch.einsum("bhqd,bkd->bhqk", qs.float(), kk[:, 0].float()) * sc
        if causal:
            s = s.masked_fill(torch.arange(S, device=cuda_device)[None, None, None, :] > rows[None, None, :, None], float("-inf"))
        ref = torch.einsum("bhqk,bkd->bhqd", torch.softmax(s, -1), vv[:, 0].float())
        assert_close(out[:, hs].index_select(2, rows), ref, torch.float16, f"C4 S=32K Hkv={Hkv}")
    else:
        _sampled_units_close(out, q, k, v, causal, torch.float16, [(0, 0), (B - 1, Hkv - 1)], f"C4 S={S} Hkv={Hkv}")


def test_prefill_is_graph_capturable(cuda_device):
    """The forward launch allocates nothing and synchronises nothing: it can be captured into a CUDA
    graph and replayed (the captured launch bakes in one slot of the scheduler-counter ring)."""
    from flash_attention import flash_attn_func

    q, k, v = make_qkv_fast(2, 8, 2, 700, 700, 128, torch.bfloat16, 31, cuda_device)
    eager = flash_attn_func(q, k, v, causal=True)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            out = flash_attn_func(q, k, v, causal=True)
    for _ in range(3):
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager)
    from flash_attention_cute_b200 import _lib
    _lib.check_device(0)


def test_watchdog_abort_is_contained(cuda_device):
    """a wait that gives up must not turn later waits into no-ops.  With the
    watchdog at 1 ms a work tile of 524288 keys (≈ 6 ms) makes the epilogue's wait for its row sums time out
    legitimately: the launch must end by itself, report through fa_sm100_check_device, make every
    further entry-point call fail loudly until it is cleared, and leave the device healthy."""
    import ctypes

    from flash_attention import flash_attn_func
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import fill_params

    lib = _lib.load()
    qs, ks, vs = make_qkv_fast(1, 4, 4, 256, 256, 128, torch.bfloat16, 77, cuda_device)
    want = flash_attn_func(qs, ks, vs, causal=True)
    _lib.check_device(0)
    q, k, v = make_qkv_fast(1, 2, 2, 256, 524288, 128, torch.bfloat16, 78, cuda_device)
    lib.fa_sm100_set_watchdog_ms(1)
    try:
        flash_attn_func(q, k, v, causal=False)          # asynchronous: returns, then times out in the kernel
        flash_attn_func(qs, ks, vs, causal=True)        # queued behind it: must do nothing
        torch.cuda.synchronize()
        # the entry points refuse to launch while the condition is pending (host-mapped status word)
        o = torch.empty_like(qs)
        p = fill_params(qs, ks, vs, o, 128 ** -0.5, True)
        rc = lib.fa_sm100_fwd(ctypes.byref(p), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 4 and "fa_sm100_check_device" in _lib.last_error()
        with pytest.raises(RuntimeError, match="watchdog fired"):
            _lib.check_device(0)
    finally:
        lib.fa_sm100_set_watchdog_ms(2000)
    _lib.check_device(0)                                 # cleared: healthy again
    again = flash_attn_func(qs, ks, vs, causal=True)
    _lib.check_device(0)
    assert torch.equal(again, want)
    full = flash_attn_func(q, k, v, causal=False)       # the long tile itself, with the default watchdog
    _lib.check_device(0)
    assert torch.isfinite(full.float()).all()
