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
    assert p["workspace_bytes"] >= 128 * 2 * 16 * (128 + 2) * 4


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
