"""fa_sm100_fwd_decode_static: the split-KV decode entry point that allocates nothing
(include/fa_sm100.h; no counterpart in the reference).

Its scratch — split partials and arrival counters — is a file-scope device array of
FA_SM100_STATIC_SCRATCH_BYTES inside the library, shared by every call, holding whatever the last
call left.  The tests call it through the C ABI, FIRST (before any other entry point has set the
device up, so the file also passes alone in a fresh process), on several calls in a row with
different splits, and compare with fa_sm100_fwd_decode on the same inputs: same kernel, same plan,
so the results must be bit-equal, which is tighter than the tolerance the parity tests use; one case
is also held against the fp32 eager oracle with that tolerance.

This file sorts last on purpose (a failure here hides nothing)."""
import ctypes

import pytest
import torch

from common import assert_close, make_qkv_fast
from oracle.eager_attention import eager_attention


def _static_call(q, k, v, lib, _lib, fill_params):
    o = torch.empty_like(q)
    params = fill_params(q, k, v, o, q.size(-1) ** -0.5, False, 0)
    stream = torch.cuda.current_stream(q.device).cuda_stream
    before = _lib.launch_count()
    rc = lib.fa_sm100_fwd_decode_static(ctypes.byref(params), ctypes.c_void_p(stream))
    _lib.raise_for_status(rc)
    assert _lib.launch_count() == before + 1
    return o


def _workspace_call(q, k, v, lib, _lib, fill_params):
    """The existing entry point, straight through the C ABI with a caller workspace."""
    o = torch.empty_like(q)
    params = fill_params(q, k, v, o, q.size(-1) ** -0.5, False, 0)
    nbytes = ctypes.c_int64()
    _lib.raise_for_status(lib.fa_sm100_decode_plan(ctypes.byref(params), None, None, ctypes.byref(nbytes)))
    ws = torch.empty(max(nbytes.value, 256), dtype=torch.uint8, device=q.device)
    stream = torch.cuda.current_stream(q.device).cuda_stream
    rc = lib.fa_sm100_fwd_decode(ctypes.byref(params), ctypes.c_void_p(ws.data_ptr()),
                                 ctypes.c_int64(ws.numel()), ctypes.c_void_p(stream))
    _lib.raise_for_status(rc)
    torch.cuda.synchronize(q.device)
    return o, nbytes.value


def test_static_symbols_and_limit():
    from flash_attention_cute_b200 import _lib

    lib = _lib.load()
    assert lib.fa_sm100_static_scratch_bytes() == 4 * 1024 * 1024
    # the largest one-wave split on a 148-SM part (296 CTAs of partials, D = 128) fits
    assert 296 * 16 * (128 + 2) * 4 <= lib.fa_sm100_static_scratch_bytes() - 2048
    assert lib.fa_sm100_fwd_decode_static(None, None) == 1
    assert "NULL" in _lib.last_error()


@pytest.mark.gpu
def test_static_decode_matches_workspace_decode_on_calls_in_a_row(cuda_device):
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import fill_params

    lib = _lib.load()
    # (B, Hq, Hkv, Sk, D, dtype): many splits, then few, then none, then many again — every call
    # finds the previous call's partials and counters in the array
    cases = [
        (1, 32, 8, 6000, 128, torch.bfloat16),
        (16, 64, 8, 1024, 128, torch.bfloat16),   # the reference's decode example
        (16, 64, 8, 1024, 128, torch.bfloat16),   # the same shape again, other data
        (3, 14, 2, 777, 64, torch.float16),
        (64, 32, 8, 512, 128, torch.float16),     # 512 units: one CTA each, no scratch
        (1, 8, 8, 65536, 128, torch.bfloat16),    # 296 CTAs: the largest request on this part
        (2, 4, 4, 1, 128, torch.float16),
        (1, 32, 8, 6000, 128, torch.bfloat16),
    ]
    inputs = [make_qkv_fast(B, Hq, Hkv, 1, Sk, D, dt, 900 + i, cuda_device)
              for i, (B, Hq, Hkv, Sk, D, dt) in enumerate(cases)]
    # 1. the static entry point first, back to back on one stream, nothing else in between
    outs = [_static_call(q, k, v, lib, _lib, fill_params) for q, k, v in inputs]
    torch.cuda.synchronize(cuda_device)
    # 2. the existing entry point on the same inputs
    split_seen = False
    for i, ((q, k, v), o_static) in enumerate(zip(inputs, outs)):
        o_ws, nbytes = _workspace_call(q, k, v, lib, _lib, fill_params)
        split_seen |= nbytes > 0
        assert torch.equal(o_static, o_ws), f"case {i} {cases[i]}: static and workspace decode differ"
    assert split_seen
    q, k, v = inputs[1]
    assert_close(outs[1], eager_attention(q, k, v, None, False, out_dtype=torch.float32), torch.bfloat16,
                 "static decode vs oracle")
    _lib.check_device(cuda_device.index or 0)


@pytest.mark.gpu
def test_static_decode_allocates_nothing(cuda_device):
    """After the first call (module load brings the array in), further calls leave the device's
    free memory untouched, and the call records under a capture mode in which cudaMalloc is an
    error; replaying the graph twice gives the eager result."""
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import fill_params

    lib = _lib.load()
    q, k, v = make_qkv_fast(4, 32, 8, 1, 5000, 128, torch.bfloat16, 77, cuda_device)
    eager = _static_call(q, k, v, lib, _lib, fill_params)
    torch.cuda.synchronize(cuda_device)
    free0, _ = torch.cuda.mem_get_info(cuda_device)
    o = torch.empty_like(q)
    params = fill_params(q, k, v, o, q.size(-1) ** -0.5, False, 0)
    for _ in range(3):
        stream = torch.cuda.current_stream(cuda_device).cuda_stream
        _lib.raise_for_status(lib.fa_sm100_fwd_decode_static(ctypes.byref(params), ctypes.c_void_p(stream)))
    torch.cuda.synchronize(cuda_device)
    free1, _ = torch.cuda.mem_get_info(cuda_device)
    assert free1 == free0, f"free device memory changed across calls: {free0} -> {free1}"
    assert torch.equal(o, eager)

    o.zero_()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, capture_error_mode="thread_local"):
        stream = torch.cuda.current_stream(cuda_device).cuda_stream
        _lib.raise_for_status(lib.fa_sm100_fwd_decode_static(ctypes.byref(params), ctypes.c_void_p(stream)))
    for _ in range(2):
        o.zero_()
        graph.replay()
        torch.cuda.synchronize(cuda_device)
        assert torch.equal(o, eager)
