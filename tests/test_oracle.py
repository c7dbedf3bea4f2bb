"""CPU: the oracle restatement against the golden vectors produced from the REFERENCE's own
eager_attention (scripts/benchmark_kernel.py:15-44) by oracle/make_golden.py."""
from pathlib import Path

import numpy as np
import pytest
import torch

from common import make_qkv
from oracle.eager_attention import attention_flops, eager_attention, eager_attention_chunked

GOLDEN = sorted((Path(__file__).parent / "golden").glob("eager_*.npz"))


def test_golden_present():
    assert len(GOLDEN) >= 6


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
def test_oracle_matches_reference_golden(path):
    g = np.load(path)
    B, Hq, Hkv, Sq, Sk, D = (int(x) for x in g["shape"])
    dtype = getattr(torch, str(g["dtype"]))
    q, k, v = make_qkv(B, Hq, Hkv, Sq, Sk, D, dtype, int(g["seed"]))
    out = eager_attention(q, k, v, None, bool(g["causal"]))
    assert out.dtype == dtype
    rows = out.float()[:, :, :: int(g["row_step"]), :].numpy()
    # generated single-threaded; allow reduction-order noise from a different core count
    np.testing.assert_allclose(rows, g["rows"], rtol=0, atol=2e-6 if dtype == torch.float32 else 0)
    assert abs(float(out.float().double().sum()) - float(g["checksum"])) < 1e-2


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("shape", [(1, 8, 8, 64, 64, 32), (2, 6, 2, 40, 72, 16), (1, 4, 1, 72, 40, 16)])
def test_chunked_equals_plain(shape, causal):
    B, Hq, Hkv, Sq, Sk, D = shape
    q, k, v = make_qkv(B, Hq, Hkv, Sq, Sk, D, torch.float32, 7)
    a = eager_attention(q, k, v, None, causal)
    b = eager_attention_chunked(q, k, v, None, causal, max_score_elems=Hq // Hkv * Sk * 16)
    torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)


def test_causal_is_bottom_right_aligned():
    # csrc/mask.cuh:37-39: column n visible to row m iff n <= m + (Sk - Sq)
    q, k, v = make_qkv(1, 1, 1, 3, 5, 8, torch.float32, 3)
    out = eager_attention(q, k, v, None, True)
    w = torch.softmax((q[0, 0] @ k[0, 0].T) / 8 ** 0.5 + torch.tensor(
        [[0, 0, 0, -np.inf, -np.inf], [0, 0, 0, 0, -np.inf], [0, 0, 0, 0, 0]], dtype=torch.float32), -1)
    torch.testing.assert_close(out[0, 0], w @ v[0, 0], rtol=1e-5, atol=1e-6)


def test_sq1_causal_equals_noncausal():
    # the reference drops `causal` for decode (csrc/flash_attention_api.cpp:79)
    q, k, v = make_qkv(2, 8, 2, 1, 33, 16, torch.float32, 5)
    torch.testing.assert_close(eager_attention(q, k, v, None, True), eager_attention(q, k, v, None, False))


def test_fully_masked_rows_are_zero():
    q, k, v = make_qkv(1, 2, 2, 6, 4, 8, torch.float32, 9)
    out = eager_attention(q, k, v, None, True)
    assert torch.all(out[:, :, :2] == 0) and torch.isfinite(out).all()


def test_flops_formula():
    assert attention_flops(4, 32, 4096, 4096, 128, False) == 2 ** 40
    assert attention_flops(1, 32, 8192, 8192, 128, True) == 2 ** 39
