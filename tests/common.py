"""Shared helpers for the parity tests (seeded inputs, error metrics, tolerances)."""
import numpy as np
import torch

# BASELINE.json north_star: max-abs <= 2e-3 (fp16) / 1e-2 (bf16); SURVEY.md §8(c): relative-L2
# <= 1e-3 (fp16) / 5e-3 (bf16) against the fp32 eager oracle.
TOL = {
    torch.float16: dict(max_abs=2e-3, rel_l2=1e-3),
    torch.bfloat16: dict(max_abs=1e-2, rel_l2=5e-3),
}


def make_qkv(B, Hq, Hkv, Sq, Sk, D, dtype, seed, device="cpu"):
    """randn in fp32 then cast, like scripts/benchmark_kernel.py:75-77 but seeded (numpy's frozen
    RandomState so CPU fixtures and GPU tests see identical values)."""
    rs = np.random.RandomState(seed)
    q = torch.from_numpy(rs.standard_normal((B, Hq, Sq, D)).astype(np.float32)).to(dtype)
    k = torch.from_numpy(rs.standard_normal((B, Hkv, Sk, D)).astype(np.float32)).to(dtype)
    v = torch.from_numpy(rs.standard_normal((B, Hkv, Sk, D)).astype(np.float32)).to(dtype)
    return q.to(device), k.to(device), v.to(device)


def make_qkv_fast(B, Hq, Hkv, Sq, Sk, D, dtype, seed, device):
    g = torch.Generator(device=device).manual_seed(seed)
    q = torch.randn(B, Hq, Sq, D, generator=g, device=device, dtype=torch.float32).to(dtype)
    k = torch.randn(B, Hkv, Sk, D, generator=g, device=device, dtype=torch.float32).to(dtype)
    v = torch.randn(B, Hkv, Sk, D, generator=g, device=device, dtype=torch.float32).to(dtype)
    return q, k, v


def errors(out, ref):
    """max-abs error (measured in units of max(1, |ref|): a 16-bit output cannot be closer than
    half an ulp, 2^-9·|o| for bf16, so for |o| > 1 the absolute bound scales with the magnitude —
    this only matters for the first causal rows, whose output is a copy of a V row) and the
    relative L2 error."""
    out = out.float()
    ref = ref.float()
    diff = out - ref
    max_abs = (diff.abs() / ref.abs().clamp_min(1.0)).max().item()
    rel_l2 = (diff.norm() / ref.norm().clamp_min(1e-30)).item()
    return max_abs, rel_l2


def assert_close(out, ref32, dtype, what=""):
    assert torch.isfinite(out.float()).all(), f"{what}: non-finite values in output"
    max_abs, rel_l2 = errors(out, ref32)
    tol = TOL[dtype]
    assert max_abs <= tol["max_abs"] and rel_l2 <= tol["rel_l2"], (
        f"{what}: max_abs={max_abs:.3e} (tol {tol['max_abs']:.0e}), rel_l2={rel_l2:.3e} (tol {tol['rel_l2']:.0e})")
    return max_abs, rel_l2
