"""TEST INFRASTRUCTURE ONLY — generates tests/golden/eager_*.npz from the REFERENCE itself.

Runs only in the build container (needs /root/reference).  The reference's ``eager_attention``
(scripts/benchmark_kernel.py:15-44) is lifted out of its script by AST — importing the script
would import flash_attn and JIT-build the reference CUDA extension (scripts/benchmark_kernel.py
:10-11) — and evaluated on CPU on seeded inputs.  For every case the restatement in
oracle/eager_attention.py must agree BIT-FOR-BIT before the vector is written.

Inputs are not stored: they come from numpy's frozen legacy generator
``RandomState(seed).standard_normal`` (stream guaranteed stable), cast to the case dtype.
Outputs are stored as fp32, row-subsampled for the larger cases, plus a float64 checksum.
"""
from __future__ import annotations

import ast
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle.eager_attention import eager_attention as restated  # noqa: E402

REF_SCRIPT = Path("/root/reference/scripts/benchmark_kernel.py")
GOLDEN = ROOT / "tests" / "golden"

# name, B, Hq, Hkv, Sq, Sk, D, dtype, causal, row_step, seed
CASES = [
    ("c1_fp32_noncausal", 1, 8, 8, 512, 512, 64, "float32", False, 16, 1001),
    ("c1_fp32_causal", 1, 8, 8, 512, 512, 64, "float32", True, 16, 1001),
    ("gqa_bf16_causal", 2, 14, 2, 128, 128, 64, "bfloat16", True, 4, 1004),
    ("mqa_fp16_noncausal", 1, 14, 1, 96, 96, 64, "float16", False, 4, 1005),
    ("gqa_bf16_d128_causal", 1, 32, 8, 256, 256, 128, "bfloat16", True, 8, 1003),
    ("ragged_fp16_causal", 1, 4, 2, 200, 200, 40, "float16", True, 5, 1006),
]


def make_inputs(B, Hq, Hkv, Sq, Sk, D, dtype, seed):
    rs = np.random.RandomState(seed)
    td = getattr(torch, dtype)
    q = torch.from_numpy(rs.standard_normal((B, Hq, Sq, D)).astype(np.float32)).to(td)
    k = torch.from_numpy(rs.standard_normal((B, Hkv, Sk, D)).astype(np.float32)).to(td)
    v = torch.from_numpy(rs.standard_normal((B, Hkv, Sk, D)).astype(np.float32)).to(td)
    return q, k, v


def load_reference_eager():
    tree = ast.parse(REF_SCRIPT.read_text())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "eager_attention")
    mod = ast.Module(body=[fn], type_ignores=[])
    ns = {"torch": torch}
    exec(compile(mod, str(REF_SCRIPT), "exec"), ns)
    return ns["eager_attention"]


def main():
    ref = load_reference_eager()
    GOLDEN.mkdir(parents=True, exist_ok=True)
    torch.set_num_threads(1)  # reduction order independent of the core count
    for name, B, Hq, Hkv, Sq, Sk, D, dtype, causal, step, seed in CASES:
        q, k, v = make_inputs(B, Hq, Hkv, Sq, Sk, D, dtype, seed)
        o_ref = ref(q, k, v, None, causal)
        o_new = restated(q, k, v, None, causal)
        assert o_ref.dtype == o_new.dtype
        assert torch.equal(o_ref, o_new), f"{name}: restatement differs from the reference function"
        o32 = o_ref.float().numpy()
        np.savez_compressed(
            GOLDEN / f"eager_{name}.npz",
            shape=np.array([B, Hq, Hkv, Sq, Sk, D]), dtype=dtype, causal=causal, seed=seed,
            row_step=step, rows=o32[:, :, ::step, :], checksum=np.float64(o32.astype(np.float64).sum()),
        )
        print(f"{name}: ok, restatement == reference bit-for-bit; checksum {o32.astype(np.float64).sum():.6f}")


if __name__ == "__main__":
    main()
