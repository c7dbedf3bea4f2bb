"""In-tree nvcc build of libfa_sm100.so (sm_100a only).

Replaces the reference's JIT loader flash_attention/load_cpp_extention.py:11-54: same idea
(build on first import, cache afterwards) but the product is a plain C-ABI shared library
(include/fa_sm100.h) compiled straight by nvcc — no torch headers, no CUTLASS include paths,
one architecture: ``-gencode arch=compute_100a,code=sm_100a``.  ``-lineinfo`` is kept so Nsight
Compute can map SASS to source, as the reference does (load_cpp_extention.py:37).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
ROOT = PKG.parent
LIB = PKG / "libfa_sm100.so"
STAMP = PKG / ".libfa_sm100.stamp"

SOURCES = [CSRC / "fa_sm100_host.cu", CSRC / "fa_sm100_selftest.cu"]
HEADERS = [CSRC / "fa_sm100_kernel.cuh", CSRC / "fa_sm100_fwd.cuh", CSRC / "fa_sm100_fwd4.cuh", CSRC / "fa_sm100_fwd5.cuh", CSRC / "fa_sm100_fwd6.cuh", CSRC / "fa_sm100_decode.cuh", CSRC / "fa_sm100_rope.cuh", CSRC / "fa_sm100_ptx.cuh",
           ROOT / "include" / "fa_sm100.h"]

NVCC_FLAGS = [
    "-O3",
    "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "--expt-relaxed-constexpr",
    "-DFA_SM100_TUNE",
    "-Xptxas", "-v",
    "-shared",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; cannot build libfa_sm100.so (there is no fallback path)")
    return cand


def source_hash() -> str:
    h = hashlib.sha256()
    for f in SOURCES + HEADERS:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_current() -> bool:
    return LIB.exists() and STAMP.exists() and STAMP.read_text().strip() == source_hash()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile the library if sources changed.  Returns the path of the .so."""
    if not force and is_current():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-I", str(ROOT / "include"), "-o", str(LIB), *map(str, SOURCES)]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    (PKG / "build.log").write_text(" ".join(cmd) + "\n" + log)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed building libfa_sm100.so:\n{log[-4000:]}")
    if verbose:
        print(log)
    STAMP.write_text(source_hash())
    return LIB


if __name__ == "__main__":
    import sys

    build(force="--force" in sys.argv, verbose=True)
    print(LIB)
