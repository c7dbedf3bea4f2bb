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
TUNE = os.environ.get("FA_SM100_TUNE", "") not in ("", "0")
# Development builds live next to the product library, never in its place: FA_SM100_TUNE=1 adds the
# tuning variants; FA_SM100_BUILD_TAG=<tag> with FA_SM100_NVCC_EXTRA="-D..." builds the product
# sources with other compile-time defaults (e.g. to run the whole parity suite on a candidate).
TAG = ("_tune" if TUNE else "") + ("_" + os.environ["FA_SM100_BUILD_TAG"] if os.environ.get("FA_SM100_BUILD_TAG") else "")
LIB = PKG / f"libfa_sm100{TAG}.so"
STAMP = PKG / f".libfa_sm100{TAG}.stamp"

SOURCES = [CSRC / "fa_sm100_host.cu", CSRC / "fa_sm100_selftest.cu"]
HEADERS = [CSRC / "fa_sm100_fwd.cuh", CSRC / "fa_sm100_sched.cuh", CSRC / "fa_sm100_decode.cuh",
           CSRC / "fa_sm100_rope.cuh", CSRC / "fa_sm100_ptx.cuh", ROOT / "include" / "fa_sm100.h"]

NVCC_FLAGS = [
    "-O3",
    "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-shared",
    "-Xcompiler", "-fPIC",
]
# FA_SM100_TUNE=1 in the environment adds the tuning variants (A/B of exp2 emulation, speculative
# max, MMA fence; debug timeline) — a development build, never the product (cold build ~25 s
# without, ~70 s with).
if TUNE:
    NVCC_FLAGS.insert(5, "-DFA_SM100_TUNE")
if os.environ.get("FA_SM100_BUILD_TAG") and os.environ.get("FA_SM100_NVCC_EXTRA"):
    NVCC_FLAGS[5:5] = os.environ["FA_SM100_NVCC_EXTRA"].split()


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
    """Compile the library if sources changed.  Returns the path of the .so.

    Safe under ``torchrun``: ranks serialise on a file lock, the first one compiles into a
    temporary file and renames it into place atomically, the others find it current."""
    import fcntl

    if not force and is_current():
        return LIB
    with open(PKG / ".libfa_sm100.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_current():  # another process built it while we waited
                return LIB
            tmp = PKG / f".{LIB.stem}.{os.getpid()}.tmp.so"
            cmd = [_nvcc(), *NVCC_FLAGS, "-I", str(ROOT / "include"), "-o", str(tmp), *map(str, SOURCES)]
            proc = subprocess.run(cmd, capture_output=True, text=True)
            log = proc.stdout + proc.stderr
            (PKG / f"build{TAG}.log").write_text(" ".join(cmd) + "\n" + log)
            if proc.returncode != 0:
                tmp.unlink(missing_ok=True)
                raise RuntimeError(f"nvcc failed building libfa_sm100.so:\n{log[-4000:]}")
            if verbose:
                print(log)
            os.replace(tmp, LIB)
            STAMP.write_text(source_hash())
            return LIB
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


if __name__ == "__main__":
    import sys

    build(force="--force" in sys.argv, verbose=True)
    print(LIB)
