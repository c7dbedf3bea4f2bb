"""A/B sweep of the kernel's tuning bits on a B200 (FA_SM100_TUNE builds).  Not part of the product.

    python tools/tune.py [--shapes c3b2,c2] [--emu 0,1,2,3] [--split 64,96,128] [--iters 10]

For every shape: the first-generation kernel (v1), then each (emu, split) variant of the current
kernel: ms, TFLOP/s, and max |o - o_v1| as a quick sanity check (parity proper is in tests/).
Writes gpurun_out/tune.json.  --halves 3 selects the generation-4 kernel."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

from prof_one import SHAPES  # noqa: E402


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="c3b2,c2,c3nc,c4s4k")
    ap.add_argument("--emu", default="0,1,2")
    ap.add_argument("--split", default="128")
    ap.add_argument("--halves", default="1,2")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--no-v1", action="store_true")
    args = ap.parse_args()
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    ext = load_extension()
    dev = torch.device("cuda:0")
    out = []
    for name in args.shapes.split(","):
        B, Hq, Hkv, Sq, Sk, D, causal, dt = SHAPES[name]
        g = torch.Generator(device=dev).manual_seed(7)
        q = torch.randn(B, Hq, Sq, D, generator=g, device=dev).to(dt)
        k = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
        v = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
        fl = 4.0 * B * Hq * Sq * Sk * D / (2 if causal else 1)
        scale = D ** -0.5
        o1 = None
        if not args.no_v1:
            o1 = ext.flash_attention_fwd(q, k, v, scale, causal, _lib.FLAG_KERNEL_V1)
            _lib.check_device(0)
            ms = timeit(lambda: ext.flash_attention_fwd(q, k, v, scale, causal, _lib.FLAG_KERNEL_V1), args.iters)
            rec = dict(shape=name, kernel="v1", ms=ms, tflops=fl / ms / 1e9)
            print(json.dumps(rec), flush=True)
            out.append(rec)
        combos = [(h, sp, e) for h in [int(x) for x in args.halves.split(",")]
                  for sp in ([int(x) for x in args.split.split(",")] if h in (0, 1, 3) else [128])
                  for e in [int(x) for x in args.emu.split(",")]]
        for halves, split, emu in combos:
            if True:
                flags = _lib.tune_flags(emu, split, halves if halves else None)  # halves 0 = default generation (5)
                try:
                    o = ext.flash_attention_fwd(q, k, v, scale, causal, flags)
                    _lib.check_device(0)
                    err = (o.float() - o1.float()).abs().max().item() if o1 is not None else None
                    nan = int(torch.isnan(o.float()).sum())
                    ms = timeit(lambda: ext.flash_attention_fwd(q, k, v, scale, causal, flags), args.iters)
                    _lib.check_device(0)
                    rec = dict(shape=name, kernel={0: "default(v5)", 2: "v5", 3: "v4"}.get(halves, "v2"), halves=halves, emu=emu, split=split, ms=ms, tflops=fl / ms / 1e9,
                               max_abs_vs_v1=err, nan=nan)
                except Exception as e:  # keep sweeping unless the device is gone
                    rec = dict(shape=name, kernel={0: "default(v5)", 2: "v5", 3: "v4"}.get(halves, "v2"), halves=halves, emu=emu, split=split, error=repr(e)[:300])
                    print(json.dumps(rec), flush=True)
                    out.append(rec)
                    if "ABORT" in repr(e) or "CUDA" in repr(e):
                        (ROOT / "gpurun_out").mkdir(exist_ok=True)
                        (ROOT / "gpurun_out" / "tune.json").write_text(json.dumps(out, indent=1))
                        raise
                    continue
                print(json.dumps(rec), flush=True)
                out.append(rec)
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / "tune.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
