"""Interleaved A/B timing of tuning variants (FA_SM100_TUNE=1 builds): every configuration is timed
`--rounds` times in round-robin order so that clock / thermal drift hits all of them alike; prints
median and best TFLOP/s per configuration.  Not part of the product.

    FA_SM100_TUNE=1 python tools/ab.py --shapes c2,c3 --configs 3:0:1,3:0:0,2:0:1 [--iters 20] [--rounds 5]
config = emu:spec:fence (exp2 pairs of 8 on the FMA pipe : speculative-max softmax : MMA fence);
"pack" / "nopack" time the default variant with GQA packing forced on / off."""
import argparse
import json
import statistics
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
from prof_one import SHAPES  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="c2,c3")
    ap.add_argument("--configs", default="1:128,2:128,3:128")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--rounds", type=int, default=5)
    args = ap.parse_args()
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    ext = load_extension()
    dev = torch.device("cuda:0")
    cfgs = [c if not c[0].isdigit() else tuple(int(x) for x in c.split(":"))
            for c in args.configs.split(",")]
    for name in args.shapes.split(","):
        B, Hq, Hkv, Sq, Sk, D, causal, dt = SHAPES[name]
        g = torch.Generator(device=dev).manual_seed(7)
        q = torch.randn(B, Hq, Sq, D, generator=g, device=dev).to(dt)
        k = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
        v = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
        fl = 4.0 * B * Hq * Sq * Sk * D / (2 if causal else 1)
        named = {"pack": _lib.FLAG_FORCE_PACK_GQA, "nopack": _lib.FLAG_NO_PACK_GQA, "default": 0}
        flags = {c: named[c] if isinstance(c, str) else _lib.tune_flags(c[0], bool(c[1]), bool(c[2])) for c in cfgs}
        res = {c: [] for c in flags}
        for c, f in flags.items():
            for _ in range(3):
                ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, f)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(args.rounds):
            for c, f in flags.items():
                e0.record()
                for _ in range(args.iters):
                    ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, f)
                e1.record()
                torch.cuda.synchronize()
                res[c].append(fl / (e0.elapsed_time(e1) / args.iters) / 1e9)
        _lib.check_device(0)
        for c in flags:
            print(json.dumps(dict(shape=name, config=str(c), median=round(statistics.median(res[c]), 1),
                                  best=round(max(res[c]), 1), worst=round(min(res[c]), 1))), flush=True)


if __name__ == "__main__":
    main()
