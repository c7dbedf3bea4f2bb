"""Short driver for ncu captures: a few launches of the forward op on one named shape.

    python tools/prof_one.py c3b1 [--iters 3] [--flags 0]

Prints the CUDA-event time of the last launches (not a bench value when run under ncu).
Not part of the product."""
import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

SHAPES = {
    # name: B, Hq, Hkv, Sq, Sk, D, causal, dtype
    "c2": (4, 32, 32, 4096, 4096, 128, False, torch.bfloat16),
    "c3b1": (1, 32, 8, 8192, 8192, 128, True, torch.bfloat16),
    "c3b2": (2, 32, 8, 8192, 8192, 128, True, torch.bfloat16),
    "c3": (8, 32, 8, 8192, 8192, 128, True, torch.bfloat16),
    "c3nc": (2, 32, 8, 8192, 8192, 128, False, torch.bfloat16),
    "c4s4k": (8, 14, 2, 4096, 4096, 64, True, torch.float16),
    "c4s32k": (1, 14, 2, 32768, 32768, 64, True, torch.float16),
    "c4s4knc": (8, 14, 2, 4096, 4096, 64, False, torch.float16),
    "s512": (128, 32, 8, 512, 512, 128, True, torch.bfloat16),      # many short work tiles per CTA
    "s1k": (64, 32, 8, 1024, 1024, 128, True, torch.bfloat16),
    "c4s512": (64, 14, 2, 512, 512, 64, True, torch.float16),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("shape", choices=sorted(SHAPES))
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--flags", type=int, default=0)
    args = ap.parse_args()
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    B, Hq, Hkv, Sq, Sk, D, causal, dt = SHAPES[args.shape]
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(7)
    q = torch.randn(B, Hq, Sq, D, generator=g, device=dev).to(dt)
    k = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
    v = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
    ext = load_extension()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = 0.0
    for _ in range(args.iters):
        e0.record()
        ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, args.flags)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    _lib.check_device(0)
    fl = 4.0 * B * Hq * Sq * Sk * D / (2 if causal else 1)
    print(f"{args.shape}: last launch {ms:.4f} ms = {fl / ms / 1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    main()
