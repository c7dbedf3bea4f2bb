"""A few decode launches of one shape (for ncu): python tools/prof_decode.py b16_s8k --iters 3"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
from bench_decode import SHAPES  # noqa: E402
from flash_attention_cute_b200.load_cpp_extention import load_extension  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("shape")
ap.add_argument("--iters", type=int, default=3)
args = ap.parse_args()
B, Hq, Hkv, Sk, D = SHAPES[args.shape]
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
q = torch.randn(B, Hq, 1, D, generator=g, device=dev).bfloat16()
k = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).bfloat16()
v = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).bfloat16()
ext = load_extension()
for _ in range(args.iters):
    o = ext.flash_attention_fwd(q, k, v, D ** -0.5, False)
torch.cuda.synchronize()
print("ok", float(o.float().abs().mean()))
