"""Packed variable-length forward (fa_sm100_fwd_varlen) against (i) the same tokens run as one
padded fixed-length batch and (ii) flash-attn 2.8's varlen call (library comparator).

FLOPs counted = sum over sequences of 4 * Hq * Sq_i * Sk_i * D (halved when causal) — padding rows
of the fixed-length run are not counted.  CUDA-event timing on the current stream, warm-up first;
total q+k+v+o of every mix exceeds the 126 MB L2."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from flash_attention import flash_attn_func, flash_attn_varlen_func  # noqa: E402

MIXES = {
    # name: sequence lengths (Llama-3-8B heads: 32 q / 8 kv, D = 128)
    "8x8k": [8192] * 8,
    "long_tail": [16384, 8192, 8192, 4096, 4096, 2048, 2048, 1024, 1024, 512, 512, 256, 256, 128],
    "many_short": [512] * 128,
    "one_32k_plus_32x1k": [32768] + [1024] * 32,
}


def time_fn(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default=None)
    ap.add_argument("--mixes", default=None, help="comma-separated subset of the mixes")
    ap.add_argument("--no-comparators", action="store_true")
    args = ap.parse_args()
    hq, hkv, d, dtype = 32, 8, 128, torch.bfloat16
    rows = []
    for name, lens in MIXES.items():
        if args.mixes and name not in args.mixes.split(","):
            continue
        g = torch.Generator(device="cuda").manual_seed(7)
        total = sum(lens)
        q = torch.randn(total, hq, d, generator=g, device="cuda").to(dtype)
        k = torch.randn(total, hkv, d, generator=g, device="cuda").to(dtype)
        v = torch.randn(total, hkv, d, generator=g, device="cuda").to(dtype)
        cu = [0]
        for n in lens:
            cu.append(cu[-1] + n)
        flops = sum(4.0 * hq * n * n * d / 2 for n in lens)
        cu_dev = torch.tensor(cu, dtype=torch.int32, device="cuda")   # device offsets: no host work per call
        ms = time_fn(lambda: flash_attn_varlen_func(q, k, v, cu_dev, cu_dev, max(lens), max(lens), causal=True),
                     args.iters, args.warmup)
        row = {"mix": name, "num_seqs": len(lens), "tokens": total, "flops": flops,
               "varlen_ms": ms, "varlen_tflops": flops / ms / 1e9}
        # the same tokens as a padded fixed-length batch (max length), padding FLOPs not counted
        smax = max(lens)
        if len(lens) * smax <= 4 * total and not args.no_comparators:
            if min(lens) == smax:   # equal lengths: the very same tokens as a [B, H, S, D] view
                qp = q.view(len(lens), smax, hq, d).transpose(1, 2)
                kp = k.view(len(lens), smax, hkv, d).transpose(1, 2)
                vp = v.view(len(lens), smax, hkv, d).transpose(1, 2)
            else:                   # random data of the padded shape (zeros would run cooler and faster)
                qp = torch.randn(len(lens), hq, smax, d, generator=g, device="cuda").to(dtype)
                kp = torch.randn(len(lens), hkv, smax, d, generator=g, device="cuda").to(dtype)
                vp = torch.randn(len(lens), hkv, smax, d, generator=g, device="cuda").to(dtype)
            ms_p = time_fn(lambda: flash_attn_func(qp, kp, vp, causal=True), args.iters, args.warmup)
            row.update(padded_ms=ms_p, padded_tflops_useful=flops / ms_p / 1e9)
        try:
            if args.no_comparators:
                raise RuntimeError("skipped")
            from flash_attn import flash_attn_varlen_func as fa2_varlen

            cu_d = torch.tensor(cu, dtype=torch.int32, device="cuda")
            ms_l = time_fn(lambda: fa2_varlen(q, k, v, cu_d, cu_d, smax, smax, causal=True),
                           args.iters, args.warmup)
            row.update(flash_attn2_ms=ms_l, flash_attn2_tflops=flops / ms_l / 1e9)
        except Exception as exc:  # comparator only
            row.update(flash_attn2_error=str(exc)[:120])
        rows.append(row)
        print(json.dumps(row), flush=True)
    if args.out:
        Path(args.out).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
