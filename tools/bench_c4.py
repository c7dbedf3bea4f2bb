"""BASELINE config 4, measured: fp16, D = 64, Qwen2-0.5B heads (14 q / 2 kv) and MQA (14 / 1),
S in {512 ... 32768}, B = 32768 / S, causal and non-causal (the reference suggests the
D = 64 case at scripts/benchmark_kernel.py:181-182 and runs it inside its D = 128 kernel,
csrc/kernel_dispatcher.h:45-52).

Per point: ms, TFLOP/s (F = 4·B·Hq·S²·D, halved for causal), fraction of the measured cuBLAS peak
(MEASURED_PEAKS.json) and of nominal 2250, and on the SAME inputs the reference's own sm80 kernel
compiled for sm_100a (oracle/_ref, comparator only) and flash-attn 2.8 (library) when importable.
CUDA events, warm-up first, inputs + outputs of every point are 2 x 32768 x 14 x 64 x 2 B x ... well
below L2 for K/V at small B·S, so every timed call runs on ROTATING input sets larger than L2.

    python tools/bench_c4.py --out profiles/r02/c4_sweep.json
"""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def time_rotating(fn, sets, iters, warmup=3):
    for i in range(warmup):
        fn(*sets[i % len(sets)])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        fn(*sets[i % len(sets)])
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--seqlens", default="512,1024,2048,4096,8192,16384,32768")
    ap.add_argument("--no-comparators", action="store_true")
    args = ap.parse_args()
    from flash_attention import flash_attn_func
    from flash_attention_cute_b200 import _lib

    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    peak = peaks.get("bf16_tflops", 1590.0)
    ref_kernel = None
    fa2 = None
    if not args.no_comparators:
        try:
            from oracle import ref_kernel as rk  # comparator only (bench tool, not the product)
            ref_kernel = rk if rk.available() else None
        except Exception:
            ref_kernel = None
        try:
            from flash_attn import flash_attn_func as fa2_func
            fa2 = fa2_func
        except Exception:
            fa2 = None
    dev = torch.device("cuda:0")
    Hq, D, dt = 14, 64, torch.float16
    rows = []
    for S in (int(x) for x in args.seqlens.split(",")):
        B = 32768 // S
        for Hkv in (2, 1):
            # q + k + v + o of one set = 32768 * (2*14 + 2*Hkv) * 64 * 2 B = 117-134 MB: two sets exceed the 126 MB L2
            g = torch.Generator(device=dev).manual_seed(1004)
            sets = []
            for _ in range(3):
                q = torch.randn(B, Hq, S, D, generator=g, device=dev).to(dt)
                k = torch.randn(B, Hkv, S, D, generator=g, device=dev).to(dt)
                v = torch.randn(B, Hkv, S, D, generator=g, device=dev).to(dt)
                sets.append((q, k, v))
            for causal in (False, True):
                flops = 4.0 * B * Hq * S * S * D / (2 if causal else 1)
                ms = time_rotating(lambda q, k, v: flash_attn_func(q, k, v, causal=causal), sets, args.iters)
                _lib.check_device(0)
                row = {"S": S, "B": B, "Hq": Hq, "Hkv": Hkv, "D": D, "dtype": "fp16", "causal": causal,
                       "flops": flops, "ms": ms, "tflops": flops / ms / 1e9,
                       "frac_of_measured_peak": flops / ms / 1e9 / peak, "pct_of_nominal_2250": 100 * flops / ms / 1e9 / 2250}
                if ref_kernel is not None:
                    try:
                        ms_r = time_rotating(lambda q, k, v: ref_kernel.ref_flash_attn_func(q, k, v, None, causal),
                                             sets, max(3, args.iters // 4), warmup=1)
                        row.update(reference_sm80_kernel_ms=ms_r, reference_sm80_kernel_tflops=flops / ms_r / 1e9)
                    except Exception as exc:
                        row.update(reference_sm80_kernel_error=str(exc)[:100])
                if fa2 is not None:
                    try:
                        sets_t = [(q.transpose(1, 2).contiguous(), k.transpose(1, 2).contiguous(),
                                   v.transpose(1, 2).contiguous()) for q, k, v in sets]
                        ms_l = time_rotating(lambda q, k, v: fa2(q, k, v, causal=causal), sets_t,
                                             max(3, args.iters // 2), warmup=1)
                        row.update(flash_attn2_ms=ms_l, flash_attn2_tflops=flops / ms_l / 1e9)
                        del sets_t
                    except Exception as exc:
                        row.update(flash_attn2_error=str(exc)[:100])
                rows.append(row)
                print(json.dumps(row), flush=True)
            del sets
            torch.cuda.empty_cache()
    if args.out:
        Path(args.out).write_text(json.dumps({"peak_tflops_measured": peak, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
