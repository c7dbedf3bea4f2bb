#!/usr/bin/env python
"""Kernel benchmark + accuracy check with the command line of the reference's
scripts/benchmark_kernel.py:125-136 (same flags and defaults: --batch-size 16 --num-heads-q 64
--num-heads-kv 8 --seqlen-q 1024 --seqlen-kv 1024 --dim 128 --iter 100 --dtype half --device cuda
[--causal]), extended with --seed and --json.

Differences from the reference script, on purpose: inputs are seeded; timing uses CUDA events
after a warm-up instead of time.time() around a cold loop; every implementation that is not
installed is reported as unavailable instead of failing the import; the accuracy section uses the
acceptance bounds of BASELINE.json (max-abs and relative L2 against an fp32 eager
softmax(QK^T)V evaluated on the GPU per (batch, kv head) — the comparator role eager_attention has
in the reference script; nothing under oracle/ is used here) next to the reference's own
MSE / allclose(atol=1e-3) lines; a JSON record with TFLOP/s and the fraction
of the measured / nominal B200 tensor peak is printed last.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def time_cuda(fn, iters: int, warmup: int = 3) -> float:
    """Average milliseconds per call, CUDA events on the current stream."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def eager_attention(q, k, v, causal: bool, out_dtype=None) -> torch.Tensor:
    """fp32 softmax(q kᵀ / sqrt(D) + causal bias) v, one (batch, kv head) at a time; GQA head h uses
    kv head h // G; causal alignment bottom-right (identical to top-left when Sq == Sk)."""
    B, Hq, Sq, D = q.shape
    Hkv, Sk = k.shape[1], k.shape[2]
    G = Hq // Hkv
    out = torch.empty((B, Hq, Sq, D), dtype=out_dtype or q.dtype, device=q.device)
    keep = None
    if causal:
        keep = torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).tril(diagonal=Sk - Sq)
    for b in range(B):
        for h in range(Hkv):
            s = torch.matmul(q[b, h * G:(h + 1) * G].float(), k[b, h].float().t()) * (D ** -0.5)
            if keep is not None:
                s = s.masked_fill(~keep, float("-inf"))
            p = torch.softmax(s, dim=-1).nan_to_num(0.0)
            out[b, h * G:(h + 1) * G] = torch.matmul(p, v[b, h].float()).to(out.dtype)
    return out


def comparators(q, k, v, causal):
    """name -> callable returning [B,H,S,D]; only what is importable on this machine."""
    out = {}
    try:
        from flash_attn import flash_attn_func as fa_official  # layout [B,S,H,D]

        out["Official"] = lambda: fa_official(q.permute(0, 2, 1, 3), k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3),
                                              causal=causal).permute(0, 2, 1, 3)
    except Exception:
        pass
    return out


def main() -> int:
    ap = argparse.ArgumentParser(description="Flash Attention Benchmark Tool (B200)")
    ap.add_argument("--batch-size", type=int, default=16)
    ap.add_argument("--num-heads-q", type=int, default=64)
    ap.add_argument("--num-heads-kv", type=int, default=8)
    ap.add_argument("--seqlen-q", type=int, default=1024)
    ap.add_argument("--seqlen-kv", type=int, default=1024)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--iter", type=int, default=100)
    ap.add_argument("--dtype", type=str, default="half")
    ap.add_argument("--device", type=str, default="cuda")
    ap.add_argument("--causal", action="store_true")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--json", type=str, default=None, help="also write the record to this file")
    ap.add_argument("--no-eager", action="store_true", help="skip the fp32 eager timing/accuracy (large shapes)")
    args = ap.parse_args()

    dtype = getattr(torch, args.dtype)
    assert isinstance(dtype, torch.dtype)
    dev = torch.device(args.device)
    if dev.type != "cuda" or not torch.cuda.is_available():
        raise SystemExit("benchmark_kernel: a CUDA (sm_100) device is required — the operator has no CPU path")
    from flash_attention import flash_attn_func

    B, Hq, Hkv, Sq, Sk, D = (args.batch_size, args.num_heads_q, args.num_heads_kv, args.seqlen_q,
                             args.seqlen_kv, args.dim)
    g = torch.Generator(device=dev).manual_seed(args.seed)
    q = torch.randn((B, Hq, Sq, D), generator=g, device=dev, dtype=torch.float32).to(dtype)
    k = torch.randn((B, Hkv, Sk, D), generator=g, device=dev, dtype=torch.float32).to(dtype)
    v = torch.randn((B, Hkv, Sk, D), generator=g, device=dev, dtype=torch.float32).to(dtype)
    print("\nBenchmark Configuration:")
    print(f"Batch size: {B}")
    print(f"Query heads: {Hq}, Key/Value heads: {Hkv}")
    print(f"Query length: {Sq}, Key/Value length: {Sk}")
    print(f"Dimension: {D}, Causal: {args.causal}")
    print(f"Data type: {dtype}, Device: {dev}, Seed: {args.seed}")
    print(f"Iterations: {args.iter}\n")

    flops = 4.0 * B * Hq * Sq * Sk * D / (2 if args.causal else 1)
    rec = {"config": vars(args), "flops": flops, "ms": {}, "tflops": {}, "accuracy": {}}
    print("Running benchmarks...")
    ms = time_cuda(lambda: flash_attn_func(q, k, v, causal=args.causal), args.iter)
    rec["ms"]["Custom"] = ms
    print(f"[Custom] Total: {ms * args.iter:.3f}ms | Per iter: {ms:.3f}ms | {flops / ms / 1e9:.1f} TFLOP/s")
    others = comparators(q, k, v, args.causal)
    for name, fn in others.items():
        try:
            ms_o = time_cuda(fn, max(1, min(args.iter, 20)))
            rec["ms"][name] = ms_o
            print(f"[{name}] Per iter: {ms_o:.3f}ms | {flops / ms_o / 1e9:.1f} TFLOP/s")
        except Exception as e:  # a comparator must never take the benchmark down
            print(f"[{name}] unavailable: {repr(e)[:120]}")
    if not args.no_eager:
        ms_e = time_cuda(lambda: eager_attention(q, k, v, args.causal), max(1, min(args.iter, 3)), 1)
        rec["ms"]["Eager"] = ms_e
        print(f"[Eager] Per iter: {ms_e:.3f}ms")
    rec["tflops"] = {n: flops / t / 1e9 for n, t in rec["ms"].items()}

    print("\nChecking accuracy...")
    with torch.no_grad():
        o = flash_attn_func(q, k, v, causal=args.causal)
        if not args.no_eager:
            ref32 = eager_attention(q, k, v, args.causal, out_dtype=torch.float32)
            diff = o.float() - ref32
            rec["accuracy"]["eager"] = {
                "mse": float((diff ** 2).mean()), "max_abs": float(diff.abs().max()),
                "rel_l2": float(diff.norm() / ref32.norm().clamp_min(1e-30)),
                "allclose_atol_1e-3": bool(torch.allclose(o, ref32.to(dtype), atol=1e-3)),
            }
            a = rec["accuracy"]["eager"]
            print(f"MSE between implementations (custom, eager): {a['mse']:.4e}")
            print(f"max-abs {a['max_abs']:.3e}, relative L2 {a['rel_l2']:.3e} (bounds: "
                  f"{'2e-3 / 1e-3' if dtype == torch.float16 else '1e-2 / 5e-3'})")
            print(f"AllClose check (custom, eager): {a['allclose_atol_1e-3']}")
        for name, fn in others.items():
            try:
                oo = fn()
                rec["accuracy"][name] = {"mse": float(((o.float() - oo.float()) ** 2).mean()),
                                         "allclose_atol_1e-3": bool(torch.allclose(o, oo, atol=1e-3))}
                print(f"MSE between implementations (custom, {name}): {rec['accuracy'][name]['mse']:.4e}")
            except Exception as e:
                print(f"[{name}] accuracy unavailable: {repr(e)[:120]}")

    peaks = {}
    pf = ROOT / "MEASURED_PEAKS.json"
    if pf.exists():
        peaks = json.loads(pf.read_text())
    measured = peaks.get("bf16_tflops", 1590.0)
    rec["roofline"] = {"bound": "tensor", "achieved": rec["tflops"]["Custom"], "peak_measured": measured,
                       "frac_measured": rec["tflops"]["Custom"] / measured,
                       "frac_nominal_2250": rec["tflops"]["Custom"] / 2250.0}
    line = json.dumps(rec, default=str)
    print("\n" + line)
    if args.json:
        Path(args.json).write_text(line + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
