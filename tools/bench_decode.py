"""Decode (seqlen_q == 1) bandwidth benchmark: split-KV kernel vs the prefill kernel on the same
inputs vs flash-attn 2.8 (library comparator), as achieved HBM GB/s against MEASURED_PEAKS.json.

Algorithmic bytes per call = K + V read once + q + o.  Every iteration uses a
different copy of the inputs (ring of copies larger than the 126 MB L2), so the K/V bytes really
come from HBM.  CUDA-event timing on the current stream, warm-up first."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from flash_attention_cute_b200.load_cpp_extention import FLAG_NO_DECODE, load_extension  # noqa: E402

SHAPES = {
    # name: B, Hq, Hkv, Sk, D
    "ref_decode": (16, 64, 8, 1024, 128),   # scripts/benchmark_kernel.py usage example (seqlen_q 1)
    "b16_s8k": (16, 64, 8, 8192, 128),
    "b64_s4k": (64, 32, 8, 4096, 128),
    "b1_s32k": (1, 32, 8, 32768, 128),
    "b4_s64k": (4, 32, 8, 65536, 128),
    "b8_s2k_d64": (8, 14, 2, 2048, 64),
}


def time_fn(fn, sets, iters, warmup):
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


def time_graph(fn, sets, iters, warmup):
    """Same loop captured into one CUDA graph and replayed: no per-call host time in the number."""
    for i in range(warmup):
        fn(*sets[i % len(sets)])
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        for i in range(3):
            fn(*sets[i % len(sets)])  # per-stream workspace is created outside the capture
        side.synchronize()
        with torch.cuda.graph(g, stream=side):
            for i in range(iters):
                fn(*sets[i % len(sets)])
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default=",".join(SHAPES))
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--dtype", default="bfloat16")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dtype = getattr(torch, args.dtype)
    dev = torch.device("cuda:0")
    ext = load_extension()
    peaks = {}
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        peaks = json.loads(p.read_text())
    hbm = float(peaks.get("hbm_gbs", 6553.0))
    try:
        from flash_attn import flash_attn_func as fa_official
    except Exception:  # library comparator only
        fa_official = None
    rows = []
    for name in args.shapes.split(","):
        B, Hq, Hkv, Sk, D = SHAPES[name]
        esz = 2
        nbytes = (2 * B * Hkv * Sk * D + 2 * B * Hq * D) * esz
        ncopies = max(2, min(64, int(3e8 // nbytes) + 1))
        g = torch.Generator(device=dev).manual_seed(0)
        sets = []
        for _ in range(ncopies):
            q = torch.randn(B, Hq, 1, D, generator=g, device=dev).to(dtype)
            k = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dtype)
            v = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dtype)
            sets.append((q, k, v))
        scale = D ** -0.5
        plan = ext.decode_plan(*sets[0])
        ms_dec = time_fn(lambda q, k, v: ext.flash_attention_fwd(q, k, v, scale, False), sets, args.iters, args.warmup)
        ms_pre = time_fn(lambda q, k, v: ext.flash_attention_fwd(q, k, v, scale, False, FLAG_NO_DECODE), sets,
                         max(10, args.iters // 4), 5)
        ms_host = ms_dec
        ms_dec = time_graph(lambda q, k, v: ext.flash_attention_fwd(q, k, v, scale, False), sets, args.iters, args.warmup)
        ms_pre = time_graph(lambda q, k, v: ext.flash_attention_fwd(q, k, v, scale, False, FLAG_NO_DECODE), sets,
                            max(10, args.iters // 4), 5)
        rec = {"shape": name, "timing": "CUDA graph replay of the op calls", "eager_call_us": ms_host * 1e3, "B": B, "Hq": Hq, "Hkv": Hkv, "Sk": Sk, "D": D, "dtype": args.dtype,
               "nsplit": plan["nsplit"], "chunk": plan["chunk"], "bytes": nbytes, "input_copies": ncopies,
               "decode_us": ms_dec * 1e3, "decode_gbs": nbytes / ms_dec / 1e6, "frac_hbm": nbytes / ms_dec / 1e6 / hbm,
               "prefill_kernel_us": ms_pre * 1e3, "prefill_kernel_gbs": nbytes / ms_pre / 1e6, "hbm_peak_gbs": hbm}
        if fa_official is not None:
            fsets = [(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)) for q, k, v in sets]
            fsets = [(q.contiguous(), k.contiguous(), v.contiguous()) for q, k, v in fsets]
            ms_fa = time_graph(lambda q, k, v: fa_official(q, k, v, softmax_scale=scale, causal=False), fsets,
                               args.iters, args.warmup)
            rec["flash_attn_us"] = ms_fa * 1e3
            rec["flash_attn_gbs"] = nbytes / ms_fa / 1e6
            del fsets
        a = ext.flash_attention_fwd(*sets[0], scale, False).float()
        b = ext.flash_attention_fwd(*sets[0], scale, False, FLAG_NO_DECODE).float()
        rec["max_abs_vs_prefill_kernel"] = (a - b).abs().max().item()
        rows.append(rec)
        print(json.dumps(rec), flush=True)
        del sets
        torch.cuda.empty_cache()
    if args.json:
        Path(args.json).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
