"""Kernel-level comparator: this repo's prefill kernel next to cuDNN's fused attention (the
Blackwell-tuned library kernel PyTorch reaches through scaled_dot_product_attention with
SDPBackend.CUDNN_ATTENTION) and PyTorch's flash backend, on the BASELINE shapes.  Library kernels
are comparators only; nothing here is on the product path.  CUDA events over `iters` back-to-back
calls after warm-up; the inputs of consecutive calls rotate through > 126 MB."""
import argparse
import json
import sys
from pathlib import Path

import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from flash_attention import flash_attn_func  # noqa: E402

SHAPES = {
    # name: B, Hq, Hkv, S, D, dtype, causal
    "c3_llama8b_causal": (8, 32, 8, 8192, 128, torch.bfloat16, True),
    "c3_llama8b_full": (8, 32, 8, 8192, 128, torch.bfloat16, False),
    "c2_mha_causal": (8, 32, 32, 4096, 128, torch.float16, True),
    "c4_qwen_s4k_causal": (8, 14, 2, 4096, 64, torch.float16, True),
    "c4_qwen_s512_causal": (64, 14, 2, 512, 64, torch.float16, True),
    "short_b128_s512_causal": (128, 32, 8, 512, 128, torch.bfloat16, True),
}


def timeit(fn, sets, iters, warmup=5):
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
    ap.add_argument("--shapes", default=",".join(SHAPES))
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    rows = []
    for name in args.shapes.split(","):
        B, Hq, Hkv, S, D, dtype, causal = SHAPES[name]
        flops = 4.0 * B * Hq * S * S * D * (0.5 if causal else 1.0)
        nbytes = 2 * B * (2 * Hq + 2 * Hkv) * S * D
        ncopies = max(2, min(6, int(3e8 // nbytes) + 1))
        g = torch.Generator(device=dev).manual_seed(0)
        sets = [tuple(torch.randn(B, h, S, D, generator=g, device=dev).to(dtype) for h in (Hq, Hkv, Hkv))
                for _ in range(ncopies)]
        rec = {"shape": name, "B": B, "Hq": Hq, "Hkv": Hkv, "S": S, "D": D, "dtype": str(dtype), "causal": causal,
               "flops": flops}
        ms = timeit(lambda q, k, v: flash_attn_func(q, k, v, causal=causal), sets, args.iters)
        rec.update(ours_ms=ms, ours_tflops=flops / ms / 1e9)
        ref = flash_attn_func(*sets[0], causal=causal).float()
        for label, backend in (("cudnn", SDPBackend.CUDNN_ATTENTION), ("torch_flash", SDPBackend.FLASH_ATTENTION)):
            try:
                def call(q, k, v):
                    with sdpa_kernel(backend):
                        return F.scaled_dot_product_attention(q, k, v, is_causal=causal, enable_gqa=Hq != Hkv)
                ms = timeit(call, sets, args.iters)
                out = call(*sets[0]).float()
                diff = (out - ref).abs()
                rec.update({f"{label}_ms": ms, f"{label}_tflops": flops / ms / 1e9,
                            f"{label}_max_abs_diff_to_ours": diff.max().item(),
                            f"{label}_mean_abs_diff_to_ours": diff.mean().item()})
            except Exception as exc:  # backend not available for this shape on this build
                rec[f"{label}_error"] = str(exc).splitlines()[0][:160]
        rows.append(rec)
        print(json.dumps(rec), flush=True)
        del sets
        torch.cuda.empty_cache()
    if args.json:
        Path(args.json).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
