"""Times the prefill op on the BASELINE shapes with whatever library FA_SM100_LIB_OVERRIDE names
(development aid: same-box A/B of two builds, run alternately from a shell loop)."""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from flash_attention import flash_attn_func  # noqa: E402

SHAPES = {
    "c3": (8, 32, 8, 8192, 128, torch.bfloat16, True),
    "c3nc": (8, 32, 8, 8192, 128, torch.bfloat16, False),
    "c2": (4, 32, 32, 4096, 128, torch.bfloat16, False),
    "c4s4k": (8, 14, 2, 4096, 64, torch.float16, True),
    "c4s4knc": (8, 14, 2, 4096, 64, torch.float16, False),
    "s512": (128, 32, 8, 512, 128, torch.bfloat16, True),
}


def main():
    dev = torch.device("cuda:0")
    out = {}
    for name, (B, Hq, Hkv, S, D, dtype, causal) in SHAPES.items():
        flops = 4.0 * B * Hq * S * S * D * (0.5 if causal else 1.0)
        g = torch.Generator(device=dev).manual_seed(0)
        sets = [tuple(torch.randn(B, h, S, D, generator=g, device=dev).to(dtype) for h in (Hq, Hkv, Hkv))
                for _ in range(3)]
        for i in range(5):
            flash_attn_func(*sets[i % 3], causal=causal)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(30):
            flash_attn_func(*sets[i % 3], causal=causal)
        b.record()
        torch.cuda.synchronize()
        out[name] = round(flops / (a.elapsed_time(b) / 30) / 1e9, 1)
        del sets
    print(json.dumps(out))


if __name__ == "__main__":
    main()
