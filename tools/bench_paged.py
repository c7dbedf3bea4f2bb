"""Paged-KV decode bandwidth (fa_sm100_fwd_decode_paged): achieved HBM GB/s against
MEASURED_PEAKS.json, next to the contiguous split-KV kernel on the same keys and flash-attn 2.8's
flash_attn_with_kvcache (library comparator) on the same page pool.

Algorithmic bytes per call = K + V pages of the live keys read once + q + o + the block-table
entries touched.  Pages are handed out in random order; every iteration uses another copy of the
pool (ring of copies larger than the 126 MB L2).  CUDA-graph replay of the op calls."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

from bench_decode import time_graph  # noqa: E402
from flash_attention import flash_attn_decode_paged  # noqa: E402
from flash_attention_cute_b200.load_cpp_extention import load_extension  # noqa: E402

SHAPES = {
    # name: B, Hq, Hkv, Sk, D, page
    "b16_s8k_p128": (16, 64, 8, 8192, 128, 128),
    "b16_s8k_p16": (16, 64, 8, 8192, 128, 16),
    "b64_s4k_p256": (64, 32, 8, 4096, 128, 256),
    "b4_s64k_p128": (4, 32, 8, 65536, 128, 128),
    "b16_ragged_p128": (16, 64, 8, 8192, 128, 128),   # lengths 1 K ... 8 K
    "b64_ragged_p256": (64, 32, 8, 4096, 128, 256),   # lengths 512 ... 4 K
    "b4_ragged_p128": (4, 32, 8, 65536, 128, 128),    # lengths 4 K ... 64 K
    "b32_skewed_p128": (32, 64, 8, 16384, 128, 128),  # one sequence of 16 K, the rest 512 keys
    "b256_short_p64": (256, 32, 8, 1024, 128, 64),    # 2048 units of 64 ... 1024 keys
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default=",".join(SHAPES))
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--json", default=None)
    ap.add_argument("--modes", default="flat,one_wave", help="ways to cut the work to time: flat, one_wave")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    dtype = torch.bfloat16
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = float(peaks.get("hbm_gbs", 6553.0))
    try:
        from flash_attn import flash_attn_with_kvcache
    except Exception:
        flash_attn_with_kvcache = None
    ext = load_extension()
    rows = []
    from flash_attention_cute_b200.paged import prefer_one_wave
    modes = [m for m in args.modes.split(",") if m]
    for name, mode in [(n, m) for n in args.shapes.split(",") for m in modes]:
        one_wave = mode == "one_wave"
        B, Hq, Hkv, Sk, D, page = SHAPES[name]
        g = torch.Generator(device=dev).manual_seed(0)
        max_pages = Sk // page
        n_pages = B * max_pages
        lens = [Sk] * B
        if "ragged" in name:
            lo = Sk // 8 if B > 4 else Sk // 16
            lens = [lo + (Sk - lo) * i // (B - 1) for i in range(B)]
        if "skewed" in name:
            lens = [Sk] + [512] * (B - 1)
        if "short" in name:
            lens = [64 + (Sk - 64) * ((i * 37) % B) // (B - 1) for i in range(B)]
        uniform = len(set(lens)) == 1
        live = sum(lens)
        nbytes = (2 * live * Hkv * D + 2 * B * Hq * D) * 2 + 4 * sum((n + page - 1) // page for n in lens)
        ncopies = max(2, min(16, int(3e8 // nbytes) + 1))
        sets = []
        for _ in range(ncopies):
            kc = torch.randn(n_pages, page, Hkv, D, generator=g, device=dev).to(dtype)
            vc = torch.randn(n_pages, page, Hkv, D, generator=g, device=dev).to(dtype)
            bt = torch.randperm(n_pages, generator=g, device=dev).to(torch.int32).view(B, max_pages).contiguous()
            q = torch.randn(B, Hq, D, generator=g, device=dev).to(dtype)
            sets.append((q, kc, vc, bt))
        sl = torch.tensor(lens, dtype=torch.int32, device=dev)
        ms = time_graph(lambda q, kc, vc, bt: flash_attn_decode_paged(q, kc, vc, bt, sl, max_seqlen_k=Sk,
                                                                      one_wave=one_wave), sets, args.iters, 10)
        rec = {"shape": name, "flat": mode, "host_lengths_would_pick": "one_wave" if prefer_one_wave(lens, Hkv) else "flat",
               "B": B, "Hq": Hq, "Hkv": Hkv, "Sk": Sk, "D": D, "page": page, "live_keys": live,
               "bytes": nbytes, "paged_us": ms * 1e3, "paged_gbs": nbytes / ms / 1e6,
               "frac_hbm": nbytes / ms / 1e6 / hbm, "hbm_peak_gbs": hbm, "input_copies": ncopies,
               "timing": "CUDA graph replay"}
        if uniform and mode in (None, modes[0]):  # the same keys, contiguous, through the non-paged kernel
            dsets = []
            for q, kc, vc, bt in sets[:max(2, ncopies // 2)]:
                dk = kc[bt.long()].reshape(B, Sk, Hkv, D).transpose(1, 2).contiguous()
                dv = vc[bt.long()].reshape(B, Sk, Hkv, D).transpose(1, 2).contiguous()
                dsets.append((q.unsqueeze(2), dk, dv))
            ms_c = time_graph(lambda q, k, v: ext.flash_attention_fwd(q, k, v, D ** -0.5, False), dsets, args.iters, 10)
            rec.update(contiguous_us=ms_c * 1e3, contiguous_gbs=nbytes / ms_c / 1e6)
            a = flash_attn_decode_paged(*sets[0], sl, max_seqlen_k=Sk, one_wave=one_wave)
            b = ext.flash_attention_fwd(*dsets[0], D ** -0.5, False)[:, :, 0]
            rec["max_abs_diff_to_contiguous"] = float((a.float() - b.float()).abs().max())
            del dsets
        if flash_attn_with_kvcache is not None and page % 256 == 0 and mode in (None, modes[0]):
            try:
                ms_l = time_graph(lambda q, kc, vc, bt: flash_attn_with_kvcache(
                    q.unsqueeze(1), kc, vc, cache_seqlens=sl, block_table=bt, softmax_scale=D ** -0.5), sets,
                    args.iters, 10)
                rec.update(flash_attn2_us=ms_l * 1e3, flash_attn2_gbs=nbytes / ms_l / 1e6)
            except Exception as exc:
                rec["flash_attn2_error"] = str(exc)[:100]
        rows.append(rec)
        print(json.dumps(rec), flush=True)
        del sets
        torch.cuda.empty_cache()
    if args.json:
        Path(args.json).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
