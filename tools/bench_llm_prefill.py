"""BASELINE config 5: Llama-3-8B-shaped random-init prefill through models.patch_llama.patch_attn(),
B = 8, S = 8192, on 1 / 2 / 4 / 8 B200 (batch-sharded replicas: 8/N sequences per GPU — the batch
axis of the (batch, kv-head) partition, no collective on the path).

Timing follows the reference's prefill loop (scripts/benchmark_llm.py:44-63): CUDA events around
``model(input_ids)`` under no_grad after a warm-up; there is no checkpoint (no network), so the
weights are random (torch.manual_seed(0)) and the logits are limited to the last position to avoid
a 16 GB logits tensor.  Reported: prefill tokens/s (box = all tokens / slowest rank),
the share of the step spent in the attention and RoPE kernels (torch.profiler kernel times), and
the same model with HF's ``sdpa`` attention as the comparator.

    python tools/bench_llm_prefill.py [--layers 32] [--batch 8] [--seqlen 8192] [--out x.json]
    torchrun --nproc-per-node N ... tools/bench_llm_prefill.py
"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def build_model(layers, dev, attn_impl):
    from transformers import LlamaConfig, LlamaForCausalLM

    cfg = LlamaConfig(hidden_size=4096, num_hidden_layers=layers, num_attention_heads=32, num_key_value_heads=8,
                      intermediate_size=14336, vocab_size=128256, rope_theta=500000.0,
                      max_position_embeddings=8192, attn_implementation=attn_impl)
    torch.manual_seed(0)
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device(dev):
            model = LlamaForCausalLM(cfg)
    finally:
        torch.set_default_dtype(old)
    return model.eval()


def time_prefill(model, ids, trials, warmup):
    with torch.no_grad():
        for _ in range(warmup):
            model(ids, use_cache=False, logits_to_keep=1)
        torch.cuda.synchronize()
        times = []
        for _ in range(trials):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = model(ids, use_cache=False, logits_to_keep=1)
            b.record()
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b) / 1e3)
    return sum(times) / len(times), out.logits


def kernel_shares(model, ids):
    """Fraction of one prefill's GPU kernel time per kernel family (CUPTI through torch.profiler)."""
    try:
        from torch.profiler import ProfilerActivity, profile

        with torch.no_grad(), profile(activities=[ProfilerActivity.CUDA]) as prof:
            model(ids, use_cache=False, logits_to_keep=1)
            torch.cuda.synchronize()
        tot, fam = 0.0, {"attention (fa_fwd_sm100)": 0.0, "rope (fa_rope_sm100)": 0.0, "gemm": 0.0}
        for e in prof.key_averages():
            t = getattr(e, "self_device_time_total", None)
            if t is None:
                t = getattr(e, "self_cuda_time_total", 0.0)
            if not t:
                continue
            tot += t
            n = e.key
            if "fa_fwd_sm100" in n:
                fam["attention (fa_fwd_sm100)"] += t
            elif "fa_rope_sm100" in n:
                fam["rope (fa_rope_sm100)"] += t
            elif "gemm" in n.lower() or "cutlass" in n.lower() or "nvjet" in n.lower() or "xmma" in n.lower():
                fam["gemm"] += t
        return {k: v / tot for k, v in fam.items()} | {"kernel_time_ms": tot / 1e3} if tot else None
    except Exception as exc:
        return {"unavailable": str(exc)[:120]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--seqlen", type=int, default=8192)
    ap.add_argument("--trials", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--no-sdpa", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch.distributed as dist

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert args.batch % world == 0, "the batch must split evenly over the ranks"
    b_local = args.batch // world
    g = torch.Generator().manual_seed(1005)
    ids = torch.randint(0, 128256, (args.batch, args.seqlen), generator=g)[rank * b_local:(rank + 1) * b_local].to(dev)

    from flash_attention_cute_b200 import _lib
    from transformers.models.llama.modeling_llama import LlamaAttention

    def box_time(t):
        x = torch.tensor([t], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(x, op=dist.ReduceOp.MAX)
        return float(x.item())

    res = {"config": f"LlamaConfig(hidden 4096, {args.layers} layers, 32q/8kv, inter 14336, vocab 128256, rope_theta 5e5), "
                     f"bf16 random-init, input_ids [{args.batch},{args.seqlen}] ({b_local} per GPU), logits of the last position",
           "n_gpus": world, "tokens": args.batch * args.seqlen}
    # comparator first: HF sdpa attention, stock forward
    sdpa_logits = None
    if not args.no_sdpa:
        model = build_model(args.layers, dev, "sdpa")
        t, sdpa_logits = time_prefill(model, ids, args.trials, args.warmup)
        t = box_time(t)
        res["hf_sdpa"] = {"prefill_s": t, "tokens_per_s": args.batch * args.seqlen / t}
        sdpa_logits = sdpa_logits.float().clone()
        del model
        torch.cuda.empty_cache()
    # ours: the reference's integration point (models/patch_llama.py:4-5)
    original = LlamaAttention.forward
    from models.patch_llama import patch_attn

    model = build_model(args.layers, dev, "eager")
    patch_attn()
    try:
        n0 = _lib.launch_count()
        t, logits = time_prefill(model, ids, args.trials, args.warmup)
        launches = (_lib.launch_count() - n0) // (args.trials + args.warmup)
        _lib.check_device(local)
        t = box_time(t)
        res["patched"] = {"prefill_s": t, "tokens_per_s": args.batch * args.seqlen / t,
                          "fa_sm100_launches_per_step": launches}
        if rank == 0:
            res["patched"]["kernel_shares"] = kernel_shares(model, ids)
        attn_flops = 4.0 * b_local * 32 * args.seqlen * args.seqlen * 128 / 2 * args.layers
        res["attention_flops_per_gpu_per_step"] = attn_flops
        if sdpa_logits is not None:
            d = (logits.float() - sdpa_logits)
            res["logits_rel_l2_vs_sdpa"] = (d.norm() / sdpa_logits.norm()).item()
            res["speedup_vs_hf_sdpa"] = res["hf_sdpa"]["prefill_s"] / t
    finally:
        LlamaAttention.forward = original
    if rank == 0:
        print(json.dumps(res), flush=True)
        if args.out:
            Path(args.out).write_text(json.dumps(res, indent=1))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
