"""(batch, kv-head) sharding of the forward operator over the GPUs of one box.

One unit = batch b, kv-head h with its G = Hq / Hkv query heads: it reads q[b, h*G:(h+1)*G],
k[b, h], v[b, h] and writes o[b, h*G:(h+1)*G]; nothing is shared between units (the reference
kernel's addressing, csrc/flash_attention_template.cuh:157-160), so ranks own disjoint contiguous
ranges of units and the hot path needs no collective.  torch.distributed (NCCL over NVLink on the
box, gloo in the CPU tests) is used only to assemble O afterwards.
"""
from __future__ import annotations

from typing import Callable, List, Tuple

import torch


def unit_range(num_units: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [begin, end) of units owned by `rank`; sizes differ by at most one."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(num_units, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def unit_slices(batch: int, heads_kv: int, world: int, rank: int) -> List[Tuple[int, int, int]]:
    """The rank's units as (batch index, first kv head, one-past-last kv head) runs."""
    begin, end = unit_range(batch * heads_kv, world, rank)
    runs = []
    u = begin
    while u < end:
        b, h0 = divmod(u, heads_kv)
        h1 = min(heads_kv, h0 + (end - u))
        runs.append((b, h0, h1))
        u += h1 - h0
    return runs


def sharded_forward(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, world: int, rank: int,
                    attn: Callable[..., torch.Tensor], **kw) -> torch.Tensor:
    """Run `attn` (flash_attn_func on a GPU box) on this rank's units only.  Returns an O of full
    shape in which only the rank's units are written (zeros elsewhere): ranks' results are
    disjoint, so a sum-reduce or an all-gather + select assembles the box-level O."""
    B, Hq, _, _ = q.shape
    Hkv = k.shape[1]
    G = Hq // Hkv
    o = torch.zeros_like(q)
    for b, h0, h1 in unit_slices(B, Hkv, world, rank):
        o[b:b + 1, h0 * G:h1 * G] = attn(q[b:b + 1, h0 * G:h1 * G], k[b:b + 1, h0:h1], v[b:b + 1, h0:h1], **kw)
    return o


def assemble(o_local: torch.Tensor, group=None) -> torch.Tensor:
    """Box-level O on every rank: all-gather of the per-rank tensors, summed (units are disjoint)."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    parts = [torch.empty_like(o_local) for _ in range(world)]
    dist.all_gather(parts, o_local.contiguous(), group=group)
    out = parts[0]
    for p in parts[1:]:
        out = out + p
    return out
