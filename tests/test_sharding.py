"""Multi-GPU host logic on CPU: world_size-2 gloo run of the (batch, kv-head) sharding.
The per-unit compute is the oracle here (a checker standing in for the GPU op — the sharding code
itself never imports it)."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from flash_attention_cute_b200 import sharding  # noqa: E402


def test_unit_ranges_cover_disjointly():
    for units in (1, 7, 8, 64, 128):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                b, e = sharding.unit_range(units, world, r)
                assert 0 <= b <= e <= units
                seen += list(range(b, e))
            assert seen == list(range(units))
            sizes = [sharding.unit_range(units, world, r) for r in range(world)]
            assert max(e - b for b, e in sizes) - min(e - b for b, e in sizes) <= 1


def test_unit_slices_runs():
    runs = sharding.unit_slices(batch=2, heads_kv=8, world=4, rank=1)  # units 4..7 -> batch 0, heads 4..8
    assert runs == [(0, 4, 8)]
    runs = sharding.unit_slices(batch=3, heads_kv=2, world=2, rank=0)  # units 0..2
    assert runs == [(0, 0, 2), (1, 0, 1)]
    with pytest.raises(ValueError):
        sharding.unit_range(8, 2, 2)


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank: int, world: int, port: int):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle.eager_attention import eager_attention  # checker only

        g = torch.Generator().manual_seed(5)  # same inputs on every rank
        B, Hq, Hkv, S, D = 3, 8, 2, 64, 32
        q = torch.randn(B, Hq, S, D, generator=g)
        k = torch.randn(B, Hkv, S, D, generator=g)
        v = torch.randn(B, Hkv, S, D, generator=g)

        def attn(q_, k_, v_, causal=False):
            return eager_attention(q_, k_, v_, None, causal)

        for causal in (False, True):
            o_local = sharding.sharded_forward(q, k, v, world, rank, attn, causal=causal)
            # the rank wrote its own units only
            G = Hq // Hkv
            mine = torch.zeros(B, Hq, dtype=torch.bool)
            for b, h0, h1 in sharding.unit_slices(B, Hkv, world, rank):
                mine[b, h0 * G:h1 * G] = True
            assert torch.all(o_local[~mine] == 0)
            o = sharding.assemble(o_local)
            ref = eager_attention(q, k, v, None, causal)
            assert torch.allclose(o, ref, atol=1e-6, rtol=0), (o - ref).abs().max()
        # max-over-ranks timing plumbing used by bench.py
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == float(world)
    finally:
        dist.destroy_process_group()


def test_sharded_forward_gloo_world2():
    port = _free_port()
    mp.spawn(_worker, args=(2, port), nprocs=2, join=True)
