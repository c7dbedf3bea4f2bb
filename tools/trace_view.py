"""Record and summarise the in-kernel timeline of CTA 0 (FA_SM100_TUNE builds).  Debug tool.

    FA_SM100_TUNE=1 python tools/trace_view.py c3nc --emu 3

Roles: 0/1 softmax warp 0 of M-tile 0/1, 2 epilogue warp 0, 3 MMA issuer of M-tile 0, 4 TMA producer.
Softmax events: 1 S ready, 2 S loaded, 3 first chunk done, 4 P published, 7 lazy rescale.
MMA events: 2 prologue issued, 3 P seen + stage acquired, 5 batch issued, 23 tile done.
Epilogue: 8 O ready, 9 tile stored.  Producer: 1 tile start, 6 stage issued."""
import argparse
import ctypes
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
from prof_one import SHAPES  # noqa: E402

ROLES, MAXN = 5, 4096


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("shape")
    ap.add_argument("--emu", type=int, default=None)
    ap.add_argument("--spec", type=int, default=None)
    ap.add_argument("--dump", default=None)
    args = ap.parse_args()
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    ext = load_extension()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    B, Hq, Hkv, Sq, Sk, D, causal, dt = SHAPES[args.shape]
    g = torch.Generator(device=dev).manual_seed(7)
    q = torch.randn(B, Hq, Sq, D, generator=g, device=dev).to(dt)
    k = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
    v = torch.randn(B, Hkv, Sk, D, generator=g, device=dev).to(dt)
    flags = _lib.tune_flags(args.emu, None if args.spec is None else bool(args.spec))
    for _ in range(2):
        ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, flags)
    torch.cuda.synchronize()
    buf = torch.zeros(ROLES * MAXN, dtype=torch.int64, device=dev)
    lib.fa_sm100_internal_set_trace.argtypes = [ctypes.c_void_p]
    lib.fa_sm100_internal_set_trace(ctypes.c_void_p(buf.data_ptr()))
    ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, flags)
    torch.cuda.synchronize()
    lib.fa_sm100_internal_set_trace(ctypes.c_void_p(0))
    _lib.check_device(0)
    raw = buf.cpu().numpy().reshape(ROLES, MAXN)
    ev = {}
    for r in range(ROLES):
        x = raw[r][raw[r] != 0]
        ev[r] = [(int(e >> 56), int(e & ((1 << 56) - 1))) for e in x]
    t0 = min(t for r in ev for _, t in ev[r]) if any(ev.values()) else 0
    if args.dump:
        Path(args.dump).write_text(json.dumps({str(r): [(c, t - t0) for c, t in ev[r]] for r in ev}))

    def deltas(role, code):
        ts = [t for c, t in ev[role] if c == code]
        return np.diff(ts) if len(ts) > 1 else np.array([])

    def med(a):
        return float(np.median(a)) if len(a) else float("nan")

    print(f"shape {args.shape} flags emu={args.emu} spec={args.spec}; events per role:",
          {r: len(ev[r]) for r in ev})
    for role in (0, 1):
        e = ev[role]
        # per-iteration phases: 1 -> 2 -> 3 -> 4 -> (5) -> next 1
        seg = {"wait(prev end->S ready)": [], "ld": [], "max+scale": [], "expA": [], "expB": [], "period": []}
        last_end = None
        cur = {}
        for c, t in e:
            if c == 1:
                if last_end is not None:
                    seg["wait(prev end->S ready)"].append(t - last_end)
                if 1 in cur:
                    seg["period"].append(t - cur[1])
                cur = {1: t}
            elif c in (2, 3, 4, 5):
                cur[c] = t
                if c == 2:
                    seg["ld"].append(t - cur[1])
                if c == 3 and 2 in cur:
                    seg["max+scale"].append(t - cur[2])
                if c == 4 and 3 in cur:
                    seg["expA"].append(t - cur[3])
                    last_end = t
                if c == 5 and 4 in cur:
                    seg["expB"].append(t - cur[4])
                    last_end = t
        print(f"softmax tile {role}: " + ", ".join(f"{k}={med(np.array(v)):.0f}" for k, v in seg.items()),
              f"(n={len(seg['period'])})")
    e = ev[3]
    for code, name in ((3, "P1Full tile0 seen"), (4, "P1Full tile1 seen"), (5, "tile0 issued"), (6, "tile1 issued")):
        print(f"mma {name}: median period {med(deltas(3, code)):.0f}")
    # MMA: time from P1 seen to issue done
    d = []
    last = None
    for c, t in e:
        if c in (3, 4):
            last = t
        elif c in (5, 6) and last is not None:
            d.append(t - last)
    print(f"mma P1 seen -> issue done: median {med(np.array(d)):.0f}")
    # cross-role latency: softmax P part A published -> MMA sees it
    for tile in (0, 1):
        pub = [t for c, t in ev[tile] if c == 4]
        seen = [t for c, t in ev[3] if c == 3 + tile]
        n = min(len(pub), len(seen))
        if n:
            print(f"tile {tile}: publish A -> mma seen median {med(np.array(seen[:n]) - np.array(pub[:n])):.0f}")
        sready = [t for c, t in ev[tile] if c == 1]
        issued = [t for c, t in ev[3] if c == 5 + tile]
        # S(j+1) issue done at event 5/6 number j -> S ready event number j+1 (event 0 is prologue)
        n = min(len(issued), len(sready) - 1)
        if n > 0:
            print(f"tile {tile}: mma issue done -> S ready at softmax median "
                  f"{med(np.array(sready[1:n + 1]) - np.array(issued[:n])):.0f}")
    tot = max(t for r in ev for _, t in ev[r]) - t0
    print(f"CTA 0 traced span: {tot} cycles")


if __name__ == "__main__":
    main()
