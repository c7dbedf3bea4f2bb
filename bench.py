#!/usr/bin/env python
"""bench.py — forward-attention throughput on B200, BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2]

A step = one pass of the hot path (`flash_attn_func`, i.e. torch op -> C ABI -> sm_100a kernel)
over one batch of synthetic q/k/v.  Default workload = the configuration the metric is quoted on
("bf16 hd128 causal"): BASELINE config 3, Llama-3-8B attention shape, causal GQA 32q/8kv, D=128,
S=8192, with the batch of config 5 (B=8) so that one step's inputs (1.25 GiB) exceed the 126 MB L2.
FLOPs = 4·B·Hq·Sq·Sk·D, halved for causal.  N>1: the SAME workload is sharded — its B·Hkv = 64
(batch, kv-head) units are independent (reference csrc/flash_attention_template.cuh:157-160), so
rank r owns a contiguous run of 64/N units (flash_attention_cute_b200/sharding.py) and the hot
path needs no collective: strong scaling, box TFLOP/s = total FLOPs / slowest rank's time.  NCCL
is used for the barrier, the max-over-ranks time and — after the timed region — one all-gather
of the ranks' own O shards (the box-level O north_star names), whose time is reported.

`--impl reference` times the reference's CPU implementation of the path: its fp32 eager
softmax(QKᵀ)V (scripts/benchmark_kernel.py:15-44, restated in oracle/eager_attention.py because a
Python reference cannot travel to the GPU box) on the host cores, on a bounded sample.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # name: (B, Hq, Hkv, Sq, Sk, D, causal, dtype)
    "c3": dict(B=8, Hq=32, Hkv=8, Sq=8192, Sk=8192, D=128, causal=True, dtype="bf16",
               desc="bf16 causal GQA fwd, Llama-3-8B shape (32q/8kv, D=128), S=8192, B=8"),
    "c2": dict(B=4, Hq=32, Hkv=32, Sq=4096, Sk=4096, D=128, causal=False, dtype="bf16",
               desc="bf16 MHA fwd B=4,H=32,S=4096,D=128 non-causal"),
}
METRIC = "fwd_attn_tflops_bf16_hd128_causal"
UNIT = "TFLOP/s"


def flops_of(w) -> float:
    f = 4.0 * w["B"] * w["Hq"] * w["Sq"] * w["Sk"] * w["D"]
    return f / 2 if w["causal"] else f


def measured_peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return json.loads(p.read_text())
        except Exception:
            pass
    return {}


class ClockSampler:
    """Polls NVML for SM clock and throttle reasons while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.02):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self.index, self.period = index, period_s
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self) -> dict:
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# --------------------------------------------------------------------------------------------
# CPU baseline: the reference's eager attention (restated), one (batch, kv-head) unit at a time
# --------------------------------------------------------------------------------------------
_CPU_INPUTS = {}


def cpu_unit_inputs(w, seed=0):
    import torch
    key = (w["desc"], seed)
    if key in _CPU_INPUTS:
        return _CPU_INPUTS[key]
    g = torch.Generator().manual_seed(seed)
    G = w["Hq"] // w["Hkv"]
    dt = torch.bfloat16 if w["dtype"] == "bf16" else torch.float16
    q = torch.randn(1, G, w["Sq"], w["D"], generator=g).to(dt)
    k = torch.randn(1, 1, w["Sk"], w["D"], generator=g).to(dt)
    v = torch.randn(1, 1, w["Sk"], w["D"], generator=g).to(dt)
    _CPU_INPUTS[key] = (q, k, v)
    return q, k, v


def cpu_time_sample(w, rows: int, repeats: int = 1):
    """Times the oracle on `rows` query rows of one (batch, kv-head) unit (all G heads, all keys).
    Eager attention does not skip masked keys, so its cost is proportional to the rows taken; the
    sample is credited rows/Sq of the unit's algorithmic FLOPs.  Returns (seconds, flops)."""
    import torch
    from oracle.eager_attention import eager_attention  # bench's cpu_baseline leg may use the oracle

    torch.set_num_threads(os.cpu_count() or 1)
    q, k, v = cpu_unit_inputs(w)
    rows = max(1, min(rows, w["Sq"]))
    r0 = (w["Sq"] - rows) // 2
    qs = q[:, :, r0:r0 + rows]
    unit_flops = flops_of(w) / (w["B"] * w["Hkv"])
    with torch.no_grad():
        t0 = time.perf_counter()
        for _ in range(repeats):
            eager_attention(qs, k, v, None, w["causal"])
        dt = time.perf_counter() - t0
    return dt, unit_flops * repeats * rows / w["Sq"]


def run_reference(args, w):
    """--impl reference: the reference's CPU path on this box's host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    budget_s = 150.0
    cpu_time_sample(w, 256)  # thread pool / allocator warm-up, untimed
    probe_t, _ = cpu_time_sample(w, 1024)
    n = args.warmup + args.steps
    rows = int(min(w["Sq"], max(128, 1024 * budget_s / max(1e-6, probe_t * n))))
    rows -= rows % 128
    times, flops = [], 0.0
    for i in range(n):
        dt, f = cpu_time_sample(w, rows)
        if i >= args.warmup:
            times.append(dt)
            flops += f
    total_t = sum(times)
    value = flops / total_t / 1e12
    sample = (f"each step = {rows} of {w['Sq']} query rows of one (batch, kv-head) unit "
              f"(G={w['Hq'] // w['Hkv']} q heads, all {w['Sk']} keys), fp32 eager softmax(QK^T)V, "
              f"torch CPU, {os.cpu_count()} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / max(1, len(times)),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": w["desc"], "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline_leg(w, target_s: float = 12.0) -> dict:
    """Rank 0, N=1 only: ~10-30 s of CPU work on a bounded sample of the same workload."""
    cpu_time_sample(w, 256)
    probe_t, _ = cpu_time_sample(w, 1024)
    rows = int(min(w["Sq"], max(128, 1024 * (target_s / 3) / max(1e-6, probe_t))))
    rows -= rows % 128
    reps = int(max(3, min(200, round(target_s / max(1e-6, probe_t * rows / 1024)))))
    dt, f = cpu_time_sample(w, rows, repeats=reps)
    return {"value": f / dt / 1e12, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
            "sample": (f"{reps} x {rows} of {w['Sq']} query rows of one (batch, kv-head) unit (G={w['Hq'] // w['Hkv']} "
                       f"heads, all keys) = {dt:.1f} s of fp32 eager attention on {os.cpu_count()} host threads")}


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def bind_to_gpu_numa(local: int) -> dict:
    """Pin this rank to the CPUs of its GPU's NUMA node BEFORE any pinned allocation (first touch
    then puts the staging buffers next to the GPU's PCIe root).  Returns what was found."""
    info = {"numa_node": None, "cpus": None}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:   # nvml prints an 8-digit domain, sysfs a 4-digit one
            bus = bus[4:]
        node = int(Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text().strip())
        info["numa_node"] = node
        if node >= 0:
            cpus = set()
            for part in Path(f"/sys/devices/system/node/node{node}/cpulist").read_text().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            allowed = os.sched_getaffinity(0) & cpus
            if allowed:
                os.sched_setaffinity(0, allowed)
                info["cpus"] = len(allowed)
    except Exception as exc:  # a container may hide sysfs / nvml: report, do not fail
        info["error"] = str(exc)[:80]
    info["affinity"] = len(os.sched_getaffinity(0))
    return info


def local_block(w, world: int, rank: int):
    """The rank's (batch, kv-head) units as one rectangular [b0:b1, h0:h1] block of the workload
    (sharding.unit_slices; 64 units over 1/2/4/8 ranks are whole batch entries)."""
    from flash_attention_cute_b200 import sharding

    runs = sharding.unit_slices(w["B"], w["Hkv"], world, rank)
    whole = all(h0 == 0 and h1 == w["Hkv"] for _, h0, h1 in runs)
    if whole:
        return runs[0][0], runs[-1][0] + 1, 0, w["Hkv"]
    if len(runs) == 1:
        b, h0, h1 = runs[0]
        return b, b + 1, h0, h1
    raise SystemExit(f"bench.py: {w['B'] * w['Hkv']} units do not split into one block per rank at N={world}")


def run_ours(args, w):
    import torch
    import torch.distributed as dist

    from flash_attention import flash_attn_func
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import fill_params, load_extension

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this operator has no CPU path")
    numa = bind_to_gpu_numa(local)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    dt = torch.bfloat16 if w["dtype"] == "bf16" else torch.float16
    B, Hq, Hkv, Sq, Sk, D, causal = (w[k] for k in ("B", "Hq", "Hkv", "Sq", "Sk", "D", "causal"))
    G = Hq // Hkv
    b0, b1, h0, h1 = local_block(w, world, rank)
    Bl, Hkvl = b1 - b0, h1 - h0
    units_local = Bl * Hkvl
    g = torch.Generator(device=dev).manual_seed(1000 + 3 + rank)
    q = torch.randn(Bl, Hkvl * G, Sq, D, generator=g, device=dev, dtype=torch.float32).to(dt)
    k = torch.randn(Bl, Hkvl, Sk, D, generator=g, device=dev, dtype=torch.float32).to(dt)
    v = torch.randn(Bl, Hkvl, Sk, D, generator=g, device=dev, dtype=torch.float32).to(dt)
    total_flops = flops_of(w)                                  # the whole job, all ranks
    local_flops = total_flops * units_local / (B * Hkv)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident throughput (`value`) ----
    for _ in range(max(3, args.warmup)):
        o = flash_attn_func(q, k, v, causal=causal)
    _lib.check_device(local)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _lib.launch_count()
    with ClockSampler(local) as clk:
        e0.record()
        for _ in range(args.steps):
            o = flash_attn_func(q, k, v, causal=causal)
        e1.record()
        torch.cuda.synchronize()
    launches = _lib.launch_count() - n0
    ms = e0.elapsed_time(e1)
    _lib.check_device(local)
    ms_max = max_over_ranks(ms)
    value = total_flops * args.steps / (ms_max * 1e-3) / 1e12
    kernel_ms = ms / max(1, launches)          # one kernel per step: average launch duration
    kernel_tflops = local_flops / (kernel_ms * 1e-3) / 1e12
    kernel_name = _lib.load().fa_sm100_kernel_name(
        ctypes.byref(fill_params(q, k, v, o, D ** -0.5, causal))).decode()

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timing ----
    hq, hk, hv = (x.cpu().pin_memory() for x in (q, k, v))
    ho = torch.empty(q.shape, dtype=dt).pin_memory()
    # two sets of device staging buffers: the copies of step k+1 run under the attention and the
    # read-back of step k (a serving pipeline); every copy of every step is inside the timed region
    dq, dk, dv = [torch.empty_like(q), torch.empty_like(q)], [torch.empty_like(k), torch.empty_like(k)], \
        [torch.empty_like(v), torch.empty_like(v)]
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    # chunks of the rank's block that pipeline H2D | attention | D2H: (batch entry, kv-head range)
    # pieces, at least eight per rank whatever N is, so the pipeline's fill and drain stay short
    per = max(1, min(Hkvl, (Bl * Hkvl) // 8))       # kv heads per chunk
    chunks = [(slice(b, b + 1), slice(h * G, min(Hkvl, h + per) * G), slice(h, min(Hkvl, h + per)))
              for b in range(Bl) for h in range(0, Hkvl, per)]
    consumed = [[None] * len(chunks), [None] * len(chunks)]   # attention of (set, chunk) has read its inputs

    def e2e_step(step: int):
        cur = torch.cuda.current_stream(dev)
        j = step & 1
        for c, (bs, qh, kh) in enumerate(chunks):
            with torch.cuda.stream(s_in):
                if consumed[j][c] is not None:
                    s_in.wait_event(consumed[j][c])   # the step before last is done with this staging piece
                dq[j][bs, qh].copy_(hq[bs, qh], non_blocking=True)
                dk[j][bs, kh].copy_(hk[bs, kh], non_blocking=True)
                dv[j][bs, kh].copy_(hv[bs, kh], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(s_in)
            cur.wait_event(ev)
            ob = flash_attn_func(dq[j][bs, qh], dk[j][bs, kh], dv[j][bs, kh], causal=causal)
            ev2 = torch.cuda.Event()
            ev2.record(cur)
            consumed[j][c] = ev2
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev2)
                ho[bs, qh].copy_(ob, non_blocking=True)
                ob.record_stream(s_out)

    def e2e_drain():
        cur = torch.cuda.current_stream(dev)
        cur.wait_stream(s_in)
        cur.wait_stream(s_out)

    e2e_steps = max(2, min(args.steps, 10))
    s_in.wait_stream(torch.cuda.current_stream(dev))
    s_out.wait_stream(torch.cuda.current_stream(dev))
    for i in range(2):
        e2e_step(i)
    e2e_drain()
    barrier()
    e0.record()
    for i in range(e2e_steps):
        e2e_step(i)
    e2e_drain()       # the last read-back is inside the timed region
    e1.record()
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = total_flops * e2e_steps / (e2e_ms * 1e-3) / 1e12
    e2e_ok = torch.equal(ho[0], o[0].cpu())
    h2d_local = sum(x.numel() * x.element_size() for x in (hq, hk, hv))
    d2h_local = ho.numel() * ho.element_size()
    h2d = int(h2d_local * (B * Hkv) / units_local)      # whole job, per step
    d2h = int(d2h_local * (B * Hkv) / units_local)

    # ---- what the host link gives with every rank copying at once (no compute): pins the e2e limiter ----
    def copy_probe(dst, src, stream):
        barrier()
        with torch.cuda.stream(stream):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(3):
                dst.copy_(src, non_blocking=True)
            b.record(stream)
        torch.cuda.synchronize()
        return 3 * src.numel() * src.element_size() / (a.elapsed_time(b) * 1e-3) / 1e9

    h2d_gbs = copy_probe(dq[0], hq, s_in)
    d2h_gbs = copy_probe(ho, dq[0], s_out)
    # both directions at once, as the e2e pipeline drives the link
    barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cur = torch.cuda.current_stream(dev)
    a0.record(cur)
    s_in.wait_stream(cur)
    s_out.wait_stream(cur)
    for _ in range(3):
        with torch.cuda.stream(s_in):
            dq[0].copy_(hq, non_blocking=True)
        with torch.cuda.stream(s_out):
            ho.copy_(o, non_blocking=True)
    cur.wait_stream(s_in)
    cur.wait_stream(s_out)
    a1.record(cur)
    torch.cuda.synchronize()
    bidir_gbs = 3 * (hq.numel() + ho.numel()) * hq.element_size() / (a0.elapsed_time(a1) * 1e-3) / 1e9
    link = torch.tensor([h2d_gbs, d2h_gbs, bidir_gbs], device=dev, dtype=torch.float64)
    link_min = link.clone()
    if world > 1:
        dist.all_reduce(link, op=dist.ReduceOp.SUM)
        dist.all_reduce(link_min, op=dist.ReduceOp.MIN)

    # ---- box-level O: one NCCL all-gather of the ranks' OWN shards after the timed region ----
    allgather_ms = None
    if world > 1:
        gathered = torch.empty((world,) + tuple(o.shape), dtype=o.dtype, device=dev)  # = the full [B, Hq, S, D]
        oc = o.contiguous()
        dist.all_gather_into_tensor(gathered, oc)   # warm-up (communicator set-up)
        barrier()
        e0.record()
        dist.all_gather_into_tensor(gathered, oc)
        e1.record()
        torch.cuda.synchronize()
        allgather_ms = max_over_ranks(e0.elapsed_time(e1))

    # ---- ablation SURVEY 8(d) asks for: GQA q-head packing on / off on the same inputs (rank 0, N = 1) ----
    pack = None
    if world == 1 and G > 1:
        ext = load_extension()
        res = {}
        for name, fl in (("off", _lib.FLAG_NO_PACK_GQA), ("on", _lib.FLAG_FORCE_PACK_GQA)):
            for _ in range(2):
                ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, fl)
        for rnd in range(3):  # interleaved, so clock drift hits both alike
            for name, fl in (("off", _lib.FLAG_NO_PACK_GQA), ("on", _lib.FLAG_FORCE_PACK_GQA)):
                torch.cuda.synchronize()
                e0.record()
                for _ in range(5):
                    ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, fl)
                e1.record()
                torch.cuda.synchronize()
                res.setdefault(name, []).append(total_flops * 5 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
        plan = ext.plan(q, k, v, causal)
        pack = {"off_tflops": sorted(res["off"])[1], "on_tflops": sorted(res["on"])[1],
                "default": "on" if plan["pack_g"] > 1 else "off",
                "note": "median of 3 interleaved rounds x 5 launches; on = 4 q heads x 32 positions per 128-row M tile"}
        _lib.check_device(local)

    if rank == 0:
        peaks = measured_peaks()
        region_s = ms * 1e-3
        if peaks.get("bf16_tflops"):
            use_sustained = region_s > 1.0 and peaks.get("bf16_tflops_sustained")
            peak = peaks["bf16_tflops_sustained"] if use_sustained else peaks["bf16_tflops"]
            peak_src = "MEASURED_PEAKS.json bf16_tflops" + ("_sustained" if use_sustained else " (burst)")
        else:
            peak, peak_src = 1590.0, "fallback (B200_PROFILING.md)"
        traffic, traffic_src = None, None
        tfile = ROOT / "profiles" / "traffic.json"
        if tfile.exists() and world == 1:
            try:
                tj = json.loads(tfile.read_text())
                traffic, traffic_src = tj.get(args.workload), tj.get("source")
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": w["dtype"], "data": "synthetic",
            "config": {"workload": w["desc"], "flops_per_step": total_flops,
                       "l2": f"per-rank q+k+v+o = {(h2d_local + d2h_local) / 2**20:.0f} MiB per step, above the 126 MB L2; no flush",
                       "parallelism": (f"(batch,kv-head) units: {B * Hkv} in the job, {units_local} per GPU "
                                       f"(rank 0 owns batch {b0}:{b1}, kv heads {h0}:{h1}); no data-path collective")},
            "pct_of_nominal_2250": 100.0 * (value / world) / 2250.0,
            "roofline": {"bound": "tensor", "achieved": kernel_tflops, "peak": peak, "unit": "TFLOP/s",
                         "frac": kernel_tflops / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "kernel": kernel_name, "kernel_ms": kernel_ms,
                         "flops_per_launch": local_flops},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "result_matches_device_run": bool(e2e_ok),
                    "host_link": {"h2d_gbs_sum": float(link[0]), "d2h_gbs_sum": float(link[1]),
                                  "h2d_gbs_min_rank": float(link_min[0]), "d2h_gbs_min_rank": float(link_min[1]),
                                  "both_directions_gbs_sum": float(link[2]), "both_directions_gbs_min_rank": float(link_min[2]),
                                  "e2e_bytes_over_time_gbs": (h2d + d2h) / (e2e_ms / e2e_steps * 1e-3) / 1e9,
                                  "note": "pinned-memory copies alone (no compute), all ranks at once: each direction by itself, then both together"},
                    "chunks_per_rank": len(chunks),
                    "numa": numa},
            "gpu_launches": int(launches),
            "clocks": clk.summary(),
        }
        if allgather_ms is not None:
            line["allgather_o_ms"] = allgather_ms
            line["allgather_o_bytes_per_rank"] = int(o.numel() * o.element_size())
        if pack is not None:
            line["gqa_pack"] = pack
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg(w)
        if world == 1 and not args.no_ref_kernel:
            line["reference_sm80_kernel_on_b200"] = time_reference_kernel(w, q, k, v, causal, total_flops)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def time_reference_kernel(w, q, k, v, causal, step_flops):
    """The reference's own sm80 CuTe kernel, compiled unmodified for sm_100a (oracle/_ref) — the
    comparator on this GPU, timed outside the timed region of our arm (checker code: bench only)."""
    import torch
    try:
        from oracle import ref_kernel
        if not ref_kernel.available():
            return {"unavailable": "oracle/_ref/libref_fa.so not built"}
        for _ in range(2):
            ref_kernel.ref_flash_attn_func(q, k, v, None, causal)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 5
        for _ in range(n):
            ref_kernel.ref_flash_attn_func(q, k, v, None, causal)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        return {"ms_per_step": ms, "value": step_flops / (ms * 1e-3) / 1e12, "unit": UNIT,
                "what": "reference csrc/flash_attention_impl.cu compiled unmodified for sm_100a (128x64 tile branch, mma.sync)"}
    except Exception as exc:
        return {"unavailable": str(exc)[:120]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-kernel", action="store_true", help="skip timing the reference's sm80 kernel on this GPU")
    ap.add_argument("--ref-kernel", action="store_true", help=argparse.SUPPRESS)  # r01 flag, now the default
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference(args, w)
    return run_ours(args, w)


if __name__ == "__main__":
    sys.exit(main())
