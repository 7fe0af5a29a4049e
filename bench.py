#!/usr/bin/env python
"""bench.py — the PS aggregate+apply path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step = one pass of the hot path over one batch of synthetic gradients:
  N = 1  BASELINE configs[1]: single-PS fused aggregate + SGD apply, 8 workers x
         100 M fp32 parameters, inputs resident in HBM (`value`), and the same step
         through the host-buffer C-ABI call with H2D/D2H inside the timed region
         (`e2e`).
  N > 1  the same per-GPU load (8 local workers x 100 M fp32 per GPU), variables
         sharded over N PS tasks = N GPUs: push (local 8-way sum + reduce-scatter),
         shard apply, pull (all-gather). Weak scaling: per-GPU gradient bytes fixed.
Besides the headline the line carries `configs`: the other BASELINE configs as
kernel-level (N = 1) or whole-step (N > 1) measurements, and at N > 1 `parity`: one
step re-run from known inputs and compared with the CPU restatement on every rank.
`--impl reference` times the reference's CPU PS path (the threaded C restatement in
oracle/, TF-1.4 is not installable here) on the host cores, same metric, unit and
config. One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "aggregated+applied gradient GB/s"
UNIT = "GB/s"
WORKERS = 8
ELEMS = 100_000_000
LR = 0.01
FALLBACK_HBM_GBS = 6650.0   # /opt/skills/guides/B200_PROFILING.md fallback
NVLINK_MEASURED = 770.0     # GB/s per direction per GPU, measured peer copy (B200_PROFILING.md)
NVLINK_NOMINAL = 900.0
L2_BYTES = 126 << 20
RESNET50_ELEMS = 25_557_032     # BASELINE configs[2]: 161 tensors
BERT_LARGE_ELEMS = 335_141_888  # BASELINE configs[3]: 391 tensors
# How the headline multi-GPU step runs per world size (measured on B200, profiles/r02_*):
# (mode, slices). "fused" = unicast peer stores, bit-exact; "switch" = in-switch reduction +
# multicast pull (rel 1e-5).
STEP_SHAPE = {2: ("fused", 0), 4: ("fused", 0), 8: ("fused", 0)}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("dram_bytes_per_launch")
    return None


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6
                          for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def workload_config(n_gpus: int):
    """The workload, identically worded for this arm and the reference arm."""
    cfg = {"workload": ("single-PS fused aggregate+SGD apply, 8 workers x 100M fp32 params on "
                        "1 B200 (BASELINE configs[1])"),
           "workers_per_gpu": WORKERS, "elems": ELEMS, "optimizer": "sgd", "aggregate": "sum",
           "lr": LR, "l2": "inputs (3.6 GB) exceed L2 (126 MB); no flush needed"}
    if n_gpus > 1:
        cfg["workload"] = (f"{n_gpus} PS tasks (one per B200), each aggregating 8 local workers x "
                           f"100M fp32 and applying SGD to its shard: push (8-way sum + "
                           f"reduce-scatter) -> shard apply -> pull (all-gather); "
                           f"{n_gpus} x 3.2 GB of gradient per step")
        cfg["ps_shards"] = n_gpus
    return cfg


# ---- the reference arm / cpu baseline: the CPU PS path on the host cores --------
def cpu_ps_rate(sample_elems: int, steps: int, warmup: int, n_tasks: int = 1):
    """Gradient GB/s of the CPU restatement on a bounded sample of the workload.
    n_tasks > 1: that many PS tasks run back to back per step (the host's cores are shared by
    all PS tasks of the box), each with its own variables; the 8 gradient buffers are re-used."""
    import numpy as np
    from oracle import ps_oracle as O
    cores = os.cpu_count() or 1
    orc = O.COracle(threads=cores, native=True)
    rng = np.random.default_rng(1234)
    block = 1 << 20   # a random 4 MiB block tiled to size: init cost stays off the clock

    def tiled(scale):
        b = rng.standard_normal(block, dtype=np.float32) * np.float32(scale)
        return np.resize(b, sample_elems)
    var = [tiled(1.0) for _ in range(n_tasks)]
    grads = [tiled(0.01) for _ in range(WORKERS)]
    acc = np.empty(sample_elems, dtype=np.float32)
    hp = O.HParams(lr=LR)

    def step():
        for t in range(n_tasks):
            orc.ps_step(O.SGD, var[t], None, None, grads, hp, acc=acc)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    gbs = n_tasks * WORKERS * sample_elems * 4 / dt / 1e9
    # host-memory traffic of the statement-per-pass form: acc = g0 (8 B), 7 x acc += g (12 B),
    # var -= acc * lr (12 B) = 104 B per element
    mem_gbs = n_tasks * sample_elems * 104 / dt / 1e9
    frac = "the whole workload" if sample_elems == ELEMS else \
        f"1/{ELEMS // sample_elems} of the workload"
    tasks = "" if n_tasks == 1 else f"{n_tasks} PS tasks back to back per step, each "
    sample = (f"{tasks}{WORKERS} workers x {sample_elems} fp32 ({frac}; working set "
              f"{(WORKERS + 2) * sample_elems * 4 / 1e9:.1f} GB, beyond any host cache) x {steps} "
              f"steps, C restatement of TF-1.4 ConditionalAccumulator + ApplyGradientDescent, "
              f"one pass per statement, persistent pool of {cores} threads, -march={orc.build}, "
              f"host memory traffic {mem_gbs:.0f} GB/s (104 B/elem)")
    return gbs, dt, cores, sample, mem_gbs


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = max(args.gpus, 1)
    steps = max(args.steps, 1)
    # full-size steps: ~0.06 s per PS task on the box's host cores
    gbs, dt, cores, desc, mem_gbs = cpu_ps_rate(ELEMS, steps, min(max(args.warmup, 1), 3), n)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(n),
        "cpu_baseline": {"value": round(gbs, 3), "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": desc, "host_mem_gbs": round(mem_gbs, 1)},
        "e2e": {"value": round(gbs, 3), "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "steps_per_s": round(1.0 / dt, 3),
    }
    print(json.dumps(line), flush=True)


# ---- N = 1: kernel-level table of the other BASELINE configs ------------------------
def kernel_case(torch, pkg, ctx, label, opt, bf16, workers, n, flush, peak, iters=None):
    """One fused aggregate+apply launch shape, each launch timed on its own with CUDA events;
    working sets under 2x L2 are measured cold (a 512 MiB memset runs between launches)."""
    dev = "cuda:0"
    var = torch.randn(n, device=dev)
    s0 = torch.zeros(n, device=dev) if opt >= 1 else None
    s1 = torch.full((n,), 1e-4, device=dev) if opt == 2 else None
    grads = []
    for _ in range(workers):
        g = torch.randn(n, device=dev).mul_(0.01)
        grads.append(g.to(torch.bfloat16) if bf16 else g)
        del g
    hp = pkg.HParams(lr=0.01, momentum=0.9)
    alg = n * (workers * (2 if bf16 else 4) + 8 * (1 + opt))
    cold = alg < 2 * L2_BYTES
    iters = iters or (7 if alg > (2 << 30) else 15)
    st = torch.cuda.current_stream()
    for _ in range(3):
        ctx.apply(opt, var, s0, s1, grads, hp, stream=st)
    times = []
    for _ in range(iters):
        if cold:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        ctx.apply(opt, var, s0, s1, grads, hp, stream=st)
        b.record(st)
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    times.sort()
    med = times[len(times) // 2]
    gbs = alg / med / 1e6
    row = {"workload": label, "opt": ["sgd", "momentum", "adam"][opt],
           "grad": "bf16" if bf16 else "f32", "workers": workers, "elems": n,
           "ms": round(med, 5), "ms_min": round(times[0], 5), "algorithmic_bytes": alg,
           "hbm_gbs": round(gbs, 1), "frac": round(gbs / peak, 4), "l2": "flushed" if cold else
           "exceeded", "launches": iters + 3}
    if alg < (64 << 20):
        row["bound"] = "latency (launch + one wave; a bandwidth fraction is not meaningful)"
    del var, s0, s1, grads
    torch.cuda.empty_cache()
    return row


def single_gpu_configs(torch, pkg, ctx, peak):
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")
    rows, launches = [], 0

    def add(*a, **k):
        nonlocal launches
        try:
            r = kernel_case(torch, pkg, ctx, *a, flush=flush, peak=peak, **k)
            launches += r["launches"]
        except Exception as e:   # noqa: BLE001 — one failing case must not lose the line
            r = {"workload": a[0], "error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()
        rows.append(r)
    # config 2's Momentum / Adam siblings and the bf16-gradient forms (SURVEY §8 d)
    for opt in (1, 2):
        add("cfg2 sibling: 8 workers x 100M fp32", opt, False, 8, ELEMS)
    for opt in (0, 1, 2):
        add("cfg2 sibling: 8 workers x 100M bf16 gradients", opt, True, 8, ELEMS)
    # post-collective shard applies of configs 3 and 4
    for p in (2, 4, 8):
        add(f"cfg3 shard apply (ResNet-50 / {p} PS)", 1, False, 1, -(-RESNET50_ELEMS // (8 * p)) * 8)
    add("cfg4 shard apply (BERT-large / 8 PS, bf16 -> Adam)", 2, True, 1, 41_892_736)
    # config 5 on one GPU: fp32 variable sizes 1 MiB - 4 GiB, one gradient
    for mib in (1, 4, 16, 64, 256, 1024, 4096):
        for opt in (0, 2):
            add(f"cfg5 {mib} MiB of fp32 variables", opt, False, 1, mib << 18)
    del flush
    torch.cuda.empty_cache()
    return rows, launches


# ---- N = 1 -------------------------------------------------------------------------
def run_single(args):
    import torch

    import kubeflow_controller_b200 as pkg
    pkg.load_library()     # raises when the CUDA extension is missing: no fallback
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; the product has no CPU path")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ctx = pkg.Context(1)
    if args.kernel:
        ctx.set_kernel(getattr(pkg, "KERNEL_" + args.kernel.upper()), args.threads, args.ctas,
                       args.unroll)
    gen = torch.Generator(device=dev)
    gen.manual_seed(42)
    var = torch.randn(ELEMS, device=dev, generator=gen)
    grads = []
    for w in range(WORKERS):
        gen.manual_seed(1234 + w)
        grads.append(torch.randn(ELEMS, device=dev, generator=gen) * 0.01)
    hp = pkg.HParams(lr=LR)
    stream = torch.cuda.current_stream()

    for _ in range(args.warmup):
        ctx.apply_sgd(var, grads, hp, stream=stream)
    torch.cuda.synchronize()
    clocks = ClockSampler(0)
    clocks.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    torch.cuda.synchronize()
    ev[0].record(stream)
    for i in range(args.steps):
        ctx.apply_sgd(var, grads, hp, stream=stream)
        ev[i + 1].record(stream)
    torch.cuda.synchronize()
    per_launch = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    total_ms = ev[0].elapsed_time(ev[-1])
    clk = clocks.stop()

    ms = total_ms / args.steps
    grad_bytes = WORKERS * ELEMS * 4
    alg_bytes = ELEMS * (WORKERS * 4 + 8)
    value = grad_bytes / (ms * 1e-3) / 1e9
    peak, peak_src = peaks()
    avg_launch_ms = sum(per_launch) / len(per_launch)
    achieved = alg_bytes / (avg_launch_ms * 1e-3) / 1e9

    # ---- e2e: the reference-facing call with HOST buffers -------------------------
    e2e = cpu = None
    if not args.no_e2e:
        e2e = measure_e2e(args, pkg, ctx, var, grads, hp, grad_bytes)
    del var, grads
    torch.cuda.empty_cache()
    configs, cfg_launches = ([], 0) if args.no_configs else single_gpu_configs(torch, pkg, ctx, peak)
    if not args.no_cpu:
        cpu_gbs, cpu_dt, cores, desc, mem_gbs = cpu_ps_rate(ELEMS, 5, 2)
        cpu = {"value": round(cpu_gbs, 3), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": desc, "host_mem_gbs": round(mem_gbs, 1)}
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(1),
        "steps_per_s": round(1e3 / ms, 2),
        "hbm_gbs": round(alg_bytes / (ms * 1e-3) / 1e9, 2),
        "roofline": {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": ncu_traffic(),
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                     "kernel": "ps_apply (fused 8-way sum + SGD)",
                     "launch_ms_min": round(min(per_launch), 4),
                     "launch_ms_avg": round(avg_launch_ms, 4)},
        "cpu_baseline": cpu,
        "e2e": e2e,
        "gpu_launches": args.steps,
        "clocks": clk,
        "configs": configs,
        "configs_note": ("kernel-level rows outside the timed region of `value`: one fused "
                         "aggregate+apply launch each, CUDA events per launch, median; frac = "
                         "algorithmic HBM bytes / time / measured HBM peak; "
                         f"{cfg_launches} more launches of the library's kernels"),
    }
    print(json.dumps(line), flush=True)


def measure_e2e(args, pkg, ctx, var, grads, hp, grad_bytes):
    import torch
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    h_grads = [torch.empty(ELEMS, dtype=torch.float32).pin_memory() for _ in range(WORKERS)]
    for h, g in zip(h_grads, grads):
        h.copy_(g)
    h_out = torch.empty(ELEMS, dtype=torch.float32).pin_memory()
    for _ in range(min(args.warmup, 3)):
        ctx.ps_step_host(pkg.OPT_SGD, var, None, None, h_grads, hp, h_out)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.ps_step_host(pkg.OPT_SGD, var, None, None, h_grads, hp, h_out)
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_value = grad_bytes / e2e_s / 1e9
    return {"value": round(e2e_value, 3), "unit": UNIT, "h2d_bytes_per_step": grad_bytes,
            "d2h_bytes_per_step": ELEMS * 4, "ms_per_step": round(e2e_s * 1e3, 3),
            "steps": e2e_steps,
            "path": "accel_ps_step_host: pinned host gradients -> chunked H2D -> fused "
                    "apply -> D2H of updated variables",
            "bound": "PCIe: 3.2 GB in per step over one Gen5 x16 link (~55.6 GB/s measured)"}


# ---- N > 1 ---------------------------------------------------------------------------
def timed_steps(torch, dist, fn, steps, warmup, stream, dev):
    """CUDA events around `steps` calls after `warmup`, barrier + sync both sides, max over ranks."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def parity_check(torch, dist, pkg, sps, grads, world, rank, dev, exact):
    """One step re-run from known inputs on the LIVE buffers of the timed run, compared on every
    rank with the CPU restatement (hierarchical order: worker order inside a rank, then rank
    order) over two windows of the replica — the first 512 Ki elements and 512 Ki elements across
    the boundary of the last two shards. The checker is test infrastructure (oracle/); nothing
    timed touches it."""
    import numpy as np
    from oracle import ps_oracle as O
    half = 1 << 19
    lo2 = max(half, (world - 1) * sps.shard - half // 2)
    lo2 -= lo2 % 8
    windows = [(0, half), (min(lo2, sps.padded - half), half)]
    idx = torch.cat([torch.arange(a, a + n, device=dev) for a, n in windows])
    before = sps.var[idx].cpu().numpy()
    mine = torch.stack([g[idx].float() for g in grads])            # [W, elems]
    allg = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allg, mine)
    sps.step(grads, torch.cuda.current_stream())
    if sps.mode != "nccl":
        sps.check(torch.cuda.current_stream())
    torch.cuda.synchronize()
    dist.barrier()
    got = sps.var[idx].cpu().numpy()
    partials = [O.numpy_aggregate([w.cpu().numpy() for w in r], False) for r in allg]
    want = before.copy()
    O.numpy_apply(O.SGD, want, None, None, O.numpy_aggregate(partials, False), O.HParams(lr=LR))
    if exact:
        bad = int((got != want).sum())
        tol = "bit-exact"
    else:
        bad = int((np.abs(got - want) > 1e-5 * np.maximum(np.abs(want), 1e-3)).sum())
        tol = "rel 1e-5 (abs floor 1e-8)"
    t = torch.tensor([bad], device=dev, dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {"checked": True, "mismatch": int(t.item()), "elems": int(idx.numel()),
            "ranks": world, "tolerance": tol,
            "against": "oracle/ps_oracle.py, per-rank worker order then rank order"}


def sharded_case(torch, dist, pkg, ShardedPS, rank, world, local, dev, stream, label, elems, opt,
                 bf16, hpkw, mode, slices, pull="f32", steps=20, warmup=4, in_place=False):
    # every rank must take the same decision (a rank that fails alone would leave the others
    # waiting in the rendezvous): agree on the smallest free memory first
    padded_est = -(-elems // (8 * world)) * 8 * world
    need = padded_est * (4 + 4 + (2 if bf16 else 4) + (4 if bf16 else 0) + (2 if pull == "bf16" else 0)) \
        + padded_est // world * 4 * opt
    gc.collect()
    torch.cuda.empty_cache()
    dist.barrier()      # every rank has dropped (and unmapped) the previous case's buffers
    free = torch.tensor([torch.cuda.mem_get_info()[0]], device=dev, dtype=torch.int64)
    dist.all_reduce(free, op=dist.ReduceOp.MIN)
    if need * 1.08 > free.item():
        return {"workload": label, "ps_shards": world, "mode": mode,
                "skipped": f"needs {need / 2**30:.1f} GiB per GPU, {free.item() / 2**30:.1f} GiB free"}
    sps = ShardedPS(rank, world, local, elems, opt, pkg.HParams(**hpkw), mode=mode,
                    n_chunks=slices or None, pull_dtype=pull)
    g = torch.randn(sps.padded, device=dev).mul_(0.01)
    if in_place:     # zero-copy in-switch form: the worker's gradient lives in the stage buffer
        view = sps.stage_grad(torch.bfloat16 if bf16 else torch.float32)
        view.copy_(g)
        grads = [view]
    else:
        grads = [g.to(torch.bfloat16) if bf16 else g]
    ms = timed_steps(torch, dist, lambda: sps.step(grads, stream), steps, warmup, stream, dev)
    if mode != "nccl":
        sps.check(stream)
    shard = sps.shard
    gb, pb = (2 if bf16 else 4), (2 if pull == "bf16" else 4)
    if mode == "switch":      # ld_reduce: arena out, shard in; multimem.st: shard out, arena in
        link = shard * world * max(gb, pb) + shard * min(gb, pb)
    elif mode == "nccl":
        link = (world - 1) * shard * (4 + pb)
    else:
        link = (world - 1) * shard * (gb + pb)
    row = {"workload": label, "ps_shards": world, "mode": mode + (" in place" if in_place else ""),
           "slices": sps.n_chunks,
           "pull": pull, "opt": ["sgd", "momentum", "adam"][opt], "grad": "bf16" if bf16 else "f32",
           "elems": elems, "shard_mib": round(shard * 4 / 2**20, 2), "ms": round(ms, 4),
           "steps_per_s": round(1e3 / ms, 1), "nvlink_bytes_per_gpu_per_dir": link,
           "nvlink_gbs_per_dir": round(link / ms / 1e6, 1),
           "frac_of_770_measured": round(link / ms / 1e6 / NVLINK_MEASURED, 3),
           "frac_of_900_nominal": round(link / ms / 1e6 / NVLINK_NOMINAL, 3),
           "launches": (steps + warmup) * sps.launches_per_step}
    if shard * 4 <= (16 << 20):
        row["bound"] = "latency (cross-GPU signalling dominates; not a bandwidth measurement)"
    sps.close()
    del sps, g, grads
    gc.collect()
    torch.cuda.empty_cache()
    return row


def sharded_configs(torch, dist, pkg, ShardedPS, rank, world, local, dev, stream, big):
    rows = []

    def add(*a, **k):
        try:
            rows.append(sharded_case(torch, dist, pkg, ShardedPS, rank, world, local, dev, stream,
                                     *a, **k))
        except Exception as e:   # noqa: BLE001 — one failing case must not lose the line
            rows.append({"workload": a[0], "mode": a[5], "error": f"{type(e).__name__}: {e}"[:300]})
            gc.collect()
            torch.cuda.empty_cache()
    modes = ["fused", "switch", "nccl"]
    # config 3: ResNet-50-shaped, Momentum, one gradient per rank, PS sharded over the ranks
    for mode in modes:
        add(f"cfg3 ResNet-50-shaped (25.6M fp32) Momentum, {world} PS shards", RESNET50_ELEMS,
            pkg.OPT_MOMENTUM, False, dict(lr=0.1, momentum=0.9), mode, 1 if mode != "switch" else 2)
    add(f"cfg3 ResNet-50-shaped (25.6M fp32) Momentum, {world} PS shards", RESNET50_ELEMS,
        pkg.OPT_MOMENTUM, False, dict(lr=0.1, momentum=0.9), "switch", 2, in_place=True)
    # the same model placed the reference's way (variable k -> PS task k mod P, buckets padded to
    # equal length): same kernels, a larger and unbalanced arena
    try:
        import numpy as np
        from kubeflow_controller_b200.sharded import BucketLayout
        with open(os.path.join(ROOT, "tests", "golden", "model_shapes.json")) as f:
            sizes = [int(np.prod(sh)) if sh else 1 for _, sh in json.load(f)["resnet50"]]
        lay = BucketLayout(sizes, world)
        add(f"cfg3 ResNet-50-shaped, round-robin-by-variable buckets (arena {lay.arena} = "
            f"{lay.arena / sum(sizes):.3f} x flat), Momentum, {world} PS shards", lay.arena,
            pkg.OPT_MOMENTUM, False, dict(lr=0.1, momentum=0.9), "fused", 1)
    except Exception as e:   # noqa: BLE001
        rows.append({"workload": "cfg3 round-robin buckets", "error": f"{type(e).__name__}: {e}"[:300]})
    # config 4: BERT-large-shaped, bf16 gradients -> fp32 Adam (named for 8 shards; run at any N)
    for mode, pull in (("fused", "f32"), ("switch", "f32"), ("fused", "bf16"), ("nccl", "f32")):
        add(f"cfg4 BERT-large-shaped (335M) bf16 grads -> fp32 Adam, {world} PS shards",
            BERT_LARGE_ELEMS, pkg.OPT_ADAM, True, dict(lr=1e-4), mode,
            4 if mode == "switch" else 2, pull=pull, steps=10, warmup=3)
    add(f"cfg4 BERT-large-shaped (335M) bf16 grads -> fp32 Adam, {world} PS shards",
        BERT_LARGE_ELEMS, pkg.OPT_ADAM, True, dict(lr=1e-4), "switch", 4, steps=10, warmup=3,
        in_place=True)
    # config 5: fp32 variable shards of 1 MiB - 4 GiB per GPU, one gradient per rank
    sizes = (1, 4, 16, 64, 256, 1024) + ((4096,) if big else ())
    for mib in sizes:
        for opt, hpkw in ((pkg.OPT_SGD, dict(lr=0.01)), (pkg.OPT_ADAM, dict(lr=0.01))):
            for mode in ("fused", "switch"):
                if mib >= 4096 and mode == "switch":
                    continue     # symmetric-memory allocations of 32 GiB per buffer: not attempted
                n = 10 if mib >= 1024 else 20
                add(f"cfg5 {mib} MiB fp32 shard per GPU, {world} PS shards", (mib << 18) * world, opt,
                    False, hpkw, mode, 1 if mode == "fused" else (1 if mib <= 16 else 4),
                    steps=n if mib < 4096 else 4, warmup=3)
            if mib < 4096 and opt == pkg.OPT_SGD:
                add(f"cfg5 {mib} MiB fp32 shard per GPU, {world} PS shards", (mib << 18) * world, opt,
                    False, hpkw, "switch", 1 if mib <= 16 else 4, steps=n, warmup=3, in_place=True)
    return rows


def run_sharded(args):
    import torch
    import torch.distributed as dist

    import kubeflow_controller_b200 as pkg
    from kubeflow_controller_b200.sharded import ShardedPS
    pkg.load_library()
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    if world != args.gpus:
        raise RuntimeError(f"--gpus {args.gpus} but WORLD_SIZE={world}; launch with torchrun")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    mode, slices = STEP_SHAPE.get(world, ("fused", 0))
    if args.mode:
        mode, slices = args.mode, 0
    if args.chunks:
        slices = args.chunks
    sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=LR), mode=mode,
                    n_chunks=slices or None, rendezvous=args.rendezvous,
                    multicast=args.multicast, pull_dtype=args.pull,
                    overlap=None if args.overlap < 0 else args.overlap)
    if args.reducers or args.no_coop:
        sps.ctx.set_step_shape(args.reducers, not args.no_coop)
    gen = torch.Generator(device=dev)
    gen.manual_seed(42)
    sps.var.copy_(torch.randn(sps.padded, device=dev, generator=gen))
    grads = []
    for w in range(WORKERS):
        gen.manual_seed(1234 + rank * WORKERS + w)
        grads.append(torch.randn(sps.padded, device=dev, generator=gen) * 0.01)
    stream = torch.cuda.current_stream()
    for _ in range(args.warmup):
        sps.step(grads, stream)
    if sps.mode != "nccl":
        sps.check(stream)      # a rank that never arrived surfaces here, not as a hang
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        sps.step(grads, stream)
    e1.record(stream)
    torch.cuda.synchronize()
    dist.barrier()
    ms_t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms = ms_t.item() / args.steps
    clk = clocks.stop() if rank == 0 else None

    # ---- parity on every rank, on the buffers the timed loop just used -----------------
    parity = parity_check(torch, dist, pkg, sps, grads, world, rank, dev,
                          exact=sps.mode in ("fused", "streams"))

    # ---- e2e: every rank's workers deliver their gradients in pinned HOST memory; the C-ABI
    # host-buffer step (accel_shard_step_host) copies slice k+1 in over the rank's own PCIe link
    # while slice k runs push/apply/pull over NVLink and slice k-1's part of the replica goes
    # back out. Wall clock, barrier on both sides, max over ranks.
    e2e = None
    if not args.no_e2e:
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        h_grads = [torch.empty(sps.padded, dtype=torch.float32).pin_memory() for _ in range(WORKERS)]
        for h, g in zip(h_grads, grads):
            h.copy_(g)
        h_out = torch.empty(sps.padded, dtype=torch.float32).pin_memory()
        for _ in range(3):
            sps.step_host(h_grads, h_out, n_slices=args.host_slices)
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            sps.step_host(h_grads, h_out, n_slices=args.host_slices)
        dist.barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / e2e_steps], device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        ok = torch.tensor([int(torch.equal(h_out.to(dev), sps.var))], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        e2e = {"value": round(world * WORKERS * ELEMS * 4 / dt.item() / 1e9, 3), "unit": UNIT,
               "h2d_bytes_per_step": world * WORKERS * sps.padded * 4,
               "d2h_bytes_per_step": world * sps.padded * 4,
               "ms_per_step": round(dt.item() * 1e3, 3), "steps": e2e_steps,
               "host_copy_equals_replica": bool(ok.item()),
               "path": f"accel_shard_step_host per rank, {args.host_slices} slices: 8 pinned host "
                       "gradients -> H2D over the rank's own PCIe link || push + apply + pull over "
                       "NVLink || D2H of the updated replica",
               "bound": "PCIe: 3.2 GB in per rank per step (~55.6 GB/s per Gen5 x16 link), then "
                        "the host's memory system at higher N"}
        del h_grads, h_out
    n_chunks, overlap, mc = sps.n_chunks, sps.overlap, bool(sps.multicast_ptr)
    launches_per_step, padded = sps.launches_per_step, sps.padded
    sps.close()
    del sps, grads
    gc.collect()
    torch.cuda.empty_cache()
    configs = [] if args.no_configs else sharded_configs(
        torch, dist, pkg, ShardedPS, rank, world, local, dev, stream, big=not args.no_big)
    if rank == 0:
        grad_bytes = world * WORKERS * ELEMS * 4
        peak, peak_src = peaks()
        shard = padded // world
        # per GPU: read 8 local gradients + write partial (push), read `world` slots + var r/w
        # (apply), write the full replica (pull)
        hbm_bytes = padded * (WORKERS * 4 + 4) + shard * (world * 4 + 8) + padded * 4
        pull_b = 2 if args.pull == "bf16" else 4
        if mode == "switch":
            nvl_bytes = padded * 4 + shard * 4      # (1 + 1/P) of the arena per direction
        else:
            nvl_bytes = (world - 1) * shard * (4 + pull_b)   # push out + pull out, per direction
        line = {
            "metric": METRIC, "value": round(grad_bytes / (ms * 1e-3) / 1e9, 2), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(world),
            "variant": {"mode": mode, "slices": n_chunks, "rendezvous": args.rendezvous
                        if mode != "switch" else "symm", "multicast_pull": mc, "pull": args.pull,
                        "overlap": ["inline", "roles", "switch"][overlap]},
            "steps_per_s": round(1e3 / ms, 2),
            "roofline": {"bound": "hbm", "achieved": round(hbm_bytes / (ms * 1e-3) / 1e9, 2),
                         "peak": peak, "unit": "GB/s",
                         "frac": round(hbm_bytes / (ms * 1e-3) / 1e9 / peak, 4), "traffic": None,
                         "peak_source": peak_src,
                         "note": "per-GPU HBM bytes of the whole step / step time; at N > 1 the step "
                                 "is NVLink-bound as well, see `nvlink`"},
            "nvlink": {"bytes_per_gpu_per_dir_per_step": nvl_bytes,
                       "achieved": round(nvl_bytes / (ms * 1e-3) / 1e9, 2),
                       "peak": NVLINK_MEASURED, "unit": "GB/s per direction per GPU",
                       "frac": round(nvl_bytes / (ms * 1e-3) / 1e9 / NVLINK_MEASURED, 4),
                       "frac_of_nominal_900": round(nvl_bytes / (ms * 1e-3) / 1e9 / NVLINK_NOMINAL, 4),
                       "peak_source": "measured peer copy, B200_PROFILING.md",
                       "note": "unicast: push sends (P-1)/P of the arena and the pull another "
                               "(P-1)/P per direction; switch: (1 + 1/P) of the arena per direction"},
            "parity": parity,
            "cpu_baseline": None,
            "e2e": e2e,
            "gpu_launches": args.steps * launches_per_step,
            "clocks": clk,
            "configs": configs,
            "configs_note": "whole sharded steps of the other BASELINE configs on these N GPUs, one "
                            "worker gradient per rank, CUDA events, max over ranks; outside the "
                            "timed region of `value`",
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if parity["mismatch"]:
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (sweeps)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (sweeps)")
    ap.add_argument("--no-configs", action="store_true",
                    help="skip the `configs` rows (profiling runs, sweeps)")
    ap.add_argument("--no-big", action="store_true", help="skip the 4 GiB-per-shard config-5 rows")
    ap.add_argument("--mode", default="", choices=["", "fused", "streams", "nccl", "switch"],
                    help="multi-GPU step form (default: the measured-best per world size)")
    ap.add_argument("--chunks", type=int, default=0,
                    help="pipeline slices of the sharded step (0 = the harness default per world)")
    ap.add_argument("--host-slices", type=int, default=32,
                    help="slices of the host-buffer sharded step (e2e leg)")
    ap.add_argument("--rendezvous", default="ipc", choices=["ipc", "symm"])
    ap.add_argument("--overlap", type=int, default=-1, choices=[-1, 0, 1],
                    help="one-kernel step: 0 = every CTA pushes then applies, 1 = split roles, "
                         "-1 = the harness default per world")
    ap.add_argument("--reducers", type=int, default=0,
                    help="switch mode: percent of CTAs that reduce+apply+pull (0 = default: 25 with one local gradient, 8 with several)")
    ap.add_argument("--no-coop", action="store_true",
                    help="one-kernel step: plain launch when the grid is <= one CTA per SM")
    ap.add_argument("--pull", default="f32", choices=["f32", "bf16"],
                    help="dtype of the replicas the pull writes (bf16 halves the pull bytes)")
    ap.add_argument("--multicast", action="store_true",
                    help="pull through an NVSwitch multicast mapping (symm rendezvous only)")
    ap.add_argument("--kernel", default="", help="ldg128 | ldg256 | tma (default: auto)")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--ctas", type=int, default=0)
    ap.add_argument("--unroll", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus == 1 and int(os.environ.get("WORLD_SIZE", "1")) == 1:
        return run_single(args)
    return run_sharded(args)


if __name__ == "__main__":
    main()
