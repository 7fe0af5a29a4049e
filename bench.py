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
`--impl reference` times the reference's CPU PS path (the threaded C restatement in
oracle/, TF-1.4 is not installable here) on the host cores, same metric and unit.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
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


# ---- the reference arm / cpu baseline: the CPU PS path on the host cores --------
def cpu_ps_rate(sample_elems: int, steps: int, warmup: int):
    """Gradient GB/s of the CPU restatement on a bounded sample of the workload."""
    import numpy as np
    from oracle import ps_oracle as O
    cores = os.cpu_count() or 1
    orc = O.COracle(threads=cores)
    rng = np.random.default_rng(1234)
    block = 1 << 20   # a random 4 MiB block tiled to size: init cost stays off the clock

    def tiled(scale):
        b = rng.standard_normal(block, dtype=np.float32) * np.float32(scale)
        return np.resize(b, sample_elems)
    var = tiled(1.0)
    grads = [tiled(0.01) for _ in range(WORKERS)]
    acc = np.empty(sample_elems, dtype=np.float32)
    hp = O.HParams(lr=LR)
    for _ in range(warmup):
        orc.ps_step(O.SGD, var, None, None, grads, hp, acc=acc)
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.ps_step(O.SGD, var, None, None, grads, hp, acc=acc)
    dt = (time.perf_counter() - t0) / steps
    gbs = WORKERS * sample_elems * 4 / dt / 1e9
    frac = "the whole workload" if sample_elems == ELEMS else \
        f"1/{ELEMS // sample_elems} of the workload"
    sample = (f"{WORKERS} workers x {sample_elems} fp32 ({frac}; working set "
              f"{(WORKERS + 2) * sample_elems * 4 / 1e9:.1f} GB, beyond any host cache) x {steps} "
              f"steps, C restatement of TF-1.4 ConditionalAccumulator + "
              f"ApplyGradientDescent, one pass per statement, {cores} threads")
    return gbs, dt, cores, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = ELEMS     # full-size steps: ~0.1-0.3 s each on the box's host cores
    gbs, dt, cores, desc = cpu_ps_rate(sample, max(args.steps, 1), min(max(args.warmup, 1), 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt * 1e3 * (ELEMS / sample), 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(1),
        "cpu_baseline": {"value": round(gbs, 3), "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": desc},
        "e2e": {"value": round(gbs, 3), "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "steps_per_s": round(1.0 / (dt * (ELEMS / sample)), 3),
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus: int):
    cfg = {"workload": ("single-PS fused aggregate+SGD apply, 8 workers x 100M fp32 params on "
                        "1 B200 (BASELINE configs[1])"),
           "workers_per_gpu": WORKERS, "elems": ELEMS, "optimizer": "sgd", "aggregate": "sum",
           "lr": LR, "l2": "inputs (3.6 GB) exceed L2 (126 MB); no flush needed"}
    if n_gpus > 1:
        cfg["workload"] = (f"PS sharded over {n_gpus} B200: per GPU 8 local workers x 100M fp32, "
                           f"push (8-way sum + reduce-scatter) -> shard SGD apply -> pull "
                           f"(all-gather)")
        cfg["ps_shards"] = n_gpus
    return cfg


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
    if not args.no_cpu:
        cpu_gbs, cpu_dt, cores, desc = cpu_ps_rate(ELEMS, 5, 1)
        cpu = {"value": round(cpu_gbs, 3), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": desc}
    emit_single(args, ms, value, alg_bytes, achieved, peak, peak_src, per_launch, avg_launch_ms,
                cpu, e2e, clk)


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
                    "apply -> D2H of updated variables"}


def emit_single(args, ms, value, alg_bytes, achieved, peak, peak_src, per_launch, avg_launch_ms,
                cpu, e2e, clk):
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
    }
    print(json.dumps(line), flush=True)


# ---- N > 1 ---------------------------------------------------------------------------
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
    sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=LR), mode=args.mode,
                    n_chunks=args.chunks or None, rendezvous=args.rendezvous,
                    multicast=args.multicast, pull_dtype=args.pull, overlap=None if args.overlap < 0 else args.overlap)
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

    # ---- e2e: every rank's workers deliver their gradients in pinned HOST memory; per step each
    # rank copies them in over its own PCIe link, runs the sharded step, and copies its updated
    # replica back out. Wall clock, barrier on both sides, max over ranks.
    e2e = None
    if not args.no_e2e:
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        h_grads = [torch.empty(sps.padded, dtype=torch.float32).pin_memory() for _ in range(WORKERS)]
        for h, g in zip(h_grads, grads):
            h.copy_(g)
        h_out = torch.empty(sps.padded, dtype=torch.float32).pin_memory()

        def host_step():
            for h, g in zip(h_grads, grads):
                g.copy_(h, non_blocking=True)
            sps.step(grads, stream)
            h_out.copy_(sps.var, non_blocking=True)
            torch.cuda.synchronize()
        for _ in range(2):
            host_step()
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            host_step()
        dist.barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / e2e_steps], device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": round(world * WORKERS * ELEMS * 4 / dt.item() / 1e9, 3), "unit": UNIT,
               "h2d_bytes_per_step": world * WORKERS * sps.padded * 4,
               "d2h_bytes_per_step": world * sps.padded * 4,
               "ms_per_step": round(dt.item() * 1e3, 3), "steps": e2e_steps,
               "path": "per rank: 8 pinned host gradients -> H2D over the rank's own PCIe link -> "
                       "sharded fused step -> D2H of the updated replica"}
    if rank == 0:
        grad_bytes = world * WORKERS * ELEMS * 4
        peak, peak_src = peaks()
        shard = sps.padded // world
        # per GPU: read 8 local gradients + write partial (push), read `world` slots + var r/w
        # (apply), write the full replica (pull)
        hbm_bytes = sps.padded * (WORKERS * 4 + 4) + shard * (world * 4 + 8) + sps.padded * 4
        pull_b = 2 if args.pull == "bf16" else 4
        nvl_bytes = (world - 1) * shard * (4 + pull_b)   # push out + pull out, per GPU per direction
        line = {
            "metric": METRIC, "value": round(grad_bytes / (ms * 1e-3) / 1e9, 2), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(world), mode=sps.mode, chunks=sps.n_chunks,
                           rendezvous=args.rendezvous, multicast=bool(sps.multicast_ptr),
                           pull=args.pull, overlap=["inline", "roles"][sps.overlap]),
            "steps_per_s": round(1e3 / ms, 2),
            "roofline": {"bound": "hbm", "achieved": round(hbm_bytes / (ms * 1e-3) / 1e9, 2),
                         "peak": peak, "unit": "GB/s",
                         "frac": round(hbm_bytes / (ms * 1e-3) / 1e9 / peak, 4), "traffic": None,
                         "peak_source": peak_src,
                         "note": "per-GPU HBM bytes of the whole step / step time; at N > 1 the step "
                                 "is NVLink-bound as well, see `nvlink`"},
            "nvlink": {"bytes_per_gpu_per_dir_per_step": nvl_bytes,
                       "achieved": round(nvl_bytes / (ms * 1e-3) / 1e9, 2),
                       "peak": 770.0, "unit": "GB/s per direction per GPU",
                       "frac": round(nvl_bytes / (ms * 1e-3) / 1e9 / 770.0, 4),
                       "frac_of_nominal_900": round(nvl_bytes / (ms * 1e-3) / 1e9 / 900.0, 4),
                       "peak_source": "measured peer copy, B200_PROFILING.md",
                       "note": "push sends (P-1)/P of the arena and the pull another (P-1)/P, in "
                               "each direction, over the whole step time"},
            "cpu_baseline": None,
            "e2e": e2e,
            "gpu_launches": args.steps * sps.launches_per_step,
            "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (sweeps)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (sweeps)")
    ap.add_argument("--mode", default="fused", choices=["fused", "streams", "nccl"])
    ap.add_argument("--chunks", type=int, default=0,
                    help="pipeline slices of the sharded step (0 = the harness default per world)")
    ap.add_argument("--rendezvous", default="ipc", choices=["ipc", "symm"])
    ap.add_argument("--overlap", type=int, default=-1, choices=[-1, 0, 1],
                    help="one-kernel step: 0 = every CTA pushes then applies, 1 = split roles, "
                         "-1 = the harness default per world")
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
