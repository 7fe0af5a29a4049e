#!/usr/bin/env python
"""Kernel-level table for the BASELINE configs on one B200: every optimizer / gradient
dtype / worker count / size the configs name, as ms, HBM GB/s and fraction of the measured
HBM peak. Each launch is timed on its own with CUDA events; an L2 flush (a 512 MiB memset)
runs between launches when the working set is under 2x L2, so small shards are measured cold.

    python tools/bench_configs.py [--quick] [--kernel ldg128|ldg256|tma] > table.jsonl
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_controller_b200 as pkg  # noqa: E402

L2_BYTES = 126 << 20


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6650.0


def run(ctx, opt, bf16, workers, n, iters, flush_buf, hp=None):
    dev = "cuda:0"
    var = torch.randn(n, device=dev)
    s0 = torch.zeros(n, device=dev) if opt >= 1 else None
    s1 = torch.full((n,), 1e-4, device=dev) if opt == 2 else None
    grads = [(torch.randn(n, device=dev) * 0.01).to(torch.bfloat16 if bf16 else torch.float32)
             for _ in range(workers)]
    hp = hp or pkg.HParams(lr=0.01, momentum=0.9)
    gb = 2 if bf16 else 4
    alg = n * (workers * gb + 8 * (1 + opt))
    cold = alg < 2 * L2_BYTES
    st = torch.cuda.current_stream()
    for _ in range(3):
        ctx.apply(opt, var, s0, s1, grads, hp, stream=st)
    times = []
    for _ in range(iters):
        if cold:
            flush_buf.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        ctx.apply(opt, var, s0, s1, grads, hp, stream=st)
        b.record(st)
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    times.sort()
    med = times[len(times) // 2]
    return {"ms_median": round(med, 5), "ms_min": round(times[0], 5), "alg_bytes": alg,
            "hbm_gbs": round(alg / med / 1e6, 1), "cold_l2": cold}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--kernel", default="")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--ctas", type=int, default=0)
    ap.add_argument("--unroll", type=int, default=0)
    args = ap.parse_args()
    pkg.load_library()
    ctx = pkg.Context(1)
    if args.kernel:
        ctx.set_kernel(getattr(pkg, "KERNEL_" + args.kernel.upper()), args.threads, args.ctas,
                       args.unroll)
    pk = peak()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")
    names = ["sgd", "momentum", "adam"]
    cases = []
    # config 2 and its Momentum / Adam siblings (SURVEY §8 d), fp32 and bf16 gradients
    for opt in (0, 1, 2):
        for bf16 in (False, True):
            cases.append(("cfg2 8x100M", opt, bf16, 8, 100_000_000))
    # post-collective shard applies: config 3 (ResNet-50, Momentum, P=2/4/8), config 4 (BERT-large)
    for p in (2, 4, 8):
        cases.append((f"cfg3 shard P={p}", 1, False, 1, -(-25_557_032 // (8 * p)) * 8))
    cases.append(("cfg4 shard P=8", 2, True, 1, 41_892_736))
    cases.append(("cfg4 whole 1 GPU", 2, True, 8, 335_141_888 // 8))
    # config 5: size sweep, single-gradient apply
    sizes = [1 << 18, 1 << 20, 1 << 22, 1 << 24, 1 << 26, 1 << 28] + ([] if args.quick else [1 << 30])
    for n in sizes:
        for opt in (0, 2):
            cases.append((f"cfg5 {n * 4 >> 20} MiB", opt, False, 1, n))
    # worker-count scan at 100 M
    if not args.quick:
        for w in (1, 2, 3, 4, 16):
            cases.append((f"workers={w} 100M", 0, False, w, 100_000_000))
    # config 1 (plumbing): the example's model (79,510 fp32), 1 PS + 2 workers, through the
    # host-buffer call, wall clock per step including copies and synchronisation
    import time
    for opt in (0, 2):
        n, workers = 79_510, 2
        var = torch.randn(n, device="cuda:0")
        s0 = torch.zeros(n, device="cuda:0") if opt else None
        s1 = torch.zeros(n, device="cuda:0") if opt == 2 else None
        hg = [(torch.randn(n) * 0.01).pin_memory() for _ in range(workers)]
        hout = torch.empty(n).pin_memory()
        hp = pkg.HParams(lr=0.5 if opt == 0 else 0.01)
        for _ in range(20):
            ctx.ps_step_host(opt, var, s0, s1, hg, hp, hout)
        t0 = time.perf_counter()
        for _ in range(200):      # the example's train_steps (mnist_replica.py:68)
            ctx.ps_step_host(opt, var, s0, s1, hg, hp, hout)
        dt = (time.perf_counter() - t0) / 200
        print(json.dumps({"case": "cfg1 mnist 1PS+2W host step", "opt": names[opt], "grad": "f32",
                          "workers": workers, "elems": n, "us_per_step": round(dt * 1e6, 1),
                          "steps_per_s": round(1 / dt, 1),
                          "reference_doc_log_global_steps_per_s": 356}), flush=True)
    # the sum-of-squares pre-pass of global-norm clipping (reads the N gradients once)
    for bf16 in (False, True):
        n, workers = 100_000_000, 8
        g = [(torch.randn(n, device="cuda:0") * 0.01).to(torch.bfloat16 if bf16 else torch.float32)
             for _ in range(workers)]
        out = torch.zeros(1, device="cuda:0")
        st = torch.cuda.current_stream()
        for _ in range(3):
            ctx.grad_sumsq(g, out, stream=st)
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            ctx.grad_sumsq(g, out, stream=st)
            b.record(st)
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        alg = n * workers * (2 if bf16 else 4)
        print(json.dumps({"case": "sumsq 8x100M", "grad": "bf16" if bf16 else "f32",
                          "ms_median": round(ts[5], 4), "hbm_gbs": round(alg / ts[5] / 1e6, 1),
                          "frac_of_measured_peak": round(alg / ts[5] / 1e6 / pk, 4)}), flush=True)
        del g
        torch.cuda.empty_cache()
    for label, opt, bf16, workers, n in cases:
        r = run(ctx, opt, bf16, workers, n, 7 if n > (1 << 26) else 15, flush)
        r.update({"case": label, "opt": names[opt], "grad": "bf16" if bf16 else "f32",
                  "workers": workers, "elems": n, "frac_of_measured_peak": round(r["hbm_gbs"] / pk, 4)})
        print(json.dumps(r), flush=True)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
