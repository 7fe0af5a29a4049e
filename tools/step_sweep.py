#!/usr/bin/env python
"""Variants of the multi-GPU step in ONE torchrun session (start-up paid once): the headline
workload (8 local workers x 100 M fp32 per GPU, SGD) under every (mode, slices, reducer share,
launch kind) of interest, the small-shard latency cases, and the host-buffer step at several
slice counts. One JSON line per row on rank 0.

    torchrun ... tools/step_sweep.py [headline] [small] [host]
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_controller_b200 as pkg  # noqa: E402
from kubeflow_controller_b200.sharded import ShardedPS  # noqa: E402

ELEMS, WORKERS = 100_000_000, 8


def timed(fn, iters, warm, stream, dev):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(iters):
        fn()
    b.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    return t.item()


def main():
    what = set(sys.argv[1:]) or {"headline", "small", "host"}
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream()

    def emit(row):
        if rank == 0:
            print(json.dumps(row), flush=True)

    if "headline" in what:
        padded = -(-ELEMS // (8 * world)) * 8 * world
        grads = [torch.randn(padded, device=dev) * 0.01 for _ in range(WORKERS)]
        variants = [("fused", 1, 0, 0, True), ("fused", 2, 0, 0, True), ("fused", 4, 1, 0, True)]
        for s, pct in ((4, 25), (8, 25), (16, 25), (4, 12), (8, 12), (16, 12), (8, 6), (8, 40)):
            variants.append(("switch", s, 2, pct, True))
        for mode, slices, overlap, pct, coop in variants:
            try:
                sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                                mode=mode, n_chunks=slices,
                                overlap=None if mode == "switch" else overlap)
                sps.ctx.set_step_shape(pct, coop)
                ms = timed(lambda: sps.step(grads, stream), 20, 4, stream, dev)
                sps.check(stream)
                emit({"case": "headline 8 x 100M fp32 SGD per GPU", "world": world, "mode": mode,
                      "slices": slices, "overlap": overlap, "reducer_pct": pct, "coop": coop,
                      "ms": round(ms, 4),
                      "grad_gbs": round(world * WORKERS * ELEMS * 4 / ms / 1e6, 1)})
                del sps
            except Exception as e:   # noqa: BLE001
                emit({"case": "headline", "mode": mode, "slices": slices, "error": str(e)[:300]})
            torch.cuda.empty_cache()
        del grads
        torch.cuda.empty_cache()

    if "roles" in what:
        # unicast step with split roles: share of the CTAs that apply + pull (the rest push)
        padded = -(-ELEMS // (8 * world)) * 8 * world
        grads = [torch.randn(padded, device=dev) * 0.01 for _ in range(WORKERS)]
        for slices, pct in ((4, 50), (4, 40), (4, 30), (4, 25), (4, 20), (4, 15), (8, 30), (8, 25),
                            (8, 20), (2, 25), (16, 25)):
            sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                            mode="fused", n_chunks=slices, overlap=1)
            sps.ctx.set_step_shape(pct, True)
            ms = timed(lambda: sps.step(grads, stream), 20, 4, stream, dev)
            sps.check(stream)
            emit({"case": "headline, unicast step, split roles", "world": world, "slices": slices,
                  "applier_pct": pct, "ms": round(ms, 4)})
            del sps
            torch.cuda.empty_cache()
        del grads
        torch.cuda.empty_cache()

    if "tune" in what:
        # second pass over the in-switch step: reducer share x quads in flight x slices
        padded = -(-ELEMS // (8 * world)) * 8 * world
        grads = [torch.randn(padded, device=dev) * 0.01 for _ in range(WORKERS)]
        for slices, pct, unroll in ((8, 6, 4), (8, 8, 4), (8, 10, 4), (6, 6, 4), (12, 6, 4), (8, 4, 4)):
            try:
                sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                                mode="switch", n_chunks=slices)
                sps.ctx.set_step_shape(pct, True)
                ms = timed(lambda: sps.step(grads, stream), 20, 4, stream, dev)
                sps.check(stream)
                emit({"case": "headline, in-switch step", "world": world, "slices": slices,
                      "reducer_pct": pct, "quads_in_flight": unroll, "ms": round(ms, 4)})
                del sps
            except Exception as e:   # noqa: BLE001
                emit({"case": "tune", "slices": slices, "pct": pct, "error": str(e)[:300]})
            torch.cuda.empty_cache()
        del grads
        torch.cuda.empty_cache()

    if "l2" in what:
        # ACCEL_STEP_L2HINT: evict-first gradient loads and evict-last stage stores in the summers,
        # so that the slice the switch reads next is still in L2; off / on, several slice counts
        padded = -(-ELEMS // (8 * world)) * 8 * world
        grads = [torch.randn(padded, device=dev) * 0.01 for _ in range(WORKERS)]
        for hint in ("0", "1"):
            os.environ["ACCEL_STEP_L2HINT"] = hint
            for slices, pct in ((8, 6), (8, 12), (16, 12), (32, 12), (16, 25)):
                try:
                    sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                                    mode="switch", n_chunks=slices)
                    sps.ctx.set_step_shape(pct, True)
                    ms = timed(lambda: sps.step(grads, stream), 20, 4, stream, dev)
                    sps.check(stream)
                    emit({"case": "headline, in-switch step, L2 policies", "world": world,
                          "l2hint": int(hint), "slices": slices, "reducer_pct": pct,
                          "ms": round(ms, 4)})
                    del sps
                except Exception as e:   # noqa: BLE001
                    emit({"case": "l2", "l2hint": hint, "slices": slices, "error": str(e)[:300]})
                torch.cuda.empty_cache()
        os.environ.pop("ACCEL_STEP_L2HINT", None)
        del grads
        torch.cuda.empty_cache()

    if "trace" in what:
        # where the time of one in-switch step goes (ACCEL_STEP_TRACE stamps, microseconds from the
        # first reducer's start), ranks 0 and world-1
        os.environ["ACCEL_STEP_TRACE"] = "1"
        padded = -(-ELEMS // (8 * world)) * 8 * world
        grads = [torch.randn(padded, device=dev) * 0.01 for _ in range(WORKERS)]
        for slices, pct, unroll in ((8, 6, 4), (8, 25, 4)):
            sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                            mode="switch", n_chunks=slices)
            sps.ctx.set_step_shape(pct, True)
            for _ in range(6):
                sps.step(grads, stream)
            sps.check(stream)
            torch.cuda.synchronize()
            dist.barrier()
            t = sps.flags[128:256].view(torch.int64).cpu().tolist()
            t0 = t[0]
            row = {"case": "trace of one in-switch step (us from reducer start)", "rank": rank,
                   "slices": slices, "reducer_pct": pct, "quads_in_flight": unroll,
                   "summed": [round((x - t0) / 1e3, 1) for x in t[1:1 + slices]],
                   "reduce_start": [round((x - t0) / 1e3, 1) for x in t[17:17 + slices]],
                   "reduce_end": [round((x - t0) / 1e3, 1) for x in t[33:33 + slices]],
                   "closing_ticket": round((t[49] - t0) / 1e3, 1),
                   "closed": round((t[50] - t0) / 1e3, 1)}
            rows = [None] * world
            dist.all_gather_object(rows, row)
            if rank == 0:
                for r in (rows[0], rows[world - 1]):
                    print(json.dumps(r), flush=True)
            del sps
            torch.cuda.empty_cache()
        os.environ.pop("ACCEL_STEP_TRACE", None)
        del grads
        torch.cuda.empty_cache()

    if "grid" in what:
        # small shards: how many items per thread before the grid shrinks (ACCEL_STEP_ITEMS is read
        # once per process, so this section takes the value from the environment of the launch)
        ipt = os.environ.get("ACCEL_STEP_ITEMS", "2")
        for mib in (1, 4):
            elems = (mib << 18) * world
            g = [torch.randn(elems, device=dev) * 0.01]
            for mode, coop in (("fused", True), ("fused", False)):
                sps = ShardedPS(rank, world, local, elems, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                                mode=mode, n_chunks=1, overlap=0)
                sps.ctx.set_step_shape(0, coop)
                ms = timed(lambda: sps.step(g, stream), 100, 10, stream, dev)
                sps.check(stream)
                emit({"case": f"grid sizing, {mib} MiB shard, SGD, 1 gradient per rank",
                      "world": world, "items_per_thread": int(ipt), "coop": coop,
                      "us": round(ms * 1e3, 2)})
                del sps
                torch.cuda.empty_cache()
            del g

    if "small" in what:
        # small shards: what the launch kind and the grid size cost (one gradient per rank)
        for mib in (1, 16):
            elems = (mib << 18) * world
            g = [torch.randn(elems, device=dev) * 0.01]
            for mode, coop in (("fused", True), ("fused", False), ("switch", True), ("switch", False)):
                try:
                    sps = ShardedPS(rank, world, local, elems, pkg.OPT_SGD, pkg.HParams(lr=0.01),
                                    mode=mode, n_chunks=1,
                                    overlap=None if mode == "switch" else 0)
                    sps.ctx.set_step_shape(0, coop)
                    ms = timed(lambda: sps.step(g, stream), 50, 5, stream, dev)
                    sps.check(stream)
                    emit({"case": f"small shard {mib} MiB per GPU, SGD, 1 gradient per rank",
                          "world": world, "mode": mode, "coop": coop, "us": round(ms * 1e3, 2)})
                    del sps
                except Exception as e:   # noqa: BLE001
                    emit({"case": f"small {mib}", "mode": mode, "coop": coop, "error": str(e)[:300]})
                torch.cuda.empty_cache()
            del g

    if "host" in what:
        sps = ShardedPS(rank, world, local, ELEMS, pkg.OPT_SGD, pkg.HParams(lr=0.01), mode="fused")
        hg = [torch.empty(sps.padded, dtype=torch.float32).pin_memory() for _ in range(WORKERS)]
        for h in hg:
            h.normal_(0, 0.01)
        out = torch.empty(sps.padded, dtype=torch.float32).pin_memory()
        for slices, with_out in ((4, True), (8, True), (16, True), (32, True), (64, True), (16, False)):
            for _ in range(2):
                sps.step_host(hg, out if with_out else None, n_slices=slices)
            dist.barrier()
            t0 = time.perf_counter()
            for _ in range(4):
                sps.step_host(hg, out if with_out else None, n_slices=slices)
            dist.barrier()
            dt = torch.tensor([(time.perf_counter() - t0) / 4], device=dev)
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            emit({"case": "host-buffer sharded step, 8 x 100M fp32 per rank", "world": world,
                  "slices": slices, "d2h": with_out, "ms": round(dt.item() * 1e3, 2),
                  "grad_gbs": round(world * WORKERS * ELEMS * 4 / dt.item() / 1e9, 1)})
        # raw link check: the 8 H2D copies alone on every rank at once
        dg = [torch.empty(sps.padded, device=dev) for _ in range(WORKERS)]
        for _ in range(2):
            for d, h in zip(dg, hg):
                d.copy_(h, non_blocking=True)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            for d, h in zip(dg, hg):
                d.copy_(h, non_blocking=True)
            torch.cuda.synchronize()
        dist.barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / 3], device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        emit({"case": "raw H2D of the 8 gradients on every rank at once", "world": world,
              "ms": round(dt.item() * 1e3, 2),
              "gbs_all_ranks": round(world * WORKERS * sps.padded * 4 / dt.item() / 1e9, 1)})
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
