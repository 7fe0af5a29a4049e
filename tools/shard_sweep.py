#!/usr/bin/env python
"""BASELINE config 5 on real peers (torchrun, one process per GPU): variable shards of
1 MiB - 1 GiB (fp32) per GPU, one worker gradient per rank, SGD and Adam. For every size:
the push alone (a pure reduce-scatter: sum of ONE local gradient, partials to the owners),
the apply_pull alone (apply + all-gather) with fp32 and bf16 pulls, and the whole fused step.
Bus bandwidth as nccl-tests defines it: (P-1)/P * arena_bytes / t per collective, against the
measured 770 GB/s peer-copy rate and the nominal 900 GB/s per direction per GPU.
One JSON line per (size, optimizer) on rank 0. Phases repeated in isolation reuse the last
epoch (their waits are already satisfied): results are garbage, timings are not."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_controller_b200 as pkg  # noqa: E402
from kubeflow_controller_b200.sharded import ShardedPS  # noqa: E402


def timed(fn, iters, stream, dev):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
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
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream()
    mibs = [int(x) for x in os.environ.get("SHARD_MIB", "1,4,16,64,256,1024").split(",")]
    for mib in mibs:
        shard = mib << 18
        elems = shard * world
        for opt, name in ((pkg.OPT_SGD, "sgd"), (pkg.OPT_ADAM, "adam")):
            row = {"world": world, "shard_mib": mib, "opt": name, "arena_mb": elems * 4 / 1e6}
            for pull in ("f32", "bf16"):
                sps = ShardedPS(rank, world, local, elems, opt, pkg.HParams(lr=0.01), mode="fused",
                                n_chunks=1, pull_dtype=pull)
                if os.environ.get("SHARD_CTAS"):
                    sps.ctx.set_kernel(ctas_per_sm=int(os.environ["SHARD_CTAS"]))
                    row["ctas_per_sm"] = int(os.environ["SHARD_CTAS"])
                g = [torch.randn(sps.padded, device=dev) * 0.01]
                iters = 30 if mib <= 64 else 10
                if pull == "f32":
                    row["push_ms"] = timed(lambda: sps.push(g, stream), iters, stream, dev)
                ep = sps.push(g, stream)
                torch.cuda.synchronize()
                dist.barrier()
                row[f"apply_pull_{pull}_ms"] = timed(
                    lambda: sps.apply_pull(world, None, stream, epoch=ep), iters, stream, dev)
                row[f"step_{pull}_ms"] = timed(lambda: sps.step(g, stream), iters, stream, dev)
                sps.check(stream)
                del sps, g
                torch.cuda.empty_cache()
            moved = (world - 1) / world * elems * 4 / 1e9      # GB per GPU per direction
            row["rs_busbw_gbs"] = round(moved / (row["push_ms"] * 1e-3), 1)
            row["ag_busbw_gbs"] = round(moved / (row["apply_pull_f32_ms"] * 1e-3), 1)
            row["ag_bf16_busbw_gbs"] = round(moved / 2 / (row["apply_pull_bf16_ms"] * 1e-3), 1)
            for k in ("rs", "ag"):
                row[f"{k}_frac_of_770_measured"] = round(row[f"{k}_busbw_gbs"] / 770, 3)
                row[f"{k}_frac_of_900_nominal"] = round(row[f"{k}_busbw_gbs"] / 900, 3)
            for k in list(row):
                if k.endswith("_ms"):
                    row[k] = round(row[k], 4)
            if rank == 0:
                print(json.dumps(row), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
