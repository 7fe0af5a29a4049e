#!/usr/bin/env python
"""Per-phase timing of the sharded PS step on real peers (torchrun, one process per GPU):
push only, apply_pull only (unicast and multicast), and the whole step for several slice
counts. Max over ranks of CUDA-event times. Phases repeated in isolation reuse the last
epoch, so their waits are already satisfied — results are garbage, timings are not."""
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
    elems, workers = 100_000_000, int(os.environ.get("WORKERS", "8"))
    stream = torch.cuda.current_stream()
    out = {}
    for mc in ((False, True) if os.environ.get('MC') else (False,)):
        sps = ShardedPS(rank, world, local, elems, pkg.OPT_SGD, pkg.HParams(lr=0.01), mode="fused",
                        rendezvous="symm", multicast=mc, n_chunks=1)
        grads = [torch.randn(sps.padded, device=dev) * 0.01 for _ in range(workers)]
        tag = "mc" if (mc and sps.multicast_ptr) else "uc"
        if mc and not sps.multicast_ptr:
            continue
        if not mc:
            out["push"] = timed(lambda: sps.push(grads, stream), 20, stream, dev)
        ep = sps.push(grads, stream)
        torch.cuda.synchronize()
        dist.barrier()
        out[f"apply_pull_{tag}"] = timed(
            lambda: sps.apply_pull(workers * world, None, stream, epoch=ep), 20, stream, dev)
        for ch in (1, 2, 4, 8):
            sps.n_chunks = ch
            out[f"step_{tag}_chunks{ch}"] = timed(lambda: sps.step(grads, stream), 20, stream, dev)
        sps.check(stream)
        del sps, grads
        torch.cuda.empty_cache()
    if rank == 0:
        for k, v in out.items():
            print(f"{k:24s} {v:.4f} ms")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
