#!/usr/bin/env python
"""BASELINE configs 3 and 4 as whole sharded steps on real peers (torchrun, one process per GPU):
  config 3  ResNet-50-shaped gradients (25,557,032 fp32), Momentum lr 0.1 / 0.9, one worker
            gradient per rank, PS sharded over the P ranks of the launch;
  config 4  BERT-large-shaped gradients (335,141,888), bf16 gradients -> fp32 Adam (lr 1e-4),
            one gradient per rank, fp32 pull and bf16 pull.
Per config: ms per step (CUDA events, max over ranks), per-GPU NVLink bytes per direction and the
rate they imply, and the NCCL-based step beside it. One JSON line per row on rank 0."""
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
    for _ in range(5):
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
    cases = [("cfg3 resnet50 momentum", 25_557_032, pkg.OPT_MOMENTUM, False,
              dict(lr=0.1, momentum=0.9)),
             ("cfg4 bert-large bf16->adam", 335_141_888, pkg.OPT_ADAM, True, dict(lr=1e-4))]
    for name, elems, opt, bf16, hpkw in cases:
        for mode, pull, slices in (("fused", "f32", 1), ("fused", "f32", 2), ("fused", "bf16", 1),
                                   ("nccl", "f32", 1)):
            if mode == "nccl" and bf16:
                pass  # the baseline pre-sums bf16 into fp32 locally, then fp32 collectives
            sps = ShardedPS(rank, world, local, elems, opt, pkg.HParams(**hpkw), mode=mode,
                            n_chunks=slices, pull_dtype=pull, overlap=0)
            g = torch.randn(sps.padded, device=dev) * 0.01
            grads = [g.to(torch.bfloat16) if bf16 else g]
            iters = 50 if elems < 10**8 else 20
            ms = timed(lambda: sps.step(grads, stream), iters, stream, dev)
            if mode != "nccl":
                sps.check(stream)
            shard = sps.shard
            link = (world - 1) * shard * (4 + (2 if pull == "bf16" else 4))
            if rank == 0:
                print(json.dumps({
                    "case": name, "world": world, "mode": mode, "pull": pull, "slices": slices,
                    "elems": elems, "ms_per_step": round(ms, 4),
                    "steps_per_s": round(1e3 / ms, 1),
                    "nvlink_bytes_per_gpu_per_dir": link,
                    "nvlink_gbs_per_dir": round(link / ms / 1e6, 1),
                    "frac_of_770_measured": round(link / ms / 1e6 / 770, 3),
                    "frac_of_900_nominal": round(link / ms / 1e6 / 900, 3)}), flush=True)
            del sps, g, grads
            torch.cuda.empty_cache()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
