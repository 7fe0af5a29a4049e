#!/usr/bin/env python
"""Real multi-GPU parity check of the sharded PS step, one process per GPU (torchrun):
fused mode must match the hierarchical-order oracle bit for bit on every rank's replica;
nccl mode within rel 1e-5 (NCCL's reduction order is its own). Prints one line per rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_controller_b200 as pkg  # noqa: E402
from kubeflow_controller_b200.sharded import ShardedPS  # noqa: E402
from oracle import ps_oracle as O  # noqa: E402

F = np.float32


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    elems, workers, steps = 400_003, 3, 3
    ok = True
    variants = [("fused", "ipc", 1), ("fused", "ipc", 16), ("fused-roles", "ipc", 8),
                ("fused", "symm", 8), ("streams", "ipc", 4), ("nccl", "ipc", 1)]
    for mode, rdv, chunks in variants:
        for opt, bf16 in ((pkg.OPT_SGD, False), (pkg.OPT_ADAM, True), (pkg.OPT_MOMENTUM, False)):
            sps = ShardedPS(rank, world, local, elems, opt, pkg.HParams(lr=0.01, momentum=0.9),
                            mode=mode.split("-")[0], rendezvous=rdv, n_chunks=chunks,
                            multicast=False, overlap=int(mode.endswith("roles")))
            padded = sps.padded
            rng = np.random.default_rng(7)          # same stream on every rank
            var = rng.standard_normal(padded).astype(F)
            s0, s1 = np.zeros(padded, F), np.zeros(padded, F)
            sps.var.copy_(torch.from_numpy(var).to(dev))
            ohp = O.HParams(lr=0.01, momentum=0.9)
            for _ in range(steps):
                allg = [[(0.01 * rng.standard_normal(padded)).astype(F) for _ in range(workers)]
                        for _ in range(world)]
                if bf16:
                    allg = [[O.f32_to_bf16_bits(g) for g in gr] for gr in allg]
                mine = [torch.from_numpy(g.view(np.int16)).to(dev).view(torch.bfloat16) if bf16
                        else torch.from_numpy(g).to(dev) for g in allg[rank]]
                sps.step(mine)
                partials = [O.numpy_aggregate(gr, False) for gr in allg]
                O.numpy_apply(opt, var, s0, s1, O.numpy_aggregate(partials, False), ohp)
                ohp.advance()
            if mode != "nccl":
                sps.check()
            torch.cuda.synchronize()
            dist.barrier()
            got = sps.var.cpu().numpy()
            if mode != "nccl":
                good = np.array_equal(got, var)
            else:
                good = np.allclose(got, var, rtol=1e-5, atol=1e-6)
            ok &= bool(good)
            bad = torch.tensor([0 if good else int((got != var).sum()) or 1], device=dev)
            dist.all_reduce(bad, op=dist.ReduceOp.MAX)      # one line per case, from rank 0
            if rank == 0:
                print(f"world={world} mode={mode} rdv={rdv} mc={int(bool(sps.multicast_ptr))} "
                      f"chunks={chunks} opt={opt} bf16={bf16} "
                      f"{'OK (all ranks)' if bad.item() == 0 else 'MISMATCH ' + str(bad.item())}",
                      flush=True)
            del sps
            torch.cuda.empty_cache()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
