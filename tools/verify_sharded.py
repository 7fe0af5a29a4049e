#!/usr/bin/env python
"""Real multi-GPU parity check of the sharded PS step, one process per GPU (torchrun):
the unicast forms (fused / streams) must match the hierarchical-order oracle bit for bit on every
rank's replica — also with raw-bf16 stage slots (one bf16 gradient per rank); nccl mode and the
in-switch form (switch: the NVSwitch adds in its own order) within rel 1e-5; the in-switch form
with raw-bf16 stage (the switch returns the sum rounded to bf16) within the bf16 path's rel 1e-2.
The host-buffer sharded step (accel_shard_step_host) is checked the same way. One line per case."""
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
    elems, steps = 400_003, 3
    ok = True
    variants = [("fused", "ipc", 1), ("fused", "ipc", 16), ("fused-roles", "ipc", 8),
                ("fused", "symm", 8), ("streams", "ipc", 4), ("nccl", "ipc", 1),
                ("switch", "symm", 1), ("switch", "symm", 5), ("switch-inplace", "symm", 4),
                ("host", "ipc", 3)]
    if os.environ.get("VERIFY_ONLY"):
        variants = [v for v in variants if v[0] in os.environ["VERIFY_ONLY"].split(",")]
    for mode, rdv, chunks in variants:
        for opt, bf16, workers in ((pkg.OPT_SGD, False, 3), (pkg.OPT_ADAM, True, 3),
                                   (pkg.OPT_MOMENTUM, False, 8), (pkg.OPT_ADAM, True, 1),
                                   (pkg.OPT_SGD, False, 1)):
            host = mode == "host"
            inplace = mode == "switch-inplace"
            if inplace and workers != 1:
                continue                      # the zero-copy form takes exactly one gradient
            sps = ShardedPS(rank, world, local, elems, opt, pkg.HParams(lr=0.01, momentum=0.9),
                            mode="fused" if host else mode.split("-")[0], rendezvous=rdv,
                            n_chunks=chunks, multicast=False,
                            overlap=int(mode.endswith("roles")) if not mode.startswith("switch")
                            else None)
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
                if host:
                    mine = [torch.from_numpy(g.view(np.int16)).view(torch.bfloat16).pin_memory()
                            if bf16 else torch.from_numpy(g).pin_memory() for g in allg[rank]]
                    sps.step_host(mine, None, n_slices=chunks)
                elif inplace:     # the worker writes its gradient straight into the stage buffer
                    view = sps.stage_grad(torch.bfloat16 if bf16 else torch.float32)
                    g0 = allg[rank][0]
                    view.copy_(torch.from_numpy(g0.view(np.int16)).to(dev).view(torch.bfloat16)
                               if bf16 else torch.from_numpy(g0).to(dev))
                    sps.step([view])
                else:
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
            if mode.startswith("switch") and bf16 and workers == 1:
                good = np.allclose(got, var, rtol=1e-2, atol=1e-4)     # bf16-rounded sums
                tol = "rel 1e-2"
            elif mode == "nccl" or mode.startswith("switch"):
                good = np.allclose(got, var, rtol=1e-5, atol=1e-6)
                tol = "rel 1e-5"
            else:
                good = np.array_equal(got, var)
                tol = "bit-exact"
            ok &= bool(good)
            bad = torch.tensor([0 if good else int((got != var).sum()) or 1], device=dev)
            dist.all_reduce(bad, op=dist.ReduceOp.MAX)      # one line per case, from rank 0
            if rank == 0:
                print(f"world={world} mode={mode} rdv={rdv} mc={int(bool(sps.multicast_ptr))} "
                      f"chunks={chunks} opt={opt} bf16={bf16} workers={workers} tol={tol} "
                      f"{'OK (all ranks)' if bad.item() == 0 else 'MISMATCH ' + str(bad.item())}",
                      flush=True)
            del sps
            torch.cuda.empty_cache()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
