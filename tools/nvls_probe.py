#!/usr/bin/env python
"""Raw rates of NVSwitch multicast stores / in-switch reductions (development probe; torchrun).
Every rank works on its own shard of a symmetric buffer at the same time, as a fused step would."""
import ctypes
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_controller_b200 as pkg  # noqa: E402
from kubeflow_controller_b200.sharded import SymmBuffer  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    lib = pkg.load_library()
    fn = lib.accel_dev_probe_multimem
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                   ctypes.c_int, ctypes.c_void_p]
    fn.restype = ctypes.c_int
    shard = 64 << 20                      # elements per rank (256 MB)
    buf = SymmBuffer(shard * world * 4, local)
    if not buf.multicast_ptr:
        if rank == 0:
            print("no multicast support on this fabric")
        return
    loc = torch.randn(2 * shard, device=dev)
    stream = torch.cuda.current_stream()
    peer = buf.peer_ptrs[(rank + 1) % world]
    for mode, name in ((2, "local copy"), (3, "peer load (right neighbour)"),
                       (4, "peer store (right neighbour)"), (0, "multimem.st (all-gather)"),
                       (1, "multimem.ld_reduce (reduce-scatter)")):
        for ctas in (296, 592, 1184):
            def go():
                base = peer if mode in (3, 4) else buf.multicast_ptr
                pkg.accel.check(fn(base, loc.data_ptr(), rank * shard, shard, mode,
                                   ctas, stream.cuda_stream))
            for _ in range(3):
                go()
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(10):
                go()
            b.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b) / 10], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.barrier()
            if rank == 0:
                gb = shard * 4 / 1e9
                print(f"{name:38s} ctas={ctas:5d} {t.item():.4f} ms  shard {gb / t.item() * 1e3:7.1f} GB/s"
                      f"  (x{world - 1} fan = {gb * (world - 1) / t.item() * 1e3:7.1f} GB/s)", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
