#!/usr/bin/env python
"""Raw rates of the cross-GPU memory operations a fused step can use (development probe;
torchrun, one process per GPU; tools/probe/libaccel_probe.so — not the product library).
Every rank works on its own shard of a symmetric buffer at the same time, as a fused step would.

  peer stores: 128-bit, 256-bit (what the step kernels issue) and bulk-async 4 KiB
  (cp.async.bulk shared -> peer global); peer loads; multimem.st; multimem.ld_reduce; and the
  in-switch steady state (ld_reduce of the own shard + multimem.st of the result, both link
  directions busy).
One JSON line per (operation, grid) on rank 0."""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubeflow_controller_b200.sharded import SymmBuffer  # noqa: E402

MODES = [(2, "local copy (HBM reference)", "none"),
         (3, "peer load 128-bit (right neighbour)", "uni"),
         (4, "peer store 128-bit (right neighbour)", "uni"),
         (5, "peer store 256-bit st.global.v8.f32 (right neighbour)", "uni"),
         (6, "peer store bulk-async 4 KiB cp.async.bulk (right neighbour)", "uni"),
         (0, "multimem.st own shard (all-gather pattern)", "mc"),
         (1, "multimem.ld_reduce own shard (reduce-scatter pattern)", "mc"),
         (7, "ld_reduce own shard + multimem.st of the result (in-switch steady state)", "mc"),
         (8, "ld_reduce (even CTAs) and multimem.st (odd CTAs) of the own shard, decoupled", "mc")]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    lib = ctypes.CDLL(os.path.join(ROOT, "tools", "probe", "libaccel_probe.so"))
    fn = lib.accel_probe_run
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                   ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    fn.restype = ctypes.c_int
    shard = int(os.environ.get("PROBE_SHARD_MIB", "256")) << 18     # elements per rank
    buf = SymmBuffer(shard * world * 4, local)
    buf2 = SymmBuffer(shard * world * 4, local)
    loc = torch.randn(2 * shard, device=dev)
    stream = torch.cuda.current_stream()
    peer = buf.peer_ptrs[(rank + 1) % world]
    only = os.environ.get("PROBE_MODES")
    modes = [m for m in MODES if not only or str(m[0]) in only.split(",")]
    for mode, name, kind in modes:
        if kind == "mc" and not (buf.multicast_ptr and buf2.multicast_ptr):
            if rank == 0:
                print(json.dumps({"op": name, "skipped": "no multicast mapping on this fabric"}))
            continue
        for ctas in [int(c) for c in os.environ.get("PROBE_CTAS", "148,296,592,1184").split(",")]:
            def go():
                base = peer if kind == "uni" else buf.multicast_ptr
                rc = fn(base, loc.data_ptr(), buf2.multicast_ptr or None, rank * shard, shard, mode,
                        ctas, stream.cuda_stream)
                assert rc == 0, rc
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
                row = {"op": name, "world": world, "ctas": ctas, "shard_mb": round(gb * 1e3, 1),
                       "ms": round(t.item(), 4), "shard_gbs": round(gb / t.item() * 1e3, 1)}
                if kind == "mc":
                    # per GPU: the switch moves `world` shards through one direction of its link
                    row["link_gbs_busy_dir"] = round(gb * world / t.item() * 1e3, 1)
                print(json.dumps(row), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
