"""World-size-2 gloo test of the host-side logic of the sharded PS (no GPU): both ranks
derive the same shard plan and placement, the rendezvous exchange returns every rank's
handles in rank order, and the hierarchical aggregation order the GPU path uses stays
within the path's tolerance of the flat worker order."""
import os
import socket
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import kubeflow_controller_b200 as pkg
        from kubeflow_controller_b200 import sharded
        pkg.load_library()
        # 1. every rank computes the same plan; shards tile the padded arena in rank order
        padded, shard = sharded.shard_layout(25_557_032, world)
        plans = sharded.exchange((padded, shard), world)
        assert all(p == plans[0] for p in plans)
        shards, _ = pkg.plan_flat(25_557_032, world, 8)
        assert shards[rank] == (rank * shard, shard)
        # 2. PS task r <-> GPU r <-> rank r, from the argv the controller would emit
        argv = pkg.placement.cluster_spec(world, world, [f"w{i}" for i in range(world)],
                                          [f"p{i}" for i in range(world)], "PS", rank)
        job, task, n_w, n_ps = pkg.placement.parse_argv(argv)
        assert (job, task, n_w, n_ps) == ("ps", rank, world, world)
        assert pkg.Context(n_ps).gpu_of(task) == rank
        # 3. the handle exchange is an all-gather in rank order
        mine = (bytes([rank]) * 64, bytes([rank + 16]) * 64, bytes([rank + 32]) * 64)
        got = sharded.exchange(mine, world)
        assert [g[0][0] for g in got] == list(range(world))
        assert [g[2][0] for g in got] == [r + 32 for r in range(world)]
        # 4. round-robin plan agrees across ranks
        rr = pkg.plan_round_robin([78400, 100, 1000, 10], world, 1, 8)
        assert all(x == rr for x in sharded.exchange(rr, world))
        # 5. the TF-faithful bucket layout of the sharded arena: same on every rank, variable k in
        #    the shard of rank k mod world
        lay = sharded.BucketLayout([78400, 100, 1000, 10, 4096, 7], world)
        views = sharded.exchange((lay.shard, lay.pos), world)
        assert all(v == views[0] for v in views)
        assert [p // lay.shard for p in lay.pos] == [k % world for k in range(6)]
        q.put((rank, "ok"))
    except Exception as e:   # surface the failure in the parent
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_host_logic():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
    assert res == {0: "ok", 1: "ok"}, res


def test_hierarchical_order_within_tolerance():
    """Per-rank partial sums then rank order (GPU sharded path) vs one flat worker-order sum:
    both are valid accumulator arrival orders; they agree to fp32 rounding."""
    from oracle import ps_oracle as O
    rng = np.random.default_rng(0)
    g = [(0.01 * rng.standard_normal(50_000)).astype(np.float32) for _ in range(16)]
    flat = O.numpy_aggregate(g, False)
    parts = [O.numpy_aggregate(g[i:i + 4], False) for i in range(0, 16, 4)]
    hier = O.numpy_aggregate(parts, False)
    np.testing.assert_allclose(hier, flat, rtol=1e-5, atol=1e-7)
