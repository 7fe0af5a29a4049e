"""The multi-GPU PS step (push -> apply_pull -> wait_pull) on ONE GPU: `world` ranks are
emulated in turn inside one process (their "peer" buffers are local allocations), which
exercises the whole data path and the flag protocol; real NVLink runs are bench.py --gpus N.
Oracle: per-rank partial sums in worker order, partials added in rank order, then the
TF-1.4 apply — bit-exact."""
import numpy as np
import pytest
import torch

from oracle import ps_oracle as O

pytestmark = pytest.mark.gpu
F = np.float32
DEV = "cuda:0"


def _dev(a):
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16)).to(DEV).view(torch.bfloat16)
    return torch.from_numpy(a).to(DEV)


def _oracle_hier(grads_per_rank, mean):
    """aggregate per rank (worker order), then across ranks (rank order), divide by the total."""
    partials = [O.numpy_aggregate(g, False) for g in grads_per_rank]
    total = partials[0].copy()
    for p in partials[1:]:
        total = total + p
    if mean:
        total = total / F(sum(len(g) for g in grads_per_rank))
    return total


@pytest.mark.parametrize("world,chunks", [(1, 1), (2, 1), (2, 3), (4, 4), (8, 1), (8, 5)])
@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
@pytest.mark.parametrize("workers,bf16,mean", [(1, False, False), (8, False, True), (3, True, True),
                                               (1, True, False)])   # last: raw-bf16 stage slots
def test_sharded_step_matches_oracle(pkg, world, chunks, opt, workers, bf16, mean):
    from kubeflow_controller_b200 import sharded
    elems = 100_003
    agg = pkg.AGG_MEAN if mean else pkg.AGG_SUM
    ranks = sharded.local_group(world, elems, opt,
                                lambda: pkg.HParams(lr=0.01, momentum=0.9, aggregate=agg),
                                n_chunks=chunks)
    assert sum(n for _, n in ranks[0].slices()) == ranks[0].shard
    padded, shard = ranks[0].padded, ranks[0].shard
    assert padded % (8 * world) == 0 and padded >= elems
    rng = np.random.default_rng(world * 10 + workers)
    var = rng.standard_normal(padded).astype(F)
    s0 = np.zeros(padded, F)
    s1 = np.zeros(padded, F)
    for x in ranks:
        x.var.copy_(_dev(var))
    ohp = O.HParams(lr=0.01, momentum=0.9)
    for step in range(3):
        gpr = []
        for r in range(world):
            g = [(0.01 * rng.standard_normal(padded)).astype(F) for _ in range(workers)]
            if bf16:
                g = [O.f32_to_bf16_bits(x) for x in g]
            gpr.append(g)
        sharded.local_group_step(ranks, [[_dev(g) for g in gr] for gr in gpr])
        O.numpy_apply(opt, var, s0, s1, _oracle_hier(gpr, mean), ohp)
        ohp.advance()
    torch.cuda.synchronize()
    for x in ranks:
        # every replica holds the whole updated arena (the pull is an all-gather)
        assert np.array_equal(x.var.cpu().numpy(), var), f"rank {x.rank} replica differs"
        lo = x.rank * shard
        if opt >= O.MOMENTUM:
            assert np.array_equal(x.slot0.cpu().numpy(), s0[lo:lo + shard])
        if opt == O.ADAM:
            assert np.array_equal(x.slot1.cpu().numpy(), s1[lo:lo + shard])
        assert x.global_step == 3


def test_sharded_clip_needs_global_sumsq(pkg):
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 4096, pkg.OPT_SGD, lambda: pkg.HParams(lr=0.1, clip_norm=0.5))
    g = [[torch.ones(ranks[0].padded, device=DEV)] for _ in ranks]
    for x, gr in zip(ranks, g):
        x.push(gr)
    with pytest.raises(pkg.AccelError, match="grad_sumsq_dev"):
        ranks[0].apply_pull(2)
    # with the global sum of squares the clip is the TF formula on the summed gradient
    ss = torch.tensor([float(ranks[0].padded) * 4.0], device=DEV)   # (1+1)^2 per element
    for x in ranks:
        x.apply_pull(2, sumsq=ss)
    for x in ranks:
        x.wait_pull()
        x.check()
    scale = O.numpy_clip_scale(F(ss.item()), 0.5)
    want = F(0.0) - (F(2.0) * scale) * F(0.1)
    assert torch.all(ranks[1].var == float(want))


def test_wait_times_out_instead_of_hanging(pkg):
    """A rank whose peer never pushes must come back with an error, not hang the GPU."""
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 4096, pkg.OPT_SGD, lambda: pkg.HParams(lr=0.1))
    ranks[0].push([torch.ones(ranks[0].padded, device=DEV)])
    ranks[0].apply_pull(2)            # rank 1 never pushed: bounded wait (2 s), then bail out
    with pytest.raises(pkg.AccelError, match="timed out"):
        ranks[0].check()
    assert not ranks[0].var.any(), "a timed-out step must not touch the variables"


def test_shard_argument_errors(pkg):
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 4096, pkg.OPT_ADAM, lambda: pkg.HParams())
    x = ranks[0]
    with pytest.raises(pkg.AccelError, match="before any"):
        x.wait_pull()
    with pytest.raises(pkg.AccelError, match="aligned"):
        x.push([torch.ones(x.padded + 1, device=DEV)[1:]])
    bad = pkg.accel.Peers(0, 2, 12, x._stage_arr, x._var_arr, x._flags_arr, None, None)
    with pytest.raises(pkg.AccelError, match="multiple of 8"):
        x.push([torch.ones(x.padded, device=DEV)], peers=bad)
    with pytest.raises(pkg.AccelError, match="sub-range"):
        x.push([torch.ones(x.padded, device=DEV)], sub=(8, x.shard))
    with pytest.raises(pkg.AccelError, match="sub-range"):
        x.push([torch.ones(x.padded, device=DEV)], sub=(4, 8))


def test_single_rank_pipelined_step_call(pkg):
    """accel_shard_step itself (two streams, slices) with a world of one: no peer to wait for."""
    from kubeflow_controller_b200 import sharded
    n = 300_000
    x = sharded.local_group(1, n, pkg.OPT_MOMENTUM,
                            lambda: pkg.HParams(lr=0.1, momentum=0.9), n_chunks=3)[0]
    rng = np.random.default_rng(1)
    var = rng.standard_normal(x.padded).astype(F)
    acc = np.zeros(x.padded, F)
    x.var.copy_(_dev(var))
    for _ in range(3):
        g = [(0.01 * rng.standard_normal(x.padded)).astype(F) for _ in range(2)]
        x.step([_dev(a) for a in g])
        O.numpy_ps_step(O.MOMENTUM, var, acc, None, g, O.HParams(lr=0.1, momentum=0.9))
    x.check()
    assert np.array_equal(x.var.cpu().numpy(), var)
    assert np.array_equal(x.slot0.cpu().numpy(), acc)


@pytest.mark.parametrize("world,slices", [(1, 1), (2, 1), (2, 4), (4, 7), (8, 16), (8, 64)])
@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
@pytest.mark.parametrize("workers,bf16,mean", [(1, False, False), (8, False, True), (5, True, True),
                                               (1, True, True)])    # last: raw-bf16 stage slots
@pytest.mark.parametrize("overlap", [0, 1, 2])
def test_one_kernel_step_matches_oracle(pkg, world, slices, opt, workers, bf16, mean, overlap):
    """The ONE-kernel step (push, apply, pull and closing wait overlapped inside a cooperative
    launch), all ranks played by one launch on this GPU: bit-exact against the oracle.
    overlap 2 = the in-switch protocol (summers / reducers) in its peer-load form: a single
    GPU has no NVSwitch, so the reduce is `world` loads in rank order — same bits."""
    from kubeflow_controller_b200 import sharded
    elems = 100_003
    agg = pkg.AGG_MEAN if mean else pkg.AGG_SUM
    ranks = sharded.local_group(world, elems, opt,
                                lambda: pkg.HParams(lr=0.01, momentum=0.9, aggregate=agg))
    padded, shard = ranks[0].padded, ranks[0].shard
    rng = np.random.default_rng(world * 100 + slices)
    var = rng.standard_normal(padded).astype(F)
    s0, s1 = np.zeros(padded, F), np.zeros(padded, F)
    for x in ranks:
        x.var.copy_(_dev(var))
    ohp = O.HParams(lr=0.01, momentum=0.9)
    for step in range(3):
        gpr = []
        for r in range(world):
            g = [(0.01 * rng.standard_normal(padded)).astype(F) for _ in range(workers)]
            if bf16:
                g = [O.f32_to_bf16_bits(x) for x in g]
            gpr.append(g)
        sharded.local_group_step_fused(ranks, [[_dev(g) for g in gr] for gr in gpr],
                                       n_slices=slices, overlap=overlap)
        O.numpy_apply(opt, var, s0, s1, _oracle_hier(gpr, mean), ohp)
        ohp.advance()
    torch.cuda.synchronize()
    for x in ranks:
        assert np.array_equal(x.var.cpu().numpy(), var), f"rank {x.rank} replica differs"
        lo = x.rank * shard
        if opt >= O.MOMENTUM:
            assert np.array_equal(x.slot0.cpu().numpy(), s0[lo:lo + shard])
        if opt == O.ADAM:
            assert np.array_equal(x.slot1.cpu().numpy(), s1[lo:lo + shard])


def test_one_kernel_step_then_three_phase_step_share_epochs(pkg):
    """Both step forms advance the same epoch counter, so they can be mixed."""
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 40_000, pkg.OPT_SGD, lambda: pkg.HParams(lr=1.0))
    g = [[torch.ones(ranks[0].padded, device=DEV)] for _ in ranks]
    sharded.local_group_step_fused(ranks, g, n_slices=3)
    sharded.local_group_step(ranks, g)
    sharded.local_group_step_fused(ranks, g, n_slices=5)
    assert torch.all(ranks[0].var == -6.0) and torch.all(ranks[1].var == -6.0)


@pytest.mark.parametrize("world,one_kernel", [(2, False), (4, True), (8, True)])
@pytest.mark.parametrize("opt", [O.SGD, O.ADAM])
def test_bf16_pull(pkg, world, one_kernel, opt):
    """Pull in bf16 (SURVEY 8 f, rank 2): every rank's bf16 replica holds the updated variables
    rounded to nearest-even; the fp32 master exists only in the owner's shard."""
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(world, 200_003, opt, lambda: pkg.HParams(lr=0.01),
                                n_chunks=2, pull_dtype="bf16")
    padded, shard = ranks[0].padded, ranks[0].shard
    rng = np.random.default_rng(world)
    var = rng.standard_normal(padded).astype(F)
    s0, s1 = np.zeros(padded, F), np.zeros(padded, F)
    for x in ranks:
        x.var.copy_(_dev(var))
    ohp = O.HParams(lr=0.01)
    for step in range(2):
        gpr = [[(0.01 * rng.standard_normal(padded)).astype(F) for _ in range(2)]
               for _ in range(world)]
        dg = [[_dev(g) for g in gr] for gr in gpr]
        if one_kernel:
            sharded.local_group_step_fused(ranks, dg, n_slices=3)
        else:
            sharded.local_group_step(ranks, dg)
        O.numpy_apply(opt, var, s0, s1, _oracle_hier(gpr, False), ohp)
        ohp.advance()
    want16 = O.f32_to_bf16_bits(var)
    for x in ranks:
        got16 = x.var_bf16.view(torch.int16).cpu().numpy().view(np.uint16)
        assert np.array_equal(got16, want16), f"rank {x.rank} bf16 replica differs"
        lo, hi = x.rank * shard, (x.rank + 1) * shard
        assert np.array_equal(x.var[lo:hi].cpu().numpy(), var[lo:hi])   # fp32 master, own shard


def test_sharded_global_norm_clip(pkg):
    """clip_by_global_norm across shards: per-shard sums of squares of the aggregated gradient
    (wait_push + accel_grad_sumsq over the stage slots), added across ranks, then apply_pull."""
    from kubeflow_controller_b200 import sharded
    world, workers = 4, 2
    ranks = sharded.local_group(world, 120_000, pkg.OPT_MOMENTUM,
                                lambda: pkg.HParams(lr=0.1, momentum=0.9, clip_norm=0.05))
    padded = ranks[0].padded
    rng = np.random.default_rng(4)
    var = rng.standard_normal(padded).astype(F)
    for x in ranks:
        x.var.copy_(_dev(var))
    gpr = [[(0.01 * rng.standard_normal(padded)).astype(F) for _ in range(workers)]
           for _ in range(world)]
    epochs = [x.push([_dev(g) for g in gr]) for x, gr in zip(ranks, gpr)]
    parts = torch.zeros(world, device=DEV)
    for x in ranks:
        x.wait_push(epoch=epochs[0])
        x.shard_sumsq(parts[x.rank:x.rank + 1])
    total = parts.sum().reshape(1)                    # the all-reduce of a real group
    for x in ranks:
        x.apply_pull(world * workers, total, epoch=epochs[0])
    for x in ranks:
        x.wait_pull(epoch=epochs[0])
        x.check()
    g = _oracle_hier(gpr, False)
    ref_ss = float(O.numpy_sumsq(g))
    assert abs(total.item() - ref_ss) <= 1e-5 * ref_ss
    acc = np.zeros(padded, F)
    O.numpy_apply(O.MOMENTUM, var, acc, None, g * O.numpy_clip_scale(F(total.item()), 0.05),
                  O.HParams(lr=0.1, momentum=0.9))
    for x in ranks:     # with the SAME sum of squares the rest is bit-exact
        assert np.array_equal(x.var.cpu().numpy(), var)


def test_raw_bf16_stage_is_what_travels(pkg):
    """One bf16 gradient per rank: the stage slots hold the raw bf16 items (2 B/elem), and the
    result equals the fp32-slot path bit for bit."""
    from kubeflow_controller_b200 import sharded
    world, elems = 4, 50_000
    rng = np.random.default_rng(11)
    outs = []
    for b16 in (True, False):
        ranks = sharded.local_group(world, elems, pkg.OPT_ADAM, lambda: pkg.HParams(lr=0.01),
                                    bf16_stage=b16)
        padded, shard = ranks[0].padded, ranks[0].shard
        rng = np.random.default_rng(11)
        var = rng.standard_normal(padded).astype(F)
        for x in ranks:
            x.var.copy_(_dev(var))
        gpr = [[O.f32_to_bf16_bits((0.01 * rng.standard_normal(padded)).astype(F))]
               for _ in range(world)]
        sharded.local_group_step(ranks, [[_dev(g) for g in gr] for gr in gpr])
        if b16:
            # slot s of rank r = rank s's gradient for shard r, as bf16, packed `shard` apart
            r = 1
            raw = ranks[r].stage.view(torch.int16).cpu().numpy().view(np.uint16)
            for s_ in range(world):
                assert np.array_equal(raw[s_ * shard:(s_ + 1) * shard],
                                      gpr[s_][0][r * shard:(r + 1) * shard])
        outs.append(ranks[0].var.cpu().numpy())
    assert np.array_equal(outs[0], outs[1])


def test_reset_after_timeout(pkg):
    """After a provoked timeout the flag block is unusable until accel_shard_reset; after the
    reset (every rank) the next step succeeds and is correct."""
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 4096, pkg.OPT_SGD, lambda: pkg.HParams(lr=1.0))
    ones = [[torch.ones(ranks[0].padded, device=DEV)] for _ in ranks]
    ranks[0].push(ones[0])
    ranks[0].apply_pull(2)                      # rank 1 never pushed
    with pytest.raises(pkg.AccelError, match="timed out"):
        ranks[0].check()
    with pytest.raises(pkg.AccelError, match="timed out"):
        ranks[0].check()                        # sticky until reset
    for x in ranks:
        x.reset()
    sharded.local_group_step(ranks, ones)
    sharded.local_group_step_fused(ranks, ones, n_slices=2)
    for x in ranks:
        assert torch.all(x.var == -4.0)


def test_one_kernel_timeout_leaves_tickets_clean(pkg):
    """A one-kernel step in which a rank never shows up times out, and afterwards every ticket
    word is back at zero (all CTAs reach the closing ticket even when their waits failed)."""
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 40_000, pkg.OPT_SGD, lambda: pkg.HParams(lr=1.0))
    x = ranks[0]
    g = [torch.ones(x.padded, device=DEV)]
    # rank 0 alone runs the real one-kernel entry point against a peer that never launches
    pkg.accel.check(x.lib.accel_shard_step_fused(
        x.ctx.handle, __import__("ctypes").byref(x.peers), x.optimizer, None, None,
        pkg.accel._grad_array(g), 1, pkg.F32, 2, __import__("ctypes").byref(x.hp), None, 2, 0,
        pkg.accel._stream_handle(None)))
    with pytest.raises(pkg.AccelError, match="timed out"):
        x.check()
    words = x.flags.cpu().numpy()
    assert words[32] == 0 and words[33] == 0 and not words[64:128].any(), "tickets not released"
    for r in ranks:
        r.reset()
    ones = [[torch.ones(x.padded, device=DEV)] for _ in ranks]
    sharded.local_group_step_fused(ranks, ones, n_slices=2, overlap=2)
    assert torch.all(ranks[1].var[ranks[1].shard:] == -2.0)


@pytest.mark.parametrize("opt,bf16,workers,slices", [(O.SGD, False, 3, 1), (O.ADAM, False, 2, 5),
                                                     (O.MOMENTUM, True, 1, 4), (O.ADAM, True, 2, 16)])
def test_shard_step_host_world_of_one(pkg, opt, bf16, workers, slices):
    """accel_shard_step_host (pinned host gradients -> sliced H2D -> push/apply_pull -> D2H of
    the replica) with a world of one: the whole pipeline, no peer to wait for."""
    from kubeflow_controller_b200 import sharded
    n = 300_007
    x = sharded.local_group(1, n, opt, lambda: pkg.HParams(lr=0.01, momentum=0.9))[0]
    rng = np.random.default_rng(5)
    var = rng.standard_normal(x.padded).astype(F)
    s0, s1 = np.zeros(x.padded, F), np.zeros(x.padded, F)
    x.var.copy_(_dev(var))
    ohp = O.HParams(lr=0.01, momentum=0.9)
    out = torch.empty(x.padded, dtype=torch.float32).pin_memory()
    for _ in range(3):
        g = [(0.01 * rng.standard_normal(x.padded)).astype(F) for _ in range(workers)]
        if bf16:
            g = [O.f32_to_bf16_bits(a) for a in g]
            hg = [torch.from_numpy(a.view(np.int16)).view(torch.bfloat16).pin_memory() for a in g]
        else:
            hg = [torch.from_numpy(a).pin_memory() for a in g]
        x.step_host(hg, out, n_slices=slices)
        O.numpy_apply(opt, var, s0, s1, O.numpy_aggregate(g, False), ohp)
        ohp.advance()
        assert np.array_equal(out.numpy(), var), "host copy of the replica differs"
    assert np.array_equal(x.var.cpu().numpy(), var)
    with pytest.raises(pkg.AccelError, match="clip"):
        x.hp.clip_norm = 1.0
        x.step_host(hg, out)


def test_rank_runs_on_the_device_it_was_given(pkg):
    """ShardedPS(rank=1, device_index=0): the default context must map task 1 to GPU 0 (local
    rank != rank on multi-node jobs), not to GPU 1."""
    from kubeflow_controller_b200 import sharded
    x = sharded.ShardedPS(1, 2, 0, 4096, pkg.OPT_SGD, pkg.HParams(lr=1.0),
                          peer_ptrs=[(0, 0, 0, 0)] * 2)
    assert x.ctx.gpu_of(1) == 0


@pytest.mark.parametrize("world,slices", [(2, 1), (4, 3), (8, 8)])
@pytest.mark.parametrize("bf16", [False, True])
def test_in_switch_step_in_place(pkg, world, slices, bf16):
    """Zero-copy form of the in-switch protocol: every rank's single gradient is written straight
    into its own stage buffer and passed as grads[0]; nothing is copied, the reduce reads it
    there. Same bits as the copying form (peer-load emulation on one GPU)."""
    from kubeflow_controller_b200 import sharded
    opt = O.ADAM
    ranks = sharded.local_group(world, 60_001, opt, lambda: pkg.HParams(lr=0.01))
    padded, shard = ranks[0].padded, ranks[0].shard
    rng = np.random.default_rng(world + slices)
    var = rng.standard_normal(padded).astype(F)
    s0, s1 = np.zeros(padded, F), np.zeros(padded, F)
    for x in ranks:
        x.var.copy_(_dev(var))
    ohp = O.HParams(lr=0.01)
    for step in range(2):
        gpr = [[(0.01 * rng.standard_normal(padded)).astype(F)] for _ in range(world)]
        if bf16:
            gpr = [[O.f32_to_bf16_bits(g[0])] for g in gpr]
        dg = []
        for x, g in zip(ranks, gpr):
            view = x.stage_grad(torch.bfloat16 if bf16 else torch.float32)
            view.copy_(_dev(g[0]))
            dg.append([view])
        sharded.local_group_step_fused(ranks, dg, n_slices=slices, overlap=2)
        O.numpy_apply(opt, var, s0, s1, _oracle_hier(gpr, False), ohp)
        ohp.advance()
    for x in ranks:
        assert np.array_equal(x.var.cpu().numpy(), var), f"rank {x.rank} replica differs"


def test_gradient_inside_stage_is_rejected_by_unicast_forms(pkg):
    from kubeflow_controller_b200 import sharded
    ranks = sharded.local_group(2, 4096, pkg.OPT_SGD, lambda: pkg.HParams(lr=1.0))
    dg = [[x.stage_grad()] for x in ranks]
    with pytest.raises(pkg.AccelError, match="stage buffer"):
        ranks[0].push(dg[0])
    with pytest.raises(pkg.AccelError, match="in place"):
        sharded.local_group_step_fused(ranks, dg, n_slices=1, overlap=0)
    with pytest.raises(pkg.AccelError, match="stage buffer"):
        ranks[0].push([ranks[0].stage[ranks[0].shard:]])     # anywhere inside it, not just the base
