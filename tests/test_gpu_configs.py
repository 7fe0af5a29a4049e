"""BASELINE configs 3, 4 and 5 at their full sizes through the C ABI on a B200, against the
threaded C oracle (bit-exact) and through size-independent properties."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import ps_oracle as O

pytestmark = pytest.mark.gpu
F = np.float32
DEV = "cuda:0"


def _sizes(name):
    shapes = json.load(open(os.path.join(os.path.dirname(__file__), "golden",
                                         "model_shapes.json")))[name]
    return [int(np.prod(s)) if s else 1 for _, s in shapes]


def _randn(n, seed, scale=1.0, bf16=False):
    gen = torch.Generator(device=DEV)
    gen.manual_seed(seed)
    t = torch.randn(n, device=DEV, generator=gen) * scale
    return t.to(torch.bfloat16) if bf16 else t


def _host(t):
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).cpu().numpy().view(np.uint16)
    return t.cpu().numpy()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_config3_resnet50_momentum_sharded(pkg, world):
    """ResNet-50-shaped gradients (161 tensors, 25,557,032 fp32), Momentum lr 0.1 / 0.9, PS
    sharded over P ranks with one worker per rank, flat arena; two steps, one-kernel path."""
    from kubeflow_controller_b200 import sharded
    elems = sum(_sizes("resnet50"))
    ranks = sharded.local_group(world, elems, pkg.OPT_MOMENTUM,
                                lambda: pkg.HParams(lr=0.1, momentum=0.9))
    padded, shard = ranks[0].padded, ranks[0].shard
    assert padded - elems < 8 * world and shard % 8 == 0
    var = _randn(padded, 42)
    var[elems:] = 0                      # padding: zero state ...
    for x in ranks:
        x.var.copy_(var)
    h_var, h_acc = _host(var), np.zeros(padded, F)
    orc = O.COracle()
    for step in range(2):
        grads = []
        for r in range(world):
            g = _randn(padded, 1234 + r + 100 * step, 0.01)
            g[elems:] = 0                # ... and zero gradient, so it never moves
            grads.append([g])
        sharded.local_group_step_fused(ranks, grads, n_slices=2)
        orc.ps_step(O.MOMENTUM, h_var, h_acc, None, [_host(g[0]) for g in grads],
                    O.HParams(lr=0.1, momentum=0.9))
    for x in ranks:
        assert np.array_equal(_host(x.var), h_var)
        assert np.array_equal(_host(x.slot0), h_acc[x.rank * shard:(x.rank + 1) * shard])
    assert not h_var[elems:].any()


def test_config3_round_robin_by_variable(pkg):
    """The same model placed the reference's way: variable k on PS task k mod 8, each task's
    bucket applied with one launch; every variable must match the oracle."""
    sizes = _sizes("resnet50")
    n_ps, workers = 8, 2
    rng = np.random.default_rng(3)
    full = [rng.standard_normal(s).astype(F) for s in sizes]
    grads = [[(0.01 * rng.standard_normal(s)).astype(F) for s in sizes] for _ in range(workers)]
    want, acc = [v.copy() for v in full], [np.zeros(s, F) for s in sizes]
    for k in range(len(sizes)):
        O.numpy_ps_step(O.MOMENTUM, want[k], acc[k], None, [g[k] for g in grads],
                        O.HParams(lr=0.1, momentum=0.9))
    seen = 0
    for t in range(n_ps):
        ps = pkg.ParameterServer(task_index=t, n_ps=n_ps, var_elems=sizes,
                                 optimizer=pkg.OPT_MOMENTUM,
                                 hparams=pkg.HParams(lr=0.1, momentum=0.9), align_elems=8,
                                 gpu_ordinals=[0] * n_ps)
        arenas = [torch.zeros(ps.count, device=DEV) for _ in range(workers)]
        for k, off, n in ps.layout:
            ps.var[off:off + n].copy_(torch.from_numpy(full[k]))
            for w in range(workers):
                arenas[w][off:off + n].copy_(torch.from_numpy(grads[w][k]))
        ps.apply_gradients(arenas)
        for k, off, n in ps.layout:
            assert k % n_ps == t
            assert np.array_equal(ps.variable(k).cpu().numpy(), want[k])
            seen += 1
    assert seen == len(sizes)


@pytest.mark.parametrize("t_step", [1, 1000])
def test_config4_bert_large_bf16_adam_8_shards(pkg, t_step):
    """BERT-large-shaped gradients (391 tensors, 335,141,888), bf16 gradients -> fp32 Adam
    (lr 1e-4, 0.9 / 0.999 / 1e-8), 8 PS shards, at step t = 1 and t = 1000."""
    from kubeflow_controller_b200 import sharded
    world = 8
    elems = sum(_sizes("bert_large"))
    assert elems == 335_141_888
    b1p, b2p = F(0.9), F(0.999)
    for _ in range(t_step - 1):
        b1p, b2p = b1p * F(0.9), b2p * F(0.999)
    hp = dict(lr=1e-4, beta1=0.9, beta2=0.999, epsilon=1e-8, beta1_power=float(b1p),
              beta2_power=float(b2p))
    ranks = sharded.local_group(world, elems, pkg.OPT_ADAM, lambda: pkg.HParams(**hp))
    padded, shard = ranks[0].padded, ranks[0].shard
    assert shard == 41_892_736
    var = _randn(padded, 42)
    m = _randn(padded, 43, 0.01)
    v = _randn(padded, 44, 0.01).square_()
    for x in ranks:
        x.var.copy_(var)
        x.slot0.copy_(m[x.rank * shard:(x.rank + 1) * shard])
        x.slot1.copy_(v[x.rank * shard:(x.rank + 1) * shard])
    grads = [[_randn(padded, 1234 + r, 0.01, bf16=True)] for r in range(world)]
    sharded.local_group_step_fused(ranks, grads, n_slices=1)
    h_var, h_m, h_v = _host(var), _host(m), _host(v)
    O.COracle().ps_step(O.ADAM, h_var, h_m, h_v, [_host(g[0]) for g in grads], O.HParams(**hp))
    for x in (ranks[0], ranks[5]):       # every replica is written by the same stores
        assert np.array_equal(_host(x.var), h_var)
    for x in ranks:
        lo = x.rank * shard
        assert np.array_equal(_host(x.slot0), h_m[lo:lo + shard])
        assert np.array_equal(_host(x.slot1), h_v[lo:lo + shard])
    assert ranks[0].hp.beta1_power == float(b1p * F(0.9))    # advanced once per step


@pytest.mark.parametrize("mib", [1, 4, 16, 64, 256, 1024, 4096])
@pytest.mark.parametrize("opt", [O.SGD, O.ADAM])
def test_config5_size_sweep(pkg, mib, opt):
    """1 MiB - 4 GiB of fp32 variables per shard, one gradient (the post-collective apply).
    Up to 1 GiB the whole result is compared with the C oracle; at 4 GiB a 64 Mi-element
    window is, plus the properties: zero lr is the identity, and the update's checksum equals
    lr times the gradient's checksum (SGD)."""
    n = mib << 18
    ctx = pkg.Context(1)
    var, g = _randn(n, 7), _randn(n, 8, 0.01)
    m = torch.zeros(n, device=DEV) if opt == O.ADAM else None
    v = torch.zeros(n, device=DEV) if opt == O.ADAM else None
    var0 = var.clone() if mib <= 1024 else None
    hp_kw = dict(lr=0.01)
    if opt == O.SGD:
        ctx.apply_sgd(var, [g], pkg.HParams(lr=0.0))
        if var0 is not None:
            assert torch.equal(var, var0)
    win = slice(0, n) if mib <= 1024 else slice(n - (64 << 20), n)
    h_var = _host(var[win])
    h_m = np.zeros(h_var.size, F)
    h_v = np.zeros(h_var.size, F)
    before = var.double().sum().item() if opt == O.SGD else None
    ctx.apply(opt, var, m, v, [g], pkg.HParams(**hp_kw))
    torch.cuda.synchronize()
    O.COracle().ps_step(opt, h_var, h_m, h_v, [_host(g[win])], O.HParams(**hp_kw))
    assert np.array_equal(_host(var[win]), h_var)
    if opt == O.ADAM:
        assert np.array_equal(_host(m[win]), h_m) and np.array_equal(_host(v[win]), h_v)
    else:
        delta = before - var.double().sum().item()
        want = 0.01 * g.double().sum().item()
        assert abs(delta - want) <= 1e-4 * max(1.0, abs(want)) + 1e-3 * (n / 1e8)


@pytest.mark.parametrize("world", [2, 8])
def test_config3_round_robin_buckets_through_the_sharded_step(pkg, world):
    """Config 3 with the reference's placement INSIDE the multi-GPU step: variable k on PS task
    k mod P (mnist_replica.py:137-141), buckets padded to equal length so bucket r = shard r,
    one worker gradient per rank, the one-kernel step. Every variable must match the oracle
    applied to that variable alone, and must live in the shard of rank k mod P."""
    from kubeflow_controller_b200 import sharded
    sizes = _sizes("resnet50")
    lay = sharded.BucketLayout(sizes, world)
    assert lay.arena == lay.shard * world and lay.shard >= max(lay.fill)
    for k in range(len(sizes)):
        assert lay.pos[k] // lay.shard == k % world
        assert (lay.pos[k] + sizes[k] - 1) // lay.shard == k % world      # never straddles
    ranks = sharded.local_group(world, lay.arena, pkg.OPT_MOMENTUM,
                                lambda: pkg.HParams(lr=0.1, momentum=0.9))
    assert ranks[0].padded == lay.arena and ranks[0].shard == lay.shard
    rng = np.random.default_rng(8)
    var = [rng.standard_normal(s).astype(F) for s in sizes]
    acc = [np.zeros(s, F) for s in sizes]
    arena = lay.pack([torch.from_numpy(v).to(DEV) for v in var])
    for x in ranks:
        x.var.copy_(arena)
    for step in range(2):
        per_rank = [[(0.01 * rng.standard_normal(s)).astype(F) for s in sizes]
                    for _ in range(world)]
        dg = [[lay.pack([torch.from_numpy(g).to(DEV) for g in gr])] for gr in per_rank]
        sharded.local_group_step_fused(ranks, dg, n_slices=2)
        for k in range(len(sizes)):
            O.numpy_ps_step(O.MOMENTUM, var[k], acc[k], None, [gr[k] for gr in per_rank],
                            O.HParams(lr=0.1, momentum=0.9))
    for x in ranks:
        for k in range(len(sizes)):
            assert np.array_equal(lay.view(x.var, k).cpu().numpy(), var[k]), f"variable {k}"
    # optimizer slots stay with the owner: variable k's accumulator is in rank (k mod P)'s slot0
    for k in (0, 1, len(sizes) - 1):
        owner = ranks[k % world]
        got = owner.slot0[lay.offset[k]:lay.offset[k] + sizes[k]].cpu().numpy()
        assert np.array_equal(got, acc[k])
