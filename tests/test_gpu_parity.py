"""Parity of the CUDA path with the CPU oracle, through the C ABI, on a B200.

Bar (BASELINE.json): fp32 within rel 1e-5, bf16 gradients within rel 1e-2. The
kernels evaluate the TF-1.4 expressions with the same operation order and
round-to-nearest per operation as the oracle, so these tests assert the stronger
property — bit-identical fp32 results — wherever the reduction order is defined
(everything except the sum of squares for clipping, compared at rel 1e-5).
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import ps_oracle as O

pytestmark = pytest.mark.gpu
F = np.float32
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ctx(pkg):
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device; there is no CPU fallback")
    return pkg.Context(1)


def _inputs(n, workers, bf16, seed):
    rng = np.random.default_rng(seed)
    var = rng.standard_normal(n).astype(F)
    s0 = (0.1 * rng.standard_normal(n)).astype(F)
    s1 = np.abs(0.01 * rng.standard_normal(n)).astype(F)
    grads = [(0.01 * rng.standard_normal(n)).astype(F) for _ in range(workers)]
    if bf16:
        grads = [O.f32_to_bf16_bits(g) for g in grads]
    return var, s0, s1, grads


def _to_dev(a):
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16)).to(DEV).view(torch.bfloat16)
    return torch.from_numpy(a).to(DEV)


def _hp_pair(pkg, **kw):
    o = O.HParams(**kw)
    a = pkg.HParams(lr=o.lr, momentum=o.momentum, beta1=o.beta1, beta2=o.beta2, epsilon=o.epsilon,
                    beta1_power=o.beta1_power, beta2_power=o.beta2_power, clip_norm=o.clip_norm,
                    use_nesterov=o.use_nesterov, aggregate=o.aggregate)
    return o, a


def _run_both(pkg, ctx, opt, n, workers, bf16, seed, sumsq_given=None, **hpkw):
    var, s0, s1, grads = _inputs(n, workers, bf16, seed)
    ohp, ahp = _hp_pair(pkg, **hpkw)
    d_var, d_s0, d_s1 = _to_dev(var), _to_dev(s0), _to_dev(s1)
    d_g = [_to_dev(g) for g in grads]
    d_ss = None
    if sumsq_given is not None:
        d_ss = torch.tensor([sumsq_given], dtype=torch.float32, device=DEV)
    ctx.apply(opt, d_var, d_s0, d_s1, d_g, ahp, sumsq=d_ss)
    torch.cuda.synchronize()
    O.numpy_ps_step(opt, var, s0, s1, grads, ohp, sumsq=sumsq_given)
    return (d_var.cpu().numpy(), d_s0.cpu().numpy(), d_s1.cpu().numpy()), (var, s0, s1)


def _assert_bits(got, want, opt):
    assert np.array_equal(got[0], want[0]), f"var differs in {(got[0] != want[0]).sum()} elems"
    if opt >= O.MOMENTUM:
        assert np.array_equal(got[1], want[1])
    if opt == O.ADAM:
        assert np.array_equal(got[2], want[2])


# ---- known answers -------------------------------------------------------------
def test_kat_through_abi(pkg, ctx, kat):
    g = [np.array(x, F) for x in kat["inputs"]["grads"]]
    d_g = [_to_dev(x) for x in g]
    var0 = np.array(kat["inputs"]["var"], F)

    v = _to_dev(var0)
    ctx.apply_sgd(v, d_g, pkg.HParams(lr=kat["sgd"]["lr"]))
    assert np.array_equal(v.cpu().numpy(), np.array(kat["sgd"]["var"], F))

    k = kat["momentum"]
    for nest, key in ((False, "var"), (True, "var_nesterov")):
        v, a = _to_dev(var0), _to_dev(np.array(k["accum0"], F))
        ctx.apply_momentum(v, a, d_g, pkg.HParams(lr=k["lr"], momentum=k["momentum"],
                                                  use_nesterov=nest))
        assert np.array_equal(a.cpu().numpy(), np.array(k["accum"], F))
        assert np.array_equal(v.cpu().numpy(), np.array(k[key], F))

    k = kat["adam"]
    v, m, vv = _to_dev(var0), _to_dev(np.zeros(4, F)), _to_dev(np.zeros(4, F))
    ctx.apply_adam(v, m, vv, d_g, pkg.HParams(lr=k["lr"], beta1=k["beta1"], beta2=k["beta2"],
                                              epsilon=k["epsilon"]))
    assert np.array_equal(m.cpu().numpy(), np.array(k["m"], F))
    assert np.array_equal(vv.cpu().numpy(), np.array(k["v"], F))
    assert np.array_equal(v.cpu().numpy(), np.array(k["var"], F))

    # mean aggregation of the same vectors: SGD on the mean
    v = _to_dev(var0)
    ctx.apply_sgd(v, d_g, pkg.HParams(lr=0.5, aggregate=pkg.AGG_MEAN))
    want = var0 - np.array(kat["mean"], F) * F(0.5)
    assert np.array_equal(v.cpu().numpy(), want)


def test_golden_fixture_through_abi(pkg, ctx):
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ps_step_golden.npz"))
    for name in ("sgd", "momentum", "adam_bf16_mean_clip"):
        opt = int(z[f"{name}.opt"])
        grads = [z[f"{name}.g{w}"] for w in range(int(z[f"{name}.workers"]))]
        kw = {k: (float(x) if k not in ("use_nesterov", "aggregate") else int(x))
              for k, x in zip(z[f"{name}.hp_keys"], z[f"{name}.hp_vals"])}
        _, ahp = _hp_pair(pkg, **kw)
        d = [_to_dev(z[f"{name}.{k}"].copy()) for k in ("var", "s0", "s1")]
        ss = torch.tensor([float(z[f"{name}.sumsq"])], dtype=torch.float32, device=DEV)
        ctx.apply(opt, d[0], d[1], d[2], [_to_dev(g) for g in grads], ahp,
                  sumsq=ss if ahp.clip_norm > 0 else None)
        assert np.array_equal(d[0].cpu().numpy(), z[f"{name}.var_out"])
        if opt >= 1:
            assert np.array_equal(d[1].cpu().numpy(), z[f"{name}.s0_out"])
        if opt == 2:
            assert np.array_equal(d[2].cpu().numpy(), z[f"{name}.s1_out"])


# ---- sweep: optimizers x dtypes x worker counts x sizes x kernel variants ----------
SIZES = [1, 3, 4, 7, 8, 1023, 4096, 100_003, (1 << 20) + 5]


@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
@pytest.mark.parametrize("bf16", [False, True])
@pytest.mark.parametrize("workers", [1, 2, 3, 4, 8, 9, 17])
def test_bit_exact_sizes(pkg, ctx, opt, bf16, workers):
    ctx.set_kernel(pkg.KERNEL_AUTO)
    for n in SIZES:
        got, want = _run_both(pkg, ctx, opt, n, workers, bf16, seed=n + workers, lr=0.01,
                              momentum=0.9)
        _assert_bits(got, want, opt)


@pytest.mark.parametrize("kernel", ["LDG128", "LDG256", "TMA"])
@pytest.mark.parametrize("threads,ctas,unroll", [(0, 0, 0), (128, 1, 1), (256, 2, 2), (512, 1, 4)])
@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
def test_bit_exact_kernel_variants(pkg, ctx, kernel, threads, ctas, unroll, opt):
    ctx.set_kernel(getattr(pkg, "KERNEL_" + kernel), threads, ctas, unroll)
    try:
        for workers, bf16 in ((8, False), (8, True), (5, False), (1, True)):
            for n in (8, 200_000, (1 << 21) + 8 * 13):
                got, want = _run_both(pkg, ctx, opt, n, workers, bf16, seed=n, lr=0.05,
                                      momentum=0.5, use_nesterov=True, aggregate=O.AGG_MEAN)
                _assert_bits(got, want, opt)
    finally:
        ctx.set_kernel(pkg.KERNEL_AUTO)


def test_max_workers_and_limits(pkg, ctx):
    got, want = _run_both(pkg, ctx, O.SGD, 10_000, 64, False, seed=64, lr=0.01)
    _assert_bits(got, want, O.SGD)
    v = torch.zeros(64, device=DEV)
    g = [torch.zeros(64, device=DEV)] * 65
    with pytest.raises(pkg.AccelError, match="n_workers"):
        ctx.apply_sgd(v, g, pkg.HParams())
    with pytest.raises(pkg.AccelError, match="aligned"):
        ctx.apply_sgd(v[1:], [torch.zeros(63, device=DEV)], pkg.HParams())
    with pytest.raises(pkg.AccelError, match="aligned"):
        ctx.apply_sgd(v, [torch.zeros(65, device=DEV)[1:]], pkg.HParams())
    with pytest.raises(pkg.AccelError, match="slot"):
        ctx.apply(pkg.OPT_ADAM, v, v.clone(), None, [v.clone()], pkg.HParams())
    # empty input: a no-op, not an error
    ctx.apply_sgd(v, [torch.ones(64, device=DEV)], pkg.HParams(lr=1.0), count=0)
    torch.cuda.synchronize()
    assert not v.any()


@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
@pytest.mark.parametrize("mean", [False, True])
@pytest.mark.parametrize("nesterov", [False, True])
def test_modes(pkg, ctx, opt, mean, nesterov):
    got, want = _run_both(pkg, ctx, opt, 50_001, 3, False, seed=7, lr=0.1, momentum=0.9,
                          use_nesterov=nesterov, aggregate=O.AGG_MEAN if mean else O.AGG_SUM)
    _assert_bits(got, want, opt)


# ---- global-norm clipping ----------------------------------------------------------
@pytest.mark.parametrize("bf16", [False, True])
@pytest.mark.parametrize("workers,n", [(1, 5), (3, 100_003), (8, (1 << 22) + 3)])
def test_grad_sumsq(pkg, ctx, bf16, workers, n):
    _, _, _, grads = _inputs(n, workers, bf16, seed=n)
    d_g = [_to_dev(g) for g in grads]
    out = torch.zeros(2, device=DEV)
    for agg in (pkg.AGG_SUM, pkg.AGG_MEAN):
        ctx.grad_sumsq(d_g, out, aggregate=agg)
        first = out[0].item()
        ctx.grad_sumsq(d_g, out[1:], aggregate=agg)
        assert out[1].item() == first, "sum of squares is not reproducible run to run"
        ref = float(O.numpy_sumsq(O.numpy_aggregate(grads, agg == pkg.AGG_MEAN)))
        assert abs(first - ref) <= 1e-5 * abs(ref)


@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
@pytest.mark.parametrize("clip", [1e-3, 10.0])
def test_clip_with_given_sumsq_is_bit_exact(pkg, ctx, opt, clip):
    n, workers = 70_001, 4
    _, _, _, grads = _inputs(n, workers, False, seed=3)
    ss = float(O.numpy_sumsq(O.numpy_aggregate(grads, False)))
    got, want = _run_both(pkg, ctx, opt, n, workers, False, seed=3, sumsq_given=ss, lr=0.1,
                          momentum=0.9, clip_norm=clip)
    _assert_bits(got, want, opt)


@pytest.mark.parametrize("bf16", [False, True])
def test_clip_with_internal_sumsq(pkg, ctx, bf16):
    """clip_norm > 0 and no sum of squares given: the library runs its own reduction."""
    tol = 1e-2 if bf16 else 1e-5
    got, want = _run_both(pkg, ctx, O.ADAM, 300_007, 8, bf16, seed=5, lr=0.01, clip_norm=0.01)
    for a, b in zip(got, want):
        np.testing.assert_allclose(a, b, rtol=tol, atol=1e-9)


# ---- the host-buffer step (the reference-facing call) ------------------------------
@pytest.mark.parametrize("opt,bf16,workers,n", [
    (O.SGD, False, 2, 1000), (O.ADAM, True, 3, (9 << 20) + 77), (O.MOMENTUM, False, 8, (4 << 20))])
def test_host_step(pkg, ctx, opt, bf16, workers, n):
    var, s0, s1, grads = _inputs(n, workers, bf16, seed=n % 1000)
    ohp, ahp = _hp_pair(pkg, lr=0.01, momentum=0.9, aggregate=O.AGG_MEAN)
    d_var, d_s0, d_s1 = _to_dev(var), _to_dev(s0), _to_dev(s1)
    h_g = []
    for g in grads:
        t = torch.from_numpy(g.view(np.int16) if bf16 else g).pin_memory()
        h_g.append(t.view(torch.bfloat16) if bf16 else t)
    h_out = torch.empty(n, dtype=torch.float32).pin_memory()
    for _ in range(2):   # two steps: staging buffers and events are reused
        ctx.ps_step_host(opt, d_var, d_s0, d_s1, h_g, ahp, h_out)
        O.numpy_ps_step(opt, var, s0, s1, grads, ohp)
    assert np.array_equal(h_out.numpy(), var)
    assert np.array_equal(d_var.cpu().numpy(), var)
    with pytest.raises(pkg.AccelError, match="clip"):
        ctx.ps_step_host(opt, d_var, d_s0, d_s1, h_g, pkg.HParams(clip_norm=1.0), h_out)


# ---- BASELINE config 1: the example's model on 1 PS (and on 2 PS tasks) -------------
@pytest.mark.parametrize("n_ps", [1, 2])
def test_config1_mnist_plumbing(pkg, n_ps):
    """1 PS + 2 workers (and dist.yml's 2 PS), Adam lr 0.01 as mnist_replica.py:170,
    20 steps; variables placed by the round-robin of replica_device_setter."""
    sizes = [784 * 100, 100, 100 * 10, 10]
    workers, steps = 2, 20
    rng = np.random.default_rng(0)
    full_var = [rng.standard_normal(s).astype(F) * F(0.05) for s in sizes]
    servers = []
    for t in range(n_ps):
        argv = pkg.placement.cluster_spec(workers, n_ps, [f"w{i}" for i in range(workers)],
                                          [f"p{i}" for i in range(n_ps)], "PS", t)
        ps = pkg.ParameterServer(argv, var_elems=sizes, optimizer=pkg.OPT_ADAM,
                                 hparams=pkg.HParams(lr=0.01), first_task=1, align_elems=4,
                                 gpu_ordinals=[0] * n_ps)
        for k, off, n in ps.layout:
            ps.var[off:off + n].copy_(torch.from_numpy(full_var[k]))
        servers.append(ps)
    o_var = [v.copy() for v in full_var]
    o_m = [np.zeros_like(v) for v in full_var]
    o_v = [np.zeros_like(v) for v in full_var]
    ohp = O.HParams(lr=0.01)
    for step in range(steps):
        g = [[(0.01 * rng.standard_normal(s)).astype(F) for s in sizes] for _ in range(workers)]
        for ps in servers:
            arenas = []
            for w in range(workers):
                a = torch.zeros(max(ps.count, 1), dtype=torch.float32, device=ps.device)
                for k, off, n in ps.layout:
                    a[off:off + n].copy_(torch.from_numpy(g[w][k]))
                arenas.append(a)
            ps.apply_gradients(arenas)
        for k in range(len(sizes)):
            O.numpy_ps_step(O.ADAM, o_var[k], o_m[k], o_v[k], [g[w][k] for w in range(workers)], ohp)
        ohp.advance()
    torch.cuda.synchronize()
    seen = set()
    for ps in servers:
        assert ps.global_step == steps
        for k, off, n in ps.layout:
            assert np.array_equal(ps.variable(k).cpu().numpy(), o_var[k])
            seen.add(k)
    assert seen == set(range(len(sizes)))


# ---- BASELINE config 2 at full size --------------------------------------------------
def test_config2_full_size(pkg, ctx):
    """8 workers x 100 M fp32, SGD: the whole result against the threaded C oracle, bit for
    bit, plus size-independent properties (zero learning rate is the identity; the
    checksum of the update equals the checksum of the gradients)."""
    n, workers, lr = 100_000_000, 8, 0.01
    gen = torch.Generator(device=DEV)
    gen.manual_seed(42)
    var = torch.randn(n, device=DEV, generator=gen)
    grads = []
    for w in range(workers):
        gen.manual_seed(1234 + w)
        grads.append(torch.randn(n, device=DEV, generator=gen) * 0.01)
    var0 = var.clone()

    ctx.apply_sgd(var, grads, pkg.HParams(lr=0.0))
    assert torch.equal(var, var0), "lr = 0 must leave the variables untouched"

    ctx.apply_sgd(var, grads, pkg.HParams(lr=lr))
    torch.cuda.synchronize()
    gsum = sum(g.double().sum().item() for g in grads)
    delta = (var0.double() - var.double()).sum().item()
    assert abs(delta - lr * gsum) <= 1e-4 * max(1.0, abs(lr * gsum)) + 1e-3

    h_var = var0.cpu().numpy()
    h_g = [g.cpu().numpy() for g in grads]
    O.COracle().ps_step(O.SGD, h_var, None, None, h_g, O.HParams(lr=lr))
    got = var.cpu().numpy()
    assert np.array_equal(got, h_var), f"{(got != h_var).sum()} of {n} elements differ"


# ---- asynchronous applies, checkpoint / restore -------------------------------
def test_async_mode_applies_each_push_alone(pkg):
    """The example's default (sync_replicas=False): each worker push is applied by itself, in
    arrival order, Adam's beta powers advancing with every apply."""
    sizes = [784 * 100, 100, 100 * 10, 10]
    ps = pkg.ParameterServer(task_index=0, n_ps=1, var_elems=sizes, optimizer=pkg.OPT_ADAM,
                             hparams=pkg.HParams(lr=0.01), align_elems=4)
    rng = np.random.default_rng(0)
    var = (0.05 * rng.standard_normal(ps.count)).astype(F)
    ps.var.copy_(torch.from_numpy(var))
    m, v = np.zeros(ps.count, F), np.zeros(ps.count, F)
    ohp = O.HParams(lr=0.01)
    for push in range(12):                      # 4 workers x 3 local steps, interleaved
        g = (0.01 * rng.standard_normal(ps.count)).astype(F)
        ps.apply_async(torch.from_numpy(g).to(DEV))
        O.numpy_ps_step(O.ADAM, var, m, v, [g], ohp)
        ohp.advance()
    assert ps.global_step == 12
    assert np.array_equal(ps.pull().cpu().numpy(), var)


@pytest.mark.parametrize("device_state", [False, True])
def test_checkpoint_restore_resumes_bit_exact(pkg, tmp_path, device_state):
    sizes = [1000, 24, 4096]

    def make():
        return pkg.ParameterServer(task_index=0, n_ps=1, var_elems=sizes, optimizer=pkg.OPT_ADAM,
                                   hparams=pkg.HParams(lr=0.01), align_elems=8,
                                   device_state=device_state)
    rng = np.random.default_rng(1)
    a = make()
    a.var.copy_(torch.from_numpy(rng.standard_normal(a.count).astype(F)))
    grads = [[torch.from_numpy((0.01 * rng.standard_normal(a.count)).astype(F)).to(DEV)
              for _ in range(2)] for _ in range(6)]
    for g in grads[:3]:
        a.apply_gradients(g)
    a.save(str(tmp_path))
    for g in grads[3:]:
        a.apply_gradients(g)
    b = make()                                   # a fresh replica resumes from the checkpoint
    b.restore(str(tmp_path))
    assert b.global_step == 3 and b.hp.beta1_power == pytest.approx(0.9 ** 4, rel=1e-6)
    for g in grads[3:]:
        b.apply_gradients(g)
    assert torch.equal(a.var, b.var) and torch.equal(a.slot0, b.slot0) and torch.equal(a.slot1, b.slot1)
    if device_state:
        assert torch.equal(a.dev_state, b.dev_state) and b.device_global_step() == 6
    else:
        assert a.hp.beta2_power == b.hp.beta2_power and b.global_step == 6


def test_host_step_with_registered_numpy_buffers(pkg, ctx):
    """Gradients living in ordinary (numpy) host memory, page-locked through the ABI."""
    n, workers = 300_000, 3
    var, _, _, grads = _inputs(n, workers, False, seed=21)
    out = np.empty(n, F)
    for a in grads + [out]:
        pkg.accel.host_register(a)
    try:
        d_var = _to_dev(var)
        arr = (ctypes.c_void_p * workers)(*[g.ctypes.data for g in grads])
        hp = pkg.HParams(lr=0.01)
        pkg.accel.check(pkg.load_library().accel_ps_step_host(
            ctx.handle, 0, pkg.OPT_SGD, d_var.data_ptr(), None, None, arr, workers, pkg.F32, n,
            ctypes.byref(hp), out.ctypes.data))
        O.numpy_ps_step(O.SGD, var, None, None, grads, O.HParams(lr=0.01))
        assert np.array_equal(out, var)
    finally:
        for a in grads + [out]:
            pkg.accel.host_unregister(a)


# ---- optimizer state that advances on the device ------------------------
@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
def test_device_resident_optimizer_state(pkg, opt):
    sizes = [5000, 3, 70_000]
    ps = pkg.ParameterServer(task_index=0, n_ps=1, var_elems=sizes, optimizer=opt,
                             hparams=pkg.HParams(lr=0.01, momentum=0.9), align_elems=8,
                             device_state=True)
    rng = np.random.default_rng(2)
    var = rng.standard_normal(ps.count).astype(F)
    ps.var.copy_(torch.from_numpy(var))
    s0, s1 = np.zeros(ps.count, F), np.zeros(ps.count, F)
    ohp = O.HParams(lr=0.01, momentum=0.9)
    for step in range(25):
        g = [(0.01 * rng.standard_normal(ps.count)).astype(F) for _ in range(2)]
        ps.apply_gradients([torch.from_numpy(x).to(DEV) for x in g])
        O.numpy_ps_step(opt, var, s0, s1, g, ohp)
        ohp.advance()
    assert np.array_equal(ps.pull().cpu().numpy(), var)
    assert ps.device_global_step() == 25
    if opt == O.ADAM:
        b = ps.dev_state[:2].cpu().numpy()
        assert b[0] == F(ohp.beta1_power) and b[1] == F(ohp.beta2_power)


def test_recorded_step_replays_correctly(pkg):
    """With the optimizer state on the device a whole Adam step (apply + finish) is a fixed
    sequence of launches: record it once in a CUDA graph, replay it with new gradients."""
    n, workers = 200_000, 4
    ps = pkg.ParameterServer(task_index=0, n_ps=1, var_elems=[n], optimizer=pkg.OPT_ADAM,
                             hparams=pkg.HParams(lr=0.01), align_elems=8, device_state=True)
    rng = np.random.default_rng(3)
    var = rng.standard_normal(n).astype(F)
    ps.var.copy_(torch.from_numpy(var))
    m, v = np.zeros(n, F), np.zeros(n, F)
    ohp = O.HParams(lr=0.01)
    static_g = [torch.zeros(ps.count, device=DEV) for _ in range(workers)]

    def next_grads():
        g = [(0.01 * rng.standard_normal(n)).astype(F) for _ in range(workers)]
        for dst, src in zip(static_g, g):
            dst[:n].copy_(torch.from_numpy(src))
        O.numpy_ps_step(O.ADAM, var, m, v, g, ohp)
        ohp.advance()

    next_grads()
    ps.apply_gradients(static_g)                 # first call outside the recording: set-up
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    next_grads()
    with torch.cuda.graph(graph):
        ps.apply_gradients(static_g, stream=torch.cuda.current_stream())
    graph.replay()                               # capture does not execute: this is step 2
    for _ in range(6):
        next_grads()
        graph.replay()
    torch.cuda.synchronize()
    assert np.array_equal(ps.var[:n].cpu().numpy(), var)
    assert ps.device_global_step() == 8


def test_device_state_advances_once_per_step_with_many_variables(pkg, ctx):
    """A task that applies its variables ONE BY ONE (round-robin buckets) must run TF's finish
    step once per step, not once per variable: accel_apply_dev_state_ex(advance=0) on each
    variable, then accel_opt_state_advance. Result = the flat-arena step, bit for bit."""
    sizes = [4096, 24, 70_000]
    n = sum(sizes)
    rng = np.random.default_rng(9)
    var = rng.standard_normal(n).astype(F)
    m, v = np.zeros(n, F), np.zeros(n, F)
    d_var, d_m, d_v = _to_dev(var), _to_dev(m), _to_dev(v)
    hp = pkg.HParams(lr=0.01)
    ohp = O.HParams(lr=0.01)
    state = torch.zeros(4, dtype=torch.float32, device=DEV)
    ctx.opt_state_init(state, hp, 0)
    offs = np.cumsum([0] + sizes)
    for step in range(4):
        g = [(0.01 * rng.standard_normal(n)).astype(F) for _ in range(2)]
        dg = [_to_dev(x) for x in g]
        for k, sz in enumerate(sizes):
            lo = int(offs[k])
            ctx.apply_dev_state_ex(pkg.OPT_ADAM, d_var[lo:lo + sz], d_m[lo:lo + sz], d_v[lo:lo + sz],
                                   [x[lo:lo + sz] for x in dg], hp, state, advance=False)
        ctx.opt_state_advance(state, hp, True)
        O.numpy_ps_step(O.ADAM, var, m, v, g, ohp)
        ohp.advance()
    assert np.array_equal(d_var.cpu().numpy(), var)
    assert int(state[2:4].view(torch.int64).item()) == 4
    b = state[:2].cpu().numpy()
    assert b[0] == F(ohp.beta1_power) and b[1] == F(ohp.beta2_power)
