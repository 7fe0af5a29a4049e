"""The CPU oracle against the known-answer vectors and against independent
restatements. No GPU. (The oracle is parity-unpinned: the reference holds no
fixture for this path — see oracle/cpu_ps.c.)"""
import numpy as np
import pytest
import torch

from oracle import ps_oracle as O

F = np.float32


def _f(x):
    return np.array(x, dtype=F)


def test_kat_aggregate(kat):
    g = [_f(x) for x in kat["inputs"]["grads"]]
    assert np.array_equal(O.numpy_aggregate(g, False), _f(kat["sum"]))
    assert np.array_equal(O.numpy_aggregate(g, True), _f(kat["mean"]))


def test_kat_sgd(kat):
    g = [_f(x) for x in kat["inputs"]["grads"]]
    var = _f(kat["inputs"]["var"])
    O.numpy_ps_step(O.SGD, var, None, None, g, O.HParams(lr=kat["sgd"]["lr"]))
    assert np.array_equal(var, _f(kat["sgd"]["var"]))


@pytest.mark.parametrize("nesterov", [False, True])
def test_kat_momentum(kat, nesterov):
    k = kat["momentum"]
    g = [_f(x) for x in kat["inputs"]["grads"]]
    var, acc = _f(kat["inputs"]["var"]), _f(k["accum0"])
    O.numpy_ps_step(O.MOMENTUM, var, acc, None, g,
                    O.HParams(lr=k["lr"], momentum=k["momentum"], use_nesterov=nesterov))
    assert np.array_equal(acc, _f(k["accum"]))
    assert np.array_equal(var, _f(k["var_nesterov" if nesterov else "var"]))


def test_kat_adam(kat):
    k = kat["adam"]
    g = [_f(x) for x in kat["inputs"]["grads"]]
    var, m, v = _f(kat["inputs"]["var"]), np.zeros(4, F), np.zeros(4, F)
    hp = O.HParams(lr=k["lr"], beta1=k["beta1"], beta2=k["beta2"], epsilon=k["epsilon"],
                   beta1_power=k["beta1_power"], beta2_power=k["beta2_power"])
    assert O.numpy_adam_alpha(hp) == F(k["alpha"])
    O.numpy_ps_step(O.ADAM, var, m, v, g, hp)
    assert np.array_equal(m, _f(k["m"]))
    assert np.array_equal(v, _f(k["v"]))
    assert np.array_equal(var, _f(k["var"]))


def _case(rng, n, workers, bf16=False):
    var = rng.standard_normal(n).astype(F)
    s0 = (0.1 * rng.standard_normal(n)).astype(F)
    s1 = np.abs(0.01 * rng.standard_normal(n)).astype(F)
    grads = [(0.01 * rng.standard_normal(n)).astype(F) for _ in range(workers)]
    if bf16:
        grads = [O.f32_to_bf16_bits(g) for g in grads]
    return var, s0, s1, grads


@pytest.mark.parametrize("opt", [O.SGD, O.MOMENTUM, O.ADAM])
@pytest.mark.parametrize("n,workers", [(1, 1), (7, 3), (4099, 8), (100_003, 5)])
@pytest.mark.parametrize("bf16", [False, True])
@pytest.mark.parametrize("mean,nesterov,clip", [(False, False, 0.0), (True, True, 0.0),
                                                (True, False, 0.05)])
def test_c_oracle_equals_numpy_oracle(opt, n, workers, bf16, mean, nesterov, clip):
    rng = np.random.default_rng(n * 31 + workers)
    var, s0, s1, grads = _case(rng, n, workers, bf16)
    hp = O.HParams(lr=0.01, momentum=0.9, use_nesterov=nesterov,
                   aggregate=O.AGG_MEAN if mean else O.AGG_SUM, clip_norm=clip)
    hp.advance()
    C = O.COracle(threads=3)
    cv, c0, c1 = var.copy(), s0.copy(), s1.copy()
    ss = C.ps_step(opt, cv, c0, c1, grads, hp)
    # hand the C oracle's sum of squares to the numpy twin so that the clip scale
    # is the same number and everything downstream can be compared bit for bit
    O.numpy_ps_step(opt, var, s0, s1, grads, hp, sumsq=ss if clip else None)
    assert np.array_equal(cv, var)
    if opt >= O.MOMENTUM:
        assert np.array_equal(c0, s0)
    if opt == O.ADAM:
        assert np.array_equal(c1, s1)
    if clip:
        ref = O.numpy_sumsq(O.numpy_aggregate(grads, mean))
        assert abs(ss - ref) <= 1e-5 * abs(ref)


def test_c_oracle_thread_count_does_not_change_results():
    rng = np.random.default_rng(5)
    var, s0, s1, grads = _case(rng, 50_001, 4)
    out = []
    for t in (1, 2, 8):
        v, a, b = var.copy(), s0.copy(), s1.copy()
        O.COracle(threads=t).ps_step(O.ADAM, v, a, b, grads, O.HParams(lr=0.01))
        out.append((v, a, b))
    for v, a, b in out[1:]:
        assert np.array_equal(v, out[0][0]) and np.array_equal(a, out[0][1]) \
            and np.array_equal(b, out[0][2])


def test_bf16_round_trip_matches_torch():
    rng = np.random.default_rng(9)
    x = rng.standard_normal(10_000).astype(F)
    bits = O.f32_to_bf16_bits(x)
    t = torch.from_numpy(x).to(torch.bfloat16)
    assert np.array_equal(bits, t.view(torch.int16).numpy().view(np.uint16))
    assert np.array_equal(O.bf16_bits_to_f32(bits), t.float().numpy())


def test_momentum_cross_check_torch_sgd():
    """torch.optim.SGD(momentum, dampening 0) has ApplyMomentum's form."""
    rng = np.random.default_rng(11)
    n = 10_000
    var = rng.standard_normal(n).astype(F)
    grads = [(0.01 * rng.standard_normal(n)).astype(F) for _ in range(3)]
    p = torch.nn.Parameter(torch.from_numpy(var.copy()))
    opt = torch.optim.SGD([p], lr=0.1, momentum=0.9)
    acc = np.zeros(n, F)
    for _ in range(5):
        p.grad = torch.from_numpy(O.numpy_aggregate(grads, False).copy())
        opt.step()
        O.numpy_ps_step(O.MOMENTUM, var, acc, None, grads, O.HParams(lr=0.1, momentum=0.9))
    np.testing.assert_allclose(var, p.detach().numpy(), rtol=1e-5, atol=1e-6)


def test_adam_cross_check_float64_formula():
    """TF's alpha/epsilon-hat form against a float64 evaluation of the same update,
    over 20 steps with the beta powers advanced as adam.py _finish does."""
    rng = np.random.default_rng(13)
    n = 5_000
    var = rng.standard_normal(n).astype(F)
    m, v = np.zeros(n, F), np.zeros(n, F)
    v64, m64, s64 = var.astype(np.float64), np.zeros(n), np.zeros(n)
    hp = O.HParams(lr=0.01)
    b1, b2, eps, lr = 0.9, 0.999, 1e-8, 0.01
    for t in range(1, 21):
        g = (0.01 * rng.standard_normal(n)).astype(F)
        O.numpy_ps_step(O.ADAM, var, m, v, [g], hp)
        hp.advance()
        g64 = g.astype(np.float64)
        m64 = m64 + (g64 - m64) * (1 - b1)
        s64 = s64 + (g64 * g64 - s64) * (1 - b2)
        alpha = lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
        v64 = v64 - m64 * alpha / (np.sqrt(s64) + eps)
    np.testing.assert_allclose(var, v64, rtol=2e-5, atol=1e-6)


def test_clip_scale_formula():
    # norm below the threshold: scale is exactly clip * (1/clip)
    assert O.numpy_clip_scale(F(0.25), 1.0) == F(1.0)
    # norm above: clip / norm
    assert abs(float(O.numpy_clip_scale(F(16.0), 1.0)) - 0.25) < 1e-7


def test_golden_fixture_regression():
    """Committed fixture (tests/golden/make_golden.py) still reproduced by the oracle."""
    import os
    path = os.path.join(os.path.dirname(__file__), "golden", "ps_step_golden.npz")
    z = np.load(path)
    for name in ("sgd", "momentum", "adam_bf16_mean_clip"):
        opt = int(z[f"{name}.opt"])
        grads = [z[f"{name}.g{w}"] for w in range(int(z[f"{name}.workers"]))]
        hp = O.HParams(**{k: (float(x) if k not in ("use_nesterov", "aggregate") else int(x))
                          for k, x in zip(z[f"{name}.hp_keys"], z[f"{name}.hp_vals"])})
        var, s0, s1 = z[f"{name}.var"].copy(), z[f"{name}.s0"].copy(), z[f"{name}.s1"].copy()
        ss = float(z[f"{name}.sumsq"])
        O.numpy_ps_step(opt, var, s0, s1, grads, hp, sumsq=ss if hp.clip_norm > 0 else None)
        assert np.array_equal(var, z[f"{name}.var_out"])
        assert np.array_equal(s0, z[f"{name}.s0_out"])
        assert np.array_equal(s1, z[f"{name}.s1_out"])


def test_nesterov_momentum_cross_check_torch_sgd():
    """TF: var -= g*lr + accum*momentum*lr (accum already updated); torch nesterov:
    p -= lr*(g + momentum*buf). Same mathematics, different rounding order."""
    rng = np.random.default_rng(17)
    n = 8_000
    var = rng.standard_normal(n).astype(F)
    p = torch.nn.Parameter(torch.from_numpy(var.copy()))
    opt = torch.optim.SGD([p], lr=0.05, momentum=0.9, nesterov=True)
    acc = np.zeros(n, F)
    for _ in range(6):
        g = (0.01 * rng.standard_normal(n)).astype(F)
        p.grad = torch.from_numpy(g.copy())
        opt.step()
        O.numpy_ps_step(O.MOMENTUM, var, acc, None, [g],
                        O.HParams(lr=0.05, momentum=0.9, use_nesterov=True))
    np.testing.assert_allclose(var, p.detach().numpy(), rtol=1e-5, atol=1e-6)


def test_adam_cross_check_torch_adam_with_zero_epsilon():
    """With epsilon = 0 TF's 'epsilon-hat' form and torch.optim.Adam's form are the same
    mathematics (they differ only in where epsilon enters), so torch serves as an independent
    implementation of m, v, the bias corrections and the beta-power bookkeeping."""
    rng = np.random.default_rng(19)
    n = 6_000
    var = rng.standard_normal(n).astype(F)
    p = torch.nn.Parameter(torch.from_numpy(var.copy()))
    opt = torch.optim.Adam([p], lr=0.01, betas=(0.9, 0.999), eps=0.0)
    m, v = np.zeros(n, F), np.zeros(n, F)
    hp = O.HParams(lr=0.01, epsilon=0.0)
    for _ in range(15):
        g = (0.01 * rng.standard_normal(n) + 0.02).astype(F)
        p.grad = torch.from_numpy(g.copy())
        opt.step()
        O.numpy_ps_step(O.ADAM, var, m, v, [g], hp)
        hp.advance()
    np.testing.assert_allclose(var, p.detach().numpy(), rtol=2e-5, atol=2e-6)
    state = opt.state[p]
    np.testing.assert_allclose(m, state["exp_avg"].numpy(), rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(v, state["exp_avg_sq"].numpy(), rtol=1e-5, atol=1e-10)
