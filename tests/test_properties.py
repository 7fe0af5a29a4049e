"""Property tests (hypothesis) of the host logic and the oracle: invariants that hold for any
input, beside the example-based tests. No GPU."""
import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import placement_oracle as PO
from oracle import ps_oracle as O

F = np.float32
NAME = st.text(alphabet="abcdefghijklmnopqrstuvwxyz0123456789-", min_size=1, max_size=40)


@settings(max_examples=150, deadline=None)
@given(n=st.integers(0, 10**12), n_ps=st.integers(1, 16), align=st.sampled_from([1, 4, 8, 64, 256]))
def test_flat_plan_invariants(pkg, n, n_ps, align):
    shards, padded = pkg.plan_flat(n, n_ps, align)
    assert padded >= n and padded - n < n_ps * align and padded % (n_ps * align) == 0
    assert [o for o, _ in shards] == [r * (padded // n_ps) for r in range(n_ps)]
    assert sum(c for _, c in shards) == padded and len({c for _, c in shards}) == 1
    assert all(c % align == 0 for _, c in shards)


@settings(max_examples=150, deadline=None)
@given(sizes=st.lists(st.integers(1, 10**7), min_size=0, max_size=60), n_ps=st.integers(1, 16),
       first=st.integers(0, 40), align=st.sampled_from([1, 8, 64]))
def test_round_robin_invariants(pkg, sizes, n_ps, first, align):
    task, off, fill = pkg.plan_round_robin(sizes, n_ps, first, align)
    assert (task, off, fill) == tuple(PO.round_robin(sizes, n_ps, first, align))
    assert task == [(first + k) % n_ps for k in range(len(sizes))]
    for t in range(n_ps):
        spans = sorted((off[k], off[k] + sizes[k]) for k in range(len(sizes)) if task[k] == t)
        assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))      # no overlap
        assert all(s % align == 0 for s, _ in spans)
        assert fill[t] == (spans[-1][1] if spans else 0)


@settings(max_examples=100, deadline=None)
@given(job=NAME, w=st.integers(0, 12), p=st.integers(0, 12), data=st.data())
def test_cluster_spec_equals_restatement(pkg, job, w, p, data):
    P = pkg.placement
    have = data.draw(st.lists(st.booleans(), min_size=w + p, max_size=w + p))
    names = {}
    for i in range(w):
        if have[i]:
            names[PO.service_key(job, "Worker", i)] = f"{job}-worker-{i}-bcdfg"[-63:]
    for i in range(p):
        if have[w + i]:
            names[PO.service_key(job, "PS", i)] = f"{job}-ps-{i}-hjklm"[-63:]
    wsvc = [names.get(PO.service_key(job, "Worker", i), "") for i in range(w)]
    psvc = [names.get(PO.service_key(job, "PS", i), "") for i in range(p)]
    typ = data.draw(st.sampled_from(["PS", "Worker"]))
    idx = data.draw(st.integers(0, 20))
    got = P.cluster_spec(w, p, wsvc, psvc, typ, idx)
    assert got == PO.cluster_spec(job, w, p, names, typ, idx)
    assert got[2] == f"--job_name={typ.lower()}" and got[3] == f"--task_index={idx}"
    if w and p:
        assert P.parse_argv(got) == (typ.lower(), idx, w, p)


@settings(max_examples=60, deadline=None)
@given(seed=st.integers(0, 2**31 - 1), log_n=st.integers(0, 1), n=st.integers(1, 300),
       opt=st.sampled_from([O.SGD, O.MOMENTUM, O.ADAM]))
def test_mean_of_identical_gradients_is_the_gradient(seed, log_n, n, opt):
    """Two workers pushing the SAME gradient, MEAN aggregation: g + g and the division by 2 are
    exact, so the step equals the single-worker step bit for bit. (With more copies the
    left-to-right sum passes through 3g, which rounds — the accumulator's order matters.)"""
    rng = np.random.default_rng(seed)
    g = (0.01 * rng.standard_normal(n)).astype(F)
    var = rng.standard_normal(n).astype(F)
    a = [var.copy(), np.zeros(n, F), np.zeros(n, F)]
    b = [var.copy(), np.zeros(n, F), np.zeros(n, F)]
    O.numpy_ps_step(opt, *a, [g] * (1 << log_n), O.HParams(lr=0.01, momentum=0.9,
                                                            aggregate=O.AGG_MEAN))
    O.numpy_ps_step(opt, *b, [g], O.HParams(lr=0.01, momentum=0.9))
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


@settings(max_examples=60, deadline=None)
@given(seed=st.integers(0, 2**31 - 1), n=st.integers(1, 500), workers=st.integers(1, 9))
def test_c_and_numpy_oracles_agree_on_any_input(seed, n, workers):
    rng = np.random.default_rng(seed)
    var = rng.standard_normal(n).astype(F)
    m = (0.1 * rng.standard_normal(n)).astype(F)
    v = np.abs(0.01 * rng.standard_normal(n)).astype(F)
    grads = [(rng.standard_normal(n) * 10.0 ** rng.integers(-6, 2)).astype(F)
             for _ in range(workers)]
    hp = O.HParams(lr=float(10.0 ** rng.integers(-4, 0)), use_nesterov=bool(rng.integers(0, 2)),
                   aggregate=int(rng.integers(0, 2)))
    for opt in (O.SGD, O.MOMENTUM, O.ADAM):
        a = [var.copy(), m.copy(), v.copy()]
        b = [var.copy(), m.copy(), v.copy()]
        O.numpy_ps_step(opt, *a, grads, hp)
        O.COracle(threads=2).ps_step(opt, *b, grads, hp)
        assert all(np.array_equal(x, y, equal_nan=True) for x, y in zip(a, b))


@settings(max_examples=100, deadline=None)
@given(x=st.lists(st.floats(width=32, allow_nan=False, allow_infinity=False), min_size=1,
                  max_size=64))
def test_bf16_rounding_is_nearest_even(x):
    import torch
    a = np.array(x, dtype=F)
    bits = O.f32_to_bf16_bits(a)
    want = torch.from_numpy(a).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    finite = np.isfinite(O.bf16_bits_to_f32(want))
    assert np.array_equal(bits[finite], want[finite])


@settings(max_examples=100, deadline=None)
@given(sizes=st.lists(st.integers(1, 10**6), min_size=1, max_size=80), world=st.integers(1, 16),
       first=st.integers(0, 20))
def test_bucket_layout_invariants(pkg, sizes, world, first):
    """Round-robin buckets as equal shards of the multi-GPU arena: every variable lies wholly in
    the shard of its round-robin task, nothing overlaps, the arena is `world` equal shards."""
    from kubeflow_controller_b200.sharded import BucketLayout
    lay = BucketLayout(sizes, world, first_task=first)
    assert lay.arena == lay.shard * world and lay.shard % 8 == 0 and lay.shard >= max(lay.fill)
    spans = sorted((lay.pos[k], lay.pos[k] + sizes[k]) for k in range(len(sizes)))
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))
    for k, n in enumerate(sizes):
        t = (first + k) % world
        assert lay.task[k] == t
        assert lay.pos[k] // lay.shard == t == (lay.pos[k] + n - 1) // lay.shard
        assert lay.pos[k] % 8 == 0
