"""Placement contract (pkg/tensorflow/distributed.go) through the C ABI against the
Python restatement and the hand-derived golden for examples/tfjob/dist.yml.
Byte-exact strings and ints. No GPU."""
import itertools

import pytest

from oracle import placement_oracle as PO


@pytest.fixture(scope="module")
def P(pkg):
    return pkg.placement


def test_dist_yml_first_sync_events(P, kat):
    g = kat["placement_dist_yml"]
    ev, composed = P.action(g["workers"], g["ps"])
    assert [list(e) for e in ev] == g["events_first_sync"]
    assert composed


def test_dist_yml_cluster_spec_golden(P, kat):
    g = kat["placement_dist_yml"]
    sfx = {f"S{k}": s for k, s in enumerate(["bcdfg", "hjklm", "npqrs", "tvwxz", "24567", "89bcd"])}
    wsvc = [PO.generate_name(P.service_key(g["job"], "Worker", i) + "-", sfx[f"S{i}"])
            for i in range(4)]
    psvc = [PO.generate_name(P.service_key(g["job"], "PS", i) + "-", sfx[f"S{4 + i}"])
            for i in range(2)]
    for i in range(2):
        want = [t.format(i=i, **sfx) for t in g["ps_argv_template"]]
        assert P.cluster_spec(4, 2, wsvc, psvc, "PS", i) == want
        lab = {k: v.format(i=i, R="rt123") for k, v in g["ps_pod_labels"].items()}
        assert P.labels(g["job"], "PS", "rt123", i) == lab
    w = P.cluster_spec(4, 2, wsvc, psvc, "Worker", 3)
    assert w[2:] == ["--job_name=worker", "--task_index=3"]
    assert w[:2] == P.cluster_spec(4, 2, wsvc, psvc, "PS", 0)[:2]


def test_config1_one_ps_two_workers(P):
    """BASELINE config 1: 1 PS + 2 Workers."""
    wsvc = ["mnist-worker-0-bbbbb", "mnist-worker-1-ccccc"]
    psvc = ["mnist-ps-0-ddddd"]
    assert P.cluster_spec(2, 1, wsvc, psvc, "PS", 0) == [
        "--worker_hosts=mnist-worker-0-bbbbb:2222,mnist-worker-1-ccccc:2222",
        "--ps_hosts=mnist-ps-0-ddddd:2222", "--job_name=ps", "--task_index=0"]
    ev, _ = P.action(2, 1)
    assert ev == [("ShouldAddWorkerService", 2), ("ShouldAddPSService", 1),
                  ("ShouldAddWorker", 2), ("ShouldAddPS", 1)]


@pytest.mark.parametrize("w,p,succ,wsvc,psvc,aw,ap", list(itertools.product(
    [0, 2, 4], [0, 1, 2], [0, 1], [0, 2], [0, 1], [0, 1, 2, 4, 5], [0, 1, 3])))
def test_action_matches_restatement(P, w, p, succ, wsvc, psvc, aw, ap):
    assert P.action(w, p, succ, wsvc, psvc, aw, ap) == PO.action(w, p, succ, wsvc, psvc, aw, ap)


def test_action_quirks(P):
    # only the PS pods are missing: no compose, workers report Nothing (distributed.go:100-112)
    ev, composed = P.action(4, 2, 0, 4, 2, 4, 1)
    assert ev == [("Nothing", 0), ("ShouldAddPS", 1)] and not composed
    # more active workers than expected: no worker event at all
    assert P.action(2, 1, 0, 2, 1, 3, 1) == ([], False)
    # succeeded workers shrink the pod count but not the host list (distributed.go:60,129)
    ev, _ = P.action(4, 2, 3, 4, 2, 0, 2)
    assert ev == [("ShouldAddWorker", 1)]
    assert P.cluster_spec(4, 2, None, None, "Worker", 0)[0].count(":2222") == 4


def test_cluster_spec_quirks(P):
    # later sync: serviceNames is empty -> hosts degenerate to ":2222" (Appendix A.5)
    assert P.cluster_spec(2, 1, None, None, "PS", 0) == [
        "--worker_hosts=:2222,:2222", "--ps_hosts=:2222", "--job_name=ps", "--task_index=0"]
    # zero replicas: the unconditional trim eats the '=' (distributed.go:137,148)
    assert P.cluster_spec(0, 0, None, None, "Worker", 7) == [
        "--worker_hosts", "--ps_hosts", "--job_name=worker", "--task_index=7"]
    # job_name is lower-cased, the job_type label keeps the CRD's case (Appendix A.9)
    assert P.labels("j", "Worker", "r")["job_type"] == "Worker"
    assert "index" not in P.labels("j", "Worker", "r")
    assert P.service_key("My-Job", "Worker", 12) == "My-Job-worker-12"


@pytest.mark.parametrize("w,p", [(1, 1), (3, 2), (8, 8), (16, 3)])
def test_cluster_spec_matches_restatement(P, w, p):
    job = "job-x"
    names = {}
    for i in range(w):
        names[PO.service_key(job, "Worker", i)] = f"{job}-worker-{i}-wwww{i % 10}"
    for i in range(p):
        names[PO.service_key(job, "PS", i)] = f"{job}-ps-{i}-pppp{i % 10}"
    wsvc = [names[PO.service_key(job, "Worker", i)] for i in range(w)]
    psvc = [names[PO.service_key(job, "PS", i)] for i in range(p)]
    for typ, n in (("PS", p), ("Worker", w)):
        for i in range(n):
            assert P.cluster_spec(w, p, wsvc, psvc, typ, i) == \
                PO.cluster_spec(job, w, p, names, typ, i)


def test_parse_argv_round_trip(P):
    argv = P.cluster_spec(4, 2, [f"w{i}" for i in range(4)], ["p0", "p1"], "PS", 1)
    assert P.parse_argv(argv) == ("ps", 1, 4, 2)
    argv = P.cluster_spec(4, 2, [f"w{i}" for i in range(4)], ["p0", "p1"], "Worker", 3)
    assert P.parse_argv(argv) == ("worker", 3, 4, 2)


def test_parse_argv_errors(P, pkg):
    with pytest.raises(pkg.AccelError, match="job_name"):
        P.parse_argv(["--task_index=0"])
    with pytest.raises(pkg.AccelError, match="task_index"):
        P.parse_argv(["--job_name=ps"])
    with pytest.raises(pkg.AccelError):
        P.parse_argv(["--job_name=chief", "--task_index=0"])


def test_task_to_gpu_map(pkg):
    ctx = pkg.Context(8)
    assert [ctx.gpu_of(i) for i in range(8)] == list(range(8))      # identity by default
    ctx = pkg.Context(2, [5, 3])
    assert (ctx.gpu_of(0), ctx.gpu_of(1)) == (5, 3)
    with pytest.raises(pkg.AccelError):
        ctx.gpu_of(2)
    with pytest.raises(pkg.AccelError):
        pkg.Context(0)
    with pytest.raises(pkg.AccelError):
        pkg.Context(17)


def test_round_robin_example_model(pkg):
    """dist.yml P=2 with the example's variables (mnist_replica.py:142-158): global_step
    takes task 0, so hid_w starts the trainable list at task 1."""
    sizes = [784 * 100, 100, 100 * 10, 10]      # hid_w, hid_b, sm_w, sm_b
    task, off, fill = pkg.plan_round_robin(sizes, 2, first_task=1)
    assert task == [1, 0, 1, 0]
    assert off == [0, 0, 78400, 100]
    assert fill == [110, 79400]
    # P = 1 (config 1): everything on task 0
    task, off, fill = pkg.plan_round_robin(sizes, 1, first_task=1)
    assert task == [0, 0, 0, 0] and fill == [79510]


@pytest.mark.parametrize("n_ps", [1, 2, 3, 8])
@pytest.mark.parametrize("align", [1, 8, 64])
def test_round_robin_matches_restatement(pkg, n_ps, align):
    sizes = [64, 9408, 64, 64, 4096, 36864, 2359296, 1000, 7, 1]
    assert pkg.plan_round_robin(sizes, n_ps, 0, align) == \
        tuple(PO.round_robin(sizes, n_ps, 0, align))


@pytest.mark.parametrize("n,n_ps,align", [(25_557_032, 8, 8), (25_557_032, 2, 8),
                                          (335_141_888, 8, 64), (10, 4, 1), (0, 2, 4)])
def test_flat_plan(pkg, n, n_ps, align):
    shards, padded = pkg.plan_flat(n, n_ps, align)
    assert (shards, padded) == tuple(PO.flat(n, n_ps, align))
    assert padded >= n and padded % (n_ps * align) == 0 and padded - n < n_ps * align
    assert all(c == padded // n_ps for _, c in shards)
    assert [o for o, _ in shards] == [r * (padded // n_ps) for r in range(n_ps)]


# ---- BASELINE configs 3 and 4: the named models' variable lists --------------------------------
def _model_sizes(name):
    import json
    import os
    path = os.path.join(os.path.dirname(__file__), "golden", "model_shapes.json")
    shapes = json.load(open(path))[name]
    sizes = []
    for _, s in shapes:
        n = 1
        for d in s:
            n *= d
        sizes.append(n)
    return sizes


@pytest.mark.parametrize("name,tensors,elems,biggest,smallest", [
    ("resnet50", 161, 25_557_032, 2_359_296, 64),
    ("bert_large", 391, 335_141_888, 31_254_528, 1_024)])
def test_model_shape_lists(name, tensors, elems, biggest, smallest):
    sizes = _model_sizes(name)
    assert (len(sizes), sum(sizes), max(sizes), min(sizes)) == (tensors, elems, biggest, smallest)


@pytest.mark.parametrize("name", ["resnet50", "bert_large"])
@pytest.mark.parametrize("n_ps", [2, 4, 8])
def test_round_robin_on_named_models(pkg, name, n_ps):
    sizes = _model_sizes(name)
    task, off, fill = pkg.plan_round_robin(sizes, n_ps, 0, 8)
    assert (task, off, fill) == tuple(PO.round_robin(sizes, n_ps, 0, 8))
    assert task == [k % n_ps for k in range(len(sizes))]
    # buckets tile without overlap and hold every variable exactly once
    for t in range(n_ps):
        spans = sorted((off[k], off[k] + sizes[k]) for k in range(len(sizes)) if task[k] == t)
        assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))
        assert spans[-1][1] == fill[t]
    # round-robin ignores size: the flat plan is the balanced one
    shards, padded = pkg.plan_flat(sum(sizes), n_ps, 8)
    assert max(fill) >= padded // n_ps


@pytest.mark.parametrize("first,second,want", [
    # the reference's own table, pkg/checker/checker_test.go:10-38
    ("Local", None, True), ("PS", "Local", False), ("PS", "Worker", False), ("Local", "PS", True)])
def test_is_local_job_table(P, first, second, want):
    specs = [first] + ([second] if second else [])
    assert P.is_local_job(specs) is want


# ---- one whole sync: events executed in order, names recorded, templates mutated ------------
SFX = ["bcdfg", "hjklm", "npqrs", "tvwxz", "24567", "89bcd", "fghjk", "lmnpq", "rstvw"]


def test_first_sync_of_dist_yml(P):
    """examples/tfjob/dist.yml, nothing exists yet: RuntimeID, 4+2 Services, 4+2 pods."""
    plan = P.sync_plan("dist-training-job", "", 4, 2, SFX[:7])
    assert plan == PO.sync_plan("dist-training-job", "", 4, 2, SFX[:7])
    assert plan["runtime_id"] == "bcdfg"                    # compose() draws first
    assert plan["events"] == [("ShouldAddWorkerService", 4), ("ShouldAddPSService", 2),
                              ("ShouldAddWorker", 4), ("ShouldAddPS", 2)]
    assert [s["name"] for s in plan["services"]] == [
        "dist-training-job-worker-0-hjklm", "dist-training-job-worker-1-npqrs",
        "dist-training-job-worker-2-tvwxz", "dist-training-job-worker-3-24567",
        "dist-training-job-ps-0-89bcd", "dist-training-job-ps-1-fghjk"]
    assert plan["services"][4]["labels"] == {
        "kubeflow.caicloud.io": "true", "job_type": "PS", "runtime_id": "bcdfg",
        "tf_job_name": "dist-training-job", "index": "0"}
    ps1 = [p for p in plan["pods"] if p["type"] == "PS" and p["index"] == 1][0]
    assert ps1["args"] == [
        "--worker_hosts=dist-training-job-worker-0-hjklm:2222,dist-training-job-worker-1-npqrs:2222,"
        "dist-training-job-worker-2-tvwxz:2222,dist-training-job-worker-3-24567:2222",
        "--ps_hosts=dist-training-job-ps-0-89bcd:2222,dist-training-job-ps-1-fghjk:2222",
        "--job_name=ps", "--task_index=1"]
    assert ps1["labels"]["index"] == "1" and ps1["labels"]["job_type"] == "PS"
    # GetSpec mutates the shared template: it is left carrying the LAST replica's index and args
    assert plan["templates"]["Worker"]["labels"]["index"] == "3"
    assert plan["templates"]["PS"]["args"][3] == "--task_index=1"


def test_later_sync_recreates_pods_with_degenerate_hosts(P):
    """Services exist, two workers died: no compose of names -> hosts are ":2222" (Appendix A.5),
    the RuntimeID changes (compose runs whenever workers are missing), indices restart at 0."""
    plan = P.sync_plan("dist-training-job", "oldid", 4, 2, ["zzzzz"[:5].replace("z", "b")],
                       worker_services=4, ps_services=2, active_workers=2, active_ps=2)
    assert plan == PO.sync_plan("dist-training-job", "oldid", 4, 2, ["bbbbb"],
                                worker_services=4, ps_services=2, active_workers=2, active_ps=2)
    assert plan["runtime_id"] == "bbbbb" and plan["services"] == []
    assert [(p["type"], p["index"]) for p in plan["pods"]] == [("Worker", 0), ("Worker", 1)]
    assert plan["pods"][0]["args"][0] == "--worker_hosts=:2222,:2222,:2222,:2222"
    assert plan["pods"][0]["args"][1] == "--ps_hosts=:2222,:2222"


def test_only_ps_missing_keeps_runtime_id(P):
    plan = P.sync_plan("j", "keepme", 2, 2, [], worker_services=2, ps_services=2,
                       active_workers=2, active_ps=1)
    assert plan == PO.sync_plan("j", "keepme", 2, 2, [], worker_services=2, ps_services=2,
                                active_workers=2, active_ps=1)
    assert plan["runtime_id"] == "keepme"
    assert plan["events"] == [("Nothing", 0), ("ShouldAddPS", 1)]
    # no compose: the PS template has no base labels yet, only the index GetSpec adds
    assert plan["pods"][0]["labels"] == {"index": "0"}


def test_generated_names_are_truncated_to_63(P):
    """simpleNameGenerator keeps at most 58 bytes of the base (generate.go:53-65)."""
    job = "x" * 70
    plan = P.sync_plan(job, "", 1, 1, SFX[:3])
    assert plan == PO.sync_plan(job, "", 1, 1, SFX[:3])
    for s in plan["services"]:
        assert len(s["name"]) == 63 and s["name"].startswith("x" * 58)


@pytest.mark.parametrize("w,p,succ,wsvc,psvc,aw,ap", [
    (4, 2, 0, 0, 0, 0, 0), (4, 2, 1, 4, 2, 1, 2), (3, 1, 0, 0, 1, 3, 0), (2, 2, 0, 2, 0, 0, 0),
    (1, 1, 0, 0, 0, 1, 1), (8, 8, 0, 0, 0, 0, 0), (0, 0, 0, 0, 0, 0, 0)])
def test_sync_plan_matches_restatement(P, w, p, succ, wsvc, psvc, aw, ap):
    sfx = (SFX * 3)[:1 + w + p]
    kw = dict(succeeded=succ, worker_services=wsvc, ps_services=psvc, active_workers=aw, active_ps=ap)
    assert P.sync_plan("job", "rid00", w, p, sfx, **kw) == PO.sync_plan("job", "rid00", w, p, sfx, **kw)


def test_sync_plan_errors(P, pkg):
    with pytest.raises(pkg.AccelError, match="exhausted"):
        P.sync_plan("j", "", 2, 1, SFX[:2])          # needs 1 + 2 + 1 suffixes
    with pytest.raises(pkg.AccelError, match="5 characters"):
        P.sync_plan("j", "", 1, 1, ["abc", "bcdfg", "bcdfg"])


# ---- LocalJob (pkg/tensorflow/local.go:34-90) ------------------------------------------------
@pytest.mark.parametrize("active", [0, 1, 2, 3])
@pytest.mark.parametrize("succeeded", [0, 1, 2])
def test_local_action_matches_restatement(P, active, succeeded):
    got = P.local_action("local-job", "oldid", active, succeeded, "bcdfg")
    assert got == PO.local_action("local-job", "oldid", active, succeeded, "bcdfg")


def test_local_action_quirks(P):
    # no pod yet: one worker to add, a fresh RuntimeID, labels carry job_type "Local"
    ev, composed, rid, lab = P.local_action("mnist", "", 0, 0, "x2456")
    assert ev == ("ShouldAddWorker", 1) and composed and rid == "x2456"
    assert lab == {"kubeflow.caicloud.io": "true", "job_type": "Local", "runtime_id": "x2456",
                   "tf_job_name": "mnist"}
    # one active pod: nothing to do, the stored RuntimeID stays
    assert P.local_action("mnist", "keep5", 1, 0, "x2456") == (("Nothing", 0), False, "keep5", {})
    # TWO active pods: NewLocalJob keeps none (local.go:41-43 only logs), so Action asks for one more
    ev, composed, rid, _ = P.local_action("mnist", "keep5", 2, 0, "x2456")
    assert ev == ("ShouldAddWorker", 1) and composed and rid == "x2456"
    # a succeeded pod wins over everything
    assert P.local_action("mnist", "keep5", 0, 1, "x2456")[0] == ("Nothing", 0)


# ---- round-robin buckets as shards of the multi-GPU arena (SURVEY 8 e, TF-faithful mode) ------
@pytest.mark.parametrize("name", ["resnet50", "bert_large"])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_bucket_layout_is_round_robin_padded_to_equal_shards(pkg, name, world):
    import json
    import os

    import numpy as np
    from kubeflow_controller_b200.sharded import BucketLayout
    shapes = json.load(open(os.path.join(os.path.dirname(__file__), "golden",
                                         "model_shapes.json")))[name]
    sizes = [int(np.prod(s)) if s else 1 for _, s in shapes]
    lay = BucketLayout(sizes, world)
    task, off, fill = PO.round_robin(sizes, world, 0, 8)
    assert lay.task == task and lay.offset == off and lay.fill == fill
    assert lay.shard % 8 == 0 and lay.shard >= max(fill) and lay.shard - max(fill) < 8
    assert lay.arena == lay.shard * world
    spans = sorted((lay.pos[k], lay.pos[k] + sizes[k]) for k in range(len(sizes)))
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:])), "variables overlap"
    for k in range(len(sizes)):
        assert lay.pos[k] // lay.shard == k % world == (lay.pos[k] + sizes[k] - 1) // lay.shard
