"""Python restatement of the reference's placement contract — TEST
INFRASTRUCTURE, NOT PRODUCT (the product's version is csrc/tfjob.cpp behind the
C ABI). Follows /root/reference/pkg/tensorflow/distributed.go:56-228,
pkg/tensorflow/util.go:19-26, vendor/k8s.io/kubernetes/pkg/api/v1/generate.go:
53-65, vendor/k8s.io/apimachinery/pkg/util/rand/rand.go:75-82 and the caller
order of pkg/controller/controller.go:369-425. The reference has no test for
this code: parity unpinned; goldens are derived by hand from the
Go source (tests/golden/kat_survey.json, "placement_dist_yml").
"""
from __future__ import annotations

ALPHANUMS = "bcdfghjklmnpqrstvwxz2456789"   # rand.go:77
PORT = 2222                                  # distributed.go:28
PORT_NAME = "kubeflow-port"                  # distributed.go:27
MAX_GENERATED = 63 - 5                       # generate.go:55-57


def generate_name(base: str, suffix: str) -> str:
    """simpleNameGenerator.GenerateName (generate.go:60-65) with the random tail given."""
    assert len(suffix) == 5
    return base[:MAX_GENERATED] + suffix


def service_key(job: str, typ: str, index: int) -> str:
    """getServiceName (distributed.go:161-163)."""
    return f"{job}-{typ.lower()}-{index}"


def labels(job: str, typ: str, runtime_id: str, index: int | None = None) -> dict:
    """getLabels (distributed.go:221-228) + the index label (:121, :167)."""
    d = {"kubeflow.caicloud.io": "true", "job_type": typ, "runtime_id": runtime_id,
         "tf_job_name": job}
    if index is not None:
        d["index"] = str(index)
    return d


def action(workers: int, ps: int, succeeded: int, worker_svcs: int, ps_svcs: int,
           active_workers: int, active_ps: int):
    """DistributedJob.Action (distributed.go:56-114) -> ([(action, number)], composed)."""
    ev, composed = [], False
    expected_w, expected_p = workers - succeeded, ps
    if worker_svcs == 0:
        ev.append(("ShouldAddWorkerService", expected_w - worker_svcs))
    if ps_svcs == 0:
        ev.append(("ShouldAddPSService", expected_p - ps_svcs))
    if active_workers < expected_w:
        composed = True
        ev.append(("ShouldAddWorker", expected_w - active_workers))
    elif active_workers == expected_w:
        ev.append(("Nothing", 0))
    if active_ps < expected_p:
        ev.append(("ShouldAddPS", expected_p - active_ps))
    return ev, composed


def local_action(job: str, runtime_id: str, active_pods: int, succeeded: int, suffix: str,
                 first_spec_type: str = "Local"):
    """LocalJob.Action (/root/reference/pkg/tensorflow/local.go:50-74) with NewLocalJob's pod
    selection (:34-48 — a pod is kept only when exactly one is active) and compose()/labels()
    (:78-90). -> ((action, number), composed, runtime id after, template labels after)."""
    if succeeded > 0:
        return ("Nothing", 0), False, runtime_id, {}
    active = 1 if active_pods == 1 else 0
    if active < 1:
        rid = generate_name("", suffix)
        return ("ShouldAddWorker", 1 - active), True, rid, labels(job, first_spec_type, rid)
    return ("Nothing", 0), False, runtime_id, {}


def cluster_spec(job: str, workers: int, ps: int, service_names: dict, typ: str, index: int):
    """generateTFClusterSpec (distributed.go:127-159). service_names: key -> generated
    name; a missing key reads as "" like a Go map."""
    def flag(name, t, n):
        s = name
        for i in range(n):
            s += f"{service_names.get(service_key(job, t, i), '')}:{PORT},"
        return s[:-1]   # unconditional: eats the '=' when n == 0
    return [flag("--worker_hosts=", "Worker", workers), flag("--ps_hosts=", "PS", ps),
            f"--job_name={typ.lower()}", f"--task_index={index}"]


def round_robin(var_elems: list[int], n_ps: int, first_task: int = 0, align: int = 1):
    """tf.train.replica_device_setter default strategy (mnist_replica.py:137-141):
    variable k -> task (first_task + k) mod n_ps; packed per-task offsets."""
    fill = [0] * n_ps
    task, off = [], []
    for k, n in enumerate(var_elems):
        t = (first_task + k) % n_ps
        start = -(-fill[t] // align) * align
        task.append(t)
        off.append(start)
        fill[t] = start + n
    return task, off, fill


def flat(n_elems: int, n_ps: int, align: int = 1):
    unit = align * n_ps
    padded = -(-n_elems // unit) * unit
    per = padded // n_ps
    return [(r * per, per) for r in range(n_ps)], padded


def sync_plan(job: str, runtime_id: str, workers: int, ps: int, suffixes, succeeded=0,
              worker_services=0, ps_services=0, active_workers=0, active_ps=0):
    """One sync of manageTFJob (pkg/controller/controller.go:369-425) restated: Action() runs to
    completion first (compose() draws the RuntimeID), then events execute in order; GetService
    draws one suffix each and records the name; GetSpec rewrites the shared template in place."""
    sfx = list(suffixes)
    events, composed = action(workers, ps, succeeded, worker_services, ps_services,
                              active_workers, active_ps)
    tmpl = {"PS": {"labels": {}, "args": []}, "Worker": {"labels": {}, "args": []}}
    if composed:
        runtime_id = generate_name("", sfx.pop(0))
        for t in tmpl:
            tmpl[t]["labels"] = labels(job, t, runtime_id)
    names, services, pods = {}, [], []
    for act, number in events:
        typ = "PS" if "PS" in act else "Worker"
        for i in range(number):
            if act.endswith("Service"):
                key = service_key(job, typ, i)
                names[key] = generate_name(key + "-", sfx.pop(0))
                services.append({"name": names[key], "labels": labels(job, typ, runtime_id, i)})
            elif act in ("ShouldAddWorker", "ShouldAddPS"):
                tmpl[typ]["labels"]["index"] = str(i)
                tmpl[typ]["args"] = cluster_spec(job, workers, ps, names, typ, i)
                pods.append({"type": typ, "index": i, "labels": dict(tmpl[typ]["labels"]),
                             "args": list(tmpl[typ]["args"])})
    return {"runtime_id": runtime_id, "events": events, "services": services, "pods": pods,
            "templates": tmpl}
