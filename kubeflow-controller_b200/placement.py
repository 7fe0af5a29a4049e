"""Placement contract through the C ABI (csrc/tfjob.cpp): the strings and ints a
TFJob's replicas are given by the reference's pkg/tensorflow, unchanged.
"""
from __future__ import annotations

import ctypes

from . import accel

_TYPE = {"PS": accel.REPLICA_PS, "Worker": accel.REPLICA_WORKER,
         "ps": accel.REPLICA_PS, "worker": accel.REPLICA_WORKER}


def action(workers, ps, succeeded=0, worker_services=0, ps_services=0, active_workers=0,
           active_ps=0):
    """DistributedJob.Action (pkg/tensorflow/distributed.go:56-114) ->
    ([(action name, number)], composed)."""
    L = accel.load_library()
    out = (accel.Event * 4)()
    n = ctypes.c_int()
    composed = ctypes.c_int()
    accel.check(L.accel_tfjob_action(workers, ps, succeeded, worker_services, ps_services,
                                     active_workers, active_ps, out, ctypes.byref(n),
                                     ctypes.byref(composed)))
    return [(accel.ACTION_NAMES[out[i].action], out[i].number) for i in range(n.value)], \
        bool(composed.value)


def is_local_job(replica_types) -> bool:
    """checker.IsLocalJob (pkg/checker/checker.go:8-14): Specs[0].TFReplicaType == Local."""
    L = accel.load_library()
    codes = {"PS": accel.REPLICA_PS, "Worker": accel.REPLICA_WORKER, "Local": accel.REPLICA_LOCAL}
    arr = (ctypes.c_int * len(replica_types))(*[codes[t] for t in replica_types])
    return bool(accel.check(L.accel_tfjob_is_local_job(arr, len(replica_types))))


def local_action(job: str, runtime_id: str, active_pods: int, succeeded: int, suffix: str = "",
                 first_spec_type: str = "Local"):
    """LocalJob.Action with NewLocalJob's pod selection and compose()/labels()
    (pkg/tensorflow/local.go:34-90) -> ((action name, number), composed, runtime id after,
    template labels after)."""
    L = accel.load_library()
    codes = {"PS": accel.REPLICA_PS, "Worker": accel.REPLICA_WORKER, "Local": accel.REPLICA_LOCAL}
    ev = accel.Event()
    composed = ctypes.c_int()
    buf = ctypes.create_string_buffer(2048)
    accel.check(L.accel_tfjob_local_action(job.encode(), runtime_id.encode(),
                                           codes[first_spec_type], active_pods, succeeded,
                                           suffix.encode(), ctypes.byref(ev),
                                           ctypes.byref(composed), buf, len(buf)))
    lines = buf.value.decode().splitlines()
    rid = lines[0].split("\t", 1)[1] if "\t" in lines[0] else ""
    labels_ = dict(ln.split("=", 1) for ln in lines[1:])
    return (accel.ACTION_NAMES[ev.action], ev.number), bool(composed.value), rid, labels_


def sync_plan(job: str, runtime_id: str, workers: int, ps: int, suffixes, succeeded=0,
              worker_services=0, ps_services=0, active_workers=0, active_ps=0):
    """One sync of manageTFJob (pkg/controller/controller.go:369-425) ->
    dict(runtime_id, events, services, pods, templates)."""
    L = accel.load_library()
    sfx = (ctypes.c_char_p * max(len(suffixes), 1))(*[s.encode() for s in suffixes])
    buf = ctypes.create_string_buffer(1 << 18)
    accel.check(L.accel_tfjob_sync_plan(job.encode(), runtime_id.encode(), workers, ps, succeeded,
                                        worker_services, ps_services, active_workers, active_ps,
                                        sfx, len(suffixes), buf, len(buf)))

    def labels(s):
        return dict(kv.split("=", 1) for kv in s.split(",")) if s else {}
    plan = {"runtime_id": None, "events": [], "services": [], "pods": [], "templates": {}}
    for line in buf.value.decode().splitlines():
        f = line.split("\t")
        if f[0] == "runtime_id":
            plan["runtime_id"] = f[1] if len(f) > 1 else ""
        elif f[0] == "event":
            plan["events"].append((f[1], int(f[2])))
        elif f[0] == "service":
            plan["services"].append({"name": f[1], "labels": labels(f[2])})
        elif f[0] == "pod":
            plan["pods"].append({"type": f[1], "index": int(f[2]), "labels": labels(f[3]),
                                 "args": f[4:8]})
        elif f[0] == "template":
            plan["templates"][f[1]] = {"labels": labels(f[2]),
                                       "args": f[3].split(" ") if len(f) > 3 and f[3] else []}
    return plan


def service_key(job: str, typ: str, index: int) -> str:
    """getServiceName (distributed.go:161-163)."""
    L = accel.load_library()
    buf = ctypes.create_string_buffer(512)
    accel.check(L.accel_tfjob_service_key(job.encode(), _TYPE[typ], index, buf, len(buf)))
    return buf.value.decode()


def cluster_spec(workers: int, ps: int, worker_services, ps_services, typ: str, index: int,
                 bufsize: int = 1 << 16):
    """generateTFClusterSpec (distributed.go:127-159) -> the four argv strings."""
    L = accel.load_library()

    def arr(names, n):
        if names is None:
            return None
        return (ctypes.c_char_p * max(n, 1))(*[(s or "").encode() for s in names])
    buf = ctypes.create_string_buffer(bufsize)
    n = accel.check(L.accel_tfjob_cluster_spec(workers, ps, arr(worker_services, workers),
                                               arr(ps_services, ps), _TYPE[typ], index, buf,
                                               bufsize))
    return [s.decode() for s in buf.raw[:n].split(b"\0")[:4]]


def labels(job: str, typ: str, runtime_id: str, index: int | None = None) -> dict:
    """getLabels (+ index) (distributed.go:121,167,221-228)."""
    L = accel.load_library()
    buf = ctypes.create_string_buffer(1024)
    accel.check(L.accel_tfjob_labels(job.encode(), _TYPE[typ], runtime_id.encode(),
                                     -1 if index is None else index, buf, len(buf)))
    return dict(line.split("=", 1) for line in buf.value.decode().splitlines())


def parse_argv(argv):
    """What examples/workdir/mnist_replica.py:99-115 derives from the four flags ->
    (job "ps"|"worker", task_index, n_workers, n_ps)."""
    L = accel.load_library()
    a = (ctypes.c_char_p * max(len(argv), 1))(*[s.encode() for s in argv])
    typ, idx, nw, nps = (ctypes.c_int() for _ in range(4))
    accel.check(L.accel_tfjob_parse_argv(a, len(argv), ctypes.byref(typ), ctypes.byref(idx),
                                         ctypes.byref(nw), ctypes.byref(nps)))
    return ("ps" if typ.value == accel.REPLICA_PS else "worker"), idx.value, nw.value, nps.value
