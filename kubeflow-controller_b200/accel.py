"""ctypes binding of libaccel_ps.so (include/accel_ps.h) — the same symbols the Go
package pkg/accel binds through cgo. Python is the executable harness because
no Go toolchain exists in this image; torch supplies device memory and streams.

There is no CPU fallback: a missing library or a failing call raises.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libaccel_ps.so")

OK = 0
F32, BF16 = 0, 1
AGG_SUM, AGG_MEAN = 0, 1
OPT_SGD, OPT_MOMENTUM, OPT_ADAM = 0, 1, 2
KERNEL_AUTO, KERNEL_LDG128, KERNEL_LDG256, KERNEL_TMA = 0, 1, 2, 3
OVERLAP_INLINE, OVERLAP_ROLES, OVERLAP_SWITCH = 0, 1, 2
REPLICA_PS, REPLICA_WORKER, REPLICA_LOCAL = 0, 1, 2
MAX_WORKERS, MAX_PS = 64, 16
ACTION_NAMES = ["ShouldAddPS", "ShouldAddPSService", "ShouldAddWorker",
                "ShouldAddWorkerService", "ShouldDelete", "Nothing"]

# every symbol include/accel_ps.h declares (tests check the library exports them all)
SYMBOLS = [
    "accel_version", "accel_last_error", "accel_ctx_create", "accel_ctx_destroy",
    "accel_ctx_gpu_of", "accel_ctx_set_kernel", "accel_plan_flat", "accel_plan_round_robin",
    "accel_apply_sgd", "accel_apply_momentum", "accel_apply_adam", "accel_adam_advance",
    "accel_opt_state_init", "accel_opt_state_advance", "accel_apply_dev_state",
    "accel_apply_dev_state_ex",
    "accel_grad_sumsq", "accel_grad_sumsq_static", "accel_ps_step_host", "accel_host_register", "accel_host_unregister",
    "accel_tfjob_action", "accel_tfjob_is_local_job", "accel_tfjob_local_action",
    "accel_tfjob_sync_plan",
    "accel_tfjob_service_key", "accel_tfjob_cluster_spec",
    "accel_tfjob_labels", "accel_tfjob_parse_argv", "accel_ipc_export", "accel_ipc_open",
    "accel_ipc_close", "accel_dev_alloc", "accel_dev_free", "accel_shard_push",
    "accel_shard_apply_pull", "accel_shard_wait_push", "accel_shard_wait_pull", "accel_shard_step", "accel_shard_step_fused",
    "accel_shard_step_local_group", "accel_shard_check", "accel_shard_reset",
    "accel_ctx_set_step_shape", "accel_shard_step_host",
]


class AccelError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"accel_ps error {status}: {message}")
        self.status = status


class HParams(ctypes.Structure):
    """accel_hparams. Defaults are TF-1.4's (Adam: lr 1e-3, 0.9, 0.999, 1e-8)."""
    _fields_ = [("lr", ctypes.c_float), ("momentum", ctypes.c_float), ("beta1", ctypes.c_float),
                ("beta2", ctypes.c_float), ("epsilon", ctypes.c_float),
                ("beta1_power", ctypes.c_float), ("beta2_power", ctypes.c_float),
                ("clip_norm", ctypes.c_float), ("use_nesterov", ctypes.c_int32),
                ("aggregate", ctypes.c_int32)]

    def __init__(self, lr=1e-3, momentum=0.0, beta1=0.9, beta2=0.999, epsilon=1e-8,
                 beta1_power=None, beta2_power=None, clip_norm=0.0, use_nesterov=False,
                 aggregate=AGG_SUM):
        super().__init__(lr, momentum, beta1, beta2, epsilon,
                         beta1 if beta1_power is None else beta1_power,
                         beta2 if beta2_power is None else beta2_power,
                         clip_norm, int(bool(use_nesterov)), int(aggregate))


class Shard(ctypes.Structure):
    _fields_ = [("offset", ctypes.c_uint64), ("count", ctypes.c_uint64)]


class Event(ctypes.Structure):
    _fields_ = [("action", ctypes.c_int32), ("number", ctypes.c_int32)]


class IpcHandle(ctypes.Structure):
    _fields_ = [("bytes", ctypes.c_ubyte * 64)]


class Peers(ctypes.Structure):
    _fields_ = [("rank", ctypes.c_int32), ("world", ctypes.c_int32),
                ("shard_elems", ctypes.c_uint64),
                ("stage", ctypes.POINTER(ctypes.c_void_p)),
                ("var", ctypes.POINTER(ctypes.c_void_p)),
                ("flags", ctypes.POINTER(ctypes.c_void_p)),
                ("var_multicast", ctypes.c_void_p),
                ("var_bf16", ctypes.POINTER(ctypes.c_void_p)),
                ("stage_multicast", ctypes.c_void_p),
                ("stage_dtype", ctypes.c_int32),
                ("reserved0", ctypes.c_int32)]


_lib = None


def load_library() -> ctypes.CDLL:
    """Load the in-tree libaccel_ps.so. Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python kubeflow-controller_b200/build.py` "
            "(or __graft_entry__.build()). There is no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, u64, i32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int
    pvp = ctypes.POINTER(ctypes.c_void_p)
    php = ctypes.POINTER(HParams)
    L.accel_version.restype = i32
    L.accel_last_error.restype = ctypes.c_char_p
    L.accel_ctx_create.argtypes = [ctypes.POINTER(i32), i32, pvp]
    L.accel_ctx_destroy.argtypes = [vp]
    L.accel_ctx_gpu_of.argtypes = [vp, i32]
    L.accel_ctx_set_kernel.argtypes = [vp, i32, i32, i32, i32]
    L.accel_plan_flat.argtypes = [u64, i32, u64, ctypes.POINTER(Shard), ctypes.POINTER(u64)]
    L.accel_plan_round_robin.argtypes = [ctypes.POINTER(u64), i32, i32, i32, u64,
                                         ctypes.POINTER(i32), ctypes.POINTER(u64),
                                         ctypes.POINTER(u64)]
    L.accel_apply_sgd.argtypes = [vp, i32, vp, pvp, i32, i32, u64, php, vp, vp]
    L.accel_apply_momentum.argtypes = [vp, i32, vp, vp, pvp, i32, i32, u64, php, vp, vp]
    L.accel_apply_adam.argtypes = [vp, i32, vp, vp, vp, pvp, i32, i32, u64, php, vp, vp]
    L.accel_adam_advance.argtypes = [php]
    L.accel_opt_state_init.argtypes = [vp, php, u64, vp]
    L.accel_apply_dev_state.argtypes = [vp, i32, i32, vp, vp, vp, pvp, i32, i32, u64, php, vp, vp, vp]
    L.accel_apply_dev_state_ex.argtypes = [vp, i32, i32, vp, vp, vp, pvp, i32, i32, u64, php, vp,
                                           vp, i32, vp]
    L.accel_opt_state_advance.argtypes = [vp, php, i32, vp]
    L.accel_grad_sumsq.argtypes = [vp, i32, pvp, i32, i32, u64, i32, vp, vp]
    L.accel_grad_sumsq_static.argtypes = [pvp, i32, i32, u64, i32, vp, vp]
    L.accel_ps_step_host.argtypes = [vp, i32, i32, vp, vp, vp, pvp, i32, i32, u64, php, vp]
    L.accel_host_register.argtypes = [vp, ctypes.c_size_t]
    L.accel_host_unregister.argtypes = [vp]
    L.accel_tfjob_action.argtypes = [i32] * 7 + [ctypes.POINTER(Event), ctypes.POINTER(i32),
                                                 ctypes.POINTER(i32)]
    L.accel_tfjob_service_key.argtypes = [ctypes.c_char_p, i32, i32, ctypes.c_char_p,
                                          ctypes.c_size_t]
    L.accel_tfjob_cluster_spec.argtypes = [i32, i32, ctypes.POINTER(ctypes.c_char_p),
                                           ctypes.POINTER(ctypes.c_char_p), i32, i32,
                                           ctypes.c_char_p, ctypes.c_size_t]
    L.accel_tfjob_labels.argtypes = [ctypes.c_char_p, i32, ctypes.c_char_p, i32, ctypes.c_char_p,
                                     ctypes.c_size_t]
    L.accel_tfjob_parse_argv.argtypes = [ctypes.POINTER(ctypes.c_char_p), i32] + \
        [ctypes.POINTER(i32)] * 4
    L.accel_ipc_export.argtypes = [vp, ctypes.POINTER(IpcHandle)]
    L.accel_ipc_open.argtypes = [ctypes.POINTER(IpcHandle), pvp]
    L.accel_ipc_close.argtypes = [vp]
    u32 = ctypes.c_uint32
    L.accel_shard_push.argtypes = [vp, ctypes.POINTER(Peers), pvp, i32, i32, u64, u64, vp,
                                   ctypes.POINTER(u32)]
    L.accel_shard_apply_pull.argtypes = [vp, ctypes.POINTER(Peers), i32, vp, vp, i32, php, vp,
                                         u64, u64, u32, vp]
    L.accel_shard_wait_pull.argtypes = [vp, ctypes.POINTER(Peers), u32, vp]
    L.accel_shard_wait_push.argtypes = [vp, ctypes.POINTER(Peers), u32, vp]
    L.accel_tfjob_sync_plan.argtypes = [ctypes.c_char_p, ctypes.c_char_p] + [i32] * 7 + \
        [ctypes.POINTER(ctypes.c_char_p), i32, ctypes.c_char_p, ctypes.c_size_t]
    L.accel_tfjob_is_local_job.argtypes = [ctypes.POINTER(i32), i32]
    L.accel_tfjob_local_action.argtypes = [ctypes.c_char_p, ctypes.c_char_p, i32, i32, i32,
                                           ctypes.c_char_p, ctypes.POINTER(Event),
                                           ctypes.POINTER(i32), ctypes.c_char_p, ctypes.c_size_t]
    L.accel_shard_step.argtypes = [vp, ctypes.POINTER(Peers), i32, vp, vp, pvp, i32, i32, i32, php,
                                   vp, i32, vp]
    L.accel_shard_step_fused.argtypes = [vp, ctypes.POINTER(Peers), i32, vp, vp, pvp, i32, i32, i32,
                                         php, vp, i32, i32, vp]
    L.accel_shard_step_local_group.argtypes = [vp, ctypes.POINTER(ctypes.POINTER(Peers)), i32, i32,
                                               pvp, pvp, ctypes.POINTER(pvp), i32, i32, i32, php,
                                               vp, i32, i32, vp]
    L.accel_shard_check.argtypes = [vp, ctypes.POINTER(Peers), vp]
    L.accel_shard_reset.argtypes = [vp, ctypes.POINTER(Peers), vp]
    L.accel_ctx_set_step_shape.argtypes = [vp, i32, i32]
    L.accel_shard_step_host.argtypes = [vp, ctypes.POINTER(Peers), i32, vp, vp, pvp, i32, i32, i32,
                                        php, vp, i32]
    L.accel_dev_alloc.argtypes = [ctypes.c_size_t, pvp]
    L.accel_dev_free.argtypes = [vp]
    for name in SYMBOLS:
        fn = getattr(L, name)
        if name != "accel_last_error":
            fn.restype = i32
    _lib = L
    return L


def check(rc: int) -> int:
    if rc < 0:
        raise AccelError(rc, load_library().accel_last_error().decode("utf-8", "replace"))
    return rc


def _ptr(t) -> int:
    """Device (or host) address of a torch tensor / raw int / None."""
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def _stream_handle(stream) -> int | None:
    if stream is None:
        import torch
        return torch.cuda.current_stream().cuda_stream
    if isinstance(stream, int):
        return stream
    return stream.cuda_stream


def _grad_dtype(t) -> int:
    import torch
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"gradient dtype {t.dtype} not supported (float32 or bfloat16)")


def _grad_array(grads):
    return (ctypes.c_void_p * len(grads))(*[_ptr(g) for g in grads])


@dataclass
class Tuning:
    kernel: int = KERNEL_AUTO
    threads: int = 0
    ctas_per_sm: int = 0
    unroll: int = 0


class Context:
    """accel_ctx: PS task_index -> GPU ordinal map plus per-GPU scratch."""

    def __init__(self, n_ps: int = 1, gpu_ordinals: list[int] | None = None):
        self.lib = load_library()
        self.n_ps = n_ps
        arr = None
        if gpu_ordinals is not None:
            arr = (ctypes.c_int * n_ps)(*gpu_ordinals)
        h = ctypes.c_void_p()
        check(self.lib.accel_ctx_create(arr, n_ps, ctypes.byref(h)))
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            self.lib.accel_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def gpu_of(self, task: int) -> int:
        return check(self.lib.accel_ctx_gpu_of(self.handle, task))

    def set_kernel(self, kernel=KERNEL_AUTO, threads=0, ctas_per_sm=0, unroll=0):
        check(self.lib.accel_ctx_set_kernel(self.handle, kernel, threads, ctas_per_sm, unroll))

    def set_step_shape(self, reducer_pct=0, coop=True):
        check(self.lib.accel_ctx_set_step_shape(self.handle, reducer_pct, int(bool(coop))))

    # ---- the hot path ---------------------------------------------------------
    def apply_sgd(self, var, grads, hp: HParams, task=0, count=None, sumsq=None, stream=None):
        n = var.numel() if count is None else count
        check(self.lib.accel_apply_sgd(self.handle, task, _ptr(var), _grad_array(grads),
                                       len(grads), _grad_dtype(grads[0]), n, ctypes.byref(hp),
                                       _ptr(sumsq), _stream_handle(stream)))

    def apply_momentum(self, var, accum, grads, hp: HParams, task=0, count=None, sumsq=None,
                       stream=None):
        n = var.numel() if count is None else count
        check(self.lib.accel_apply_momentum(self.handle, task, _ptr(var), _ptr(accum),
                                            _grad_array(grads), len(grads),
                                            _grad_dtype(grads[0]), n, ctypes.byref(hp),
                                            _ptr(sumsq), _stream_handle(stream)))

    def apply_adam(self, var, m, v, grads, hp: HParams, task=0, count=None, sumsq=None,
                   stream=None):
        n = var.numel() if count is None else count
        check(self.lib.accel_apply_adam(self.handle, task, _ptr(var), _ptr(m), _ptr(v),
                                        _grad_array(grads), len(grads), _grad_dtype(grads[0]), n,
                                        ctypes.byref(hp), _ptr(sumsq), _stream_handle(stream)))

    def apply(self, optimizer, var, slot0, slot1, grads, hp, **kw):
        if optimizer == OPT_SGD:
            return self.apply_sgd(var, grads, hp, **kw)
        if optimizer == OPT_MOMENTUM:
            return self.apply_momentum(var, slot0, grads, hp, **kw)
        if optimizer == OPT_ADAM:
            return self.apply_adam(var, slot0, slot1, grads, hp, **kw)
        raise ValueError(f"unknown optimizer {optimizer}")

    def opt_state_init(self, dev_state, hp: HParams, global_step=0, stream=None):
        """dev_state: a CUDA tensor of 4 float32 (16 bytes): beta powers + uint64 global step."""
        check(self.lib.accel_opt_state_init(_ptr(dev_state), ctypes.byref(hp), global_step,
                                            _stream_handle(stream)))

    def apply_dev_state(self, optimizer, var, slot0, slot1, grads, hp: HParams, dev_state, task=0,
                        count=None, sumsq=None, stream=None):
        n = var.numel() if count is None else count
        check(self.lib.accel_apply_dev_state(self.handle, task, optimizer, _ptr(var), _ptr(slot0),
                                             _ptr(slot1), _grad_array(grads), len(grads),
                                             _grad_dtype(grads[0]), n, ctypes.byref(hp),
                                             _ptr(sumsq), _ptr(dev_state), _stream_handle(stream)))

    def apply_dev_state_ex(self, optimizer, var, slot0, slot1, grads, hp: HParams, dev_state,
                           advance: bool, task=0, count=None, sumsq=None, stream=None):
        """apply_dev_state with the finish step under the caller's control (several variables
        per step: advance=False on each, then opt_state_advance once)."""
        n = var.numel() if count is None else count
        check(self.lib.accel_apply_dev_state_ex(
            self.handle, task, optimizer, _ptr(var), _ptr(slot0), _ptr(slot1), _grad_array(grads),
            len(grads), _grad_dtype(grads[0]), n, ctypes.byref(hp), _ptr(sumsq), _ptr(dev_state),
            int(bool(advance)), _stream_handle(stream)))

    def opt_state_advance(self, dev_state, hp: HParams, is_adam: bool, stream=None):
        check(self.lib.accel_opt_state_advance(_ptr(dev_state), ctypes.byref(hp),
                                               int(bool(is_adam)), _stream_handle(stream)))

    def grad_sumsq(self, grads, out, aggregate=AGG_SUM, task=0, count=None, stream=None):
        n = grads[0].numel() if count is None else count
        check(self.lib.accel_grad_sumsq(self.handle, task, _grad_array(grads), len(grads),
                                        _grad_dtype(grads[0]), n, aggregate, _ptr(out),
                                        _stream_handle(stream)))

    def ps_step_host(self, optimizer, var, slot0, slot1, host_grads, hp: HParams, host_var_out,
                     task=0, count=None):
        """host_grads: pinned CPU tensors; var/slots: CUDA tensors; host_var_out: pinned CPU."""
        n = var.numel() if count is None else count
        check(self.lib.accel_ps_step_host(self.handle, task, optimizer, _ptr(var), _ptr(slot0),
                                          _ptr(slot1), _grad_array(host_grads), len(host_grads),
                                          _grad_dtype(host_grads[0]), n, ctypes.byref(hp),
                                          _ptr(host_var_out)))


SUMSQ_STATIC_MAX_ELEMS = 1 << 28


def grad_sumsq_static(grads, out, aggregate=AGG_SUM, count=None, stream=None) -> None:
    """accel_grad_sumsq_static: the sum of squares of the aggregated gradient for callers
    that must not allocate device memory (no context; scratch is a static array inside
    the library; runs on the current device; count <= SUMSQ_STATIC_MAX_ELEMS)."""
    n = grads[0].numel() if count is None else count
    check(load_library().accel_grad_sumsq_static(_grad_array(grads), len(grads),
                                                 _grad_dtype(grads[0]), n, aggregate, _ptr(out),
                                                 _stream_handle(stream)))


def host_register(array) -> None:
    """Page-lock a host buffer (numpy array or CPU tensor) so ps_step_host copies at PCIe rate."""
    ptr, nbytes = (array.ctypes.data, array.nbytes) if hasattr(array, "ctypes") \
        else (array.data_ptr(), array.numel() * array.element_size())
    check(load_library().accel_host_register(ptr, nbytes))


def host_unregister(array) -> None:
    ptr = array.ctypes.data if hasattr(array, "ctypes") else array.data_ptr()
    check(load_library().accel_host_unregister(ptr))


def adam_advance(hp: HParams) -> None:
    check(load_library().accel_adam_advance(ctypes.byref(hp)))


def version() -> int:
    return load_library().accel_version()


# ---- sharding plans ------------------------------------------------------------
def plan_flat(n_elems: int, n_ps: int, align_elems: int = 1):
    out = (Shard * n_ps)()
    padded = ctypes.c_uint64()
    check(load_library().accel_plan_flat(n_elems, n_ps, align_elems, out, ctypes.byref(padded)))
    return [(s.offset, s.count) for s in out], padded.value


def plan_round_robin(var_elems, n_ps: int, first_task: int = 0, align_elems: int = 1):
    n = len(var_elems)
    ve = (ctypes.c_uint64 * max(n, 1))(*var_elems)
    task = (ctypes.c_int * max(n, 1))()
    off = (ctypes.c_uint64 * max(n, 1))()
    fill = (ctypes.c_uint64 * n_ps)()
    check(load_library().accel_plan_round_robin(ve, n, n_ps, first_task, align_elems, task, off,
                                                fill))
    return list(task)[:n], list(off)[:n], list(fill)
