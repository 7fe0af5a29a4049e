"""PS tasks sharded over the GPUs of one NVSwitch box: one process per GPU, PS
task_index r = GPU r = rank r = shard r of the flat arena (SURVEY §8 e).

A step is push -> apply_pull -> wait_pull (csrc/ps_shard.cu): the worker->PS push
is a reduce-scatter fused into the local gradient sum, the PS->worker pull an
all-gather fused into the optimizer apply, both as peer stores over NVLink.
`mode="nccl"` runs the same step with NCCL's reduce-scatter / all-gather around
the library's kernels — the baseline the fused path is measured against.

torch.distributed is used only for the rendezvous (exchange of IPC handles) and,
in nccl mode, for the two collectives.
"""
from __future__ import annotations

import ctypes

import torch

from . import accel


class DevBuffer:
    """One exportable device allocation (accel_dev_alloc), viewable as a torch tensor."""

    def __init__(self, nbytes: int, device_index: int):
        self.lib = accel.load_library()
        self.nbytes = nbytes
        self.device_index = device_index
        p = ctypes.c_void_p()
        with torch.cuda.device(device_index):
            accel.check(self.lib.accel_dev_alloc(nbytes, ctypes.byref(p)))
        self.ptr = p.value

    def tensor(self, dtype=torch.float32) -> torch.Tensor:
        item = torch.empty(0, dtype=dtype).element_size()
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]

        class _View:
            __cuda_array_interface__ = {"shape": (self.nbytes // item,), "typestr": typestr,
                                        "data": (self.ptr, False), "version": 2}
        t = torch.as_tensor(_View(), device=torch.device("cuda", self.device_index))
        t._accel_owner = self   # the allocation outlives every view of it
        return t

    def export(self) -> bytes:
        h = accel.IpcHandle()
        accel.check(self.lib.accel_ipc_export(self.ptr, ctypes.byref(h)))
        return bytes(h.bytes)

    def free(self):
        if self.ptr:
            with torch.cuda.device(self.device_index):
                self.lib.accel_dev_free(self.ptr)
            self.ptr = None


def open_peer(handle: bytes) -> int:
    h = accel.IpcHandle()
    ctypes.memmove(h.bytes, handle, len(handle))
    p = ctypes.c_void_p()
    accel.check(accel.load_library().accel_ipc_open(ctypes.byref(h), ctypes.byref(p)))
    return p.value


def shard_layout(elems: int, world: int, align: int = 8):
    """Flat plan (accel_plan_flat): (padded elements, elements per shard)."""
    _, padded = accel.plan_flat(elems, world, align)
    return padded, padded // world


def exchange(obj, world: int):
    """All-gather a picklable object over the job's rendezvous (gloo or nccl group)."""
    import torch.distributed as dist
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out


class ShardedPS:
    def __init__(self, rank: int, world: int, device_index: int, elems: int, optimizer: int,
                 hp: accel.HParams, mode: str = "fused", ctx: accel.Context | None = None,
                 peer_ptrs=None, gpu_ordinals=None):
        if mode not in ("fused", "nccl"):
            raise ValueError(f"unknown mode {mode!r}")
        self.rank, self.world, self.mode = rank, world, mode
        self.device_index = device_index
        self.device = torch.device("cuda", device_index)
        self.optimizer, self.hp = optimizer, hp
        self.elems = elems
        self.padded, self.shard = shard_layout(elems, world)
        self.ctx = ctx or accel.Context(world, gpu_ordinals)
        # the nccl-mode pre-sum is a world-of-one push on THIS rank's GPU
        self.solo_ctx = accel.Context(1, [device_index])
        self.lib = accel.load_library()
        self.global_step = 0
        # exportable buffers: full variable replica, receive slots, flag words
        self.buf_var = DevBuffer(self.padded * 4, device_index)
        self.buf_stage = DevBuffer(self.padded * 4, device_index)
        self.buf_flags = DevBuffer(256, device_index)
        self.var = self.buf_var.tensor()
        self.stage = self.buf_stage.tensor()
        self.flags = self.buf_flags.tensor(torch.int32)
        with torch.cuda.device(device_index):
            self.slot0 = torch.zeros(self.shard, device=self.device) \
                if optimizer >= accel.OPT_MOMENTUM else None
            self.slot1 = torch.zeros(self.shard, device=self.device) \
                if optimizer == accel.OPT_ADAM else None
            self.gshard = torch.empty(self.shard, device=self.device) if mode == "nccl" else None
        self.launches_per_step = 3 if mode == "fused" else 2
        if peer_ptrs is None:
            mine = (self.buf_var.export(), self.buf_stage.export(), self.buf_flags.export())
            handles = exchange(mine, world)
            peer_ptrs = []
            with torch.cuda.device(device_index):
                for r, (hv, hs, hf) in enumerate(handles):
                    if r == rank:
                        peer_ptrs.append((self.buf_var.ptr, self.buf_stage.ptr, self.buf_flags.ptr))
                    else:
                        peer_ptrs.append((open_peer(hv), open_peer(hs), open_peer(hf)))
        self.set_peers(peer_ptrs)

    def set_peers(self, peer_ptrs):
        """peer_ptrs[r] = (var, stage, flags) device addresses of rank r as seen from here."""
        w = self.world
        self._var_arr = (ctypes.c_void_p * w)(*[p[0] for p in peer_ptrs])
        self._stage_arr = (ctypes.c_void_p * w)(*[p[1] for p in peer_ptrs])
        self._flags_arr = (ctypes.c_void_p * w)(*[p[2] for p in peer_ptrs])
        self.peers = accel.Peers(self.rank, w, self.shard, self._stage_arr, self._var_arr,
                                 self._flags_arr)
        # world = 1 view onto the same stage buffer: the local pre-sum of nccl mode
        self._solo_var = (ctypes.c_void_p * 1)(self.buf_var.ptr)
        self._solo_stage = (ctypes.c_void_p * 1)(self.buf_stage.ptr)
        self._solo_flags = (ctypes.c_void_p * 1)(self.buf_flags.ptr)
        self.solo = accel.Peers(0, 1, self.padded, self._solo_stage, self._solo_var,
                                self._solo_flags)

    @property
    def var_shard(self) -> torch.Tensor:
        return self.var[self.rank * self.shard:(self.rank + 1) * self.shard]

    # ---- the three phases (exposed so a single GPU can emulate several ranks in turn) ----
    def push(self, grads, stream=None, peers=None):
        ctx = self.solo_ctx if peers is self.solo else self.ctx
        accel.check(self.lib.accel_shard_push(
            ctx.handle, ctypes.byref(peers or self.peers), accel._grad_array(grads),
            len(grads), accel._grad_dtype(grads[0]), accel._stream_handle(stream)))

    def apply_pull(self, n_contrib: int, sumsq=None, stream=None):
        accel.check(self.lib.accel_shard_apply_pull(
            self.ctx.handle, ctypes.byref(self.peers), self.optimizer, accel._ptr(self.slot0),
            accel._ptr(self.slot1), n_contrib, ctypes.byref(self.hp), accel._ptr(sumsq),
            accel._stream_handle(stream)))

    def wait_pull(self, stream=None):
        accel.check(self.lib.accel_shard_wait_pull(self.ctx.handle, ctypes.byref(self.peers),
                                                   accel._stream_handle(stream)))

    def check(self, stream=None):
        """Synchronise and raise if a cross-GPU wait of this rank timed out."""
        accel.check(self.lib.accel_shard_check(self.ctx.handle, ctypes.byref(self.peers),
                                               accel._stream_handle(stream)))

    def finish_step(self):
        if self.optimizer == accel.OPT_ADAM:
            accel.adam_advance(self.hp)
        self.global_step += 1

    def step(self, grads, stream=None, sumsq=None):
        """grads: this rank's local worker gradients, each `padded` elements."""
        n_contrib = len(grads) * self.world
        if self.mode == "fused":
            self.push(grads, stream)
            self.apply_pull(n_contrib, sumsq, stream)
            self.wait_pull(stream)
        else:
            import torch.distributed as dist
            if self.hp.aggregate != accel.AGG_SUM:
                raise ValueError("nccl baseline mode implements SUM aggregation only")
            self.push(grads, stream, peers=self.solo)       # local N-way sum -> stage
            dist.reduce_scatter_tensor(self.gshard, self.stage, op=dist.ReduceOp.SUM)
            self.ctx.apply(self.optimizer, self.var_shard, self.slot0, self.slot1, [self.gshard],
                           self.hp, task=self.rank, count=self.shard, sumsq=sumsq, stream=stream)
            dist.all_gather_into_tensor(self.var, self.var_shard.clone())
        self.finish_step()


def local_group(world: int, elems: int, optimizer: int, hp_factory, device_index: int = 0):
    """`world` ranks emulated on ONE GPU inside one process (tests): every rank's buffers are
    plain local allocations and the phases are run rank after rank on one stream, so every
    device-side wait is already satisfied when its kernel starts."""
    ctx = accel.Context(world, [device_index] * world)
    ranks = []
    for r in range(world):
        ranks.append(ShardedPS(r, world, device_index, elems, optimizer, hp_factory(), ctx=ctx,
                               peer_ptrs=[(0, 0, 0)] * world))
    ptrs = [(x.buf_var.ptr, x.buf_stage.ptr, x.buf_flags.ptr) for x in ranks]
    for x in ranks:
        x.set_peers(ptrs)
    return ranks


def local_group_step(ranks, grads_per_rank, stream=None, sumsq=None):
    n_contrib = sum(len(g) for g in grads_per_rank)
    for x, g in zip(ranks, grads_per_rank):
        x.push(g, stream)
    for x in ranks:
        x.apply_pull(n_contrib, sumsq, stream)
    for x in ranks:
        x.wait_pull(stream)
    for x in ranks:
        x.check(stream)
        x.finish_step()
