"""PS tasks sharded over the GPUs of one NVSwitch box: one process per GPU, PS
task_index r = GPU r = rank r = shard r of the flat arena (SURVEY §8 e).

A step is push -> apply_pull -> wait_pull (csrc/ps_shard.cu): the worker->PS push
is a reduce-scatter fused into the local gradient sum, the PS->worker pull an
all-gather fused into the optimizer apply, both as stores over NVLink (unicast
peer stores, or one multicast store through the NVSwitch when a multicast mapping
is available). `accel_shard_step` cuts the arena into slices and overlaps the
push of slice k+1 with the apply_pull of slice k. `mode="nccl"` runs the same
step with NCCL's reduce-scatter / all-gather around the library's kernels — the
baseline the fused path is measured against.

torch.distributed is plumbing only: the rendezvous (exchange of CUDA IPC handles,
or torch's symmetric-memory allocator, which also yields the multicast mapping)
and, in nccl mode, the two collectives.
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
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.int16: "<i2",
                   torch.uint8: "|u1"}[dtype]

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

    def __del__(self):
        # views keep a reference to their DevBuffer (t._accel_owner), so this runs only once
        # no tensor over the allocation is alive
        try:
            self.free()
        except Exception:
            pass


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


class BucketLayout:
    """TF-faithful placement for the sharded step (SURVEY §8 e): variable k (creation order)
    belongs to PS task k mod P (tf.train.replica_device_setter, the call site is
    examples/workdir/mnist_replica.py:137-141); each task's variables are packed into bucket r
    and the buckets are padded to ONE length, so bucket r is exactly shard r of the arena the
    flat step works on — the kernels are the same, only the layout differs. Unlike the flat
    plan the shards are as unbalanced as the round-robin makes them (arena = P x the largest
    bucket)."""

    def __init__(self, var_elems, world: int, first_task: int = 0, align: int = 8):
        self.var_elems = list(var_elems)
        self.world = world
        self.task, self.offset, self.fill = accel.plan_round_robin(self.var_elems, world,
                                                                   first_task, align)
        self.shard = max(8, -(-max(self.fill) // 8) * 8)
        self.arena = self.shard * world
        self.pos = [t * self.shard + o for t, o in zip(self.task, self.offset)]

    def view(self, arena, k: int):
        """Variable k inside an arena-shaped tensor."""
        return arena[self.pos[k]:self.pos[k] + self.var_elems[k]]

    def pack(self, tensors, out=None):
        """Per-variable tensors -> one arena-shaped tensor (padding zero)."""
        import torch
        if out is None:
            out = torch.zeros(self.arena, dtype=tensors[0].dtype, device=tensors[0].device)
        for k, t in enumerate(tensors):
            self.view(out, k).copy_(t.reshape(-1))
        return out


def exchange(obj, world: int):
    """All-gather a picklable object over the job's rendezvous (gloo or nccl group)."""
    import torch.distributed as dist
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out


class SymmBuffer:
    """A buffer from torch's symmetric-memory allocator, rendezvoused over the default
    process group: peers' addresses plus (when the fabric supports it) a multicast address."""

    def __init__(self, nbytes: int, device_index: int):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.nbytes, self.device_index = nbytes, device_index
        dev = torch.device("cuda", device_index)
        group = dist.group.WORLD
        import warnings
        with warnings.catch_warnings():      # newer torch: a deprecated no-op
            warnings.simplefilter("ignore")
            symm.enable_symm_mem_for_group(group.group_name)
        self._t = symm.empty(nbytes // 4, dtype=torch.float32, device=dev)
        self._t.zero_()
        self.handle = symm.rendezvous(self._t, group)
        self.ptr = self._t.data_ptr()
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        mc = getattr(self.handle, "multicast_ptr", 0)
        self.multicast_ptr = int(mc) if mc else 0

    def tensor(self, dtype=torch.float32) -> torch.Tensor:
        return self._t if dtype == torch.float32 else self._t.view(dtype)

    def free(self):
        self._t = None


class ShardedPS:
    def __init__(self, rank: int, world: int, device_index: int, elems: int, optimizer: int,
                 hp: accel.HParams, mode: str = "fused", ctx: accel.Context | None = None,
                 peer_ptrs=None, gpu_ordinals=None, n_chunks: int | None = None,
                 rendezvous: str = "ipc", multicast: bool = False, pull_dtype: str = "f32",
                 overlap: int | None = None, bf16_stage: bool = True):
        if mode not in ("fused", "streams", "nccl", "switch"):
            raise ValueError(f"unknown mode {mode!r}")
        if mode == "switch":
            # the in-switch step is the one-kernel step with ACCEL_OVERLAP_SWITCH; on real peers
            # it needs multicast mappings of stage AND var (symmetric-memory rendezvous)
            if peer_ptrs is None and world > 1:
                rendezvous, multicast = "symm", True
            overlap = accel.OVERLAP_SWITCH
        if rendezvous not in ("ipc", "symm"):
            raise ValueError(f"unknown rendezvous {rendezvous!r}")
        if pull_dtype not in ("f32", "bf16"):
            raise ValueError(f"unknown pull_dtype {pull_dtype!r}")
        if pull_dtype == "bf16" and (mode == "nccl" or multicast or rendezvous != "ipc"):
            raise ValueError("the bf16 pull is implemented for the fused paths over ipc rendezvous")
        self.pull_dtype = pull_dtype
        self.rank, self.world, self.mode = rank, world, mode
        # slices of the step: measured on B200, overlap pays at 2 ranks (4 slices: 0.817 ms vs
        # 0.874 ms) and costs at 8 (1 slice: 1.166 ms, 2: 1.184 ms, 4: 1.299 ms)
        self.n_chunks = n_chunks or (8 if mode == "switch" else
                                     4 if world <= 2 else 2 if world <= 4 else 1)
        # split roles measured faster at 2 ranks (0.773 vs 0.815 ms) and slower at 4 (1.012 vs 0.983)
        self.overlap = overlap if overlap is not None else \
            (accel.OVERLAP_ROLES if world <= 2 else accel.OVERLAP_INLINE)
        self.multicast_ptr = 0
        self.stage_multicast_ptr = 0
        self.bf16_stage = bf16_stage and mode != "nccl"
        self.device_index = device_index
        self.device = torch.device("cuda", device_index)
        self.optimizer, self.hp = optimizer, hp
        self.elems = elems
        self.padded, self.shard = shard_layout(elems, world)
        if ctx is None and gpu_ordinals is None:
            # task r <-> GPU r, except that THIS rank runs on the device it was given (local rank
            # != rank on multi-node jobs or under a CUDA_VISIBLE_DEVICES remap)
            gpu_ordinals = list(range(world))
            gpu_ordinals[rank] = device_index
        self.ctx = ctx or accel.Context(world, gpu_ordinals)
        # the nccl-mode pre-sum is a world-of-one push on THIS rank's GPU
        self.solo_ctx = accel.Context(1, [device_index])
        self.lib = accel.load_library()
        self.global_step = 0
        # exportable buffers: full variable replica, receive slots, flag words
        if rendezvous == "symm" and peer_ptrs is None:
            self.buf_var = SymmBuffer(self.padded * 4, device_index)
            self.buf_stage = SymmBuffer(self.padded * 4, device_index)
            self.buf_flags = SymmBuffer(1024, device_index)
            peer_ptrs = list(zip(self.buf_var.peer_ptrs, self.buf_stage.peer_ptrs,
                                 self.buf_flags.peer_ptrs))
            if multicast:
                self.multicast_ptr = self.buf_var.multicast_ptr
            if mode == "switch":
                self.stage_multicast_ptr = self.buf_stage.multicast_ptr
                if not (self.multicast_ptr and self.stage_multicast_ptr):
                    raise RuntimeError("mode='switch' needs NVSwitch multicast mappings; the "
                                       "symmetric-memory rendezvous returned none on this fabric")
        else:
            self.buf_var = DevBuffer(self.padded * 4, device_index)
            self.buf_stage = DevBuffer(self.padded * 4, device_index)
            self.buf_flags = DevBuffer(1024, device_index)
        # bf16 pull: workers read a bf16 replica, the owner keeps the fp32 master of its shard
        self.buf_var16 = DevBuffer(self.padded * 2, device_index) if pull_dtype == "bf16" else None
        self.var_bf16 = self.buf_var16.tensor(torch.int16).view(torch.bfloat16) \
            if self.buf_var16 else None
        self.var = self.buf_var.tensor()
        self.stage = self.buf_stage.tensor()
        self.flags = self.buf_flags.tensor(torch.int32)
        with torch.cuda.device(device_index):
            self.slot0 = torch.zeros(self.shard, device=self.device) \
                if optimizer >= accel.OPT_MOMENTUM else None
            self.slot1 = torch.zeros(self.shard, device=self.device) \
                if optimizer == accel.OPT_ADAM else None
            self.gshard = torch.empty(self.shard, device=self.device) if mode == "nccl" else None
        n_slices = -(-(self.shard // 8) // (-(-(self.shard // 8) // self.n_chunks)))
        self.launches_per_step = {"fused": 1, "switch": 1, "streams": 2 * n_slices + 1,
                                  "nccl": 2}[mode]
        self._opened = []     # peers' allocations mapped into this process (accel_ipc_open)
        if peer_ptrs is None:
            mine = (self.buf_var.export(), self.buf_stage.export(), self.buf_flags.export(),
                    self.buf_var16.export() if self.buf_var16 else None)
            handles = exchange(mine, world)
            peer_ptrs = []
            with torch.cuda.device(device_index):
                for r, (hv, hs, hf, h16) in enumerate(handles):
                    if r == rank:
                        peer_ptrs.append((self.buf_var.ptr, self.buf_stage.ptr, self.buf_flags.ptr,
                                          self.buf_var16.ptr if self.buf_var16 else 0))
                    else:
                        mapped = (open_peer(hv), open_peer(hs), open_peer(hf),
                                  open_peer(h16) if h16 else 0)
                        self._opened += [m for m in mapped if m]
                        peer_ptrs.append(mapped)
        self.set_peers(peer_ptrs)

    def close(self):
        """Unmap the peers' buffers. An exported allocation is only returned to the device once
        every process that mapped it has closed its handle, so a rank that drops its ShardedPS
        without this keeps its peers' memory alive."""
        opened, self._opened = getattr(self, "_opened", []), []
        if opened:
            with torch.cuda.device(self.device_index):
                torch.cuda.synchronize()
                for ptr in opened:
                    self.lib.accel_ipc_close(ptr)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_peers(self, peer_ptrs):
        """peer_ptrs[r] = (var, stage, flags) device addresses of rank r as seen from here."""
        w = self.world
        self._var_arr = (ctypes.c_void_p * w)(*[p[0] for p in peer_ptrs])
        self._stage_arr = (ctypes.c_void_p * w)(*[p[1] for p in peer_ptrs])
        self._flags_arr = (ctypes.c_void_p * w)(*[p[2] for p in peer_ptrs])
        self._var16_arr = None
        if self.pull_dtype == "bf16":
            self._var16_arr = (ctypes.c_void_p * w)(*[p[3] for p in peer_ptrs])
        self.peers = accel.Peers(self.rank, w, self.shard, self._stage_arr, self._var_arr,
                                 self._flags_arr, self.multicast_ptr or None, self._var16_arr,
                                 self.stage_multicast_ptr or None, accel.F32, 0)
        # the same view with raw-bf16 stage slots: used when a step pushes ONE bf16 gradient
        self.peers_b16 = accel.Peers(self.rank, w, self.shard, self._stage_arr, self._var_arr,
                                     self._flags_arr, self.multicast_ptr or None, self._var16_arr,
                                     self.stage_multicast_ptr or None, accel.BF16, 0)
        # world = 1 view onto the same stage buffer: the local pre-sum of nccl mode
        self._solo_var = (ctypes.c_void_p * 1)(self.buf_var.ptr)
        self._solo_stage = (ctypes.c_void_p * 1)(self.buf_stage.ptr)
        self._solo_flags = (ctypes.c_void_p * 1)(self.buf_flags.ptr)
        self.solo = accel.Peers(0, 1, self.padded, self._solo_stage, self._solo_var,
                                self._solo_flags, None, None, None, accel.F32, 0)

    def stage_grad(self, dtype=torch.float32) -> torch.Tensor:
        """The view of this rank's stage buffer a worker can write its ONE gradient into (arena
        layout), for the zero-copy form of the in-switch step: pass it back as the only gradient of
        step() and nothing is copied before the switch reads it. bf16 needs bf16_stage."""
        if dtype == torch.bfloat16:
            return self.stage.view(torch.bfloat16)[:self.padded]
        return self.stage[:self.padded]

    @property
    def var_shard(self) -> torch.Tensor:
        return self.var[self.rank * self.shard:(self.rank + 1) * self.shard]

    # ---- the three phases (exposed so a single GPU can emulate several ranks in turn) ----
    def peers_for(self, grads):
        """The peer view a step with these gradients uses: raw-bf16 stage slots when every rank
        pushes exactly one bf16 gradient (half the reduce-scatter bytes, same bits)."""
        import torch
        b16 = self.bf16_stage and len(grads) == 1 and grads[0].dtype == torch.bfloat16
        self._last_peers = self.peers_b16 if b16 else self.peers
        return self._last_peers

    def push(self, grads, stream=None, peers=None, sub=(0, 0)) -> int:
        """Returns the epoch this push opened."""
        ctx = self.solo_ctx if peers is self.solo else self.ctx
        epoch = ctypes.c_uint32()
        accel.check(self.lib.accel_shard_push(
            ctx.handle, ctypes.byref(peers or self.peers_for(grads)), accel._grad_array(grads),
            len(grads), accel._grad_dtype(grads[0]), sub[0], sub[1],
            accel._stream_handle(stream), ctypes.byref(epoch)))
        return epoch.value

    def apply_pull(self, n_contrib: int, sumsq=None, stream=None, sub=(0, 0), epoch=0):
        accel.check(self.lib.accel_shard_apply_pull(
            self.ctx.handle, ctypes.byref(getattr(self, "_last_peers", self.peers)),
            self.optimizer, accel._ptr(self.slot0),
            accel._ptr(self.slot1), n_contrib, ctypes.byref(self.hp), accel._ptr(sumsq),
            sub[0], sub[1], epoch, accel._stream_handle(stream)))

    def wait_push(self, stream=None, epoch=0):
        accel.check(self.lib.accel_shard_wait_push(self.ctx.handle, ctypes.byref(self.peers),
                                                   epoch, accel._stream_handle(stream)))

    def shard_sumsq(self, out, stream=None):
        """Sum of squares of THIS shard's aggregated gradient (the `world` stage slots summed in
        rank order) into out[0]; call after wait_push. Add the ranks' results for the global
        norm."""
        import torch
        if getattr(self, "_last_peers", self.peers) is self.peers_b16:
            st16 = self.stage.view(torch.bfloat16)      # raw-bf16 slots, `shard` elements apart
            slots = [st16[s * self.shard:(s + 1) * self.shard] for s in range(self.world)]
        else:
            slots = [self.stage[s * self.shard:(s + 1) * self.shard] for s in range(self.world)]
        self.ctx.grad_sumsq(slots, out, aggregate=accel.AGG_SUM, task=self.rank,
                            count=self.shard, stream=stream)

    def wait_pull(self, stream=None, epoch=0):
        accel.check(self.lib.accel_shard_wait_pull(self.ctx.handle, ctypes.byref(self.peers),
                                                   epoch, accel._stream_handle(stream)))

    def slices(self):
        """The (begin, count) slices accel_shard_step cuts every shard into."""
        per = -(-(self.shard // 8) // self.n_chunks) * 8
        return [(b, min(per, self.shard - b)) for b in range(0, self.shard, per)]

    def check(self, stream=None):
        """Synchronise and raise if a cross-GPU wait of this rank timed out."""
        accel.check(self.lib.accel_shard_check(self.ctx.handle, ctypes.byref(self.peers),
                                               accel._stream_handle(stream)))

    def reset(self, stream=None):
        """After a timeout (check() raised): zero this rank's flag words and restart its epoch.
        Every rank must call it, with a barrier before the next step."""
        accel.check(self.lib.accel_shard_reset(self.ctx.handle, ctypes.byref(self.peers),
                                               accel._stream_handle(stream)))

    def step_host(self, host_grads, host_var_out=None, n_slices=16):
        """The step with this rank's worker gradients in pinned HOST memory
        (accel_shard_step_host): H2D of slice k+1, the sharded step on slice k and the D2H of
        slice k-1's part of the replica overlap; returns when the step is complete."""
        n_contrib = len(host_grads) * self.world
        accel.check(self.lib.accel_shard_step_host(
            self.ctx.handle, ctypes.byref(self.peers_for(host_grads)), self.optimizer,
            accel._ptr(self.slot0), accel._ptr(self.slot1), accel._grad_array(host_grads),
            len(host_grads), accel._grad_dtype(host_grads[0]), n_contrib, ctypes.byref(self.hp),
            accel._ptr(host_var_out), n_slices))
        self.finish_step()

    def finish_step(self):
        if self.optimizer == accel.OPT_ADAM:
            accel.adam_advance(self.hp)
        self.global_step += 1

    def step(self, grads, stream=None, sumsq=None):
        """grads: this rank's local worker gradients, each `padded` elements."""
        n_contrib = len(grads) * self.world
        if self.hp.clip_norm > 0 and sumsq is None and self.mode != "nccl":
            # global-norm clipping needs the whole aggregated gradient's norm before any shard
            # is applied: push, reduce the per-shard sums of squares across ranks, then apply
            import torch.distributed as dist
            if self.hp.aggregate != accel.AGG_SUM:
                raise ValueError("clipping in the sharded step is implemented for SUM aggregation")
            ss = torch.zeros(1, device=self.device)
            epoch = self.push(grads, stream)
            self.wait_push(stream, epoch)
            self.shard_sumsq(ss, stream)
            if self.world > 1:
                dist.all_reduce(ss, op=dist.ReduceOp.SUM)
            self.apply_pull(n_contrib, ss, stream, epoch=epoch)
            self.wait_pull(stream, epoch)
        elif self.mode in ("fused", "streams", "switch"):
            head = (self.ctx.handle, ctypes.byref(self.peers_for(grads)), self.optimizer,
                    accel._ptr(self.slot0), accel._ptr(self.slot1), accel._grad_array(grads),
                    len(grads), accel._grad_dtype(grads[0]), n_contrib, ctypes.byref(self.hp),
                    accel._ptr(sumsq), self.n_chunks)
            if self.mode in ("fused", "switch"):      # one cooperative kernel
                accel.check(self.lib.accel_shard_step_fused(*head, self.overlap,
                                                            accel._stream_handle(stream)))
            else:                         # separate kernels, slices pipelined over two streams
                accel.check(self.lib.accel_shard_step(*head, accel._stream_handle(stream)))
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


def local_group(world: int, elems: int, optimizer: int, hp_factory, device_index: int = 0,
                n_chunks: int = 1, pull_dtype: str = "f32", bf16_stage: bool = True):
    """`world` ranks emulated on ONE GPU inside one process (tests): every rank's buffers are
    plain local allocations and the phases are run rank after rank on one stream, so every
    device-side wait is already satisfied when its kernel starts."""
    ctx = accel.Context(world, [device_index] * world)
    ranks = []
    for r in range(world):
        ranks.append(ShardedPS(r, world, device_index, elems, optimizer, hp_factory(), ctx=ctx,
                               peer_ptrs=[(0, 0, 0, 0)] * world, n_chunks=n_chunks,
                               pull_dtype=pull_dtype, bf16_stage=bf16_stage))
    ptrs = [(x.buf_var.ptr, x.buf_stage.ptr, x.buf_flags.ptr,
             x.buf_var16.ptr if x.buf_var16 else 0) for x in ranks]
    for x in ranks:
        x.set_peers(ptrs)
    return ranks


def local_group_step(ranks, grads_per_rank, stream=None, sumsq=None):
    """One step of the emulated group, slice by slice and phase by phase: all ranks push slice
    k, then all ranks apply_pull slice k (its waits are already satisfied), as the pipelined
    accel_shard_step orders them on real peers."""
    n_contrib = sum(len(g) for g in grads_per_rank)
    epoch = 0
    for sub in ranks[0].slices():
        epochs = [x.push(g, stream, sub=sub) for x, g in zip(ranks, grads_per_rank)]
        assert len(set(epochs)) == 1, "ranks must advance their epochs in lock step"
        epoch = epochs[0]
        for x in ranks:
            x.apply_pull(n_contrib, sumsq, stream, sub=sub, epoch=epoch)
    for x in ranks:
        x.wait_pull(stream, epoch=epoch)
    for x in ranks:
        x.check(stream)
        x.finish_step()


def local_group_step_fused(ranks, grads_per_rank, n_slices=None, stream=None, sumsq=None,
                           overlap=accel.OVERLAP_INLINE):
    """One step of the emulated group through the ONE-kernel path: a single cooperative
    launch plays every rank (accel_shard_step_local_group)."""
    n = len(ranks)
    x0 = ranks[0]
    n_contrib = sum(len(g) for g in grads_per_rank)
    PP = ctypes.POINTER(accel.Peers)
    peers = (PP * n)(*[ctypes.pointer(x.peers_for(g)) for x, g in zip(ranks, grads_per_rank)])
    s0 = (ctypes.c_void_p * n)(*[accel._ptr(x.slot0) for x in ranks])
    s1 = (ctypes.c_void_p * n)(*[accel._ptr(x.slot1) for x in ranks])
    garr = [accel._grad_array(g) for g in grads_per_rank]
    pvp = ctypes.POINTER(ctypes.c_void_p)
    gg = (pvp * n)(*[ctypes.cast(a, pvp) for a in garr])
    accel.check(x0.lib.accel_shard_step_local_group(
        x0.ctx.handle, peers, n, x0.optimizer, s0, s1, gg, len(grads_per_rank[0]),
        accel._grad_dtype(grads_per_rank[0][0]), n_contrib, ctypes.byref(x0.hp),
        accel._ptr(sumsq), n_slices or x0.n_chunks, overlap, accel._stream_handle(stream)))
    for x in ranks:
        x.check(stream)
        x.finish_step()
