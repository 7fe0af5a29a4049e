package accel

/*
#include <stdlib.h>
#include "accel_ps.h"
*/
import "C"

import (
	"fmt"
	"unsafe"
)

// NOTE: like the rest of this package, written against include/accel_ps.h and never compiled in
// this repository's build image (no Go toolchain). One Go function per C entry point.

// IpcHandle is the 64-byte token that lets another PS process of the box map a buffer.
type IpcHandle [C.ACCEL_IPC_HANDLE_BYTES]byte

// DevAlloc returns one exportable, zero-filled device allocation on the current device.
func DevAlloc(bytes uint64) (DevPtr, error) {
	var p unsafe.Pointer
	err := locked(func() C.int { return C.accel_dev_alloc(C.size_t(bytes), &p) })
	return DevPtr(uintptr(p)), err
}

// DevFree releases a DevAlloc allocation.
func DevFree(p DevPtr) error { return locked(func() C.int { return C.accel_dev_free(unsafe.Pointer(uintptr(p))) }) }

// IpcExport / IpcOpen / IpcClose move a buffer between the PS processes of one box.
func IpcExport(p DevPtr) (IpcHandle, error) {
	var h C.accel_ipc_handle
	err := locked(func() C.int { return C.accel_ipc_export(unsafe.Pointer(uintptr(p)), &h) })
	var out IpcHandle
	for i := range out {
		out[i] = byte(h.bytes[i])
	}
	return out, err
}

func IpcOpen(h IpcHandle) (DevPtr, error) {
	var ch C.accel_ipc_handle
	for i := range h {
		ch.bytes[i] = C.uchar(h[i])
	}
	var p unsafe.Pointer
	err := locked(func() C.int { return C.accel_ipc_open(&ch, &p) })
	return DevPtr(uintptr(p)), err
}

func IpcClose(p DevPtr) error { return locked(func() C.int { return C.accel_ipc_close(unsafe.Pointer(uintptr(p))) }) }

// Peers is one rank's view of the sharded group: entry [r] of every slice is rank r's buffer as
// mapped into THIS process (entry [Rank] is the local one). VarBF16 may be nil (fp32 pull).
type Peers struct {
	Rank, World int
	ShardElems  uint64
	Stage, Var  []DevPtr
	Flags       []DevPtr
	VarBF16     []DevPtr
	// NVSwitch multicast mappings over every rank's var / stage buffer (0 = none); both are
	// needed by the in-switch step (overlap = OverlapSwitch).
	VarMulticast, StageMulticast DevPtr
	// StageDType BF16: every rank pushes ONE bf16 gradient, which travels as raw bf16.
	StageDType DType
}

// How the one-kernel step overlaps its phases (accel_overlap).
const (
	OverlapInline = int(C.ACCEL_OVERLAP_INLINE)
	OverlapRoles  = int(C.ACCEL_OVERLAP_ROLES)
	OverlapSwitch = int(C.ACCEL_OVERLAP_SWITCH)
)

// cPeers builds the C struct in C memory; call the returned function to release it.
func (p *Peers) cPeers() (*C.accel_peers, func(), error) {
	if len(p.Stage) != p.World || len(p.Var) != p.World || len(p.Flags) != p.World {
		return nil, nil, fmt.Errorf("accel: peers need %d entries per array", p.World)
	}
	stage, f1 := gradArray(p.Stage)
	vr, f2 := gradArray(p.Var)
	fl, f3 := gradArray(p.Flags)
	frees := []func(){f1, f2, f3}
	cp := (*C.accel_peers)(C.calloc(1, C.size_t(unsafe.Sizeof(C.accel_peers{}))))
	cp.rank, cp.world = C.int32_t(p.Rank), C.int32_t(p.World)
	cp.shard_elems = C.uint64_t(p.ShardElems)
	cp.stage = (**C.float)(unsafe.Pointer(stage))
	cp._var = (**C.float)(unsafe.Pointer(vr)) // cgo renames the C field `var`
	cp.flags = (**C.uint32_t)(unsafe.Pointer(fl))
	if p.VarBF16 != nil {
		b, f4 := gradArray(p.VarBF16)
		frees = append(frees, f4)
		cp.var_bf16 = (**C.uint16_t)(unsafe.Pointer(b))
	}
	cp.var_multicast = (*C.float)(unsafe.Pointer(uintptr(p.VarMulticast)))
	cp.stage_multicast = (*C.float)(unsafe.Pointer(uintptr(p.StageMulticast)))
	cp.stage_dtype = C.int32_t(p.StageDType)
	return cp, func() {
		for _, f := range frees {
			f()
		}
		C.free(unsafe.Pointer(cp))
	}, nil
}

// ShardStepFused runs one whole sharded PS step (push, apply, pull, closing wait) as one
// cooperative kernel on this rank's GPU. nContrib = worker gradients summed over all ranks.
func (c *Context) ShardStepFused(p *Peers, optimizer int, slot0, slot1 DevPtr, grads []DevPtr,
	dt DType, nContrib int, hp *HParams, sumsq DevPtr, nSlices, overlap int, stream uintptr) error {
	cp, free, err := p.cPeers()
	if err != nil {
		return err
	}
	defer free()
	g, freeG := gradArray(grads)
	defer freeG()
	ch := hp.c()
	return locked(func() C.int { return C.accel_shard_step_fused(c.h, cp, C.int(optimizer),
		(*C.float)(unsafe.Pointer(uintptr(slot0))), (*C.float)(unsafe.Pointer(uintptr(slot1))), g,
		C.int(len(grads)), C.int(dt), C.int(nContrib), &ch,
		(*C.float)(unsafe.Pointer(uintptr(sumsq))), C.int(nSlices), C.int(overlap),
		C.accel_stream_t(unsafe.Pointer(stream))) })
}

// ShardCheck synchronises the stream and reports a peer that never arrived as an error.
func (c *Context) ShardCheck(p *Peers, stream uintptr) error {
	cp, free, err := p.cPeers()
	if err != nil {
		return err
	}
	defer free()
	return locked(func() C.int { return C.accel_shard_check(c.h, cp, C.accel_stream_t(unsafe.Pointer(stream))) })
}

// ShardReset recovers after ShardCheck reported a timeout: every rank calls it, then a barrier.
func (c *Context) ShardReset(p *Peers, stream uintptr) error {
	cp, free, err := p.cPeers()
	if err != nil {
		return err
	}
	defer free()
	return locked(func() C.int { return C.accel_shard_reset(c.h, cp, C.accel_stream_t(unsafe.Pointer(stream))) })
}

// ShardStepHost is the sharded step with this rank's worker gradients in page-locked HOST memory;
// the updated replica is copied to hostVarOut (nil to skip). Returns when the step is complete.
func (c *Context) ShardStepHost(p *Peers, optimizer int, slot0, slot1 DevPtr,
	hostGrads []unsafe.Pointer, dt DType, nContrib int, hp *HParams, hostVarOut unsafe.Pointer,
	nSlices int) error {
	cp, free, err := p.cPeers()
	if err != nil {
		return err
	}
	defer free()
	ptrs := make([]DevPtr, len(hostGrads))
	for i, q := range hostGrads {
		ptrs[i] = DevPtr(uintptr(q))
	}
	g, freeG := gradArray(ptrs)
	defer freeG()
	ch := hp.c()
	return locked(func() C.int { return C.accel_shard_step_host(c.h, cp, C.int(optimizer),
		(*C.float)(unsafe.Pointer(uintptr(slot0))), (*C.float)(unsafe.Pointer(uintptr(slot1))), g,
		C.int(len(hostGrads)), C.int(dt), C.int(nContrib), &ch, (*C.float)(hostVarOut),
		C.int(nSlices)) })
}

// GradSumSqStatic is GradSumSq for callers that must not allocate device memory: no context,
// scratch built into the library, at most 2^28 elements per call. It runs on the CURRENT CUDA
// device, which is per-OS-thread state: the caller must hold runtime.LockOSThread() from its
// cudaSetDevice until this call returns (a goroutine that migrated in between would launch on
// another thread's device with pointers of the intended one).
func GradSumSqStatic(grads []DevPtr, dt DType, count uint64, agg Aggregate, out DevPtr,
	stream uintptr) error {
	g, free := gradArray(grads)
	defer free()
	return locked(func() C.int { return C.accel_grad_sumsq_static(g, C.int(len(grads)), C.int(dt), C.uint64_t(count),
		C.int(agg), (*C.float)(unsafe.Pointer(uintptr(out))),
		C.accel_stream_t(unsafe.Pointer(stream))) })
}

// OptStateInit fills the 16-byte device block {beta1_power, beta2_power, global_step}.
func OptStateInit(devState DevPtr, hp *HParams, globalStep uint64, stream uintptr) error {
	ch := hp.c()
	return locked(func() C.int { return C.accel_opt_state_init(unsafe.Pointer(uintptr(devState)), &ch,
		C.uint64_t(globalStep), C.accel_stream_t(unsafe.Pointer(stream))) })
}

// ApplyDevState is the fused apply with the optimizer state advancing on the device.
func (c *Context) ApplyDevState(task, optimizer int, v, slot0, slot1 DevPtr, grads []DevPtr,
	dt DType, count uint64, hp *HParams, sumsq, devState DevPtr, stream uintptr) error {
	g, free := gradArray(grads)
	defer free()
	ch := hp.c()
	return locked(func() C.int { return C.accel_apply_dev_state(c.h, C.int(task), C.int(optimizer),
		(*C.float)(unsafe.Pointer(uintptr(v))), (*C.float)(unsafe.Pointer(uintptr(slot0))),
		(*C.float)(unsafe.Pointer(uintptr(slot1))), g, C.int(len(grads)), C.int(dt),
		C.uint64_t(count), &ch, (*C.float)(unsafe.Pointer(uintptr(sumsq))),
		unsafe.Pointer(uintptr(devState)), C.accel_stream_t(unsafe.Pointer(stream))) })
}

// SetStepShape tunes the one-kernel step: reducerPct = share of CTAs that reduce + apply + pull in
// the in-switch form (0 = default), coop = cooperative launch (false: plain launch for small grids).
func (c *Context) SetStepShape(reducerPct int, coop bool) error {
	k := C.int(0)
	if coop {
		k = 1
	}
	return locked(func() C.int { return C.accel_ctx_set_step_shape(c.h, C.int(reducerPct), k) })
}
