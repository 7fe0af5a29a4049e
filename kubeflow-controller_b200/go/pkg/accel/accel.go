// Package accel is the Go face of libaccel_ps.so: the B200-native parameter-server
// step (fused N-way gradient aggregation + SGD / Momentum / Adam apply) for the PS
// replicas of a TFJob. It is a thin cgo wrapper over include/accel_ps.h and adds
// nothing of its own; pkg/tensorflow and the TFJob CRD stay unchanged.
//
// NOTE: written against the C ABI but NOT compiled in this repository's build image
// (no Go toolchain there). It is kept mechanical on purpose: one Go function per C
// entry point, Go error <- negative status + accel_last_error().
package accel

/*
#cgo CFLAGS: -I${SRCDIR}/../../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../.. -laccel_ps -Wl,-rpath,${SRCDIR}/../../..
#include <stdlib.h>
#include "accel_ps.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

// DType of the worker gradients.
type DType int

const (
	F32  DType = C.ACCEL_F32
	BF16 DType = C.ACCEL_BF16
)

// Aggregate selects sum or mean of the worker gradients.
type Aggregate int

const (
	Sum  Aggregate = C.ACCEL_AGG_SUM
	Mean Aggregate = C.ACCEL_AGG_MEAN
)

// HParams mirrors accel_hparams.
type HParams struct {
	LR, Momentum, Beta1, Beta2, Epsilon float32
	Beta1Power, Beta2Power              float32
	ClipNorm                            float32
	UseNesterov                         bool
	Aggregate                           Aggregate
}

func (h *HParams) c() C.accel_hparams {
	n := C.int32_t(0)
	if h.UseNesterov {
		n = 1
	}
	return C.accel_hparams{
		lr: C.float(h.LR), momentum: C.float(h.Momentum), beta1: C.float(h.Beta1),
		beta2: C.float(h.Beta2), epsilon: C.float(h.Epsilon),
		beta1_power: C.float(h.Beta1Power), beta2_power: C.float(h.Beta2Power),
		clip_norm: C.float(h.ClipNorm), use_nesterov: n, aggregate: C.int32_t(h.Aggregate),
	}
}

// AdvanceAdam multiplies the beta powers once per step (TF adam.py _finish).
func (h *HParams) AdvanceAdam() {
	ch := h.c()
	C.accel_adam_advance(&ch)
	h.Beta1Power, h.Beta2Power = float32(ch.beta1_power), float32(ch.beta2_power)
}

func status(rc C.int) error {
	if rc >= 0 {
		return nil
	}
	return fmt.Errorf("accel: status %d: %s", int(rc), C.GoString(C.accel_last_error()))
}

// locked runs ONE C call and, if it failed, reads its message — both on the same OS thread.
// accel_last_error() is thread-local in the library and goroutines migrate between OS threads
// from one cgo call to the next, so the call and the fetch must not be separated. The library
// also sets the CUDA device per call (task -> GPU), which is per-thread state as well.
func locked(f func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	return status(f())
}

// Context maps PS task_index -> GPU ordinal and owns per-GPU scratch.
type Context struct{ h *C.accel_ctx }

// NewContext creates a context for nPS parameter-server tasks; gpuOrdinals may be nil
// (identity: PS task i on GPU i).
func NewContext(nPS int, gpuOrdinals []int) (*Context, error) {
	var p *C.int
	if gpuOrdinals != nil {
		if len(gpuOrdinals) != nPS {
			return nil, fmt.Errorf("accel: %d ordinals for %d PS tasks", len(gpuOrdinals), nPS)
		}
		arr := make([]C.int, nPS)
		for i, g := range gpuOrdinals {
			arr[i] = C.int(g)
		}
		p = &arr[0]
	}
	c := &Context{}
	if err := locked(func() C.int { return C.accel_ctx_create(p, C.int(nPS), &c.h) }); err != nil {
		return nil, err
	}
	runtime.SetFinalizer(c, func(c *Context) { c.Close() })
	return c, nil
}

// Close releases the context and its device scratch.
func (c *Context) Close() {
	if c.h != nil {
		C.accel_ctx_destroy(c.h)
		c.h = nil
	}
}

// GPUOf returns the GPU ordinal serving PS task `task`.
func (c *Context) GPUOf(task int) (int, error) {
	rc := C.accel_ctx_gpu_of(c.h, C.int(task))
	return int(rc), status(rc)
}

// DevPtr is a caller-owned device address (from the process's CUDA allocator).
type DevPtr uintptr

// gradArray copies the pointers into C memory. With no gradients it returns nil (the library
// then rejects the call with ACCEL_ERR_INVALID) instead of indexing an empty allocation.
func gradArray(grads []DevPtr) (*unsafe.Pointer, func()) {
	n := len(grads)
	if n == 0 {
		return nil, func() {}
	}
	mem := (*[1 << 20]unsafe.Pointer)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	for i, g := range grads {
		mem[i] = unsafe.Pointer(uintptr(g))
	}
	return &mem[0], func() { C.free(unsafe.Pointer(mem)) }
}

// ApplySGD: var -= lr * aggregate(grads). stream is a cudaStream_t (0 = default).
func (c *Context) ApplySGD(task int, v DevPtr, grads []DevPtr, dt DType, count uint64,
	hp *HParams, sumsq DevPtr, stream uintptr) error {
	g, free := gradArray(grads)
	defer free()
	ch := hp.c()
	return locked(func() C.int { return C.accel_apply_sgd(c.h, C.int(task), (*C.float)(unsafe.Pointer(uintptr(v))), g,
		C.int(len(grads)), C.int(dt), C.uint64_t(count), &ch,
		(*C.float)(unsafe.Pointer(uintptr(sumsq))), C.accel_stream_t(unsafe.Pointer(stream))) })
}

// ApplyMomentum: accum = accum*momentum + G; var -= accum*lr (or the Nesterov form).
func (c *Context) ApplyMomentum(task int, v, accum DevPtr, grads []DevPtr, dt DType, count uint64,
	hp *HParams, sumsq DevPtr, stream uintptr) error {
	g, free := gradArray(grads)
	defer free()
	ch := hp.c()
	return locked(func() C.int { return C.accel_apply_momentum(c.h, C.int(task),
		(*C.float)(unsafe.Pointer(uintptr(v))), (*C.float)(unsafe.Pointer(uintptr(accum))), g,
		C.int(len(grads)), C.int(dt), C.uint64_t(count), &ch,
		(*C.float)(unsafe.Pointer(uintptr(sumsq))), C.accel_stream_t(unsafe.Pointer(stream))) })
}

// ApplyAdam: TF-1.4 ApplyAdam on the aggregated gradient; call hp.AdvanceAdam() after
// the step's last apply.
func (c *Context) ApplyAdam(task int, v, m, vv DevPtr, grads []DevPtr, dt DType, count uint64,
	hp *HParams, sumsq DevPtr, stream uintptr) error {
	g, free := gradArray(grads)
	defer free()
	ch := hp.c()
	return locked(func() C.int { return C.accel_apply_adam(c.h, C.int(task), (*C.float)(unsafe.Pointer(uintptr(v))),
		(*C.float)(unsafe.Pointer(uintptr(m))), (*C.float)(unsafe.Pointer(uintptr(vv))), g,
		C.int(len(grads)), C.int(dt), C.uint64_t(count), &ch,
		(*C.float)(unsafe.Pointer(uintptr(sumsq))), C.accel_stream_t(unsafe.Pointer(stream))) })
}

// GradSumSq writes sum(aggregate(grads)^2) to the device scalar `out`.
func (c *Context) GradSumSq(task int, grads []DevPtr, dt DType, count uint64, agg Aggregate,
	out DevPtr, stream uintptr) error {
	g, free := gradArray(grads)
	defer free()
	return locked(func() C.int { return C.accel_grad_sumsq(c.h, C.int(task), g, C.int(len(grads)), C.int(dt),
		C.uint64_t(count), C.int(agg), (*C.float)(unsafe.Pointer(uintptr(out))),
		C.accel_stream_t(unsafe.Pointer(stream))) })
}

// StepHost runs one PS step with the workers' gradients in (page-locked) host memory and
// copies the updated variables back to hostVarOut (nil to skip).
func (c *Context) StepHost(task int, optimizer int, v, slot0, slot1 DevPtr,
	hostGrads []unsafe.Pointer, dt DType, count uint64, hp *HParams, hostVarOut unsafe.Pointer) error {
	ptrs := make([]DevPtr, len(hostGrads))
	for i, p := range hostGrads {
		ptrs[i] = DevPtr(uintptr(p))
	}
	g, free := gradArray(ptrs)
	defer free()
	ch := hp.c()
	return locked(func() C.int { return C.accel_ps_step_host(c.h, C.int(task), C.int(optimizer),
		(*C.float)(unsafe.Pointer(uintptr(v))), (*C.float)(unsafe.Pointer(uintptr(slot0))),
		(*C.float)(unsafe.Pointer(uintptr(slot1))), g, C.int(len(hostGrads)), C.int(dt),
		C.uint64_t(count), &ch, (*C.float)(hostVarOut)) })
}
