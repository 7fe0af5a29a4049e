package accel

/*
#include <stdlib.h>
#include "accel_ps.h"
*/
import "C"

import (
	"unsafe"

	api "github.com/caicloud/kubeflow-clientset/apis/kubeflow/v1alpha1"
)

// Placement derives, from the same fields pkg/tensorflow reads
// (Specs[getTemplateIndex(PS)].Replicas, pkg/tensorflow/distributed.go:140,198-206),
// how many PS tasks and workers a TFJob has and which GPU of the box serves each PS
// task: task_index i <-> GPU i. pkg/tensorflow's argv generation is not touched; a PS
// replica started with --task_index=i simply opens GPU gpuOf(i).
func Placement(tfJob *api.TFJob) (psCount, workerCount int, gpuOf func(taskIndex int) int) {
	for i := 0; i < len(tfJob.Spec.Specs) && i < 2; i++ { // only Specs[0..1] are ever inspected
		s := tfJob.Spec.Specs[i]
		if s.TFReplicaType == nil || s.Replicas == nil {
			continue
		}
		switch *s.TFReplicaType {
		case api.TFReplicaPS:
			psCount = int(*s.Replicas)
		case api.TFReplicaWorker:
			workerCount = int(*s.Replicas)
		}
	}
	return psCount, workerCount, func(taskIndex int) int { return taskIndex }
}

// Shard is one PS task's contiguous range of the padded flat arena.
type Shard struct{ Offset, Count uint64 }

// PlanFlat splits nElems over nPS equal shards padded to alignElems.
func PlanFlat(nElems uint64, nPS int, alignElems uint64) ([]Shard, uint64, error) {
	out := make([]C.accel_shard, nPS)
	var padded C.uint64_t
	if err := locked(func() C.int { return C.accel_plan_flat(C.uint64_t(nElems), C.int(nPS), C.uint64_t(alignElems),
		&out[0], &padded) }); err != nil {
		return nil, 0, err
	}
	shards := make([]Shard, nPS)
	for i, s := range out {
		shards[i] = Shard{uint64(s.offset), uint64(s.count)}
	}
	return shards, uint64(padded), nil
}

// PlanRoundRobin reproduces tf.train.replica_device_setter's default placement:
// variable k -> PS task (firstTask + k) mod nPS, packed per task.
func PlanRoundRobin(varElems []uint64, nPS, firstTask int, alignElems uint64) (task []int,
	offset, bucket []uint64, err error) {
	n := len(varElems)
	if n == 0 {
		return nil, nil, make([]uint64, nPS), nil
	}
	ve := make([]C.uint64_t, n)
	for i, v := range varElems {
		ve[i] = C.uint64_t(v)
	}
	ct := make([]C.int, n)
	co := make([]C.uint64_t, n)
	cb := make([]C.uint64_t, nPS)
	if err = locked(func() C.int { return C.accel_plan_round_robin(&ve[0], C.int(n), C.int(nPS), C.int(firstTask),
		C.uint64_t(alignElems), &ct[0], &co[0], &cb[0]) }); err != nil {
		return
	}
	task, offset, bucket = make([]int, n), make([]uint64, n), make([]uint64, nPS)
	for i := range ct {
		task[i], offset[i] = int(ct[i]), uint64(co[i])
	}
	for i := range cb {
		bucket[i] = uint64(cb[i])
	}
	return
}

// LocalAction restates LocalJob.Action (pkg/tensorflow/local.go:50-74) with NewLocalJob's pod
// selection (:34-48) and compose()/labels() (:78-90). suffix is the 5-character RuntimeID that
// compose() would draw. It returns the action type (accel_action order), the number of replicas
// to add, whether compose() ran, and the text block "runtime_id\t<id>\n" + "key=value\n" labels.
func LocalAction(jobName, runtimeID string, firstSpecType, activePods, succeeded int,
	suffix string) (action, number int, composed bool, text string, err error) {
	cj, cr, cs := C.CString(jobName), C.CString(runtimeID), C.CString(suffix)
	defer C.free(unsafe.Pointer(cj))
	defer C.free(unsafe.Pointer(cr))
	defer C.free(unsafe.Pointer(cs))
	var ev C.accel_event
	var comp C.int
	buf := (*C.char)(C.malloc(2048))
	defer C.free(unsafe.Pointer(buf))
	err = locked(func() C.int { return C.accel_tfjob_local_action(cj, cr, C.int(firstSpecType), C.int(activePods),
		C.int(succeeded), cs, &ev, &comp, buf, 2048) })
	if err != nil {
		return
	}
	return int(ev.action), int(ev.number), comp != 0, C.GoString(buf), nil
}
