// +build gpuacct

// UNBUILT SOURCE: neither the build image nor the GPU box has a Go toolchain.
// This is the binding a maintainer of the reference adds next to
// cmd/arena/commands/top_node.go (package commands) to route `arena top node`
// through libarena_gpuacct.so; see INTEGRATION.md for the seam-by-seam map.
package commands

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../lib -larena_gpuacct -Wl,-rpath,$ORIGIN/../lib
#include <stdlib.h>
#include "arena_top.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

func gpuacctErr(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("gpuacct: %s (%d)", C.GoString(C.arena_last_error()), int(rc))
}

// topNodeGPU replaces acquireAllActivePods + getAllNodeInfos + caculateNodeGPU
// + displayTopNode (top_node.go:38-49): podList / nodeList are the raw JSON
// bodies of the two List calls; the returned text is what the reference prints.
func topNodeGPU(podList, nodeList []byte, details bool) (string, error) {
	runtime.LockOSThread() // a CUDA context is used from one OS thread at a time
	defer runtime.UnlockOSThread()
	if err := gpuacctErr(C.arena_acct_init(0)); err != nil {
		return "", err
	}
	var host *C.arena_host_snapshot
	// Go memory is only read during the call (cgo pointer rule: nothing is retained).
	if err := gpuacctErr(C.arena_load_json(
		(*C.char)(unsafe.Pointer(&podList[0])), C.size_t(len(podList)),
		(*C.char)(unsafe.Pointer(&nodeList[0])), C.size_t(len(nodeList)),
		nil, 0, nil, 0, nil, 0, 0, &host)); err != nil {
		return "", err
	}
	defer C.arena_host_snapshot_free(host)
	var view C.arena_snapshot_view
	C.arena_host_snapshot_view(host, &view)
	n, p := int(view.n_nodes), int(view.n_pods)
	total := make([]int64, n+1)
	alloc := make([]int64, n+1)
	podGPU := make([]int64, p+1)
	var cluster [2]C.int64_t
	var pg *C.int64_t
	d := C.int(0)
	if details {
		pg = (*C.int64_t)(unsafe.Pointer(&podGPU[0]))
		d = 1
	}
	if err := gpuacctErr(C.arena_top_node_host(&view,
		(*C.int64_t)(unsafe.Pointer(&total[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		&cluster[0], pg)); err != nil {
		return "", err
	}
	var text *C.char
	var tlen C.size_t
	if err := gpuacctErr(C.arena_format_top_node(host,
		(*C.int64_t)(unsafe.Pointer(&total[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		pg, d, &text, &tlen)); err != nil {
		return "", err
	}
	defer C.arena_free_text(text)
	return C.GoStringN(text, C.int(tlen)), nil
}

// topJobGPU replaces the body of NewTopJobCommand's Run closure after the
// List calls (top_job.go:50-77): trainers, RequestedGPU/AllocatedGPU, the
// ordering and displayTrainingJobList(jobs, true).
func topJobGPU(podList, nodeList, jobList, tfjobList, helmList []byte, nowUnix int64) (string, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if err := gpuacctErr(C.arena_acct_init(0)); err != nil {
		return "", err
	}
	ptr := func(b []byte) *C.char {
		if len(b) == 0 {
			return nil
		}
		return (*C.char)(unsafe.Pointer(&b[0]))
	}
	var host *C.arena_host_snapshot
	if err := gpuacctErr(C.arena_load_json(ptr(podList), C.size_t(len(podList)), ptr(nodeList), C.size_t(len(nodeList)),
		ptr(jobList), C.size_t(len(jobList)), ptr(tfjobList), C.size_t(len(tfjobList)),
		ptr(helmList), C.size_t(len(helmList)), 1 /* group rows by job */, &host)); err != nil {
		return "", err
	}
	defer C.arena_host_snapshot_free(host)
	var view C.arena_snapshot_view
	C.arena_host_snapshot_view(host, &view)
	var dev *C.arena_snapshot
	if err := gpuacctErr(C.arena_snapshot_upload(&view, &dev)); err != nil {
		return "", err
	}
	defer C.arena_snapshot_free(dev)
	j := int(view.n_jobs)
	req := make([]int64, j+1)
	alloc := make([]int64, j+1)
	order := make([]uint32, j+1)
	latest := make([]uint32, j+1)
	var totals [2]C.int64_t
	if err := gpuacctErr(C.arena_top_job(dev,
		(*C.int64_t)(unsafe.Pointer(&req[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		(*C.uint32_t)(unsafe.Pointer(&order[0])), (*C.uint32_t)(unsafe.Pointer(&latest[0])), &totals[0])); err != nil {
		return "", err
	}
	var text *C.char
	var tlen C.size_t
	if err := gpuacctErr(C.arena_format_top_job(host,
		(*C.int64_t)(unsafe.Pointer(&req[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		(*C.uint32_t)(unsafe.Pointer(&order[0])), (*C.uint32_t)(unsafe.Pointer(&latest[0])),
		C.int64_t(nowUnix), &text, &tlen)); err != nil {
		return "", err
	}
	defer C.arena_free_text(text)
	return C.GoStringN(text, C.int(tlen)), nil
}

// topNodeSummaryGPU is topNodeGPU for the summary view with the PodList parsed on the GPU
// (arena_load_json_device): the library stages the pageable Go bytes through page-locked
// slots itself, parses the document on the device and falls back to its host loader for
// anything unusual, with identical output and error texts either way.
func topNodeSummaryGPU(podList, nodeList []byte) (string, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if err := gpuacctErr(C.arena_acct_init(0)); err != nil {
		return "", err
	}
	var nodes *C.arena_host_snapshot
	var dev *C.arena_snapshot
	if err := gpuacctErr(C.arena_load_json_device(
		(*C.char)(unsafe.Pointer(&podList[0])), C.size_t(len(podList)),
		(*C.char)(unsafe.Pointer(&nodeList[0])), C.size_t(len(nodeList)),
		&nodes, &dev, nil)); err != nil {
		return "", err
	}
	defer C.arena_host_snapshot_free(nodes)
	defer C.arena_snapshot_free(dev)
	var counts [4]C.uint64_t
	C.arena_snapshot_counts(dev, &counts[0])
	n := int(counts[2])
	total := make([]int64, n+1)
	alloc := make([]int64, n+1)
	var cluster [2]C.int64_t
	if err := gpuacctErr(C.arena_top_node(dev,
		(*C.int64_t)(unsafe.Pointer(&total[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		&cluster[0], nil)); err != nil {
		return "", err
	}
	var text *C.char
	var tlen C.size_t
	if err := gpuacctErr(C.arena_format_top_node(nodes,
		(*C.int64_t)(unsafe.Pointer(&total[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		nil, 0, &text, &tlen)); err != nil {
		return "", err
	}
	defer C.arena_free_text(text)
	return C.GoStringN(text, C.int(tlen)), nil
}

// listJobsGPU replaces the tail of NewListCommand's Run closure (list.go:50-75): the jobs `top job`
// finds, ordered by makeTrainingJobOrderdByAge and printed by displayTrainingJobList(jobs, false).
// STATUS and NODE of mpijob / standalonejob rows need the latest-pod rule, which the `top job`
// kernels evaluate (job_latest).
func listJobsGPU(podList, nodeList, jobList, tfjobList, helmList []byte, nowUnix int64) (string, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if err := gpuacctErr(C.arena_acct_init(0)); err != nil {
		return "", err
	}
	ptr := func(b []byte) *C.char {
		if len(b) == 0 {
			return nil
		}
		return (*C.char)(unsafe.Pointer(&b[0]))
	}
	var host *C.arena_host_snapshot
	// the PodList is decoded on the GPU (names, labels, owners, timestamps come back as places in the
	// document); the snapshot is the one arena_load_json builds, and so are the errors
	if err := gpuacctErr(C.arena_load_json_gpu(ptr(podList), C.size_t(len(podList)), ptr(nodeList), C.size_t(len(nodeList)),
		ptr(jobList), C.size_t(len(jobList)), ptr(tfjobList), C.size_t(len(tfjobList)),
		ptr(helmList), C.size_t(len(helmList)), 1, &host, nil)); err != nil {
		return "", err
	}
	defer C.arena_host_snapshot_free(host)
	var view C.arena_snapshot_view
	C.arena_host_snapshot_view(host, &view)
	var dev *C.arena_snapshot
	if err := gpuacctErr(C.arena_snapshot_upload(&view, &dev)); err != nil {
		return "", err
	}
	defer C.arena_snapshot_free(dev)
	j := int(view.n_jobs)
	req := make([]int64, j+1)
	alloc := make([]int64, j+1)
	order := make([]uint32, j+1)
	latest := make([]uint32, j+1)
	var totals [2]C.int64_t
	if err := gpuacctErr(C.arena_top_job(dev,
		(*C.int64_t)(unsafe.Pointer(&req[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])),
		(*C.uint32_t)(unsafe.Pointer(&order[0])), (*C.uint32_t)(unsafe.Pointer(&latest[0])), &totals[0])); err != nil {
		return "", err
	}
	if err := gpuacctErr(C.arena_job_order_by_age(host, (*C.uint32_t)(unsafe.Pointer(&order[0])))); err != nil {
		return "", err
	}
	var text *C.char
	var tlen C.size_t
	if err := gpuacctErr(C.arena_format_job_list(host, nil, nil,
		(*C.uint32_t)(unsafe.Pointer(&order[0])), (*C.uint32_t)(unsafe.Pointer(&latest[0])),
		C.int64_t(nowUnix), 0, &text, &tlen)); err != nil {
		return "", err
	}
	defer C.arena_free_text(text)
	return C.GoStringN(text, C.int(tlen)), nil
}

// topNodeMultiGPU is topNodeGPU on every GPU of the box from this ONE process (the reference CLI is
// one process, main.go:12-28): arena_acct_init_devices binds devices 0..nGPU-1, the snapshot is cut
// into node ranges of about equal pod count (one per device), every device computes its range at
// once and the library assembles one table; the cluster totals are traded inside the kernels over
// peer memory.  Output is identical for any nGPU.
func topNodeMultiGPU(podList, nodeList []byte, details bool, nGPU int) (string, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	ids := make([]C.int, nGPU)
	for i := range ids {
		ids[i] = C.int(i)
	}
	if err := gpuacctErr(C.arena_acct_init_devices(C.int(nGPU), &ids[0])); err != nil {
		return "", err
	}
	var host *C.arena_host_snapshot
	if err := gpuacctErr(C.arena_load_json(
		(*C.char)(unsafe.Pointer(&podList[0])), C.size_t(len(podList)),
		(*C.char)(unsafe.Pointer(&nodeList[0])), C.size_t(len(nodeList)),
		nil, 0, nil, 0, nil, 0, 0, &host)); err != nil {
		return "", err
	}
	defer C.arena_host_snapshot_free(host)
	var view C.arena_snapshot_view
	C.arena_host_snapshot_view(host, &view)
	var ms *C.arena_multi_snapshot
	if err := gpuacctErr(C.arena_multi_snapshot_upload(&view, 0 /* node ranges */, &ms)); err != nil {
		return "", err
	}
	defer C.arena_multi_snapshot_free(ms)
	n, p := int(view.n_nodes), int(view.n_pods)
	total := make([]int64, n+1)
	alloc := make([]int64, n+1)
	podGPU := make([]int64, p+1)
	var cluster [2]C.int64_t
	var pg *C.int64_t
	d := C.int(0)
	if details {
		pg = (*C.int64_t)(unsafe.Pointer(&podGPU[0]))
		d = 1
	}
	if err := gpuacctErr(C.arena_top_node_multi(ms,
		(*C.int64_t)(unsafe.Pointer(&total[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])), &cluster[0], pg)); err != nil {
		return "", err
	}
	var text *C.char
	var tlen C.size_t
	if err := gpuacctErr(C.arena_format_top_node(host,
		(*C.int64_t)(unsafe.Pointer(&total[0])), (*C.int64_t)(unsafe.Pointer(&alloc[0])), pg, d, &text, &tlen)); err != nil {
		return "", err
	}
	defer C.arena_free_text(text)
	return C.GoStringN(text, C.int(tlen)), nil
}
// (`top job` on several GPUs is the same shape: arena_load_json with group_by_job = 1,
// arena_multi_snapshot_upload(&view, 1, &ms), arena_top_job_multi, arena_format_top_job.)
