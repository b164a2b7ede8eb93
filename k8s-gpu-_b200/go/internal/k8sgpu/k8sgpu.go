// UNCOMPILED on this image (no Go toolchain).  cgo shim over include/k8sgpu.h.
// Every entry point of libk8sgpu.so calls cudaSetDevice itself, so no LockOSThread is needed;
// a Ctx is not re-entrant: callers serialise calls on it.
package k8sgpu

/*
#cgo CFLAGS: -I${SRCDIR}/../../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../.. -lk8sgpu -Wl,-rpath,${SRCDIR}/../../..
#include <stdlib.h>
#include "k8sgpu.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

type Ctx struct{ h *C.k8sgpu_ctx }

// Decisions of one batch pass: steps 3-5 of Reconcile for every pool (README.md:187-230).
type Decisions struct {
	PoolOf, Current, Delta, Ready, CreateCount []int32
	VictimOff, MemberOff, CreateOff            []int32
	VictimVM, MemberVM, CreateOrdinal          []uint32
}

// New opens one GPU, or several as one placement group (a controller process that owns the
// node's eight B200s passes all eight: k8sgpu_place then shards the pods of every round across
// them inside the library, over NCCL, and returns the same bytes as one GPU would).
func New(devices ...int) (*Ctx, error) {
	if len(devices) == 0 {
		devices = []int{0}
	}
	devs := make([]C.int, len(devices))
	for i, d := range devices {
		devs[i] = C.int(d)
	}
	var h *C.k8sgpu_ctx
	errbuf := (*C.char)(C.calloc(512, 1)) // per-call buffer: safe when goroutines create contexts concurrently
	defer C.free(unsafe.Pointer(errbuf))
	if rc := C.k8sgpu_create_ex(&devs[0], C.int(len(devs)), &h, errbuf, 512); rc != C.K8SGPU_OK {
		// no GPU is an error: the library has no CPU path
		return nil, fmt.Errorf("k8sgpu_create: %d: %s", int(rc), C.GoString(errbuf))
	}
	return &Ctx{h}, nil
}

func (c *Ctx) Close() { C.k8sgpu_destroy(c.h); c.h = nil }

func (c *Ctx) ReconcileBatch(desired []int32, poolOwner, vmOwner, vmFlags, vmOrdinal []uint32,
	nOwnerIDs uint32) (*Decisions, error) {
	P, V := len(desired), len(vmOwner)
	capCreate := 1
	for _, d := range desired {
		if d > 0 {
			capCreate += int(d)
		}
	}
	d := &Decisions{
		PoolOf: make([]int32, V+1), Current: make([]int32, P+1), Delta: make([]int32, P+1),
		Ready: make([]int32, P+1), CreateCount: make([]int32, P+1), VictimOff: make([]int32, P+2),
		MemberOff: make([]int32, P+2), CreateOff: make([]int32, P+2), VictimVM: make([]uint32, V+1),
		MemberVM: make([]uint32, V+1), CreateOrdinal: make([]uint32, capCreate),
	}
	// The table structs hold pointers, so they live in C memory (cgo pointer-passing rules);
	// the Go column slices are pinned for the duration of the call.
	pools := (*C.k8sgpu_pool_table)(C.calloc(1, C.sizeof_k8sgpu_pool_table))
	vms := (*C.k8sgpu_vm_table)(C.calloc(1, C.sizeof_k8sgpu_vm_table))
	out := (*C.k8sgpu_reconcile_out)(C.calloc(1, C.sizeof_k8sgpu_reconcile_out))
	defer C.free(unsafe.Pointer(pools))
	defer C.free(unsafe.Pointer(vms))
	defer C.free(unsafe.Pointer(out))
	var pin runtime.Pinner
	defer pin.Unpin()
	u32 := func(s []uint32) *C.uint32_t {
		if len(s) == 0 {
			return nil
		}
		pin.Pin(&s[0])
		return (*C.uint32_t)(unsafe.Pointer(&s[0]))
	}
	i32 := func(s []int32) *C.int32_t {
		if len(s) == 0 {
			return nil
		}
		pin.Pin(&s[0])
		return (*C.int32_t)(unsafe.Pointer(&s[0]))
	}
	pools.p, pools.desired, pools.owner_id = C.uint32_t(P), i32(desired), u32(poolOwner)
	vms.v, vms.owner_id, vms.flags, vms.ordinal = C.uint32_t(V), u32(vmOwner), u32(vmFlags), u32(vmOrdinal)
	out.pool_of, out.current, out.delta, out.ready = i32(d.PoolOf), i32(d.Current), i32(d.Delta), i32(d.Ready)
	out.create_count, out.victim_off, out.victim_vm = i32(d.CreateCount), i32(d.VictimOff), u32(d.VictimVM)
	out.member_off, out.member_vm = i32(d.MemberOff), u32(d.MemberVM)
	out.create_off, out.create_ordinal = i32(d.CreateOff), u32(d.CreateOrdinal)
	out.create_capacity = C.uint64_t(len(d.CreateOrdinal))
	if rc := C.k8sgpu_reconcile_batch(c.h, pools, vms, C.uint32_t(nOwnerIDs), out); rc != C.K8SGPU_OK {
		return nil, fmt.Errorf("k8sgpu_reconcile_batch: %d: %s", int(rc), C.GoString(C.k8sgpu_last_error(c.h)))
	}
	return d, nil
}

// Place binds k8sgpu_place: eight uint32 columns per table, assign and free columns out.
func (c *Ctx) Place(nodes, pods [8][]uint32, flags, maxRounds uint32) (assign []int32, free [3][]uint32, err error) {
	N, M := len(nodes[0]), len(pods[0])
	nt := (*C.k8sgpu_node_table)(C.calloc(1, C.sizeof_k8sgpu_node_table))
	pt := (*C.k8sgpu_pod_table)(C.calloc(1, C.sizeof_k8sgpu_pod_table))
	po := (*C.k8sgpu_place_out)(C.calloc(1, C.sizeof_k8sgpu_place_out))
	defer C.free(unsafe.Pointer(nt))
	defer C.free(unsafe.Pointer(pt))
	defer C.free(unsafe.Pointer(po))
	var pin runtime.Pinner
	defer pin.Unpin()
	col := func(s []uint32) *C.uint32_t {
		if len(s) == 0 {
			return nil
		}
		pin.Pin(&s[0])
		return (*C.uint32_t)(unsafe.Pointer(&s[0]))
	}
	nt.n, pt.m = C.uint32_t(N), C.uint32_t(M)
	nt.free_gpu, nt.free_mcpu, nt.free_mib, nt.cap_gpu = col(nodes[0]), col(nodes[1]), col(nodes[2]), col(nodes[3])
	nt.sku_id, nt.zone_id, nt.taint_mask, nt.flags = col(nodes[4]), col(nodes[5]), col(nodes[6]), col(nodes[7])
	pt.req_gpu, pt.req_mcpu, pt.req_mib, pt.sku_mask = col(pods[0]), col(pods[1]), col(pods[2]), col(pods[3])
	pt.zone_mask, pt.tol_mask, pt.pref_zone, pt.priority = col(pods[4]), col(pods[5]), col(pods[6]), col(pods[7])
	assign = make([]int32, M+1)
	for i := range free {
		free[i] = make([]uint32, N+1)
	}
	pin.Pin(&assign[0])
	po.assign = (*C.int32_t)(unsafe.Pointer(&assign[0]))
	po.free_gpu, po.free_mcpu, po.free_mib = col(free[0]), col(free[1]), col(free[2])
	pol := C.k8sgpu_policy{flags: C.uint32_t(flags), max_rounds: C.uint32_t(maxRounds)}
	if rc := C.k8sgpu_place(c.h, nt, pt, &pol, po); rc != C.K8SGPU_OK {
		return nil, free, fmt.Errorf("k8sgpu_place: %d: %s", int(rc), C.GoString(C.k8sgpu_last_error(c.h)))
	}
	return assign[:M], free, nil
}
