// Package routing binds libpkroute.so (include/pk_route.h) and provides the batching
// service.Service decorator that puts the GPU routing stage at the reference's plugin seam
// (internal/service/service.go:13-15).
//
// UNVERIFIED SOURCE: there is no Go toolchain in the environment this was written in, so
// these files were never compiled.  The compiled and tested form of the same logic is the C++
// mirror under polykey_service_b200/host/ (batching_service.cpp); keep the two in step.
package routing

/*
#cgo CFLAGS:  -I${SRCDIR}/../../third_party/pkroute/include
#cgo LDFLAGS: -L${SRCDIR}/../../third_party/pkroute -lpkroute -Wl,-rpath,$ORIGIN
#include <stdlib.h>
#include "pk_route.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"unsafe"
)

// Decision mirrors pk_decision (24 bytes, same field order).
type Decision struct {
	ByteLen   uint32
	EstTokens uint32
	RouteHash uint64
	ToolClass uint8
	Provider  uint8 // 0xFF: no provider eligible
	Flags     uint16
	Status    uint32 // always 200, as mock.go:25-28
}

const (
	FlagNoProvider   = 0x0002
	FlagEmptyPayload = 0x0004
	FlagUnknownTool  = 0x0008 // the default branch of mock.go:60-64; a success, not an error
	Align            = 16     // payload start alignment in the arena
)

// Router owns one pk_handle: one GPU.  Use from one goroutine at a time (the flusher).
type Router struct {
	h     *C.pk_handle
	slots uint32
}

// NewRouter is the start-up hook for cmd/polykey/main.go:85-86.
func NewRouter(device int, maxRequests uint32, maxPayloadBytes uint64, slots uint32) (*Router, error) {
	var tbl C.pk_table
	C.pk_default_table(&tbl) // example_tool / struct_tool / file_tool / default: mock.go:33,37,52,60
	cfg := C.pk_config{
		max_requests:      C.uint32_t(maxRequests),
		n_slots:           C.uint32_t(slots),
		max_payload_bytes: C.uint64_t(maxPayloadBytes),
		max_name_bytes:    C.uint64_t(256 * uint64(maxRequests)), // >= the front-end's MaxNameBytes per request
	}
	var h *C.pk_handle
	if rc := C.pk_init(C.int(device), &tbl, &cfg, &h); rc != 0 {
		return nil, fmt.Errorf("pk_init: %s", C.GoString(C.pk_strerror(rc)))
	}
	return &Router{h: h, slots: slots}, nil
}

// Close is the teardown hook for cmd/polykey/main.go:118-119.
func (r *Router) Close() {
	if r.h != nil {
		C.pk_destroy(r.h)
		r.h = nil
	}
}

func (r *Router) Slots() uint32 { return r.slots }

// HostAlloc returns pinned memory owned by the C side.  gRPC buffers live on the Go heap and
// cannot be DMA sources, and cgo forbids handing C any Go memory that holds Go pointers, so
// the flusher packs every batch into arenas from here (one copy per request).
func (r *Router) HostAlloc(n int) []byte {
	p := C.pk_host_alloc(r.h, C.size_t(n))
	if p == nil {
		return nil
	}
	return unsafe.Slice((*byte)(p), n)
}

// HostAllocDecisions: a pinned record array (the device writes the batch's decisions into it).
func (r *Router) HostAllocDecisions(n int) []Decision {
	p := C.pk_host_alloc(r.h, C.size_t(n)*C.size_t(unsafe.Sizeof(Decision{})))
	if p == nil {
		return nil
	}
	return unsafe.Slice((*Decision)(p), n)
}

// RouteError carries the pk_status of a failed call.
type RouteError struct {
	Status int // PK_ERR_*
	Cuda   int // cudaError_t when Status is PK_ERR_CUDA
	Where  string
}

func (e *RouteError) Error() string {
	return fmt.Sprintf("routing failed: %s: %s (cuda %d)", e.Where, C.GoString(C.pk_strerror(C.int(e.Status))), e.Cuda)
}

// IsDeviceFailure: a failure of the device (health goes NOT_SERVING), as opposed to a batch the
// device refused (argument / capacity: only its requests fail).
func IsDeviceFailure(err error) bool {
	var re *RouteError
	return errors.As(err, &re) && (re.Status == int(C.PK_ERR_CUDA) || re.Status == int(C.PK_ERR_NO_DEVICE))
}

// Submit enqueues H2D copies, the fused kernel and the D2H copy of one batch on `slot` and
// returns at once; Wait blocks until `out` holds the records.  payload and out must come from
// HostAlloc / HostAllocDecisions and stay untouched until Wait returns; the offset, length and name
// arrays are consumed before Submit returns and may be ordinary Go slices (no Go pointers in them).
func (r *Router) Submit(slot uint32, payload []byte, payOff []uint64, payLen []uint32,
	names []byte, nameOff []uint32, out []Decision) error {
	n := len(payLen)
	if n == 0 {
		return nil
	}
	rc := C.pk_submit(r.h, C.uint32_t(slot),
		(*C.uint8_t)(unsafe.Pointer(unsafe.SliceData(payload))),
		(*C.uint64_t)(unsafe.Pointer(unsafe.SliceData(payOff))),
		(*C.uint32_t)(unsafe.Pointer(unsafe.SliceData(payLen))),
		(*C.uint8_t)(unsafe.Pointer(unsafe.SliceData(names))),
		(*C.uint32_t)(unsafe.Pointer(unsafe.SliceData(nameOff))),
		C.uint32_t(n),
		(*C.pk_decision)(unsafe.Pointer(unsafe.SliceData(out))))
	if rc != 0 {
		return &RouteError{int(rc), int(C.pk_last_cuda_error(r.h)), "pk_submit"}
	}
	return nil
}

func (r *Router) Wait(slot uint32) error {
	if rc := C.pk_wait(r.h, C.uint32_t(slot)); rc != 0 {
		return &RouteError{int(rc), int(C.pk_last_cuda_error(r.h)), "pk_wait"}
	}
	return nil
}
