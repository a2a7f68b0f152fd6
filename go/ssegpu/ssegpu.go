// Package ssegpu is the cgo binding of libssegpu.so (include/ssegpu.h): the B200 batch path for
// the per-token work of src/sse-adapter and src/redis-nats-bridge.
//
// NOT COMPILED IN THIS REPOSITORY'S CI: the build image has no Go toolchain.  It is the binding a
// maintainer of the reference adds; the services must be built with CGO_ENABLED=1 (the
// reference Dockerfiles set CGO_ENABLED=0: src/sse-adapter/Dockerfile:12,
// src/redis-nats-bridge/Dockerfile:10).
package ssegpu

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -lssegpu
#include <stdlib.h>
#include "ssegpu.h"
*/
import "C"

import (
	"fmt"
	"unsafe"
)

// Op flags (SSEGPU_OP_*).
const (
	OpAdapter = 0
	OpBridge  = 1
	OpGo122   = 2
	OpOrdered = 8
	OpInspect, OpInspectInline, OpMatch = 16, 32, 64
	OpNatsPub = 128 // with OpBridge: records are complete "PUB <subject> <#bytes>\r\n<payload>\r\n" messages
	OpResp    = 256 // EncodeLaunch only: records are RESP2 PUBLISH commands
)

// Status codes (low 4 bits of the status word) and flags.
const (
	StOK, StErrSyntax, StErrType, StFiltered, StInspectDrop = 0, 1, 2, 3, 4
	StCodeMask                                              = 0xF
	StDone                                                  = 0x100
)

// Context is one GPU (ssegpu_init / ssegpu_destroy).
type Context struct{ c *C.ssegpu_ctx }

// Batch is one pinned ring slot.  Its slices alias C-owned pinned memory: no Go pointer is ever
// retained by C, and nothing here survives Release.
type Batch struct {
	ctx      *Context
	b        *C.ssegpu_batch
	In       []byte   // append payload bytes here ...
	InOff    []uint32 // ... and close each record with InOff[n+1] = len so far
	StartSeq []int64  // per record: the connection's startSequence (0 = no filter)
	n        int
	used     uint32
}

func errOf(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("ssegpu: %s (%s)", C.GoString(C.ssegpu_strerror(rc)), C.GoString(C.ssegpu_last_cuda_error()))
}

// New creates the context of one GPU.
func New(device int, maxRecords uint32, maxInBytes uint64, ringSlots uint32) (*Context, error) {
	cfg := C.ssegpu_config{device: C.int32_t(device), max_records: C.uint32_t(maxRecords),
		max_in_bytes: C.uint64_t(maxInBytes), ring_slots: C.uint32_t(ringSlots)}
	var c *C.ssegpu_ctx
	if err := errOf(C.ssegpu_init(&cfg, &c)); err != nil {
		return nil, err
	}
	return &Context{c}, nil
}

func (x *Context) Close() { C.ssegpu_destroy(x.c) }

// Shard returns the GPU that owns a conversation: fnv1a64(id) % nGPU.
func Shard(conversationID string, nGPU uint32) uint32 {
	p := unsafe.Pointer(unsafe.StringData(conversationID))
	return uint32(C.ssegpu_shard((*C.uint8_t)(p), C.size_t(len(conversationID)), C.uint32_t(nGPU)))
}

// StartSequence reproduces sse_handler.go:92-96 (Sscanf "%d" then ++).
func StartSequence(lastEventID string) int64 {
	cs := C.CString(lastEventID)
	defer C.free(unsafe.Pointer(cs))
	return int64(C.ssegpu_start_sequence(cs))
}

// Acquire takes a free slot (error "no free batch slot" when the ring is full).
func (x *Context) Acquire(maxRecords int, maxInBytes int) (*Batch, error) {
	var b *C.ssegpu_batch
	if err := errOf(C.ssegpu_batch_acquire(x.c, &b)); err != nil {
		return nil, err
	}
	return &Batch{ctx: x, b: b,
		In:       unsafe.Slice((*byte)(unsafe.Pointer(C.ssegpu_in_bytes(b))), maxInBytes),
		InOff:    unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_in_off(b))), maxRecords+1),
		StartSeq: unsafe.Slice((*int64)(unsafe.Pointer(C.ssegpu_start_seq(b))), maxRecords)}, nil
}

// Append is what the NATS callback (sse_handler.go:135-157) / the bridge loop
// (redis-nats-bridge/main.go:142-157) does instead of json.Unmarshal: copy the payload into the slot.
// It returns the record index, or -1 when the slot is full (the caller submits and acquires another).
func (b *Batch) Append(payload []byte, startSeq int64) int {
	if b.n+1 >= len(b.InOff) || int(b.used)+len(payload) > len(b.In) {
		return -1
	}
	copy(b.In[b.used:], payload)
	b.used += uint32(len(payload))
	b.StartSeq[b.n] = startSeq
	b.n++
	b.InOff[b.n] = b.used
	return b.n - 1
}

// Submit starts H2D -> kernel -> D2H and returns immediately.
func (b *Batch) Submit(op uint32, useFilter bool) error {
	f := C.int(0)
	if useFilter {
		f = 1
	}
	return errOf(C.ssegpu_batch_submit(b.ctx.c, b.b, C.uint32_t(b.n), C.uint32_t(op), f))
}

// Result is the host view of a finished batch; Frame(i) is what the handler goroutine writes with one
// w.Write + Flush instead of three Fprintf (sse_handler.go:212-215, :428-431).
type Result struct {
	Out     []byte
	Off     []uint64
	Len     []uint32
	Status  []uint32
	SubjLen []uint32
}

func (r *Result) Frame(i int) []byte { return r.Out[r.Off[i] : r.Off[i]+uint64(r.Len[i])] }

// Subject/JSON split of a bridge record: nc.Publish(subject, data) (redis-nats-bridge/main.go:164).
func (r *Result) Publish(i int) (subject string, data []byte) {
	f := r.Frame(i)
	return string(f[:r.SubjLen[i]]), f[r.SubjLen[i]:]
}

// Wait blocks until the results are in host memory.
func (b *Batch) Wait() (*Result, error) {
	if err := errOf(C.ssegpu_batch_wait(b.ctx.c, b.b, 0)); err != nil {
		return nil, err
	}
	n := b.n
	off := unsafe.Slice((*uint64)(unsafe.Pointer(C.ssegpu_out_off(b.b))), n+1)
	return &Result{
		Out:     unsafe.Slice((*byte)(unsafe.Pointer(C.ssegpu_out_bytes(b.b))), int(off[n])),
		Off:     off,
		Len:     unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_out_len(b.b))), n),
		Status:  unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_status(b.b))), n),
		SubjLen: unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_subj_len(b.b))), n)}, nil
}

// Release returns the slot to the ring; every slice obtained from it becomes invalid.
func (b *Batch) Release() error { return errOf(C.ssegpu_batch_release(b.ctx.c, b.b)) }

// EncodeLaunch is the producer side (publishToken, src/llm-stream-proxy/main.go:230-245) for a batch of messages whose
// columns already sit in this GPU's memory (device pointers as uintptr: the proxy owns their allocation and the pinned
// staging buffers it fills them through).  With OpBridge every record is preceded by its channel name, with OpResp it is
// the complete RESP2 PUBLISH command.  Asynchronous on `stream` (0 = the context's stream); EncodeResult after the
// stream has been synchronised.
func (x *Context) EncodeLaunch(dCid, dCidOff, dTok, dTokOff, dSeq, dDone, dTs uintptr, n uint32, op uint32,
	dOut uintptr, outCap uint64, dOutOff, dOutLen, dChanLen uintptr, stream unsafe.Pointer) error {
	return errOf(C.ssegpu_encode_launch(x.c, unsafe.Pointer(dCid), (*C.uint32_t)(unsafe.Pointer(dCidOff)),
		unsafe.Pointer(dTok), (*C.uint32_t)(unsafe.Pointer(dTokOff)), (*C.int64_t)(unsafe.Pointer(dSeq)),
		(*C.uint8_t)(unsafe.Pointer(dDone)), (*C.int64_t)(unsafe.Pointer(dTs)), C.uint32_t(n), C.uint32_t(op),
		unsafe.Pointer(dOut), C.uint64_t(outCap), (*C.uint64_t)(unsafe.Pointer(dOutOff)),
		(*C.uint32_t)(unsafe.Pointer(dOutLen)), (*C.uint32_t)(unsafe.Pointer(dChanLen)), stream))
}

// EncodeResult returns the bytes of the output arena the last EncodeLaunch used (error "output arena too small" on overflow).
func (x *Context) EncodeResult() (uint64, error) {
	var used C.uint64_t
	rc := C.ssegpu_encode_launch_result(x.c, &used)
	return uint64(used), errOf(rc)
}
