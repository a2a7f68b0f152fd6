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
	// OpHeaderOnly: a record that is exactly json.Marshal's output comes back as its header only (status bit
	// StHeaderOnly); FrameParts gives the three pieces the handler writes (header, the payload it already holds, "\n\n").
	OpHeaderOnly = 512
)

// Status codes (low 4 bits of the status word) and flags.
const (
	StOK, StErrSyntax, StErrType, StFiltered, StInspectDrop, StNoInterest = 0, 1, 2, 3, 4, 5
	StCodeMask                                                            = 0xF
	StDone                                                                = 0x100
	StRedacted                                                            = 0x200
	StHeaderOnly                                                          = 0x400
)

// Verdict actions (low byte of the verdict word); pattern index = verdict >> 8 & 0xFF.
const (
	ActAllow, ActRedact, ActDrop, ActAlert = 0, 1, 2, 3
	VerdictBeginUnknown                    = 0x10000
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
	ConvSlot []uint32 // per record: the conversation's inspection-state slot (streaming inspection only)
	n        int
	used     uint32
}

func errOf(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("ssegpu: %s (%s)", C.GoString(C.ssegpu_strerror(rc)), C.GoString(C.ssegpu_last_cuda_error()))
}

// New creates the context of one GPU.  convSlots > 0 reserves that many per-conversation inspection states
// (streaming inspection); 0 = none.
func New(device int, maxRecords uint32, maxInBytes uint64, ringSlots uint32, convSlots uint32) (*Context, error) {
	cfg := C.ssegpu_config{device: C.int32_t(device), max_records: C.uint32_t(maxRecords),
		max_in_bytes: C.uint64_t(maxInBytes), ring_slots: C.uint32_t(ringSlots), conv_slots: C.uint32_t(convSlots)}
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
		StartSeq: unsafe.Slice((*int64)(unsafe.Pointer(C.ssegpu_start_seq(b))), maxRecords),
		ConvSlot: unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_conv_slot(b))), maxRecords)}, nil
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
	Verdict []uint32 // OpInspect / OpInspectInline: action | pattern << 8
	Match   []int32  // 2n: begin / end offsets into the decoded token text
	SubSpan []uint32 // OpMatch, 3n: offset, count into Subs; wildcard mask over WildIDs
	Subs    []uint32 // subscriber-id arena of the table generation this batch saw (valid until Release)
	WildIDs []uint32
	Version uint64 // number of InterestApply calls this batch saw
	in      []byte
	inOff   []uint32
}

func (r *Result) Frame(i int) []byte { return r.Out[r.Off[i] : r.Off[i]+uint64(r.Len[i])] }

var frameTrailer = []byte("\n\n")

// FrameParts is Frame for OpHeaderOnly batches: the pieces to write in order.  A canonical record (StHeaderOnly) is
// header + the payload still sitting in the batch's input slot + "\n\n" (bridge mode: no trailer); every other record is
// one piece, the whole frame.  net.Buffers{...}.WriteTo(w) sends them with one writev.
func (r *Result) FrameParts(i int, bridge bool) [][]byte {
	if r.Status[i]&StHeaderOnly == 0 {
		return [][]byte{r.Frame(i)}
	}
	payload := r.in[r.inOff[i]:r.inOff[i+1]]
	if bridge {
		return [][]byte{r.Frame(i), payload}
	}
	return [][]byte{r.Frame(i), payload, frameTrailer}
}

// Subscribers of record i (OpMatch): the literal subscriptions' ids, then the matching wildcard subscriptions' ids.
func (r *Result) Subscribers(i int) []uint32 {
	off, cnt, mask := r.SubSpan[3*i], r.SubSpan[3*i+1], r.SubSpan[3*i+2]
	ids := append([]uint32(nil), r.Subs[off:off+cnt]...)
	for w := 0; mask != 0; w, mask = w+1, mask>>1 {
		if mask&1 != 0 {
			ids = append(ids, r.WildIDs[w])
		}
	}
	return ids
}

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
	r := &Result{
		Out:     unsafe.Slice((*byte)(unsafe.Pointer(C.ssegpu_out_bytes(b.b))), int(off[n])),
		Off:     off,
		Len:     unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_out_len(b.b))), n),
		Status:  unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_status(b.b))), n),
		SubjLen: unsafe.Slice((*uint32)(unsafe.Pointer(C.ssegpu_subj_len(b.b))), n),
		in:      b.In, inOff: b.InOff}
	if v := C.ssegpu_verdict(b.b); v != nil {
		r.Verdict = unsafe.Slice((*uint32)(unsafe.Pointer(v)), n)
		r.Match = unsafe.Slice((*int32)(unsafe.Pointer(C.ssegpu_match(b.b))), 2*n)
	}
	if sp := C.ssegpu_sub_span(b.b); sp != nil {
		var cnt C.uint32_t
		r.SubSpan = unsafe.Slice((*uint32)(unsafe.Pointer(sp)), 3*n)
		if p := C.ssegpu_batch_subs(b.b, &cnt); p != nil {
			r.Subs = unsafe.Slice((*uint32)(unsafe.Pointer(p)), int(cnt))
		}
		if p := C.ssegpu_batch_wild_ids(b.b, &cnt); p != nil {
			r.WildIDs = unsafe.Slice((*uint32)(unsafe.Pointer(p)), int(cnt))
		}
		r.Version = uint64(C.ssegpu_batch_table_version(b.b))
	}
	return r, nil
}

// Pattern is one inspection pattern: Kind 0 = literal (ASCII case folded), 1 = regex subset; Action = Act*.
type Pattern struct {
	Kind, Action uint32
	Text         string
}

// PatternsLoad installs the pattern set of the inspection hook (sse_handler.go:193-208); streaming = state carried per
// conversation slot across tokens.  Safe while batches are in flight (they finish on the old set).
func (x *Context) PatternsLoad(ps []Pattern, streaming bool) error {
	cp := make([]C.ssegpu_pattern, len(ps))
	for i, p := range ps {
		cs := C.CString(p.Text)
		defer C.free(unsafe.Pointer(cs))
		cp[i] = C.ssegpu_pattern{kind: C.uint32_t(p.Kind), action: C.uint32_t(p.Action), text: cs}
	}
	mode := C.uint32_t(0)
	if streaming {
		mode = 1
	}
	var first *C.ssegpu_pattern
	if len(cp) > 0 {
		first = &cp[0]
	}
	return errOf(C.ssegpu_patterns_load(x.c, first, C.uint32_t(len(cp)), mode))
}

// Subscribe / Unsubscribe are what nc.Subscribe(subject, cb) and sub.Unsubscribe() (sse_handler.go:133-135, :166-169)
// become for this path: one slot of the device table changes in place; the call does not wait for the GPU and may run
// beside Submit on other goroutines.
func (x *Context) Subscribe(subject string, subID uint32) error   { return x.interest(0, subject, subID) }
func (x *Context) Unsubscribe(subject string, subID uint32) error { return x.interest(1, subject, subID) }

func (x *Context) interest(op uint32, subject string, subID uint32) error {
	cs := C.CString(subject)
	defer C.free(unsafe.Pointer(cs))
	d := C.ssegpu_sub_delta{op: C.uint32_t(op), sub_id: C.uint32_t(subID), subject: cs, subject_len: C.uint32_t(len(subject))}
	return errOf(C.ssegpu_interest_apply(x.c, &d, 1))
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

// SplitLaunch is the receive side of the wire (nats.go's parser in front of the Subscribe callback, go-redis's reader in
// front of the bridge loop; sse_handler.go:135, redis-nats-bridge/main.go:115,127,142): many connections' unread socket
// bytes (device memory) -> payload arena + offsets, the input of FrameLaunch.  proto 0 = NATS, 1 = RESP.
func (x *Context) SplitLaunch(proto uint32, dBuf, dBufOff uintptr, nConn uint32, dArena uintptr, arenaCap uint64, dInOff uintptr,
	maxMsgs uint32, dMsgConn, dMsgSid, dMsgSubj, dConnConsumed, dConnStatus, dConnPings uintptr, stream unsafe.Pointer) error {
	return errOf(C.ssegpu_split_launch(x.c, C.uint32_t(proto), unsafe.Pointer(dBuf), (*C.uint32_t)(unsafe.Pointer(dBufOff)),
		C.uint32_t(nConn), unsafe.Pointer(dArena), C.uint64_t(arenaCap), (*C.uint32_t)(unsafe.Pointer(dInOff)), C.uint32_t(maxMsgs),
		(*C.uint32_t)(unsafe.Pointer(dMsgConn)), (*C.uint64_t)(unsafe.Pointer(dMsgSid)), (*C.uint32_t)(unsafe.Pointer(dMsgSubj)),
		(*C.uint32_t)(unsafe.Pointer(dConnConsumed)), (*C.uint32_t)(unsafe.Pointer(dConnStatus)),
		(*C.uint32_t)(unsafe.Pointer(dConnPings)), stream))
}

// SplitResult: deliveries found and arena bytes used by the last SplitLaunch (after the stream was synchronised).
func (x *Context) SplitResult() (uint32, uint64, error) {
	var n C.uint32_t
	var bytes C.uint64_t
	rc := C.ssegpu_split_result(x.c, &n, &bytes)
	return uint32(n), uint64(bytes), errOf(rc)
}

// ChunkDecodeLaunch is the loop body of streamFromLLM (src/llm-stream-proxy/main.go:188-226) over the response lines of
// many chats: "data: " filter, "[DONE]", json.Unmarshal into StreamChunk, per choice content / finish_reason.  Its token
// and done columns feed EncodeLaunch; the proxy adds conversation id, sequence++ and the timestamp per chat.
func (x *Context) ChunkDecodeLaunch(dLines, dLineOff uintptr, nLines uint32, dTok uintptr, tokCap uint64, dTokOff, dDone, dMsgLine uintptr,
	maxMsgs uint32, dLineStatus uintptr, stream unsafe.Pointer) error {
	return errOf(C.ssegpu_chunk_decode_launch(x.c, unsafe.Pointer(dLines), (*C.uint32_t)(unsafe.Pointer(dLineOff)), C.uint32_t(nLines),
		unsafe.Pointer(dTok), C.uint64_t(tokCap), (*C.uint32_t)(unsafe.Pointer(dTokOff)), (*C.uint8_t)(unsafe.Pointer(dDone)),
		(*C.uint32_t)(unsafe.Pointer(dMsgLine)), C.uint32_t(maxMsgs), (*C.uint32_t)(unsafe.Pointer(dLineStatus)), stream))
}

// ChunkDecodeResult: messages and token bytes of the last ChunkDecodeLaunch.
func (x *Context) ChunkDecodeResult() (uint32, uint64, error) {
	var n C.uint32_t
	var bytes C.uint64_t
	rc := C.ssegpu_chunk_decode_result(x.c, &n, &bytes)
	return uint32(n), uint64(bytes), errOf(rc)
}

// FanoutLaunch expands the subscriber spans OpMatch left on the device into the flat (subscriber, record) delivery list.
func (x *Context) FanoutLaunch(dSubSpan uintptr, n uint32, dPairSub, dPairRec uintptr, maxPairs uint64, dRecOff uintptr, stream unsafe.Pointer) error {
	return errOf(C.ssegpu_fanout_launch(x.c, (*C.uint32_t)(unsafe.Pointer(dSubSpan)), C.uint32_t(n), (*C.uint32_t)(unsafe.Pointer(dPairSub)),
		(*C.uint32_t)(unsafe.Pointer(dPairRec)), C.uint64_t(maxPairs), (*C.uint32_t)(unsafe.Pointer(dRecOff)), stream))
}

// FanoutResult: deliveries written by the last FanoutLaunch.
func (x *Context) FanoutResult() (uint64, error) {
	var n C.uint64_t
	rc := C.ssegpu_fanout_result(x.c, &n)
	return uint64(n), errOf(rc)
}
