// Package triage is the cgo binding of libtriage_b200.so (include/triage.h): the B200 batch triage engine behind
// llm-message-queue's preprocessor -> priority queue path.
//
// SOURCE ONLY: this image has no Go toolchain, so the package has never been compiled here.  The same C ABI is
// exercised call by call from Python (llm-message-queue_b200/_capi.py) by the test-suite; this file is the thin
// translation a maintainer drops into the reference tree (build with CGO_ENABLED=1; deployments/Dockerfile:15-17 sets 0).
//
// Ownership: Go owns every *models.Message.  cgo forbids C code to keep Go pointers, so message content is copied into
// pinned arenas allocated by the engine (triage_host_alloc) and messages are tracked by handle on the Go side.
package triage

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../llm-message-queue_b200 -ltriage_b200
#include <stdlib.h>
#include "triage.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"unsafe"
)

// Status codes 1..3 are mapped to the reference's error singletons by package priorityqueue.
const (
	OK        = 0
	ENotFound = 1
	EFull     = 2
	EEmpty    = 3
	EPattern  = 4
	EArg      = 5
)

// Error carries a status of include/triage.h and triage_last_error()'s text.
type Error struct {
	Status int
	Msg    string
}

func (e *Error) Error() string { return fmt.Sprintf("triage status %d: %s", e.Status, e.Msg) }

func check(rc C.int32_t) error {
	if rc == 0 {
		return nil
	}
	return &Error{Status: int(rc), Msg: C.GoString(C.triage_last_error())}
}

// Meta and Result mirror triage_meta / triage_result (16 bytes each).
type Meta = C.triage_meta
type Result = C.triage_result

// Engine is one triage_engine: one GPU, one set of level queues.
type Engine struct {
	h    *C.triage_engine
	ring [2]*Arena // pinned host arenas: batch k+1 is packed and uploaded while batch k is processed
	turn int
}

// New = preprocessor.NewPreprocessor + NewMultiLevelQueue/AddQueue x4 (preprocessor.go:24, queue.go:70-90).
func New(device int, rules bool, maxQueueSize int64) (*Engine, error) {
	opts := C.triage_opts{device: C.int32_t(device), max_queue_size: C.int64_t(maxQueueSize)}
	if rules {
		opts.rules_enabled = 1
	}
	e := &Engine{}
	if err := check(C.triage_create(&opts, &e.h)); err != nil {
		return nil, err
	}
	return e, nil
}

func (e *Engine) Close() {
	for _, a := range e.ring {
		if a != nil {
			a.free()
		}
	}
	C.triage_destroy(e.h)
	e.h = nil
}

// AddPattern = Preprocessor.AddKeywordPattern (preprocessor.go:176-184): literals and the supported RE2 subset.
func (e *Engine) AddPattern(prio int, pattern string) error {
	cs := C.CString(pattern)
	defer C.free(unsafe.Pointer(cs))
	return check(C.triage_dict_add_pattern(e.h, C.int32_t(prio), cs, C.uint32_t(len(pattern))))
}

func (e *Engine) SetDefaultPriority(prio int) error { return check(C.triage_set_default_priority(e.h, C.int32_t(prio))) }

func (e *Engine) SetUserPriority(userID string, prio int) error {
	cs := C.CString(userID)
	defer C.free(unsafe.Pointer(cs))
	return check(C.triage_set_user_priority(e.h, cs, C.uint32_t(len(userID)), C.int32_t(prio)))
}

func (e *Engine) PadByte() (int, error) {
	var p C.int32_t
	err := check(C.triage_pad_byte(e.h, &p))
	return int(p), err
}

// Arena is one pinned host buffer set in the padded layout (include/triage.h, "padded arena"): message i starts at
// Offsets[i] (a multiple of 16), holds Lengths[i] bytes and is followed by 1..16 pad bytes.
type Arena struct {
	bytes   unsafe.Pointer
	cap     uint64
	Offsets []uint64
	Lengths []uint32
	Meta    []Meta
	Results []Result
	n       int
	total   uint64
	pad     byte
	Tag     interface{} // what the caller packed into it (lets a batch packed ahead of time be found again)
}

func pinned(n uint64) (unsafe.Pointer, error) {
	var p unsafe.Pointer
	err := check(C.triage_host_alloc(&p, C.uint64_t(n)))
	return p, err
}

func (e *Engine) newArena(maxMsgs int, maxBytes uint64) (*Arena, error) {
	pad, err := e.PadByte()
	if err != nil {
		return nil, err
	}
	if pad < 0 {
		return nil, errors.New("triage: this dictionary has no reset pad byte; use the packed entry points")
	}
	a := &Arena{cap: maxBytes + 16*uint64(maxMsgs) + 64, pad: byte(pad)}
	if a.bytes, err = pinned(a.cap); err != nil {
		return nil, err
	}
	// the small arrays are pinned as well: every array of a batch is copied by the DMA engine
	po, _ := pinned(uint64(8 * (maxMsgs + 1)))
	pl, _ := pinned(uint64(4 * maxMsgs))
	pm, _ := pinned(uint64(16 * maxMsgs))
	pr, err := pinned(uint64(16 * maxMsgs))
	if err != nil {
		return nil, err
	}
	a.Offsets = unsafe.Slice((*uint64)(po), maxMsgs+1)
	a.Lengths = unsafe.Slice((*uint32)(pl), maxMsgs)
	a.Meta = unsafe.Slice((*Meta)(pm), maxMsgs)
	a.Results = unsafe.Slice((*Result)(pr), maxMsgs)
	return a, nil
}

func (a *Arena) free() {
	C.triage_host_free(a.bytes)
	C.triage_host_free(unsafe.Pointer(&a.Offsets[0]))
	C.triage_host_free(unsafe.Pointer(&a.Lengths[0]))
	C.triage_host_free(unsafe.Pointer(&a.Meta[0]))
	C.triage_host_free(unsafe.Pointer(&a.Results[0]))
}

// Reset starts a new batch in the arena.
func (a *Arena) Reset() { a.n, a.total, a.Tag = 0, 0, nil }

// Packed returns the ring arena tagged with `tag`, if any.
func (e *Engine) Packed(tag interface{}) *Arena {
	for _, a := range e.ring {
		if a != nil && a.Tag != nil && a.Tag == tag {
			return a
		}
	}
	return nil
}

// Append copies one message's content into the arena (the copy the shim has to make anyway) and records its meta.
func (a *Arena) Append(content string, m Meta) bool {
	slot := (uint64(len(content))/16 + 1) * 16
	if a.n >= len(a.Lengths) || a.total+slot > a.cap {
		return false
	}
	dst := unsafe.Slice((*byte)(unsafe.Add(a.bytes, a.total)), slot)
	copy(dst, content)
	for i := len(content); i < int(slot); i++ {
		dst[i] = a.pad
	}
	a.Offsets[a.n] = a.total
	a.Lengths[a.n] = uint32(len(content))
	a.Meta[a.n] = m
	a.n++
	a.total += slot
	a.Offsets[a.n] = a.total
	return true
}

func (a *Arena) Len() int { return a.n }

// Ring hands out the two pinned arenas in turn.
func (e *Engine) Ring(maxMsgs int, maxBytes uint64) (*Arena, error) {
	i := e.turn & 1
	e.turn++
	if a := e.ring[i]; a != nil && (len(a.Lengths) < maxMsgs || a.cap < maxBytes+16*uint64(maxMsgs)+64) {
		a.free() // grown on demand; steady-state batches reuse the same pinned memory
		e.ring[i] = nil
	}
	if e.ring[i] == nil {
		a, err := e.newArena(maxMsgs, maxBytes)
		if err != nil {
			return nil, err
		}
		e.ring[i] = a
	}
	e.ring[i].Reset()
	return e.ring[i], nil
}

// Prefetch starts the upload of a packed arena; the matching Submit then only waits for the copy.
func (e *Engine) Prefetch(a *Arena) error {
	return check(C.triage_prefetch(e.h, (*C.uint64_t)(&a.Offsets[0]), (*C.uint32_t)(&a.Lengths[0]), (*C.uint8_t)(a.bytes),
		&a.Meta[0], nil, C.uint32_t(a.n)))
}

// Submit = ProcessMessage + PushMessage(fmt.Sprint(Priority), msg) for the whole arena (api/handlers.go:178-202), or
// BatchPushMessages into one queue when targetQueue != 0 (queue_manager.go:246-287).  Returns the handle of message 0.
func (e *Engine) Submit(a *Arena, targetQueue int) (uint64, error) {
	var base C.uint64_t
	err := check(C.triage_submit_padded(e.h, C.int32_t(targetQueue), (*C.uint64_t)(&a.Offsets[0]), (*C.uint32_t)(&a.Lengths[0]),
		(*C.uint8_t)(a.bytes), &a.Meta[0], nil, C.uint32_t(a.n), &a.Results[0], &base))
	return uint64(base), err
}

// Process = ProcessMessage only (no queue is touched).
func (e *Engine) Process(a *Arena) error {
	return check(C.triage_process_padded(e.h, (*C.uint64_t)(&a.Offsets[0]), (*C.uint32_t)(&a.Lengths[0]), (*C.uint8_t)(a.bytes),
		&a.Meta[0], C.uint32_t(a.n), &a.Results[0]))
}

// MetaFields are the exported twins of triage_meta's fields (cgo struct fields are not visible outside this package).
type MetaFields struct{ Preset, UserPriority, UserDefault, MetaPriority, Flags uint8 }

func MakeMeta(f MetaFields) Meta {
	var m Meta
	m.preset, m.user_priority, m.user_default = C.uint8_t(f.Preset), C.uint8_t(f.UserPriority), C.uint8_t(f.UserDefault)
	m.meta_priority, m.flags = C.uint8_t(f.MetaPriority), C.uint8_t(f.Flags)
	return m
}

// accessors of triage_result
func Label(r *Result) int      { return int(r.label) }
func Reason(r *Result) int     { return int(r.reason) }
func Sentiment(r *Result) int  { return int(r.sentiment) }
func Question(r *Result) bool  { return r.question != 0 }
func Accepted(r *Result) bool  { return r.accepted != 0 }
func Admit(r *Result) int      { return int(r.admit) }
func Flags(r *Result) int      { return int(r.flags) }
func QueueLevel(r *Result) int { return int(r.queue_level) }
func Words(r *Result) uint32   { return uint32(r.words) }
