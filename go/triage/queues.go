package triage

// Queue side of the binding: queue.go:92-211, queue_manager.go:170-367, worker.go:197-239, delayed_queue.go,
// dead_letter_queue.go:187-258.  SOURCE ONLY (see triage.go).

/*
#include "triage.h"
*/
import "C"

// Stats mirrors the counters of priorityqueue.QueueStats (queue.go:60-68).
type Stats struct{ Pending, Processing, Completed, Failed int64 }

func (e *Engine) QueueCreate(maxSize int64) (int, error) {
	var q C.int32_t
	err := check(C.triage_queue_create(e.h, C.int64_t(maxSize), &q))
	return int(q), err
}

func (e *Engine) QueueDelete(q int) error { return check(C.triage_queue_delete(e.h, C.int32_t(q))) }

func (e *Engine) SetLimits(maxQueueSize int64, rules bool) error {
	r := C.int32_t(0)
	if rules {
		r = 1
	}
	return check(C.triage_set_limits(e.h, C.int64_t(maxQueueSize), r))
}

// Push = MultiLevelQueue.Push(name, msg, priority) (queue.go:92-119).
func (e *Engine) Push(q, itemPriority int) (uint64, error) {
	var h C.uint64_t
	err := check(C.triage_push(e.h, C.int32_t(q), C.int32_t(itemPriority), 0, &h))
	return uint64(h), err
}

func (e *Engine) Pop(q int) (uint64, error) {
	var h C.uint64_t
	err := check(C.triage_pop(e.h, C.int32_t(q), &h))
	return uint64(h), err
}

func (e *Engine) Peek(q int) (uint64, error) {
	var h C.uint64_t
	err := check(C.triage_peek(e.h, C.int32_t(q), &h))
	return uint64(h), err
}

func (e *Engine) Size(q int) (int, error) {
	var n C.uint64_t
	err := check(C.triage_queue_size(e.h, C.int32_t(q), &n))
	return int(n), err
}

func (e *Engine) Stats(q int) (Stats, error) {
	var s C.triage_qstats
	err := check(C.triage_stats(e.h, C.int32_t(q), &s))
	return Stats{int64(s.pending), int64(s.processing), int64(s.completed), int64(s.failed)}, err
}

func (e *Engine) Complete(q int) { C.triage_complete(e.h, C.int32_t(q)) }
func (e *Engine) Fail(q int)     { C.triage_fail(e.h, C.int32_t(q)) }

// Drain = the realtime -> high -> normal -> low loop of cmd/queue-manager/main.go:111-124, max times.
func (e *Engine) Drain(max int) (handles []uint64, levels []uint8, err error) {
	if max <= 0 {
		return nil, nil, nil
	}
	handles = make([]uint64, max)
	levels = make([]uint8, max)
	var n C.uint32_t
	err = check(C.triage_drain(e.h, C.uint32_t(max), (*C.uint64_t)(&handles[0]), (*C.uint8_t)(&levels[0]), &n))
	return handles[:n], levels[:n], err
}

// DrainQueue = QueueManager.BatchPopMessages(name, count) (queue_manager.go:326-367): one lock, one stats update.
func (e *Engine) DrainQueue(q, max int) ([]uint64, error) {
	if max <= 0 {
		return nil, nil
	}
	handles := make([]uint64, max)
	var n C.uint32_t
	err := check(C.triage_drain_queue(e.h, C.int32_t(q), C.uint32_t(max), (*C.uint64_t)(&handles[0]), &n))
	return handles[:n], err
}

func (e *Engine) CompleteBatch(q, k int) { C.triage_complete_batch(e.h, C.int32_t(q), C.uint32_t(k)) }
func (e *Engine) FailBatch(q, k int)     { C.triage_fail_batch(e.h, C.int32_t(q), C.uint32_t(k)) }

// Requeue = PushMessage for messages that already have handles (worker.go:229 retries, dead-letter requeue).
func (e *Engine) Requeue(q int, handles []uint64, prios []int32) (accepted []uint8, err error) {
	if len(handles) == 0 {
		return nil, nil
	}
	accepted = make([]uint8, len(handles))
	var n C.uint32_t
	err = check(C.triage_requeue(e.h, C.int32_t(q), (*C.uint64_t)(&handles[0]), (*C.int32_t)(&prios[0]), nil,
		C.uint32_t(len(handles)), (*C.uint8_t)(&accepted[0]), &n))
	return accepted, err
}

// Schedule / ReleaseDue = DelayedQueue.Schedule and one turn of its run loop (delayed_queue.go:98-170).
func (e *Engine) Schedule(q int, handles []uint64, prios []int32, readyAtNs []uint64) error {
	if len(handles) == 0 {
		return nil
	}
	return check(C.triage_schedule(e.h, C.int32_t(q), (*C.uint64_t)(&handles[0]), (*C.int32_t)(&prios[0]),
		(*C.uint64_t)(&readyAtNs[0]), C.uint32_t(len(handles))))
}

func (e *Engine) ReleaseDue(nowNs uint64, max int) (handles []uint64, queues []int32, rejected int, err error) {
	handles = make([]uint64, max)
	queues = make([]int32, max)
	var n, r C.uint32_t
	err = check(C.triage_release_due(e.h, C.uint64_t(nowNs), (*C.uint64_t)(&handles[0]), (*C.int32_t)(&queues[0]), C.uint32_t(max), &n, &r))
	return handles[:n], queues[:n], int(r), err
}

// BackoffNs = ExponentialBackoff.NextBackoff (worker.go:274-290).
func BackoffNs(initialNs, maxNs int64, factor float64, retryCount int) int64 {
	return int64(C.triage_backoff_ns(C.int64_t(initialNs), C.int64_t(maxNs), C.double(factor), C.int32_t(retryCount)))
}
