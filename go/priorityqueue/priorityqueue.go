// Package priorityqueue keeps the exported surface of llm-message-queue/internal/priorityqueue (queue.go:52-226,
// queue_manager.go:15-448) on top of the B200 engine.  SOURCE ONLY: not compiled in this image.
//
// Message objects, names and error singletons live here; ordering, capacity and the statistics counters live in the
// engine.  BatchTriage (manager.go) is the additive call the GPU wants: ProcessMessage + PushMessage(fmt.Sprint(Priority),
// msg) for a whole batch in one submit through the pinned ring.
package priorityqueue

import (
	"errors"
	"sync"
	"time"

	"llm-message-queue/internal/triage"
	"llm-message-queue/pkg/models"
)

// Error singletons, compared with == by the reference's tests (queue.go:213-226).
type QueueError struct{ Code, Message string }

func (e *QueueError) Error() string { return e.Message }

var (
	ErrQueueNotFound = &QueueError{Code: "QUEUE_NOT_FOUND", Message: "queue not found"}
	ErrQueueFull     = &QueueError{Code: "QUEUE_FULL", Message: "queue is full"}
	ErrQueueEmpty    = &QueueError{Code: "QUEUE_EMPTY", Message: "queue is empty"}
)

func mapErr(err error) error {
	var te *triage.Error
	if errors.As(err, &te) {
		switch te.Status {
		case triage.ENotFound:
			return ErrQueueNotFound
		case triage.EFull:
			return ErrQueueFull
		case triage.EEmpty:
			return ErrQueueEmpty
		}
	}
	return err
}

type QueueStats struct { // queue.go:60-68
	PendingCount, ProcessingCount, CompletedCount, FailedCount int64
	LastUpdateTime                                             time.Time
}

var levels = map[string]int{"realtime": 1, "high": 2, "normal": 3, "low": 4}

// registry shared by both front-ends: queue names -> engine ids, handles -> messages
type registry struct {
	mu   sync.Mutex
	e    *triage.Engine
	ids  map[string]int
	msgs map[uint64]*models.Message
}

func (r *registry) id(name string) (int, bool) {
	r.mu.Lock()
	defer r.mu.Unlock()
	q, ok := r.ids[name]
	return q, ok
}

func (r *registry) names() []string {
	r.mu.Lock()
	defer r.mu.Unlock()
	out := make([]string, 0, len(r.ids))
	for n := range r.ids {
		out = append(out, n)
	}
	return out
}

func (r *registry) take(handles []uint64) []*models.Message {
	out := make([]*models.Message, 0, len(handles))
	r.mu.Lock()
	for _, h := range handles {
		out = append(out, r.msgs[h])
		delete(r.msgs, h)
	}
	r.mu.Unlock()
	return out
}

func (r *registry) stats(name string) (*QueueStats, error) {
	q, ok := r.id(name)
	if !ok {
		return nil, ErrQueueNotFound
	}
	s, err := r.e.Stats(q)
	if err != nil {
		return nil, mapErr(err)
	}
	return &QueueStats{s.Pending, s.Processing, s.Completed, s.Failed, time.Now()}, nil
}

func (r *registry) allStats() map[string]*QueueStats {
	out := map[string]*QueueStats{}
	for _, n := range r.names() {
		if s, err := r.stats(n); err == nil {
			out[n] = s
		}
	}
	return out
}

// ---- MultiLevelQueue (queue.go:52-211) --------------------------------------------------------------------------
type MultiLevelQueue struct {
	registry
	maxSize int
}

func NewMultiLevelQueue(maxSize int) *MultiLevelQueue {
	e, err := triage.New(0, false, int64(maxSize))
	if err != nil {
		panic(err) // no CPU fallback
	}
	return &MultiLevelQueue{registry: registry{e: e, ids: map[string]int{}, msgs: map[uint64]*models.Message{}}, maxSize: maxSize}
}

func (mlq *MultiLevelQueue) AddQueue(name string) {
	mlq.mu.Lock()
	defer mlq.mu.Unlock()
	if _, ok := mlq.ids[name]; ok {
		return
	}
	if l, ok := levels[name]; ok {
		mlq.ids[name] = l
	} else if q, err := mlq.e.QueueCreate(int64(mlq.maxSize)); err == nil {
		mlq.ids[name] = q
	}
}

func (mlq *MultiLevelQueue) Push(queueName string, message *models.Message, priority int) error {
	q, ok := mlq.id(queueName)
	if !ok {
		return ErrQueueNotFound
	}
	h, err := mlq.e.Push(q, priority)
	if err != nil {
		return mapErr(err)
	}
	mlq.mu.Lock()
	mlq.msgs[h] = message
	mlq.mu.Unlock()
	return nil
}

func (mlq *MultiLevelQueue) Pop(queueName string) (*models.Message, error) {
	q, ok := mlq.id(queueName)
	if !ok {
		return nil, ErrQueueEmpty // queue.go:126-128: a missing queue reports empty
	}
	h, err := mlq.e.Pop(q)
	if err != nil {
		return nil, mapErr(err)
	}
	return mlq.take([]uint64{h})[0], nil
}

func (mlq *MultiLevelQueue) Peek(queueName string) (*models.Message, error) {
	q, ok := mlq.id(queueName)
	if !ok {
		return nil, ErrQueueEmpty
	}
	h, err := mlq.e.Peek(q)
	if err != nil {
		return nil, mapErr(err)
	}
	mlq.mu.Lock()
	defer mlq.mu.Unlock()
	return mlq.msgs[h], nil
}

func (mlq *MultiLevelQueue) Size(queueName string) (int, error) {
	q, ok := mlq.id(queueName)
	if !ok {
		return 0, ErrQueueNotFound
	}
	n, err := mlq.e.Size(q)
	return n, mapErr(err)
}

func (mlq *MultiLevelQueue) GetStats(queueName string) (*QueueStats, error) { return mlq.stats(queueName) }
func (mlq *MultiLevelQueue) GetAllStats() map[string]*QueueStats            { return mlq.allStats() }

func (mlq *MultiLevelQueue) CompleteMessage(queueName string) {
	if q, ok := mlq.id(queueName); ok {
		mlq.e.Complete(q)
	}
}

func (mlq *MultiLevelQueue) FailMessage(queueName string) {
	if q, ok := mlq.id(queueName); ok {
		mlq.e.Fail(q)
	}
}
