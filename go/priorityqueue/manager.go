package priorityqueue

// QueueManager (queue_manager.go:15-466) on the engine, plus the additive batch calls.  SOURCE ONLY.

import (
	"fmt"
	"time"

	"go.uber.org/zap"

	"llm-message-queue/internal/preprocessor"
	"llm-message-queue/internal/triage"
	"llm-message-queue/pkg/models"
)

type PriorityAdjustRule struct { // queue_manager.go:39-44
	Condition   func(*models.Message) bool
	NewPriority models.Priority
	Description string
}

// QueueManagerConfig: queue_manager.go:27-37; the monitoring knobs are accepted and unused (observability is out of scope).
type QueueManagerConfig struct {
	DefaultMaxSize      int
	EnableMetrics       bool
	PriorityAdjustRules []PriorityAdjustRule
	BuiltinRules        bool // exactly the two rules of queue_factory.go:211-233: evaluated on the device by BatchTriage
}

type QueueManager struct {
	registry
	config *QueueManagerConfig
	logger *zap.Logger
	next   uint64 // handles of re-pushed messages, disjoint from the engine's own
}

func NewQueueManager(config *QueueManagerConfig, logger *zap.Logger) *QueueManager {
	e, err := triage.New(0, config.BuiltinRules, int64(config.DefaultMaxSize))
	if err != nil {
		panic(err)
	}
	return &QueueManager{registry: registry{e: e, ids: map[string]int{}, msgs: map[uint64]*models.Message{}}, config: config,
		logger: logger, next: 1 << 61}
}

func (qm *QueueManager) Engine() *triage.Engine { return qm.e }
func (qm *QueueManager) Start()                 {}
func (qm *QueueManager) Stop()                  {}

func (qm *QueueManager) CreateQueue(name string, maxSize int) {
	qm.mu.Lock()
	defer qm.mu.Unlock()
	if _, ok := qm.ids[name]; ok {
		return
	}
	if maxSize <= 0 {
		maxSize = qm.config.DefaultMaxSize
	}
	if l, ok := levels[name]; ok {
		qm.ids[name] = l
	} else if q, err := qm.e.QueueCreate(int64(maxSize)); err == nil {
		qm.ids[name] = q
	}
}

func (qm *QueueManager) DeleteQueue(name string) error {
	qm.mu.Lock()
	q, ok := qm.ids[name]
	delete(qm.ids, name)
	qm.mu.Unlock()
	if !ok {
		return ErrQueueNotFound
	}
	if q > 4 {
		return mapErr(qm.e.QueueDelete(q))
	}
	return nil
}

func (qm *QueueManager) applyPriorityRules(message *models.Message) { // :451-466, first match wins
	for _, rule := range qm.config.PriorityAdjustRules {
		if rule.Condition(message) {
			message.Priority = rule.NewPriority
			return
		}
	}
}

func (qm *QueueManager) PushMessage(queueName string, message *models.Message) error {
	n, err := qm.BatchPushMessages(queueName, []*models.Message{message})
	if err == nil && n == 0 {
		return ErrQueueFull
	}
	return err
}

// BatchPushMessages: rules and capacity per message in order, a failed push is skipped (:246-287) - one engine call.
// Retries (worker.go:229) and dead-letter requeues (dead_letter_queue.go:187-258) come through here as well.
func (qm *QueueManager) BatchPushMessages(queueName string, messages []*models.Message) (int, error) {
	q, ok := qm.id(queueName)
	if !ok {
		return 0, ErrQueueNotFound
	}
	handles := make([]uint64, len(messages))
	prios := make([]int32, len(messages))
	qm.mu.Lock()
	for i, m := range messages {
		qm.applyPriorityRules(m)
		qm.next++
		handles[i], prios[i] = qm.next, int32(m.Priority)
	}
	qm.mu.Unlock()
	acc, err := qm.e.Requeue(q, handles, prios)
	if err != nil {
		return 0, mapErr(err)
	}
	n := 0
	qm.mu.Lock()
	for i, a := range acc {
		if a != 0 {
			qm.msgs[handles[i]] = messages[i]
			n++
		}
	}
	qm.mu.Unlock()
	return n, nil
}

func (qm *QueueManager) PopMessage(queueName string) (*models.Message, error) {
	msgs, err := qm.BatchPopMessages(queueName, 1)
	if err != nil {
		return nil, err
	}
	if len(msgs) == 0 {
		return nil, ErrQueueEmpty
	}
	return msgs[0], nil
}

// BatchPopMessages pops until the queue is empty and returns what it has (:326-367): one lock, one statistics update.
func (qm *QueueManager) BatchPopMessages(queueName string, count int) ([]*models.Message, error) {
	q, ok := qm.id(queueName)
	if !ok {
		return nil, ErrQueueNotFound
	}
	handles, err := qm.e.DrainQueue(q, count)
	if err != nil {
		return nil, mapErr(err)
	}
	return qm.take(handles), nil
}

func (qm *QueueManager) CompleteMessage(queueName string, messageID string, processingTime time.Duration) {
	if q, ok := qm.id(queueName); ok {
		qm.e.Complete(q)
	}
}

func (qm *QueueManager) FailMessage(queueName string, messageID string, err error) {
	if q, ok := qm.id(queueName); ok {
		qm.e.Fail(q)
	}
}

func (qm *QueueManager) GetQueueStats(queueName string) (*QueueStats, error) { return qm.stats(queueName) }
func (qm *QueueManager) GetAllQueueStats() map[string]*QueueStats            { return qm.allStats() }

// BatchTriage = for every message in arrival order: ProcessMessage, then PushMessage(fmt.Sprint(msg.Priority), msg)
// (api/handlers.go:178-202), as ONE submit: the batch is packed into a pinned arena of the engine's ring (the copy of
// Content a cgo shim has to make anyway), uploaded in chunks with the scan following the copy, labelled, admitted and
// ordered on the device.  `next`, when not empty, is packed into the ring's other arena and its upload started first,
// so that it travels while this batch is processed; pass it as `msgs` of the following call.  Returns the number admitted.
func (qm *QueueManager) BatchTriage(p *preprocessor.Preprocessor, msgs, next []*models.Message) (int, error) {
	if p.Engine != qm.e {
		return 0, fmt.Errorf("BatchTriage needs a Preprocessor built on this manager's engine")
	}
	for name := range levels {
		if _, ok := qm.id(name); !ok {
			return 0, ErrQueueNotFound
		}
	}
	a, err := qm.pack(p, msgs)
	if err != nil {
		return 0, err
	}
	if len(next) > 0 {
		if b, err := qm.pack(p, next); err == nil {
			qm.e.Prefetch(b) // matched to its later Submit by the arena's pointers
		}
	}
	base, err := qm.e.Submit(a, 0)
	if err != nil {
		return 0, mapErr(err)
	}
	now := time.Now()
	n := 0
	qm.mu.Lock()
	for i, m := range msgs {
		r := &a.Results[i]
		preprocessor.Apply(m, r, now)
		if l := triage.QueueLevel(r); l >= 1 && l <= 4 {
			m.Priority = models.Priority(triage.Label(r)) // the rules mutate msg.Priority before the push (:455)
		}
		if triage.Accepted(r) {
			qm.msgs[base+uint64(i)] = m
			n++
		}
	}
	qm.mu.Unlock()
	return n, nil
}

// pack fills the ring arena that belongs to this batch; a batch that was already packed as `next` of the previous call
// finds its arena again (same slice header) and is not copied twice.
func (qm *QueueManager) pack(p *preprocessor.Preprocessor, msgs []*models.Message) (*triage.Arena, error) {
	if a := qm.e.Packed(&msgs[0]); a != nil {
		return a, nil
	}
	var total uint64
	for _, m := range msgs {
		total += uint64(len(m.Content))
	}
	a, err := qm.e.Ring(len(msgs), total)
	if err != nil {
		return nil, err
	}
	for _, m := range msgs {
		a.Append(m.Content, p.MetaOf(m))
	}
	a.Tag = &msgs[0]
	return a, nil
}

// Drain = the worker loop of cmd/queue-manager/main.go:111-124: realtime, high, normal, low - first non-empty wins.
func (qm *QueueManager) Drain(max int) ([]*models.Message, error) {
	handles, _, err := qm.e.Drain(max)
	if err != nil {
		return nil, mapErr(err)
	}
	return qm.take(handles), nil
}
