"""Host-side mirror of the consumer side of internal/priorityqueue: what happens to messages after the dequeue.

  ExponentialBackoff / FixedBackoff   worker.go:266-316
  handle_failures                     Worker.handleFailure for a whole processBatch round, worker.go:128-158, 197-239
  DelayedQueue                        delayed_queue.go (Schedule :98, ScheduleAfter :108, Peek, the run loop :113-170)
  DeadLetterQueue                     dead_letter_queue.go (Push :62, Get, Size, Clear :181, Requeue :187, BatchRequeue :218)

Message objects and names live here; ordering, capacity and statistics arithmetic run in the engine behind the C ABI
(triage_drain_queue, triage_requeue, triage_schedule, triage_release_due, triage_backoff_ns).  Durations and times are
integer nanoseconds, like time.Duration / UnixNano."""
from __future__ import annotations

import time
from dataclasses import dataclass, field

from . import _capi
from .models import Message
from .priorityqueue import ErrQueueFull, QueueError, QueueManager

SECOND = 1_000_000_000
MILLISECOND = 1_000_000

ErrIndexOutOfRange = QueueError("INDEX_OUT_OF_RANGE", "index out of range")   # dead_letter_queue.go:273-275


class ExponentialBackoff:   # worker.go:266-294
    def __init__(self, initial_ns: int, max_ns: int, factor: float, max_retries: int):
        self.initial_backoff, self.max_backoff, self.factor, self._max_retries = int(initial_ns), int(max_ns), float(factor), int(max_retries)

    def next_backoff(self, retry_count: int) -> int:
        return _capi.backoff_ns(self.initial_backoff, self.max_backoff, self.factor, retry_count)

    def max_retries(self) -> int:
        return self._max_retries

    NextBackoff, MaxRetries = next_backoff, max_retries


class FixedBackoff:   # worker.go:296-316
    def __init__(self, backoff_ns: int, max_retries: int):
        self.backoff, self._max_retries = int(backoff_ns), int(max_retries)

    def next_backoff(self, retry_count: int) -> int:
        return self.backoff

    def max_retries(self) -> int:
        return self._max_retries

    NextBackoff, MaxRetries = next_backoff, max_retries


class DelayedQueue:
    """delayed_queue.go without the goroutines: poll(now) plays one turn of the run loop.  process_fn(message) is what
    the reference calls for a ready item; with a QueueManager given instead, a ready item is pushed into the queue it was
    scheduled for (what the comment at worker.go:226-228 says retries should do)."""

    def __init__(self, process_fn=None, manager: QueueManager | None = None, engine: _capi.Engine | None = None):
        self.process_fn = process_fn
        self.manager = manager
        self.engine = manager.engine if manager is not None else (engine if engine is not None else _capi.Engine())
        self._msgs: dict[int, tuple[Message, str]] = {}
        self._own = self.engine.queue_create(0) if manager is None else None   # stand-alone: a private unbounded queue
        self._next = 1 << 62                                                     # handles of items not known to a manager

    def schedule(self, message: Message, ready_at_ns: int, queue_name: str = "") -> None:
        self.schedule_batch([message], [ready_at_ns], queue_name)

    def schedule_after(self, message: Message, delay_ns: int, queue_name: str = "") -> None:
        self.schedule(message, time.time_ns() + int(delay_ns), queue_name)

    def schedule_batch(self, messages, ready_at_ns, queue_name: str = "") -> None:
        if self.manager is not None:
            qid = self.manager._ids.get(queue_name)
            if qid is None:
                raise QueueError("QUEUE_NOT_FOUND", "queue not found")
        else:
            qid = self._own
        handles = []
        for m in messages:
            handles.append(self._next)
            self._msgs[self._next] = (m, queue_name)
            self._next += 1
        self.engine.schedule(qid, handles, [int(m.priority) for m in messages], ready_at_ns)

    def peek(self):
        """(message, ready_at_ns, True) of the earliest item, or (None, 0, False)"""
        try:
            h, t = self.engine.delayed_peek()
        except _capi.TriageError as e:
            if e.status == _capi.E_EMPTY:
                return None, 0, False
            raise
        return self._msgs[h][0], t, True

    def size(self) -> int:
        return len(self._msgs)

    def poll(self, now_ns: int | None = None) -> list[Message]:
        """every item whose ReadyAt <= now, earliest first; each is handed to process_fn, or sits in its queue afterwards"""
        handles, queues, _ = self.engine.release_due(time.time_ns() if now_ns is None else int(now_ns))
        out = []
        for h in handles:
            m, qname = self._msgs.pop(int(h))
            out.append(m)
            if self.manager is not None:
                self.manager._msgs[int(h)] = m          # it is now an ordinary queued message of the manager
            elif self._own is not None:
                self.engine.drain_queue(self._own, 1)    # stand-alone: the private queue is only a conveyor
        if self.process_fn is not None:
            for m in out:
                self.process_fn(m)
        return out

    Schedule, ScheduleAfter, Peek = schedule, schedule_after, peek


@dataclass
class DeadLetterItem:   # dead_letter_queue.go:13-20
    message: Message
    fail_reason: str
    failed_at: float
    retry_count: int
    source_queue: str
    error_details: str = ""


@dataclass
class DeadLetterQueue:
    max_size: int = 0
    items: list = field(default_factory=list)

    def push(self, message: Message, fail_reason: str, source_queue: str) -> None:
        if self.max_size > 0 and len(self.items) >= self.max_size:   # :66-69
            raise ErrQueueFull
        self.items.append(DeadLetterItem(message, fail_reason, time.time(), message.retry_count, source_queue))

    def get(self, index: int) -> DeadLetterItem:
        if index < 0 or index >= len(self.items):
            raise ErrIndexOutOfRange
        return self.items[index]

    def size(self) -> int:
        return len(self.items)

    def clear(self) -> None:
        self.items = []

    def requeue(self, index: int, manager: QueueManager) -> None:   # :187-215
        item = self.get(index)
        item.message.retry_count = 0
        manager.push_message(item.source_queue, item.message)       # an error leaves the item where it is
        del self.items[index]

    def batch_requeue(self, indices: list[int], manager: QueueManager) -> int:   # :218-258
        """one triage_requeue per source queue instead of one locked push per item; same result: the items are taken
        from the highest index down, a push the queue turns away is skipped and its item stays"""
        order = sorted(set(i for i in indices if 0 <= i < len(self.items)), reverse=True)
        ok = 0
        by_queue: dict[str, list[int]] = {}
        for i in order:
            by_queue.setdefault(self.items[i].source_queue, []).append(i)
        done = []
        for qname, idxs in by_queue.items():
            msgs = [self.items[i].message for i in idxs]
            for m in msgs:
                m.retry_count = 0
            try:
                accepted = manager.requeue_batch(qname, msgs)
            except QueueError:
                continue
            for i, a in zip(idxs, accepted):
                if a:
                    done.append(i)
                    ok += 1
        for i in sorted(done, reverse=True):
            del self.items[i]
        return ok

    Push, Get, Size, Clear, Requeue, BatchRequeue = push, get, size, clear, requeue, batch_requeue


def handle_failures(manager: QueueManager, queue_name: str, failed: list[Message], backoff, delayed: DelayedQueue | None = None,
                    now_ns: int | None = None) -> tuple[list[Message], list[Message]]:
    """Worker.handleFailure (worker.go:197-239) for the failed messages of one batch: a message with retries left gets
    RetryCount+1 and ScheduledAt = now + NextBackoff(RetryCount) and is pushed again - straight back into the queue as the
    reference does (:229), or into `delayed` so that the back-off is honoured; the others are marked failed.
    Returns (retried, given up)."""
    now = time.time_ns() if now_ns is None else int(now_ns)
    retry, dead, ready = [], [], []
    for m in failed:
        if m.retry_count < backoff.max_retries():
            m.retry_count += 1
            at = now + backoff.next_backoff(m.retry_count)
            m.scheduled_at = at / 1e9
            retry.append(m)
            ready.append(at)
        else:
            dead.append(m)
    if retry:
        if delayed is not None:
            delayed.schedule_batch(retry, ready, queue_name)
        else:
            manager.requeue_batch(queue_name, retry)
    if dead:
        manager.fail_batch(queue_name, len(dead))
    return retry, dead
