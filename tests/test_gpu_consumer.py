"""GPU-box tests of the consumer side: batch pop, batch complete/fail, retries that re-enter the queue
behind their priority class, delayed items released by ReadyAt, dead-letter requeue.  Restates
/root/reference/tests/priorityqueue_test.go TestDelayedQueue (:471-560) and TestDeadLetterQueue (:563-698) against the
mirror, and checks the engine's ordering of large retry / schedule batches (sorted on the device) against the oracle's
container/heap and a numpy sort."""
import numpy as np
import pytest

import llm_message_queue_b200 as mq
from llm_message_queue_b200 import (DeadLetterQueue, DelayedQueue, ErrQueueFull, ExponentialBackoff, FixedBackoff, Priority, QueueError,
                                    QueueManager, QueueManagerConfig, handle_failures, new_message)
from llm_message_queue_b200.consumer import MILLISECOND, SECOND, ErrIndexOutOfRange
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu


def test_batch_pop_complete_fail_counters():
    qm = QueueManager(QueueManagerConfig(default_max_size=100))
    qm.create_queue("work", 100)
    msgs = [new_message("c", "u", f"m{i}", Priority.NORMAL) for i in range(7)]
    assert qm.batch_push_messages("work", msgs) == 7
    got = qm.batch_pop_messages("work", 5)                       # queue_manager.go:326-367
    assert got == msgs[:5]
    s = qm.get_queue_stats("work")
    assert (s.pending_count, s.processing_count) == (2, 5)
    qm.complete_batch("work", 3)
    qm.fail_batch("work", 2)
    s = qm.get_queue_stats("work")
    assert (s.pending_count, s.processing_count, s.completed_count, s.failed_count) == (2, 0, 3, 2)
    assert qm.batch_pop_messages("work", 10) == msgs[5:] and qm.batch_pop_messages("work", 10) == []
    with pytest.raises(QueueError):
        qm.batch_pop_messages("missing", 1)


def test_retry_goes_behind_its_priority_class_and_capacity_applies():
    qm = QueueManager(QueueManagerConfig(default_max_size=4))
    qm.create_queue("work", 4)
    a, b, c, d = (new_message("c", "u", x, p) for x, p in (("a", 3), ("b", 3), ("c", 2), ("d", 4)))
    for m in (a, b, c, d):
        qm.push_message("work", m)
    assert qm.batch_pop_messages("work", 2) == [c, a]            # (priority, time)
    back = qm.requeue_batch("work", [a, c])                       # worker.go:229: pushed again, fresh timestamps
    assert back == [True, True]
    e = new_message("c", "u", "e", 1)
    assert qm.requeue_batch("work", [e]) == [False]               # queue.go:101-103: full, skipped
    assert qm.batch_pop_messages("work", 10) == [c, b, a, d]      # a is now behind b


def test_large_retry_batch_is_ordered_on_the_device_like_the_heap():
    n = 70000
    rng = np.random.default_rng(11)
    prio = rng.integers(0, 16, n).astype(np.int32)
    ts = rng.permutation(n).astype(np.uint64) * np.uint64(13) + np.uint64(10**18)      # unique, shuffled
    handles = np.arange(1, n + 1, dtype=np.uint64) + np.uint64(1 << 40)
    e = mq.Engine(device=0)
    try:
        q = e.queue_create(0)
        acc = e.requeue(q, handles, prio, ts)
        assert acc.all()
        got = e.drain_queue(q, n + 5)
        want = handles[np.lexsort((ts, prio))]
        assert np.array_equal(got, want)
        h = co.lib().orc_heap_new()                               # the oracle's container/heap gives the same pops
        for i in range(0, n, 7):
            co.lib().orc_heap_push(h, int(prio[i]), int(ts[i]), i)
        import ctypes as C
        ident = C.c_uint32()
        pops = []
        while co.lib().orc_heap_pop(h, C.byref(ident)):
            pops.append(ident.value)
        co.lib().orc_heap_free(h)
        sub = np.arange(0, n, 7)
        assert np.array_equal(np.array(pops), sub[np.lexsort((ts[sub], prio[sub]))])
        # duplicates in (priority, time): arrival order decides; second batch merges into what is waiting
        ts2 = (rng.integers(0, 40, n).astype(np.uint64) << np.uint64(20)) + np.uint64(10**18)
        e.requeue(q, handles[:n // 2], prio[:n // 2], ts2[:n // 2])
        e.requeue(q, handles[n // 2:], prio[n // 2:], ts2[n // 2:])
        got = e.drain_queue(q, n)
        assert np.array_equal(got, handles[np.lexsort((np.arange(n), ts2, prio))])
        # priorities outside 0..15 fall back to the host sort
        wide = prio.copy(); wide[::100] = 1000; wide[5::100] = -3
        e.requeue(q, handles, wide, ts)
        assert np.array_equal(e.drain_queue(q, n), handles[np.lexsort((ts, wide))])
    finally:
        e.close()


def test_delayed_scheduling():                                    # priorityqueue_test.go:471-529
    got = []
    dq = DelayedQueue(process_fn=lambda m: got.append(m.id))
    msg1 = new_message("conv-delay", "user-delay", "Delay test message 1", Priority.NORMAL)
    msg2 = new_message("conv-delay", "user-delay", "Delay test message 2", Priority.HIGH)
    start = 1_700_000_000 * SECOND
    dq.schedule(msg1, start + 200 * MILLISECOND)
    dq.schedule(msg2, start + 500 * MILLISECOND)
    assert dq.poll(start + 199 * MILLISECOND) == []               # "Message processed too early"
    assert dq.poll(start + 200 * MILLISECOND) == [msg1] and got == [msg1.id]
    assert dq.poll(start + 499 * MILLISECOND) == []
    assert dq.poll(start + 900 * MILLISECOND) == [msg2] and got == [msg1.id, msg2.id]
    # :532-560 Peek
    msg = new_message("conv-peek", "user-peek", "Peek test message", Priority.NORMAL)
    ready = start + 2 * SECOND
    dq.schedule(msg, ready)
    pm, pt, ok = dq.peek()
    assert ok and pm is msg and pt == ready
    assert dq.poll(ready) == [msg]
    assert dq.peek() == (None, 0, False)


def test_large_schedule_batch_is_released_in_ready_order_into_the_queues():
    n = 50000
    rng = np.random.default_rng(12)
    qm = QueueManager(QueueManagerConfig(default_max_size=0))
    qm.create_queue("work")
    dq = DelayedQueue(manager=qm)
    msgs = [new_message("c", "u", str(i), int(p)) for i, p in enumerate(rng.integers(1, 5, n))]
    base = 1_700_000_000 * SECOND
    ready = base + rng.permutation(n).astype(np.uint64) * np.uint64(1000)
    dq.schedule_batch(msgs, ready, "work")
    cut = base + 30000 * 1000
    due = dq.poll(cut)
    idx = np.nonzero(ready <= cut)[0]
    idx = idx[np.argsort(ready[idx], kind="stable")]
    assert [m.content for m in due] == [msgs[i].content for i in idx]        # earliest ReadyAt first
    assert dq.size() == n - len(idx) and dq.peek()[1] == int(ready[ready > cut].min())
    # released items were pushed into "work" in that order: the queue pops them by (priority, release order)
    prios = np.array([int(msgs[i].priority) for i in idx])
    want = idx[np.argsort(prios, kind="stable")]
    got = qm.batch_pop_messages("work", n)
    assert [m.content for m in got] == [msgs[i].content for i in want]


def test_dead_letter_queue():                                     # priorityqueue_test.go:563-698
    dlq = DeadLetterQueue(100)
    manager = QueueManager(QueueManagerConfig(default_max_size=100))
    manager.create_queue("requeue-test", 100)
    msg = new_message("conv-dlq", "user-dlq", "DLQ test message", Priority.NORMAL)
    msg.retry_count = 3
    dlq.push(msg, "处理超时", "test-queue")
    item = dlq.get(0)
    assert item.message is msg and item.fail_reason == "处理超时" and item.source_queue == "test-queue" and item.retry_count == 3
    msg2 = new_message("conv-requeue", "user-requeue", "Requeue test message", Priority.HIGH)
    msg2.retry_count = 2
    dlq.push(msg2, "处理失败", "requeue-test")
    size = dlq.size()
    dlq.requeue(size - 1, manager)
    popped = manager.pop_message("requeue-test")
    assert popped is msg2 and popped.retry_count == 0
    with pytest.raises(QueueError) as ei:
        dlq.requeue(99, manager)
    assert ei.value is ErrIndexOutOfRange
    dlq.clear()
    batch = []
    for _ in range(5):
        m = new_message("conv-batch-requeue", "user-batch", "Batch requeue test", Priority.NORMAL)
        m.retry_count = 1
        dlq.push(m, "批量测试", "requeue-test")
        batch.append(m)
    assert dlq.size() == 5
    assert dlq.batch_requeue([0, 2, 4], manager) == 3 and dlq.size() == 2
    assert [it.message for it in dlq.items] == [batch[1], batch[3]]
    got = [manager.pop_message("requeue-test") for _ in range(3)]
    assert got == [batch[4], batch[2], batch[0]] and all(m.retry_count == 0 for m in got)   # taken from the highest index down
    small = DeadLetterQueue(1)
    small.push(msg, "x", "q")
    with pytest.raises(QueueError) as ei:
        small.push(msg, "x", "q")
    assert ei.value is ErrQueueFull


def test_handle_failures_backoff_retry_and_give_up():
    qm = QueueManager(QueueManagerConfig(default_max_size=0))
    qm.create_queue("work")
    dq = DelayedQueue(manager=qm)
    bo = ExponentialBackoff(1 * SECOND, 10 * SECOND, 2.0, 3)
    msgs = [new_message("c", "u", f"m{i}", Priority.NORMAL) for i in range(4)]
    for m, rc in zip(msgs, (0, 1, 2, 3)):
        m.retry_count = rc
    qm.batch_push_messages("work", msgs)
    popped = qm.batch_pop_messages("work", 4)
    now = 1_700_000_000 * SECOND
    retried, dead = handle_failures(qm, "work", popped, bo, delayed=dq, now_ns=now)
    assert retried == msgs[:3] and dead == [msgs[3]]
    assert [m.retry_count for m in msgs] == [1, 2, 3, 3]
    assert qm.get_queue_stats("work").failed_count == 1
    assert dq.poll(now + 1 * SECOND) == [msgs[0]]                 # 1 s, 2 s, 4 s
    assert dq.poll(now + 3 * SECOND) == [msgs[1]]
    assert dq.poll(now + 4 * SECOND) == [msgs[2]]
    assert qm.batch_pop_messages("work", 9) == msgs[:3]
    # the reference's own behaviour (no delayed queue): straight back into the queue
    retried, dead = handle_failures(qm, "work", [msgs[0]], FixedBackoff(5 * SECOND, 5), now_ns=now)
    assert retried == [msgs[0]] and msgs[0].retry_count == 2 and qm.pop_message("work") is msgs[0]
