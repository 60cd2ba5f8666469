"""Host-side mirror of internal/priorityqueue (reference: queue.go, queue_manager.go, queue_factory.go:211-233).

MultiLevelQueue: NewMultiLevelQueue :70, AddQueue :78, Push :92, Pop :121, Peek :141, Size :153, GetStats :165,
GetAllStats :177, CompleteMessage :197, FailMessage :205; error singletons :213-217 (compared by identity).
QueueManager: CreateQueue :170, DeleteQueue :191, PushMessage :210, BatchPushMessages :246, PopMessage :290,
BatchPopMessages :326, CompleteMessage/FailMessage, GetQueueStats, applyPriorityRules :451.
Additive: QueueManager.batch_triage(msgs, preprocessor) — ProcessMessage + PushMessage(fmt.Sprint(Priority)) for a
whole batch in one GPU submit.  Queue bookkeeping (named queues, message objects) is host logic; ordering and
admission arithmetic for batches run on the device."""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Callable

from . import _capi
from .models import Message, priority_name
from .preprocessor import Preprocessor, apply_result, pack_messages

LEVELS = {"realtime": 1, "high": 2, "normal": 3, "low": 4}


class QueueError(Exception):
    def __init__(self, code: str, message: str):
        super().__init__(message)
        self.code = code
        self.message = message


ErrQueueNotFound = QueueError("QUEUE_NOT_FOUND", "queue not found")
ErrQueueFull = QueueError("QUEUE_FULL", "queue is full")
ErrQueueEmpty = QueueError("QUEUE_EMPTY", "queue is empty")
_BY_STATUS = {_capi.E_NOT_FOUND: ErrQueueNotFound, _capi.E_FULL: ErrQueueFull, _capi.E_EMPTY: ErrQueueEmpty}


def _raise(e: _capi.TriageError):
    err = _BY_STATUS.get(e.status)
    if err is None:
        raise e
    raise err


@dataclass
class QueueStats:   # queue.go:60-68
    pending_count: int = 0
    processing_count: int = 0
    completed_count: int = 0
    failed_count: int = 0
    last_update: float = 0.0


class MultiLevelQueue:
    def __init__(self, max_size: int = 0, engine: _capi.Engine | None = None, device: int = 0):
        self.max_size = max_size
        self.engine = engine if engine is not None else _capi.Engine(device=device, max_queue_size=max_size)
        self.engine.set_limits(max_size, False)
        self._ids: dict[str, int] = {}
        self._msgs: dict[int, Message] = {}

    def add_queue(self, name: str) -> None:
        if name in self._ids:
            return
        self._ids[name] = LEVELS[name] if name in LEVELS else self.engine.queue_create(self.max_size)

    def push(self, queue_name: str, message: Message, priority: int) -> None:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        try:
            h = self.engine.push(qid, int(priority))
        except _capi.TriageError as e:
            _raise(e)
        self._msgs[h] = message

    def pop(self, queue_name: str) -> Message:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueEmpty   # queue.go:126-128: a missing queue reports empty
        try:
            h = self.engine.pop(qid)
        except _capi.TriageError as e:
            _raise(e)
        return self._msgs.pop(h)

    def peek(self, queue_name: str) -> Message:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueEmpty
        try:
            h = self.engine.peek(qid)
        except _capi.TriageError as e:
            _raise(e)
        return self._msgs[h]

    def size(self, queue_name: str) -> int:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        return self.engine.queue_size(qid)

    def get_stats(self, queue_name: str) -> QueueStats:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        s = self.engine.stats(qid)
        return QueueStats(s["pending"], s["processing"], s["completed"], s["failed"], time.time())

    def get_all_stats(self) -> dict[str, QueueStats]:
        return {name: self.get_stats(name) for name in self._ids}

    def complete_message(self, queue_name: str) -> None:
        qid = self._ids.get(queue_name)
        if qid is not None:
            self.engine.complete(qid)

    def fail_message(self, queue_name: str) -> None:
        qid = self._ids.get(queue_name)
        if qid is not None:
            self.engine.fail(qid)

    # -- batch paths (device) -------------------------------------------------
    def attach(self, base_handle: int, msgs, results) -> int:
        n = 0
        for i, (m, r) in enumerate(zip(msgs, results)):
            if r["accepted"]:
                self._msgs[base_handle + i] = m
                n += 1
        return n

    def drain(self, max_items: int) -> list[Message]:
        """cmd/queue-manager/main.go:111-124 run max_items times: realtime -> high -> normal -> low."""
        handles, _ = self.engine.drain(max_items)
        return [self._msgs.pop(int(h)) for h in handles]


@dataclass
class PriorityAdjustRule:   # queue_manager.go:47-51
    condition: Callable[[Message], bool]
    new_priority: int
    description: str = ""


def builtin_priority_rules() -> list[PriorityAdjustRule]:
    """queue_factory.go:211-233"""
    def vip(m):
        return "vip_user" in (m.metadata or {})

    def long_content(m):
        return len(m.content_bytes()) > 10000
    vip._builtin = long_content._builtin = True   # marks the predicates the device implements (batch_triage checks it)
    return [PriorityAdjustRule(vip, 2, "vip"), PriorityAdjustRule(long_content, 4, "long")]


@dataclass
class QueueManagerConfig:
    default_max_size: int = 10000   # config.go:157
    enable_metrics: bool = False    # metrics / monitor goroutine are out of scope (observability)
    priority_adjust_rules: list = field(default_factory=list)
    builtin_rules: bool = False     # True = exactly the two rules of queue_factory.go, evaluated on the device in batch_triage


class QueueManager:
    """name -> one heap ordered by (int(msg.Priority), enqueue time) (queue_manager.go:170-188)."""

    def __init__(self, config: QueueManagerConfig | None = None, logger=None, engine: _capi.Engine | None = None, device: int = 0):
        self.config = config or QueueManagerConfig()
        if self.config.builtin_rules and not self.config.priority_adjust_rules:
            self.config.priority_adjust_rules = builtin_priority_rules()
        self.engine = engine if engine is not None else _capi.Engine(device=device)
        self.engine.set_limits(self.config.default_max_size, self.config.builtin_rules)
        self._ids: dict[str, int] = {}
        self._msgs: dict[int, Message] = {}

    def start(self):
        return None

    def stop(self):
        return None

    def create_queue(self, name: str, max_size: int = 0) -> None:
        if name in self._ids:
            raise ValueError("queue already exists")   # errors.New("queue already exists"), :175
        if max_size <= 0:
            max_size = self.config.default_max_size
        if name in LEVELS:
            self._ids[name] = LEVELS[name]
        else:
            self._ids[name] = self.engine.queue_create(max_size)

    def delete_queue(self, name: str) -> None:
        qid = self._ids.pop(name, None)
        if qid is None:
            raise ErrQueueNotFound
        if qid > 4:
            self.engine.queue_delete(qid)

    def _apply_rules(self, message: Message) -> None:   # :451-466, first match wins
        for rule in self.config.priority_adjust_rules:
            if rule.condition(message):
                message.priority = int(rule.new_priority)
                break

    def push_message(self, queue_name: str, message: Message) -> None:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        self._apply_rules(message)
        try:
            h = self.engine.push(qid, int(message.priority))
        except _capi.TriageError as e:
            _raise(e)
        self._msgs[h] = message

    def batch_push_messages(self, queue_name: str, messages: list[Message]) -> int:
        if queue_name not in self._ids:
            raise ErrQueueNotFound
        ok = 0
        for m in messages:   # :256-266: a failed push is skipped
            try:
                self.push_message(queue_name, m)
                ok += 1
            except QueueError:
                continue
        return ok

    def pop_message(self, queue_name: str) -> Message:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        try:
            h = self.engine.pop(qid)
        except _capi.TriageError as e:
            _raise(e)
        return self._msgs.pop(h)

    def batch_pop_messages(self, queue_name: str, count: int) -> list[Message]:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        # :336-342 pops until the queue is empty and returns what it has: one engine call, one statistics update
        return [self._msgs.pop(int(h)) for h in self.engine.drain_queue(qid, count)]

    def requeue_batch(self, queue_name: str, messages: list[Message]) -> list[bool]:
        """PushMessage(queue_name, m) for messages that are pushed AGAIN (worker.go:229 retries, dead-letter requeue): rules
        applied per message, capacity in the given order, fresh enqueue times - one engine call.  -> accepted per message"""
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        if not messages:
            return []
        for m in messages:
            self._apply_rules(m)
        handles = [self._new_handle() for _ in messages]
        acc = self.engine.requeue(qid, handles, [int(m.priority) for m in messages])
        for h, m, a in zip(handles, messages, acc):
            if a:
                self._msgs[h] = m
        return [bool(a) for a in acc]

    def complete_batch(self, queue_name: str, k: int) -> None:
        qid = self._ids.get(queue_name)
        if qid is not None:
            self.engine.complete_batch(qid, k)

    def fail_batch(self, queue_name: str, k: int) -> None:
        qid = self._ids.get(queue_name)
        if qid is not None:
            self.engine.fail_batch(qid, k)

    def _new_handle(self) -> int:
        self._requeue_handles = getattr(self, "_requeue_handles", 1 << 61) + 1   # disjoint from the engine's own handles
        return self._requeue_handles

    def complete_message(self, queue_name: str, message_id: str = "", processing_time: float = 0.0) -> None:
        qid = self._ids.get(queue_name)
        if qid is not None:
            self.engine.complete(qid)

    def fail_message(self, queue_name: str, message_id: str = "", reason: str = "") -> None:
        qid = self._ids.get(queue_name)
        if qid is not None:
            self.engine.fail(qid)

    def get_queue_stats(self, queue_name: str) -> QueueStats:
        qid = self._ids.get(queue_name)
        if qid is None:
            raise ErrQueueNotFound
        s = self.engine.stats(qid)
        return QueueStats(s["pending"], s["processing"], s["completed"], s["failed"], time.time())

    def get_all_queue_stats(self) -> dict[str, QueueStats]:
        return {n: self.get_queue_stats(n) for n in self._ids}

    # -- the fused device path --------------------------------------------------
    def batch_triage(self, messages: list[Message], preprocessor: Preprocessor, enqueue_ns=None):
        """ProcessMessage then PushMessage(fmt.Sprint(msg.Priority), msg) for every message, in arrival order, in one
        GPU submit.  Needs the four level queues created and a Preprocessor built on this manager's engine.
        Returns the result records (numpy) so callers can see admission per message."""
        if preprocessor.engine is not self.engine:
            raise ValueError("batch_triage needs Preprocessor(engine=manager.engine)")
        for name in LEVELS:
            if name not in self._ids:
                raise ErrQueueNotFound
        if not self._rules_are_device_rules():
            # a custom rule list (arbitrary Python predicates) cannot run in the kernel: same semantics as the reference
            # loop - ProcessMessage for the batch on the device, then PushMessage per message, which applies the rules
            # (queue_manager.go:451-466) exactly as push_message does
            preprocessor.process_messages(messages)
            res = []
            for m in messages:
                try:
                    self.push_message(priority_name(m.priority), m)
                    res.append(True)
                except QueueError:
                    res.append(False)
            return res
        offsets, data, meta = pack_messages(messages, preprocessor.user_priorities)
        if self.engine.pad_byte() >= 0:   # preferred wire format: 16-byte aligned starts + reset pads
            poff, plen, pbytes = self.engine.pack_padded(offsets, data)
            res, base = self.engine.submit_padded(poff, plen, pbytes, meta, enqueue_ns, 0)
        else:
            res, base = self.engine.submit(offsets, data, meta, enqueue_ns, 0)
        now = time.time()
        for i, (m, r) in enumerate(zip(messages, res)):
            apply_result(m, r, now)
            if r["queue_level"] >= 1 and r["queue_level"] <= 4:
                m.priority = int(r["label"])   # rules mutate msg.Priority before the push (:455)
            if r["accepted"]:
                self._msgs[base + i] = m
        return res

    def _rules_are_device_rules(self) -> bool:
        """True when the configured rule list is what the device evaluates: none, or exactly the two built-in rules of
        queue_factory.go:211-233 with builtin_rules set (the engine was told so in __init__)."""
        rules = self.config.priority_adjust_rules
        if not rules:
            return not self.config.builtin_rules
        if not self.config.builtin_rules or len(rules) != 2:
            return False
        return [(r.new_priority, r.description) for r in rules] == [(r.new_priority, r.description) for r in builtin_priority_rules()] \
            and all(getattr(r.condition, "_builtin", False) for r in rules)

    def drain(self, max_items: int) -> list[Message]:
        handles, _ = self.engine.drain(max_items)
        return [self._msgs.pop(int(h)) for h in handles]

    # Go-style aliases
    CreateQueue = create_queue
    DeleteQueue = delete_queue
    PushMessage = push_message
    BatchPushMessages = batch_push_messages
    PopMessage = pop_message
    BatchPopMessages = batch_pop_messages
    CompleteMessage = complete_message
    FailMessage = fail_message
    GetQueueStats = get_queue_stats
    GetAllQueueStats = get_all_queue_stats


def queue_name_for(priority: int) -> str:
    return priority_name(priority)
