"""The reference's own black-box tests, restated against the host mirror (same names, same assertions):
/root/reference/tests/preprocessor_test.go and tests/priorityqueue_test.go (TestMultiLevelQueue, TestQueueManager).
Runs on the B200 box: the mirror computes through the C ABI."""
import numpy as np
import pytest

import llm_message_queue_b200 as mq
from llm_message_queue_b200 import (ErrQueueEmpty, ErrQueueFull, ErrQueueNotFound, MultiLevelQueue, Preprocessor, Priority,
                                    QueueError, QueueManager, QueueManagerConfig, new_message)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def proc():
    return Preprocessor(cfg=dict(DefaultMaxSize=1000, EnableMetrics=True))


# ---- tests/preprocessor_test.go ------------------------------------------------------------
def test_content_based_priority(proc):          # :25-43
    msg = new_message("conv1", "user1", "This is an urgent request that needs attention", Priority.NORMAL)
    out = proc.process_message(msg)
    assert out is msg and out.priority == Priority.HIGH
    assert out.metadata["priority_reason"] == "content_keywords"
    assert out.queue_name == "high" and out.metadata["word_count"] == 8


def test_user_priority_override(proc):          # :46-65
    msg = new_message("conv2", "user2", "Regular message content", Priority.NORMAL)
    msg.metadata = {"user_priority": "high"}
    out = proc.process_message(msg)
    assert out.priority == Priority.HIGH and out.metadata["priority_reason"] == "user_override"


def test_respect_explicit_priority(proc):       # :68-79
    msg = new_message("conv3", "user3", "This is a regular message", Priority.REALTIME)
    out = proc.process_message(msg)
    assert out.priority == Priority.REALTIME and "analyzed" not in out.metadata and out.queue_name == ""


def test_metadata_based_priority(proc):         # :82-105
    msg = new_message("conv4", "user4", "Regular message content", Priority.NORMAL)
    msg.metadata = {"conversation_type": "support", "customer_tier": "premium"}
    out = proc.process_message(msg)
    assert out.metadata["conversation_type"] == "support" and out.priority == Priority.NORMAL and out.metadata["analyzed"] is True


def test_custom_keyword_patterns(proc):         # :108-124
    msg = new_message("conv5", "user5", "I need this right now, it's immediate", Priority.NORMAL)
    out = proc.process_message(msg)
    assert out.priority == Priority.REALTIME and out.metadata["priority_reason"] == "content_keywords"


def test_content_analysis(proc):                # :127-149
    m1 = proc.process_message(new_message("conv6", "user6", "Can you help me with this problem?", Priority.NORMAL))
    assert m1.metadata["contains_question"] == "true"
    m2 = proc.process_message(new_message("conv7", "user7", "I'm very frustrated and angry with this service!", Priority.NORMAL))
    assert m2.metadata["sentiment"] == "negative" and m2.metadata["word_count"] == 8
    assert proc.analyze_message_content("How are you") == {"word_count": 3, "sentiment": "neutral", "is_question": True}
    assert proc.analyze_message_content("") == {"word_count": 0, "sentiment": "neutral", "is_question": False}


def test_setters_and_added_patterns():
    p = Preprocessor()
    p.set_user_priority("vipuser", Priority.REALTIME)
    m = p.process_message(new_message("c", "vipuser", "nothing special", Priority.NORMAL))
    assert m.priority == Priority.REALTIME and m.metadata["priority_reason"] == "user_default"
    assert p.add_keyword_pattern(Priority.LOW, "newsletter") is None          # case-sensitive, as regexp.Compile(pattern) is
    assert p.add_keyword_pattern(Priority.LOW, "(?i)unsubscribe") is None
    assert isinstance(p.add_keyword_pattern(Priority.LOW, "^dear"), ValueError)  # the stated deviation: outside the regex subset (anchors)
    assert isinstance(p.add_keyword_pattern(Priority.LOW, "x*"), ValueError)     # ... and patterns that match the empty string
    assert p.add_keyword_pattern(Priority.LOW, "haha") is None                  # self-overlapping literal: counted without overlap
    assert p.get_keyword_patterns(Priority.LOW) == ["newsletter", "(?i)unsubscribe", "haha"] and len(p.get_keyword_patterns(Priority.HIGH)) == 5
    m = p.process_message(new_message("c", "u", "weekly newsletter, UNSUBSCRIBE here. Newsletter", Priority.NORMAL))
    assert m.priority == Priority.LOW and m.queue_name == "low"
    m = p.process_message(new_message("c", "u", "urgent urgent hahahahaha", Priority.NORMAL))   # haha x2 (not 4) ties with urgent x2
    assert m.priority == Priority.HIGH
    assert p.add_keyword_pattern(Priority.REALTIME, "(?i)p[0-9]+ (incident|outage)s?") is None   # a real regexp (RE2 subset)
    m = p.process_message(new_message("c", "u", "weekly newsletter about the P1 Incident and the p23 outages", Priority.NORMAL))
    assert m.priority == Priority.REALTIME and m.metadata["priority_reason"] == "content_keywords"
    p.set_default_priority(Priority.LOW)
    m = p.process_message(new_message("c", "u", "plain text", 0))
    assert m.priority == Priority.LOW and m.metadata["priority_reason"] == "content_keywords"


# ---- tests/priorityqueue_test.go: TestMultiLevelQueue ------------------------------------------
def test_multi_level_queue_push_pop_peek_stats():
    q = MultiLevelQueue(100)
    for name in ("realtime", "high", "normal", "low"):          # :17-21
        q.add_queue(name)
    msgs = {name: new_message("conv", "user", f"{name} priority message", pr)
            for name, pr in (("low", 4), ("normal", 3), ("high", 2), ("realtime", 1))}
    for name, m in msgs.items():                                 # :27-60
        q.push(name, m, int(m.priority))
    for name in ("realtime", "high", "normal", "low"):          # :63-86 pops give back the same object
        assert q.pop(name) is msgs[name]
    with pytest.raises(QueueError) as ei:                        # :94-97 identity of the singleton
        q.pop("realtime")
    assert ei.value is ErrQueueEmpty
    with pytest.raises(QueueError) as ei:
        q.push("nope", msgs["low"], 4)
    assert ei.value is ErrQueueNotFound
    q.push("normal", msgs["normal"], 3)                           # :101-132 Peek is non-destructive
    assert q.peek("normal") is msgs["normal"] and q.size("normal") == 1
    assert q.pop("normal") is msgs["normal"]
    for i in range(5):                                            # :135-165
        q.push("high", new_message("c", "u", f"h{i}", 2), 2)
    for i in range(3):
        q.push("low", new_message("c", "u", f"l{i}", 4), 4)
    assert q.get_stats("high").pending_count == 5 and q.get_stats("low").pending_count == 3
    for _ in range(5):                                            # :192-237
        q.pop("high")
    for _ in range(3):
        q.complete_message("high")
    for _ in range(2):
        q.fail_message("high")
    s = q.get_stats("high")
    assert (s.completed_count, s.failed_count, s.processing_count, s.pending_count) == (3, 2, 1, 0)   # 1: the pop at :63-86
    assert set(q.get_all_stats()) == {"realtime", "high", "normal", "low"}


def test_queue_orders_by_priority_then_time_and_capacity():
    q = MultiLevelQueue(4)
    q.add_queue("work")
    a, b, c, d = (new_message("c", "u", x, p) for x, p in (("a", 3), ("b", 1), ("c", 3), ("d", 2)))
    for m in (a, b, c, d):
        q.push("work", m, int(m.priority))
    with pytest.raises(QueueError) as ei:                        # queue.go:101-103
        q.push("work", a, 1)
    assert ei.value is ErrQueueFull
    assert [q.pop("work") for _ in range(4)] == [b, d, a, c]      # queue.go:22-27


# ---- tests/priorityqueue_test.go: TestQueueManager -----------------------------------------------
def test_queue_manager_batch_push_pop_and_counters():
    qm = QueueManager(QueueManagerConfig(default_max_size=100))
    qm.create_queue("test-queue", 100)                            # :268
    with pytest.raises(ValueError):
        qm.create_queue("test-queue", 100)
    msgs = [new_message("conv", "user", f"Batch message {i}", Priority.NORMAL) for i in range(5)]
    assert qm.batch_push_messages("test-queue", msgs) == 5        # :296-303
    got = qm.batch_pop_messages("test-queue", 3)                  # :306-317
    assert got == msgs[:3]
    assert qm.batch_pop_messages("test-queue", 10) == msgs[3:]
    qm.complete_message("test-queue", got[0].id, 0.1)             # :327-362
    qm.fail_message("test-queue", got[1].id, "x")
    s = qm.get_queue_stats("test-queue")
    assert (s.completed_count, s.failed_count, s.processing_count) == (1, 1, 3)
    with pytest.raises(QueueError) as ei:
        qm.push_message("missing", msgs[0])
    assert ei.value is ErrQueueNotFound
    qm.delete_queue("test-queue")
    with pytest.raises(QueueError):
        qm.pop_message("test-queue")


def test_batch_triage_is_process_then_push_for_the_whole_batch():
    qm = QueueManager(QueueManagerConfig(default_max_size=0))
    for name in ("realtime", "high", "normal", "low"):
        qm.create_queue(name)
    p = Preprocessor(engine=qm.engine)
    texts = ["nothing to see", "this is urgent", "do it right now", "urgent and critical, asap", "bulk", "what is this?"]
    msgs = [new_message("c", f"u{i}", t, Priority.NORMAL) for i, t in enumerate(texts)]
    msgs[4].priority = Priority.LOW
    res = qm.batch_triage(msgs, p)
    assert [int(m.priority) for m in msgs] == [3, 2, 1, 2, 4, 3]
    assert bool(res[3]["flags"] & 2) is False and int(res[3]["score_hi"]) == 2 and int(res[3]["score_rt"]) == 1
    assert msgs[5].metadata["contains_question"] == "true" and "analyzed" not in (msgs[4].metadata or {})
    assert qm.drain(10) == [msgs[2], msgs[1], msgs[3], msgs[0], msgs[5], msgs[4]]
    assert qm.drain(1) == []


def test_batch_triage_with_rules_sets_queue_name_from_the_pre_rule_priority():
    """queue_manager.go:451-466 mutates msg.Priority inside PushMessage, after ProcessMessage has already named the queue
    from the pre-rule priority (preprocessor.go:103-105): a VIP message about something "right now" sits in queue
    "realtime" with Priority High."""
    qm = QueueManager(QueueManagerConfig(default_max_size=0, builtin_rules=True))
    for name in ("realtime", "high", "normal", "low"):
        qm.create_queue(name)
    p = Preprocessor(engine=qm.engine)
    vip = new_message("c", "u", "do it right now", Priority.NORMAL)
    vip.metadata = {"vip_user": True}
    long_msg = new_message("c", "u", "urgent " + "z" * 10001, Priority.NORMAL)
    plain = new_message("c", "u", "nothing", Priority.NORMAL)
    qm.batch_triage([vip, long_msg, plain], p)
    assert (int(vip.priority), vip.queue_name) == (2, "realtime")
    assert (int(long_msg.priority), long_msg.queue_name) == (4, "high")
    assert (int(plain.priority), plain.queue_name) == (3, "normal")
    assert qm.drain(5) == [vip, long_msg, plain]


def test_batch_triage_with_a_custom_rule_list_applies_it_like_push_message():
    """A custom PriorityAdjustRule list cannot run in the kernel; batch_triage then processes the batch on the device and
    pushes message by message, so the rules act exactly as in push_message (first match wins, queue_manager.go:451-466)."""
    from llm_message_queue_b200 import PriorityAdjustRule, builtin_priority_rules
    rules = [PriorityAdjustRule(lambda m: "escalate" in (m.metadata or {}), 1, "escalate")] + builtin_priority_rules()
    qm = QueueManager(QueueManagerConfig(default_max_size=0, priority_adjust_rules=rules, builtin_rules=True))
    for name in ("realtime", "high", "normal", "low"):
        qm.create_queue(name)
    p = Preprocessor(engine=qm.engine)
    a = new_message("c", "u", "nothing special", Priority.NORMAL)
    a.metadata = {"escalate": 1, "vip_user": 1}
    b = new_message("c", "u", "nothing special either", Priority.NORMAL)
    b.metadata = {"vip_user": 1}
    c = new_message("c", "u", "this is urgent", Priority.NORMAL)
    qm.batch_triage([a, b, c], p)
    assert [int(m.priority) for m in (a, b, c)] == [1, 2, 2]
    assert [m.queue_name for m in (a, b, c)] == ["normal", "normal", "high"]
    via_push = QueueManager(QueueManagerConfig(default_max_size=0, priority_adjust_rules=rules, builtin_rules=True))
    for name in ("realtime", "high", "normal", "low"):
        via_push.create_queue(name)
    a2 = new_message("c", "u", "nothing special", Priority.NORMAL)
    a2.metadata = {"escalate": 1, "vip_user": 1}
    via_push.push_message("normal", a2)
    assert int(a2.priority) == int(a.priority)
    assert qm.drain(5) == [c, a, b]      # queue high first; inside "normal" the item priorities 1 < 2 decide
