"""Host-side mirror of internal/preprocessor (reference: preprocessor.go).  Same method names, argument
meaning and error behaviour; the arithmetic runs on the B200 through the C ABI (no CPU path here).

  NewPreprocessor :24   ProcessMessage :56   AnalyzeMessageContent :253   SetUserPriority :171
  AddKeywordPattern :176   SetDefaultPriority :187   GetKeywordPatterns :192
Additive: process_messages(list) — the batch entry the GPU wants."""
from __future__ import annotations

import time

import numpy as np

from . import _capi
from .models import PRIORITY_BY_NAME, Message, priority_name

_GO_LOWER = {c: c + 32 for c in range(ord("A"), ord("Z") + 1)}
_GO_LOWER[0x130] = ord("i")    # unicode.ToLower(U+0130) = 'i'
_GO_LOWER[0x212A] = ord("k")   # unicode.ToLower(U+212A) = 'k'
_REASONS = {1: "user_override", 2: "user_default", 3: "content_keywords"}
_SENTIMENT = {1: "positive", 0: "neutral", -1: "negative"}


def go_lower_ascii_view(s: str) -> str:
    """strings.ToLower as far as a comparison with an ASCII word can tell."""
    return s.translate(_GO_LOWER)


def meta_record(msg: Message, user_priorities: dict, out) -> None:
    """models.Message -> the 16-byte triage_meta the engine reads (what ProcessMessage looks at)."""
    md = msg.metadata or {}
    out["preset"] = int(msg.priority) & 0xFF
    up = md.get("user_priority")
    if isinstance(up, str):  # .(string) type assertion, preprocessor.go:68
        out["user_priority"] = PRIORITY_BY_NAME.get(go_lower_ascii_view(up), _capi.UP_UNRECOGNISED)
    else:
        out["user_priority"] = 0
    out["user_default"] = int(user_priorities.get(msg.user_id, 0))
    mp = md.get("priority")
    out["meta_priority"] = PRIORITY_BY_NAME.get(go_lower_ascii_view(mp), 0) if isinstance(mp, str) else 0
    out["flags"] = _capi.META_VIP if "vip_user" in md else 0


def pack_messages(msgs, user_priorities):
    contents = [m.content_bytes() for m in msgs]
    offsets, data = _capi.pack_contents(contents)
    meta = np.zeros(len(msgs), dtype=_capi.META_DTYPE)
    for i, m in enumerate(msgs):
        meta_record(m, user_priorities, meta[i])
    return offsets, data, meta


def apply_result(msg: Message, r, now: float) -> Message:
    """Writes back what ProcessMessage writes (preprocessor.go:58-113)."""
    if msg.metadata is None:
        msg.metadata = {}
    if not (r["flags"] & _capi.RF_ANALYZED):   # :63-65 explicit priority respected, nothing else touched
        return msg
    # ProcessMessage sets Priority and QueueName from the PRE-rule priority (:103-105); the rules of PushMessage run
    # later and only change Priority.  On the submit path r["label"] is post-rule and r["queue_level"] pre-rule.
    pre_rule = int(r["queue_level"]) if 1 <= int(r["queue_level"]) <= 4 else int(r["label"])
    msg.priority = int(r["label"])
    if r["reason"]:
        msg.metadata["priority_reason"] = _REASONS[int(r["reason"])]
    if r["flags"] & _capi.RF_HAS_ANALYSIS:
        msg.metadata["word_count"] = int(r["words"])
        msg.metadata["sentiment"] = _SENTIMENT[int(r["sentiment"])]
        msg.metadata["contains_question"] = "true" if r["question"] else "false"
    msg.metadata["analyzed"] = True
    if msg.queue_name == "":
        msg.queue_name = priority_name(pre_rule)
    if not msg.created_at:
        msg.created_at = now
    msg.updated_at = now
    return msg


class Preprocessor:
    def __init__(self, cfg=None, engine: _capi.Engine | None = None, device: int = 0):
        # cfg (config.QueueConfig) is accepted and ignored exactly like the reference does (:24)
        self.cfg = cfg
        self.engine = engine if engine is not None else _capi.Engine(device=device)
        self.user_priorities: dict[str, int] = {}
        self._patterns: dict[int, list[str]] = {
            1: ["(?i)" + w for w in ("immediate", "emergency", "asap", "right now")],
            2: ["(?i)" + w for w in ("urgent", "important", "priority", "critical", "soon")],
        }
        self.engine.dict_reset(True)

    # -- reference API ------------------------------------------------------
    def process_message(self, msg: Message) -> Message:
        return self.process_messages([msg])[0]

    def process_messages(self, msgs: list[Message]) -> list[Message]:
        if not msgs:
            return msgs
        offsets, data, meta = pack_messages(msgs, self.user_priorities)
        res = self.engine.process(offsets, data, meta)
        now = time.time()
        for m, r in zip(msgs, res):
            apply_result(m, r, now)
        return msgs

    def analyze_message_content(self, content) -> dict:
        """AnalyzeMessageContent (:253-299): same analysis, key is_question (bool), computed even for ""."""
        b = content if isinstance(content, (bytes, bytearray)) else content.encode("utf-8")
        if len(b) == 0:
            return {"word_count": 0, "sentiment": "neutral", "is_question": False}
        offsets, data = _capi.pack_contents([bytes(b)])
        meta = np.zeros(1, dtype=_capi.META_DTYPE)
        meta["preset"] = 3
        r = self.engine.process(offsets, data, meta)[0]
        return {"word_count": int(r["words"]), "sentiment": _SENTIMENT[int(r["sentiment"])], "is_question": bool(r["question"])}

    def set_user_priority(self, user_id: str, priority: int) -> None:
        self.user_priorities[user_id] = int(priority)

    def add_keyword_pattern(self, priority: int, pattern: str):
        """Returns None on success or an error (like Go's `error`): anything that is not an optionally
        (?i)-prefixed literal is reported instead of compiled — the one stated API deviation."""
        try:
            self.engine.add_pattern(int(priority), pattern)
        except _capi.TriageError as e:
            if e.status == _capi.E_PATTERN:
                return ValueError(e.message)
            raise
        self._patterns.setdefault(int(priority), []).append(pattern)
        return None

    def set_default_priority(self, priority: int) -> None:
        self.engine.set_default_priority(int(priority))

    def get_keyword_patterns(self, priority: int) -> list[str]:
        return list(self._patterns.get(int(priority), []))

    # Go-style aliases
    ProcessMessage = process_message
    AnalyzeMessageContent = analyze_message_content
    SetUserPriority = set_user_priority
    AddKeywordPattern = add_keyword_pattern
    SetDefaultPriority = set_default_priority
    GetKeywordPatterns = get_keyword_patterns


def new_preprocessor(cfg=None, **kw) -> Preprocessor:
    return Preprocessor(cfg, **kw)
