"""Mirror of pkg/models/message.go (reference lines :15-37 Priority, :58-91 Message / NewMessage)."""
from __future__ import annotations

import time
import uuid
from dataclasses import dataclass, field
from enum import IntEnum
from typing import Any, Optional


class Priority(IntEnum):
    UNSET = 0
    REALTIME = 1
    HIGH = 2
    NORMAL = 3
    LOW = 4

    def __str__(self) -> str:  # Priority.String(), message.go:24-37
        return _NAMES.get(int(self), "unknown")


_NAMES = {1: "realtime", 2: "high", 3: "normal", 4: "low"}
PRIORITY_BY_NAME = {v: k for k, v in _NAMES.items()}


def priority_name(p: int) -> str:
    return _NAMES.get(int(p), "unknown")


@dataclass
class Message:
    id: str = ""
    conversation_id: str = ""
    user_id: str = ""
    content: str | bytes = ""
    priority: int = 0
    status: str = "pending"
    queue_name: str = ""
    retry_count: int = 0
    max_retries: int = 3
    timeout: float = 30.0
    created_at: float = 0.0     # 0.0 plays time.Time{}.IsZero()
    updated_at: float = 0.0
    scheduled_at: Optional[float] = None
    completed_at: Optional[float] = None
    metadata: Optional[dict[str, Any]] = field(default=None)

    def content_bytes(self) -> bytes:
        # Go strings are byte strings; str is encoded as UTF-8, bytes pass through (may be invalid UTF-8)
        return self.content if isinstance(self.content, (bytes, bytearray)) else self.content.encode("utf-8")


def new_message(conversation_id: str, user_id: str, content, priority: int) -> Message:
    """models.NewMessage (message.go:76-91)"""
    now = time.time()
    return Message(id=str(uuid.uuid4()), conversation_id=conversation_id, user_id=user_id, content=content,
                   priority=int(priority), status="pending", retry_count=0, max_retries=3, timeout=30.0,
                   created_at=now, updated_at=now, metadata={})
