"""llm-message-queue_b200 — B200-native batch triage engine for llm-message-queue's hot path
(internal/preprocessor content analysis -> priority label -> internal/priorityqueue dequeue order).

The product is the CUDA library libtriage_b200.so behind include/triage.h; this package is the thin host-side
mirror of the reference's Go API for that path (the Go toolchain is absent from this image; INTEGRATION.md
shows the cgo shim).  Import it as `llm_message_queue_b200` (see the shim module at the repo root)."""
from . import _capi
from ._capi import META_DTYPE, RESULT_DTYPE, Engine, TriageError, build
from .models import Message, Priority, new_message, priority_name
from .preprocessor import Preprocessor, new_preprocessor
from .priorityqueue import (ErrQueueEmpty, ErrQueueFull, ErrQueueNotFound, MultiLevelQueue, PriorityAdjustRule,
                            QueueError, QueueManager, QueueManagerConfig, QueueStats, builtin_priority_rules)

from .consumer import (DeadLetterItem, DeadLetterQueue, DelayedQueue, ErrIndexOutOfRange, ExponentialBackoff,  # noqa: E402
                       FixedBackoff, handle_failures)

__all__ = ["ExponentialBackoff", "FixedBackoff", "DelayedQueue", "DeadLetterQueue", "DeadLetterItem", "ErrIndexOutOfRange", "handle_failures",
           "_capi", "Engine", "TriageError", "build", "META_DTYPE", "RESULT_DTYPE", "Message", "Priority", "new_message",
           "priority_name", "Preprocessor", "new_preprocessor", "MultiLevelQueue", "QueueManager", "QueueManagerConfig",
           "QueueStats", "QueueError", "ErrQueueEmpty", "ErrQueueFull", "ErrQueueNotFound", "PriorityAdjustRule",
           "builtin_priority_rules"]
