/*
 * triage.h — C ABI of the B200 batch triage engine (libtriage_b200.so).
 *
 * This is the drop-in boundary for llm-message-queue's one data-parallel path:
 * internal/preprocessor content analysis -> priority label -> internal/priorityqueue
 * dequeue order.  The reference has no FFI of its own (it is pure Go, CGO_ENABLED=0,
 * deployments/Dockerfile:15-17); every entry point below cites the exported Go method
 * it stands in for (paths relative to the reference repo).  A thin cgo shim keeps the
 * Go signatures; plain pointers and sizes only, no torch types.
 *
 * All functions return 0 on success or a TRIAGE_E_* status; triage_last_error() gives a
 * thread-local message for the last failure.  There is no CPU fallback: every compute
 * entry point fails with TRIAGE_E_CUDA when no sm_100 device / kernel image is usable.
 */
#ifndef TRIAGE_B200_H
#define TRIAGE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* pkg/models/message.go:15-22 */
#define TRIAGE_PRIO_UNSET    0
#define TRIAGE_PRIO_REALTIME 1
#define TRIAGE_PRIO_HIGH     2
#define TRIAGE_PRIO_NORMAL   3
#define TRIAGE_PRIO_LOW      4

/* status codes; 1..3 map to the reference's error singletons (internal/priorityqueue/queue.go:213-217) */
#define TRIAGE_OK            0
#define TRIAGE_E_NOT_FOUND   1   /* ErrQueueNotFound */
#define TRIAGE_E_FULL        2   /* ErrQueueFull     */
#define TRIAGE_E_EMPTY       3   /* ErrQueueEmpty    */
#define TRIAGE_E_PATTERN     4   /* AddKeywordPattern: pattern outside the supported regexp subset */
#define TRIAGE_E_ARG         5
#define TRIAGE_E_CUDA        100 /* CUDA / NCCL failure, see triage_last_error() */

/* Metadata["priority_reason"], internal/preprocessor/preprocessor.go:72,86,92 */
#define TRIAGE_REASON_NONE             0
#define TRIAGE_REASON_USER_OVERRIDE    1
#define TRIAGE_REASON_USER_DEFAULT     2
#define TRIAGE_REASON_CONTENT_KEYWORDS 3

#define TRIAGE_UP_ABSENT        0   /* Metadata["user_priority"] absent or not a string  */
#define TRIAGE_UP_UNRECOGNISED  5   /* a string that is none of realtime/high/normal/low */
#define TRIAGE_META_VIP         1   /* Metadata has key "vip_user" (queue_factory.go:217) */
#define TRIAGE_META_ANALYZE     2   /* fill words / sentiment / question even where ProcessMessage returns early
                                       (api/handlers.go:180-190 analyses every submitted message); no Metadata is implied */

#define TRIAGE_RF_ANALYZED      1   /* Metadata["analyzed"] = true (preprocessor.go:100)  */
#define TRIAGE_RF_TIE           2   /* equal non-zero top scores: reference is random here */
#define TRIAGE_RF_HAS_ANALYSIS  4   /* word_count/sentiment/contains_question were written */

/* The fields of models.Message (pkg/models/message.go:58-74) the path reads, 16 bytes. */
typedef struct {
    uint8_t preset;        /* msg.Priority on entry */
    uint8_t user_priority; /* Metadata["user_priority"].(string) lower-cased: 1..4, or TRIAGE_UP_* */
    uint8_t user_default;  /* userPriorities[msg.UserID] resolved by the shim: 0 absent, 1..4  */
    uint8_t meta_priority; /* Metadata["priority"].(string): 0 absent/unrecognised, 1..4       */
    uint8_t flags;         /* TRIAGE_META_* */
    uint8_t reserved[11];
} triage_meta;

/* What ProcessMessage + PushMessage write back, 16 bytes. */
typedef struct {
    uint8_t label;       /* msg.Priority after the path (post-rule when rules are on) */
    uint8_t reason;      /* TRIAGE_REASON_* */
    int8_t  sentiment;   /* -1 negative / 0 neutral / +1 positive */
    uint8_t question;    /* contains_question */
    uint8_t accepted;    /* 1 = admitted to its level queue */
    uint8_t admit;       /* TRIAGE_OK / TRIAGE_E_NOT_FOUND / TRIAGE_E_FULL for this message */
    uint8_t flags;       /* TRIAGE_RF_* */
    uint8_t queue_level; /* level queue it was pushed to: fmt.Sprint(pre-rule Priority) (api/handlers.go:202) */
    uint32_t words;      /* word_count */
    uint16_t score_rt;   /* saturating realtime / high keyword scores (diagnostics) */
    uint16_t score_hi;
} triage_result;

/* internal/priorityqueue/queue.go:60-68 (counters only) */
typedef struct {
    int64_t pending, processing, completed, failed;
} triage_qstats;

typedef struct {
    int32_t  device;          /* CUDA ordinal */
    int32_t  rules_enabled;   /* 1 = PriorityQueueType rules (queue_factory.go:202-204, 211-233) */
    int64_t  max_queue_size;  /* per level queue; <= 0 unbounded (queue.go:101) */
    uint64_t max_batch_bytes; /* capacity hints; buffers grow on demand when 0 */
    uint32_t max_batch_msgs;
    uint32_t flags;           /* reserved, 0 */
} triage_opts;

typedef struct triage_engine triage_engine;

/* -- lifecycle.  Replaces preprocessor.NewPreprocessor (preprocessor.go:24) +
 *    priorityqueue.NewMultiLevelQueue / AddQueue x4 (queue.go:70-90, cmd/queue-manager/main.go:58-63). */
int32_t triage_create(const triage_opts *opts, triage_engine **out);
void    triage_destroy(triage_engine *e);
const char *triage_last_error(void);

/* -- dictionary.  Replaces the pattern table built in NewPreprocessor (preprocessor.go:28-43),
 *    AddKeywordPattern (:176-184), SetDefaultPriority (:187) and GetKeywordPatterns (:192). */
int32_t triage_dict_reset(triage_engine *e, int32_t with_defaults);
/* regexp source as AddKeywordPattern receives it (preprocessor.go:176-184 hands it to regexp.Compile).  A plain,
 * optionally "(?i)"-prefixed literal joins the scan automaton.  Anything else is compiled as a regular expression of the
 * supported RE2 subset - leading (?i)/(?s), classes, ".", \d \w \s, groups, alternation, ? * + {n,m} and their lazy
 * forms - and counted on the device with FindAllString's leftmost-first, non-overlapping semantics (csrc/regex_vm.h).
 * Outside the subset (anchors, \b, named groups, POSIX classes, patterns that can match "", non-ASCII under (?i), more
 * than 64 instructions) -> TRIAGE_E_PATTERN with the reason: the stated API deviation, never a silent approximation. */
int32_t triage_dict_add_pattern(triage_engine *e, int32_t prio, const char *regex, uint32_t len);
int32_t triage_dict_add_literal(triage_engine *e, int32_t prio, const uint8_t *lit, uint32_t len, int32_t ci);
int32_t triage_dict_commit(triage_engine *e);            /* compile + upload the automaton */
int32_t triage_dict_count(triage_engine *e, int32_t prio, uint32_t *n);
int32_t triage_set_default_priority(triage_engine *e, int32_t prio);
/* automaton facts for tests / DESIGN.md: states, warm-up bytes, table bytes */
int32_t triage_dict_info(triage_engine *e, uint32_t *n_states, uint32_t *warm_bytes, uint32_t *table_bytes);

/* -- pinned host arenas for the shim's ring buffers */
int32_t triage_host_alloc(void **p, uint64_t bytes);
int32_t triage_host_free(void *p);

/* -- the hot path, host buffers.  One call = for every message i in arrival order:
 *    ProcessMessage (preprocessor.go:56-114) then QueueManager.PushMessage(fmt.Sprint(Priority), msg)
 *    (queue_manager.go:210-243 -> queue.go:92-119), i.e. BatchPushMessages semantics (:246-287):
 *    a rejected message is skipped, not an error.  offsets has n+1 entries; message i is
 *    bytes[offsets[i] .. offsets[i+1]).  enqueue_ns may be NULL: the engine then stamps strictly
 *    increasing times (queue.go:108).  out (n entries) may be NULL.  handle_base receives the
 *    handle of message 0; message i has handle_base + i.                                      */
int32_t triage_submit(triage_engine *e, int32_t target_queue, const uint64_t *offsets, const uint8_t *bytes,
                      const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n,
                      triage_result *out, uint64_t *handle_base);
/* target_queue = 0 routes every message to the level queue named after its label (the pipeline of
 * api/handlers.go:178-202).  target_queue = a queue id pushes the whole batch into that one queue with
 * item priority = int(msg.Priority): QueueManager.BatchPushMessages(queueName, msgs) (queue_manager.go:246-287). */

/* -- padded arena: the engine's preferred wire format.  Message i starts at offsets[i] (a multiple of 16), holds
 *    lengths[i] content bytes and is followed by 1..16 copies of triage_pad_byte() up to offsets[i+1]; offsets[n] is the
 *    padded total.  The pad byte resets the automaton, so the scan never has to split a 16-byte group at a message
 *    boundary.  A shim that copies Content into its arena anyway lays it out like this for free
 *    (triage_pack_padded does it for an existing packed arena).  Same semantics and results as triage_submit.      */
int32_t triage_pad_byte(triage_engine *e, int32_t *pad);      /* -1: this dictionary has none, use the packed calls */
int32_t triage_pack_padded(const uint64_t *src_offsets, const uint8_t *src_bytes, uint32_t n, int32_t pad,
                           uint64_t *dst_offsets, uint32_t *dst_lengths, uint8_t *dst_bytes, uint64_t *dst_total);
int32_t triage_submit_padded(triage_engine *e, int32_t target_queue, const uint64_t *offsets, const uint32_t *lengths,
                             const uint8_t *bytes, const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n,
                             triage_result *out, uint64_t *handle_base);
int32_t triage_process_padded(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                              const triage_meta *meta, uint32_t n, triage_result *out);   /* ProcessMessage only */
/* Ring of two: starts the host->device copy of a batch on the engine's copy stream and returns at once.  A later
 * triage_submit / triage_submit_padded with the SAME host pointers and n finds the batch on the device and only waits
 * for that copy, so the shim can upload batch k+1 while batch k is scanned, ordered, read back and queued (the "pinned
 * host ring buffers, async streams" of the design brief; the reference has no counterpart - its handlers push one
 * message at a time).  lengths = NULL: packed arena.  The host buffers (pinned, triage_host_alloc) must stay unchanged
 * until that submit returns; at most two prefetched batches are kept, a third replaces the older one.  Results are
 * identical to a submit without prefetch.                                                                           */
int32_t triage_prefetch(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                        const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n);
int32_t triage_submit_device_padded(triage_engine *e, const uint64_t *offsets_dev, const uint32_t *lengths_dev,
                                    const uint8_t *bytes_dev, uint64_t total_bytes, const triage_meta *meta_dev,
                                    const uint64_t *enqueue_ns_dev, uint32_t n, triage_result *results_dev,
                                    uint32_t *order_dev, uint32_t *n_admitted);

/* ProcessMessage only (preprocessor.go:56-114) for a batch: labels + analysis, no queue is touched. */
int32_t triage_process(triage_engine *e, const uint64_t *offsets, const uint8_t *bytes,
                       const triage_meta *meta, uint32_t n, triage_result *out);

/* -- same path with every input already resident in device memory (device pointers).  order_dev
 *    (n x u32, device) receives this batch's dequeue permutation of admitted messages, results_dev
 *    (n x triage_result, device) the labels; n_admitted is read back (4 bytes).  Does not touch the
 *    host-side queues.  Used for the device-resident throughput figure and the multi-GPU driver. */
int32_t triage_submit_device(triage_engine *e, const uint64_t *offsets_dev, const uint8_t *bytes_dev,
                             uint64_t total_bytes, const triage_meta *meta_dev,
                             const uint64_t *enqueue_ns_dev, uint32_t n, triage_result *results_dev,
                             uint32_t *order_dev, uint32_t *n_admitted);

/* -- triage_submit_device for callers that may not allocate device memory (no counterpart in the reference: the Go
 *    code has no device and no allocator to avoid).  Same arguments and results as triage_submit_device, but every byte of
 *    kernel scratch comes from one fixed 3 MiB __device__ array inside the library instead of buffers grown with cudaMalloc,
 *    so the call performs no cudaMalloc / cudaFree at all.  For serving loops and embedding applications that own all of
 *    HBM or cannot afford the allocator's implicit synchronisation.  Limits: at most 65536 messages and 256 MiB of arena
 *    per call (larger -> TRIAGE_E_ARG), the dictionary must already be committed with triage_dict_commit, and calls that
 *    overlap on one GPU race on the shared array (serialise them).  The first call on a device looks the array up with
 *    cudaGetSymbolAddress; make it before any stream capture.  It takes the packed arena and runs the packed-arena scan
 *    (scan_kernel), which is about 2.5x slower per byte than the row kernel behind the padded entry points; dictionaries
 *    that need the wide table format are refused (TRIAGE_E_ARG).                                                      */
int32_t triage_submit_device_static(triage_engine *e, const uint64_t *offsets_dev, const uint8_t *bytes_dev,
                                    uint64_t total_bytes, const triage_meta *meta_dev,
                                    const uint64_t *enqueue_ns_dev, uint32_t n, triage_result *results_dev,
                                    uint32_t *order_dev, uint32_t *n_admitted);

/* -- dequeue.  triage_pop = MultiLevelQueue.Pop(name) (queue.go:121-139); triage_drain = the
 *    realtime->high->normal->low loop of cmd/queue-manager/main.go:111-124 run max times, which is
 *    also QueueManager.BatchPopMessages per level (queue_manager.go:326-367).                   */
/* Queue ids: 1..4 are the built-in level queues realtime/high/normal/low.  QueueManager.CreateQueue(name, max)
 * (queue_manager.go:170-188) makes further single-heap queues; ids are never reused. */
int32_t triage_queue_create(triage_engine *e, int64_t max_size, int32_t *queue);
int32_t triage_queue_delete(triage_engine *e, int32_t queue);                    /* DeleteQueue, :191-207 */
int32_t triage_set_limits(triage_engine *e, int64_t max_queue_size, int32_t rules_enabled);
int32_t triage_pop(triage_engine *e, int32_t level, uint64_t *handle);
int32_t triage_peek(triage_engine *e, int32_t level, uint64_t *handle);
int32_t triage_drain(triage_engine *e, uint32_t max, uint64_t *handles, uint8_t *levels, uint32_t *n_out);
/* single push with an explicit item priority: MultiLevelQueue.Push(name, msg, priority) (queue.go:92) */
int32_t triage_push(triage_engine *e, int32_t level, int32_t item_priority, uint64_t enqueue_ns, uint64_t *handle);
int32_t triage_queue_size(triage_engine *e, int32_t level, uint64_t *size);       /* Size, queue.go:153 */
int32_t triage_stats(triage_engine *e, int32_t level, triage_qstats *st);         /* GetStats, :165 */
int32_t triage_complete(triage_engine *e, int32_t level);                         /* CompleteMessage, :197 */
int32_t triage_fail(triage_engine *e, int32_t level);                             /* FailMessage, :205 */

/* -- the wire format on either side of the path, in batches.
 *    triage_ingest_json = c.ShouldBindJSON(&message) (api/handlers.go:160-166) for n request bodies laid end to end
 *    (body i = bodies[body_offsets[i] .. body_offsets[i+1])): content goes straight into the padded arena, the fields of
 *    models.Message the path reads (pkg/models/message.go:58-74: priority, user_id -> the table of
 *    triage_set_user_priority, metadata.user_priority / priority / vip_user) into triage_meta, with TRIAGE_META_ANALYZE set;
 *    every other field is type-checked only.  status[i] = 1 where the handler would answer 400 (bad JSON, or a known
 *    field of the wrong type): that message becomes an empty slot the caller skips.  arena = NULL computes offsets,
 *    lengths, meta, status and *arena_total (the size to allocate) only.  encoding/json rules followed: csrc/ingest.cc.
 *    triage_set_user_priority = Preprocessor.SetUserPriority (preprocessor.go:171); prio 0 removes the entry.
 *    triage_emit_analysis_json = json.Marshal(AnalyzeMessageContent(content)) (api/handlers.go:180-190) for n results:
 *    {"is_question":..,"sentiment":"..","word_count":..} each, message i at out[out_offsets[i] .. out_offsets[i+1]);
 *    *needed = total bytes (call with out = NULL to size the buffer).
 *    Both run on the host; large batches are decoded / written on several threads (TRIAGE_INGEST_THREADS, default up to 16). */
int32_t triage_set_user_priority(triage_engine *e, const char *user_id, uint32_t len, int32_t prio);
int32_t triage_ingest_json(triage_engine *e, const uint8_t *bodies, const uint64_t *body_offsets, uint32_t n, uint8_t *arena,
                           uint64_t arena_cap, uint64_t *offsets, uint32_t *lengths, triage_meta *meta, uint8_t *status,
                           uint64_t *arena_total);
int32_t triage_emit_analysis_json(const triage_result *res, uint32_t n, char *out, uint64_t cap, uint64_t *out_offsets, uint64_t *needed);

/* -- batch consumer side.  triage_drain_queue = QueueManager.BatchPopMessages(queueName, count) (queue_manager.go:326-367):
 *    pops from ONE queue until it is empty or max is reached, under one lock and with one statistics update
 *    (Pending -= k, Processing += k; queue.go:132-136 k times).  triage_complete_batch / triage_fail_batch =
 *    CompleteMessage / FailMessage (queue.go:197-211) for k messages of a Worker.processBatch round (worker.go:128-158). */
int32_t triage_drain_queue(triage_engine *e, int32_t queue, uint32_t max, uint64_t *handles, uint32_t *n_out);
int32_t triage_complete_batch(triage_engine *e, int32_t queue, uint32_t k);
int32_t triage_fail_batch(triage_engine *e, int32_t queue, uint32_t k);
/* Retry of failed messages, Worker.handleFailure (worker.go:208-229), and DeadLetterQueue.Requeue / BatchRequeue
 * (dead_letter_queue.go:187-258): QueueManager.PushMessage(queueName, message) for n messages that already have handles,
 * in the given order - capacity rule per message (queue.go:101-103; accepted[i] = 0 where the queue was full, n_accepted
 * = the count BatchRequeue returns), a fresh enqueue time each (enqueue_ns NULL: stamped strictly increasing, queue.go:108),
 * so a retry goes behind everything of its priority that is already waiting.  item_priority = int(msg.Priority) after the
 * shim applied its rules.  The accepted messages are ordered by (priority, time, arrival) - on the device for batches of
 * 32768 and more (the LSD radix passes of the submit path), on the host below that - and merged into the queue as one run. */
int32_t triage_requeue(triage_engine *e, int32_t queue, const uint64_t *handles, const int32_t *item_priority, const uint64_t *enqueue_ns,
                       uint32_t n, uint8_t *accepted, uint32_t *n_accepted);
/* DelayedQueue (delayed_queue.go): Schedule(message, readyAt) (:98-111) for n messages; the engine keeps them ordered by
 * ReadyAt (Less, :37-39; equal times: arrival).  triage_release_due plays the run loop (:113-170) at time `now` (0 = the
 * wall clock): every item whose ReadyAt has passed goes, earliest first, to processFn = PushMessage into the queue it
 * was scheduled for, stamped at the release; handles/queues (nullable, room for max entries; max = 0: no limit)
 * receive what was released, n_rejected counts pushes a full or deleted queue turned away.  This is where
 * Worker.handleFailure's ScheduledAt (worker.go:220-229) becomes effective - the reference leaves that as a comment.
 * triage_delayed_peek = DelayedQueue.Peek (earliest item; TRIAGE_E_EMPTY when none). */
int32_t triage_schedule(triage_engine *e, int32_t queue, const uint64_t *handles, const int32_t *item_priority, const uint64_t *ready_at_ns,
                        uint32_t n);
int32_t triage_release_due(triage_engine *e, uint64_t now_ns, uint64_t *handles, int32_t *queues, uint32_t max, uint32_t *n_released,
                           uint32_t *n_rejected);
int32_t triage_delayed_peek(triage_engine *e, uint64_t *handle, uint64_t *ready_at_ns);
int32_t triage_delayed_size(triage_engine *e, uint64_t *n);
/* ExponentialBackoff.NextBackoff(retryCount) (worker.go:274-290) in nanoseconds: initial * factor^(retryCount-1) in float64,
 * capped at max, truncated like time.Duration(backoff). */
int64_t triage_backoff_ns(int64_t initial_ns, int64_t max_ns, double factor, int32_t retry_count);

/* -- instrumentation: device times (CUDA events on the engine's stream) of the last submit */
typedef struct {
    float scan_ms, finalize_ms, order_ms, total_ms; /* kernels only */
    float h2d_ms, d2h_ms;                           /* triage_submit only */
    uint32_t launches;                              /* kernels launched by the last submit */
    uint32_t order_passes;                          /* radix digit passes executed */
    uint64_t scan_bytes;                            /* algorithmic bytes of the scan */
    uint32_t scan_variant;                          /* kernel that ran: 1 packed-arena tiles, 2 padded mini-tiles, 3 row lanes (2-D TMA), 4 wide table */
    uint32_t scan_row_bytes;                        /* row length of the row-lane scan (1280 / 640 / 320), else 0 */
    uint32_t scan_warps;                            /* warps per CTA of that kernel */
    uint32_t reserved;
} triage_timings;
int32_t triage_last_timings(triage_engine *e, triage_timings *t);
/* CUDA events on the engine's own stream (torch.cuda.Event cannot see it): slots 0..3 */
int32_t triage_event_record(triage_engine *e, int32_t slot);
int32_t triage_event_elapsed_ms(triage_engine *e, int32_t from_slot, int32_t to_slot, float *ms);

/* -- multi-GPU: one engine per rank/GPU; the caller broadcasts the 128-byte id (rank 0 makes it).  With a communicator the
 *    scan leaves a few SMs to the exchange, which runs next to it (TRIAGE_COMM_SMS: 4, 12 from 8 ranks on), and the NCCL
 *    communicator is created with as few CTAs (TRIAGE_NCCL_CTAS: 4 / 8; 0 = NCCL's own choice). */
int32_t triage_comm_unique_id(uint8_t id[128]);
int32_t triage_comm_init(triage_engine *e, int32_t nranks, int32_t rank, const uint8_t id[128]);
/* all-gather every rank's sorted (key, global index) run of the last triage_submit_device and
 * merge them: order_global_dev (sum of admitted, u32 global indices = rank_base + local index)
 * receives the whole-job dequeue order on every rank.  rank_base[r] = first global index of rank r. */
int32_t triage_merge_allgather(triage_engine *e, const uint64_t *rank_base, uint32_t *order_global_dev,
                               uint64_t capacity, uint64_t *n_total);
/* The same merge started on the engine's merge stream without waiting for it: for contiguous sequence shards with
 * engine-stamped enqueue times (enqueue_ns_dev = NULL at the submit) the exchange is ONE all-gather of [bucket counts |
 * local order] and a concatenation that works out its geometry on the device, so the caller can submit the next batch
 * while the previous order travels.  The next submit waits (on the device) only until its predecessor's order has been
 * read.  triage_merge_wait blocks until the merge is complete and returns the merged count; order_global_dev may be read
 * after it.  Explicit enqueue times need triage_merge_allgather (key exchange).  No counterpart in the reference: its
 * queues live in one process (cmd/queue-manager/main.go:111-124 is the order being reproduced). */
int32_t triage_merge_allgather_async(triage_engine *e, const uint64_t *rank_base, uint32_t *order_global_dev, uint64_t capacity);
int32_t triage_merge_wait(triage_engine *e, uint64_t *n_total);

#ifdef __cplusplus
}
#endif
#endif
