/*
 * triage_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C) of llm-message-queue's preprocess -> priority-order
 * path, used only as the checker in tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs.  Nothing under
 * llm-message-queue_b200/ may include, link or call this.
 *
 * Parity pin: the Go reference cannot be built in this image (no Go toolchain,
 * un-vendored modules), so this restatement is pinned against the reference's
 * own known-answer tests (tests/preprocessor_test.go K1..K7, and the queue
 * counter KATs of tests/priorityqueue_test.go) and cross-checked against a
 * second, independent Python restatement (oracle/py_oracle.py).  Dequeue order
 * across >= 2 messages is pinned by NO reference test ("parity unpinned" for
 * that property): it follows internal/priorityqueue/queue.go:22-27 and
 * cmd/queue-manager/main.go:111-124 only.
 *
 * Reference lines restated (relative to /root/reference):
 *   internal/preprocessor/preprocessor.go:24-53   default dictionary
 *   internal/preprocessor/preprocessor.go:56-114  ProcessMessage decision tree
 *   internal/preprocessor/preprocessor.go:117-168 analyzePriority
 *   internal/preprocessor/preprocessor.go:197-249 performContentAnalysis
 *   internal/priorityqueue/queue.go:22-27,92-139  Less / Push / Pop
 *   internal/priorityqueue/queue_manager.go:451-466 applyPriorityRules
 *   internal/priorityqueue/queue_factory.go:211-233 built-in rules
 *   cmd/queue-manager/main.go:111-124             cross-level drain order
 * Third-party arithmetic restated from published semantics: Go 1.21 stdlib
 * (go.mod:3) regexp (RE2 leftmost-first, FindAllString non-overlap, (?i)
 * SimpleFold orbits), strings.Fields/ToLower/Contains/HasSuffix,
 * unicode.IsSpace, unicode/utf8 decoding, container/heap up/down.
 */
#ifndef TRIAGE_ORACLE_H
#define TRIAGE_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

/* pkg/models/message.go:15-22 */
enum { ORC_PRIO_UNSET = 0, ORC_PRIO_REALTIME = 1, ORC_PRIO_HIGH = 2, ORC_PRIO_NORMAL = 3, ORC_PRIO_LOW = 4 };
/* Metadata["priority_reason"], preprocessor.go:72,86,92 */
enum { ORC_REASON_NONE = 0, ORC_REASON_USER_OVERRIDE = 1, ORC_REASON_USER_DEFAULT = 2, ORC_REASON_CONTENT_KEYWORDS = 3 };
/* meta.user_priority encodings: Metadata["user_priority"].(string) lower-cased */
enum { ORC_UP_ABSENT = 0, ORC_UP_UNRECOGNISED = 5 };
enum { ORC_FLAG_VIP = 1 };
enum { ORC_RF_ANALYZED = 1, ORC_RF_TIE = 2, ORC_RF_HAS_ANALYSIS = 4 };
enum { ORC_ADMIT_OK = 0, ORC_ADMIT_NOT_FOUND = 1, ORC_ADMIT_FULL = 2 };

/* Same 16-byte wire layout as triage_meta / triage_result in include/triage.h. */
typedef struct {
    uint8_t preset;        /* msg.Priority on entry (0 unset, 1..4) */
    uint8_t user_priority; /* 0 absent / non-string, 1..4 recognised, 5 unrecognised string */
    uint8_t user_default;  /* userPriorities[UserID]: 0 absent, 1..4 */
    uint8_t meta_priority; /* Metadata["priority"] string: 0 absent/unrecognised, 1..4 */
    uint8_t flags;         /* bit0: Metadata has key "vip_user" */
    uint8_t reserved[11];
} orc_meta;

typedef struct {
    uint8_t label;       /* msg.Priority after the path (post-rule when rules are on) */
    uint8_t reason;      /* ORC_REASON_* */
    int8_t  sentiment;   /* -1 negative, 0 neutral, +1 positive (valid iff HAS_ANALYSIS) */
    uint8_t question;    /* contains_question */
    uint8_t accepted;    /* admitted to its level queue */
    uint8_t admit;       /* ORC_ADMIT_* */
    uint8_t flags;       /* ORC_RF_* */
    uint8_t queue_level; /* level queue the message was pushed to (pre-rule label) */
    uint32_t words;      /* word_count */
    uint16_t score_rt;   /* saturating keyword scores, diagnostics */
    uint16_t score_hi;
} orc_result;

typedef struct orc_dict orc_dict;

orc_dict *orc_dict_new(int with_defaults);
/* literal pattern; ci = 1 for a "(?i)" literal.  returns 0 ok, 4 unsupported */
int  orc_dict_add(orc_dict *d, int prio, const uint8_t *lit, uint32_t len, int ci);
void orc_dict_free(orc_dict *d);
/* on = 1: one-pass scoring for border-free case-sensitive literals (same numbers as the per-pattern passes; see
 * triage_oracle.c "fast mode"); call after the last orc_dict_add.  The timed CPU baseline never uses it.        */
void orc_dict_set_fast(orc_dict *d, int on);
uint32_t orc_dict_size(const orc_dict *d);
/* AddKeywordPattern(prio, pattern) with a regular expression (regex_oracle.c: RE2 subset, leftmost-first,
 * non-overlapping count on runes).  0 ok, 4 outside the restated subset. */
typedef struct orc_regex orc_regex;
orc_regex *orc_regex_compile(const uint8_t *pattern, uint32_t len);
void orc_regex_free(orc_regex *re);
uint32_t orc_regex_count(const orc_regex *re, const uint8_t *s, uint64_t n);
int orc_dict_add_regex(orc_dict *d, int prio, const uint8_t *pattern, uint32_t len);

/* a5: per-priority scores for one content; scores[1..4] */
void orc_scores(const orc_dict *d, const uint8_t *s, uint64_t n, uint32_t scores[5]);
/* a6 */
void orc_analyze(const uint8_t *s, uint64_t n, uint32_t *words, int32_t *pos, int32_t *neg, int *question);
/* a4 o a5 o a6 for a batch, nthreads >= 1 (threads shard messages) */
void orc_process_batch(const orc_dict *d, int default_prio, const uint64_t *offsets,
                       const uint8_t *bytes, const orc_meta *meta, uint32_t n,
                       orc_result *out, int nthreads);
/* a9-a12: rules (optional), admission in arrival order, container/heap push, then
 * drain realtime->high->normal->low.  ts must be the enqueue timestamps.
 * max_size <= 0 means unbounded (queue.go:101).  Returns number admitted and
 * fills order_out[0..ret) with message indices in dequeue order.             */
uint32_t orc_admit_and_order(orc_result *res, const orc_meta *meta, const uint64_t *offsets,
                             const uint64_t *ts, uint32_t n, int rules_on, int64_t max_size,
                             uint32_t *order_out);
/* one-call convenience used by the timed CPU baseline */
uint32_t orc_pipeline(const orc_dict *d, int default_prio, const uint64_t *offsets,
                      const uint8_t *bytes, const orc_meta *meta, const uint64_t *ts, uint32_t n,
                      int rules_on, int64_t max_size, orc_result *out, uint32_t *order_out,
                      int nthreads);

/* worker.go:274-290 */
int64_t orc_next_backoff_ns(int64_t initial_ns, int64_t max_ns, double factor, int retry_count);

/* container/heap exposed for the queue KATs */
typedef struct orc_heap orc_heap;
orc_heap *orc_heap_new(void);
void orc_heap_free(orc_heap *h);
void orc_heap_push(orc_heap *h, int prio, uint64_t ts, uint32_t id);
int  orc_heap_pop(orc_heap *h, uint32_t *id);
uint32_t orc_heap_len(const orc_heap *h);

/* deterministic synthetic workloads (oracle/gen.c) */
typedef struct {
    uint64_t seed;
    uint32_t n;
    uint32_t len_lo, len_hi;      /* primary length range */
    uint32_t mix;                 /* 0 uniform, 1 C3 mix, 2 C5 skew (90% low bulk + bursts) */
    uint32_t edge_permille_x10;   /* edge-slice rate in 1/10000 (default 50 = 0.5%) */
    uint32_t plain;               /* 1: no meta variation, no edge slice (bench-friendly) */
} orc_gen_cfg;
/* two-call protocol: bytes == NULL returns the total byte size needed */
uint64_t orc_generate(const orc_gen_cfg *cfg, uint64_t *offsets, uint8_t *bytes, orc_meta *meta,
                      uint64_t *ts);

#ifdef __cplusplus
}
#endif
#endif
