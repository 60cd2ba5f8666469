// scan_core.cuh — the per-lane body of the arena scan, written once for device and host.
//
// The CUDA kernel (scan_kernel.cu) calls lane_scan() from every thread; tests/csrc/ compiles the
// very same function with g++ and walks all lanes of all tiles sequentially to check the
// automaton + chunking logic against the oracle on machines without a GPU.  That host build is
// a test harness only — the shipped library has no CPU path.
//
// Model: the arena is cut into tiles of NT*C bytes; lane l of a tile owns the C bytes starting at
// tile_start + l*C.  A lane first walks `warm` bytes before its chunk (or from the message start,
// whichever is later) with events discarded so that its automaton state is exact at the chunk
// start (the automaton has finite memory `span` <= warm+1), then walks its chunk, splitting at
// message boundaries.  Per-message counters live in global memory and are bumped with atomics.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TR_HD __host__ __device__ __forceinline__
#else
#define TR_HD inline
#endif

namespace triage {

// per-message counters written by the scan, read by the finalize kernel (8 x u32 = 32 bytes)
enum { CNT_WORDS = 0, CNT_KW = 1 /* ..4 */, CNT_POS = 5, CNT_NEG = 6, CNT_Q = 7, CNT_STRIDE = 8 };

// events[id]: x = kw_rt | kw_hi << 16, y = kw_norm | kw_low << 16, z = pos | neg << 8 | question << 16
struct EventPacked { uint32_t x, y, z, w; };

TR_HD void tr_add(uint32_t *p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}
TR_HD void tr_or(uint32_t *p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    atomicOr(p, v);
#else
    *p |= v;
#endif
}
TR_HD uint32_t tr_idx(uint32_t word, uint32_t entry, uint32_t k) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(word, entry, 0x7640u | k);   // (entry & 0xFF) << 8 | byte k of word; entry < 65536
#else
    return ((entry & 0xFFu) << 8) | ((word >> (8 * k)) & 0xFFu);
#endif
}

struct LaneCtx {
    const uint16_t *tbl;        // D8 table (shared memory on device)
    const EventPacked *events;  // event id -> counters
    const uint8_t *eom;         // per state end-of-message output
    const uint8_t *stage;       // staged bytes; stage[i] is the byte at tile-relative position i - warm
    const int32_t *list;        // tile-relative message starts of messages msg_lo .. (clamped to +-2^30), or
    const uint64_t *list_abs;   // ... when the tile holds too many messages for shared memory: the global offsets
    int64_t tile_start;         //     (absolute), made relative on the fly
    uint32_t list_n;            // entries (the last one is the first boundary at or beyond the tile end)
    uint32_t msg_lo;            // global index of list[0]
    uint32_t *counters;         // global, CNT_STRIDE per message
    int32_t warm;
    uint32_t start_state, neutral_state;
};

TR_HD void apply_event(const LaneCtx &c, uint32_t id, uint32_t msg) {
    EventPacked e = c.events[id];
    uint32_t *cnt = c.counters + (uint64_t)msg * CNT_STRIDE;
    if (e.x & 0xFFFFu) tr_add(cnt + CNT_KW + 0, e.x & 0xFFFFu);
    if (e.x >> 16) tr_add(cnt + CNT_KW + 1, e.x >> 16);
    if (e.y & 0xFFFFu) tr_add(cnt + CNT_KW + 2, e.y & 0xFFFFu);
    if (e.y >> 16) tr_add(cnt + CNT_KW + 3, e.y >> 16);
    if (e.z & 0xFFu) tr_add(cnt + CNT_POS, e.z & 0xFFu);
    if ((e.z >> 8) & 0xFFu) tr_add(cnt + CNT_NEG, (e.z >> 8) & 0xFFu);
    if (e.z >> 16) tr_or(cnt + CNT_Q, 1u);
}

// one byte, full event handling
TR_HD void step_slow(const LaneCtx &c, uint32_t byte, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    e = c.tbl[((e & 0xFFu) << 8) | byte];
    if (count) {
        words += e >> 15;
        uint32_t id = (e >> 8) & 0x7Fu;
        if (id) apply_event(c, id, msg);
    }
}

// bytes [a, b) (tile-relative), byte at a time
TR_HD void run_slow(const LaneCtx &c, int32_t a, int32_t b, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    for (int32_t p = a; p < b; p++) step_slow(c, c.stage[p + c.warm], e, words, msg, count);
}

#define TR_STEP4(wd)                                                          \
    idx = tr_idx(wd, e, 0); e = c.tbl[idx]; words += e >> 15; ev |= e;        \
    idx = tr_idx(wd, e, 1); e = c.tbl[idx]; words += e >> 15; ev |= e;        \
    idx = tr_idx(wd, e, 2); e = c.tbl[idx]; words += e >> 15; ev |= e;        \
    idx = tr_idx(wd, e, 3); e = c.tbl[idx]; words += e >> 15; ev |= e;

// one aligned 16-byte group fully inside a segment; rare events are replayed byte by byte
TR_HD void run_group16(const LaneCtx &c, int32_t p, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    struct alignas(16) Quad { uint32_t a, b, c, d; };
    const Quad q = *reinterpret_cast<const Quad *>(c.stage + p + c.warm);   // one 128-bit shared load
    uint32_t w0 = q.a, w1 = q.b, w2 = q.c, w3 = q.d;
    uint32_t e0 = e, words0 = words, ev = 0, idx;
    TR_STEP4(w0) TR_STEP4(w1) TR_STEP4(w2) TR_STEP4(w3)
    if (!count) { words = words0; return; }
    if (ev & 0x7F00u) {   // some transition carried a keyword / sentiment / question event
        e = e0;
        words = words0;
        run_slow(c, p, p + 16, e, words, msg, true);
    }
}

TR_HD void run_range(const LaneCtx &c, int32_t a, int32_t b, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    int32_t p = a;
    int32_t head = (16 - ((p + c.warm) & 15)) & 15;        // stage is 16-byte aligned at index 0
    if (head > b - p) head = b - p;
    if (head) { run_slow(c, p, p + head, e, words, msg, count); p += head; }
    for (; p + 16 <= b; p += 16) run_group16(c, p, e, words, msg, count);
    if (p < b) run_slow(c, p, b, e, words, msg, count);
}

TR_HD int32_t list_at(const LaneCtx &c, uint32_t k) {
    if (c.list) return c.list[k];
    int64_t r = (int64_t)c.list_abs[c.msg_lo + k] - c.tile_start;
    const int64_t lim = 1 << 30;
    return (int32_t)(r < -lim ? -lim : (r > lim ? lim : r));
}

// index k into list with list[k] <= pos < list[k+1]  (upper_bound - 1; empty messages repeat a value,
// so this lands on the one non-empty message that contains pos)
TR_HD uint32_t find_msg(const LaneCtx &c, int32_t pos) {
    uint32_t lo = 0, hi = c.list_n;   // first index with list[idx] > pos
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (list_at(c, mid) > pos) hi = mid; else lo = mid + 1;
    }
    return lo - 1;
}

// Walks the lane's chunk [chunk, chunk_end) (tile-relative).  Every segment is flushed with
// atomics except the partial count of the message at the chunk start, which is returned in
// (*head_msg, *head_words) so the caller can combine neighbouring lanes first.
TR_HD void lane_scan(const LaneCtx &c, int32_t chunk, int32_t chunk_end, uint32_t *head_msg, uint32_t *head_words) {
    uint32_t k = find_msg(c, chunk);
    int32_t mstart = list_at(c, k), mend = list_at(c, k + 1);
    uint32_t msg = c.msg_lo + k;
    // warm-up
    int32_t ws = chunk - c.warm;
    uint32_t e;
    if (ws <= mstart) { ws = mstart; e = c.start_state; } else e = c.neutral_state;
    uint32_t words = 0;
    if (ws < chunk) run_range(c, ws, chunk, e, words, msg, false);
    words = 0;
    bool first = true;
    int32_t pos = chunk;
    *head_msg = msg;
    *head_words = 0;
    while (pos < chunk_end) {
        int32_t seg_end = mend < chunk_end ? mend : chunk_end;
        run_range(c, pos, seg_end, e, words, msg, true);
        pos = seg_end;
        if (pos == mend) {   // the message ends inside this chunk: end-of-message outputs
            uint32_t eo = c.eom[e & 0xFFu];
            words += eo >> 7;
            if (eo & 0x7Fu) apply_event(c, eo & 0x7Fu, msg);
            if (c.stage[pos - 1 + c.warm] == '?') tr_or(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_Q, 1u);  // HasSuffix "?"
        }
        if (first) { *head_words = words; first = false; }
        else if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
        words = 0;
        if (pos == mend && pos < chunk_end) {   // next non-empty message
            do { k++; } while (list_at(c, k + 1) == list_at(c, k));
            msg = c.msg_lo + k;
            mend = list_at(c, k + 1);
            e = c.start_state;
        }
    }
}

}  // namespace triage
