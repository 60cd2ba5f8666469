// scan_core.cuh — the per-lane body of the arena scan, written once for device and host.
//
// The CUDA kernel (kernels.cu) calls lane_scan*() from every thread; tests/csrc/ compiles the very same functions
// with g++ and walks all lanes of all tiles sequentially to check the automaton + chunking logic against the oracle
// on machines without a GPU.  That host build is a test harness only — the shipped library has no CPU path.
//
// Model: the arena is cut into tiles of NT*C bytes; lane l of a tile owns the C bytes starting at tile_start + l*C.
// A lane first walks `warm` bytes before its chunk (or from the message start, whichever is later) with outputs
// discarded so that its automaton state is exact at the chunk start (the automaton has finite memory `span` <= warm+1),
// then walks its chunk in aligned 16-byte groups.  Message boundaries inside a group are handled WITHOUT leaving the
// unrolled group code: the bytes that belong to a neighbouring message are overwritten in registers with pad bytes the
// dictionary compiler proved inert (head pad: START --pad--> START, no output; tail pad: reading it from any state emits
// exactly that state's end-of-message output, then nothing).  Per-message counters live in global memory (atomics).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TR_HD __host__ __device__ __forceinline__
#define TR_COLD static __host__ __device__ __noinline__
#else
#define TR_HD inline
#define TR_COLD static inline
#endif

namespace triage {

// per-message counters written by the scan, read by the finalize kernel (8 x u32 = 32 bytes)
enum { CNT_WORDS = 0, CNT_KW = 1 /* ..4 */, CNT_POS = 5, CNT_NEG = 6, CNT_Q = 7, CNT_STRIDE = 8 };

// Table rows are 256 16-bit entries + 4 bytes of padding: row r starts at bank r (mod 32), so lanes that read the
// same byte pair in different states hit different banks.  The byte offset 516*state + 2*byte is one dp4a.
constexpr uint32_t kRowBytes = 516;
constexpr uint32_t kDpMul = 0xACACAC02u;   // dot((byte, st, st, st), (2, 172, 172, 172)) = 2*byte + 516*st

// events[id]: x = kw_rt | kw_hi << 16, y = kw_norm | kw_low << 16, z = pos | neg << 8 | question << 16 | risk << 24
// CNT_Q holds flags: bit 0 question, bit 1 overlap risk (message needs the sequential exact recount)
struct EventPacked { uint32_t x, y, z, w; };
struct alignas(16) Quad { uint32_t a, b, c, d; };

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
// byte offset of entry (state of `entry`, byte k of `word`)
TR_HD uint32_t tr_off(uint32_t word, uint32_t entry, uint32_t k) {
#if defined(__CUDA_ARCH__)
    return __dp4a(__byte_perm(word, entry, 0x4440u | k), kDpMul, 0u);
#else
    return (entry & 0xFFu) * kRowBytes + 2u * ((word >> (8 * k)) & 0xFFu);
#endif
}
TR_HD uint32_t tr_entry(const uint8_t *tbl, uint32_t off) { return *reinterpret_cast<const uint16_t *>(tbl + off); }

struct LaneCtx;
// one transition: returns the table entry for (state of `entry`, byte k of `word`)
TR_HD uint32_t tr_next(const LaneCtx &c, uint32_t word, uint32_t entry, uint32_t k);

struct LaneCtx {
    const uint8_t *tbl;         // padded D8 table (shared memory on device), kRowBytes per state
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
    uint32_t head_pad4, tail_pad4;   // pad byte replicated x4 (fast path only)
    uint32_t tbl_s;                  // device: address of tbl in the shared window (the dp4a accumulates onto it)
};

TR_HD uint32_t tr_next(const LaneCtx &c, uint32_t word, uint32_t entry, uint32_t k) {
#if defined(__CUDA_ARCH__)
    // PRMT (byte, st, st, st) -> IDP.4A: table base + 2*byte + 516*st -> LDS.U16
    uint32_t addr = __dp4a(__byte_perm(word, entry, 0x4440u | k), kDpMul, c.tbl_s);
    uint16_t v;
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
#else
    return tr_entry(c.tbl, tr_off(word, entry, k));
#endif
}

// per-format accessors (D8 here; the wide format overloads them in scan_wide.cuh)
TR_HD uint32_t tr_ws(const LaneCtx &, uint32_t e) { return e >> 15; }
TR_HD uint32_t tr_evmask(const LaneCtx &) { return 0x7F00u; }
TR_HD uint32_t tr_evid(const LaneCtx &, uint32_t e) { return (e >> 8) & 0x7Fu; }

// rare: kept out of line so the 16x unrolled replay stays small
TR_COLD void apply_event_raw(const EventPacked *events, uint32_t *counters, uint32_t id, uint32_t msg) {
    EventPacked e = events[id];
    uint32_t *cnt = counters + (uint64_t)msg * CNT_STRIDE;
    if (e.x & 0xFFFFu) tr_add(cnt + CNT_KW + 0, e.x & 0xFFFFu);
    if (e.x >> 16) tr_add(cnt + CNT_KW + 1, e.x >> 16);
    if (e.y & 0xFFFFu) tr_add(cnt + CNT_KW + 2, e.y & 0xFFFFu);
    if (e.y >> 16) tr_add(cnt + CNT_KW + 3, e.y >> 16);
    if (e.z & 0xFFu) tr_add(cnt + CNT_POS, e.z & 0xFFu);
    if ((e.z >> 8) & 0xFFu) tr_add(cnt + CNT_NEG, (e.z >> 8) & 0xFFu);
    if ((e.z >> 16) & 0xFFu) tr_or(cnt + CNT_Q, 1u);
    if (e.z >> 24) tr_or(cnt + CNT_Q, 2u);
}
TR_HD void apply_event(const LaneCtx &c, uint32_t id, uint32_t msg) { apply_event_raw(c.events, c.counters, id, msg); }

// tiles with more message starts than the shared-memory list holds: from the arena's offsets (kept out of line, the
// lanes meet it once in a great while)
TR_COLD int32_t list_at_abs(const uint64_t *list_abs, uint32_t idx, int64_t tile_start) {
    int64_t r = (int64_t)list_abs[idx] - tile_start;
    const int64_t lim = 1 << 30;
    return (int32_t)(r < -lim ? -lim : (r > lim ? lim : r));
}
template <class Ctx>
TR_HD int32_t list_at(const Ctx &c, uint32_t k) {
    if (c.list) return c.list[k];
    return list_at_abs(c.list_abs, c.msg_lo + k, c.tile_start);
}

// index k into list with list[k] <= pos < list[k+1]  (upper_bound - 1; empty messages repeat a value, so this lands on
// the one non-empty message that contains pos)
template <class Ctx>
TR_HD uint32_t find_msg(const Ctx &c, int32_t pos) {
    uint32_t lo = 0, hi = c.list_n;   // first index with list[idx] > pos
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (list_at(c, mid) > pos) hi = mid; else lo = mid + 1;
    }
    return lo - 1;
}
// same, walking forward from a hint known to be <= the answer (the warp's first lane)
TR_HD uint32_t find_msg_from(const LaneCtx &c, int32_t pos, uint32_t k) {
    while (list_at(c, k + 1) <= pos) k++;
    return k;
}

// ------------------------------------------------------------------------------------------- group code ---
// Formats whose walk is bound by the latency of the table reads and whose traffic is event-dense (the wide format:
// thousands of keywords) apply an event where it is met instead of re-walking the half group afterwards: one test per
// transition costs less there than eight more dependent reads per event.  Specialised next to the format's context.
template <class Ctx> struct InlineEvents { static constexpr bool value = false; };

#define TR_STEP(wd, k)                                                                      \
    e = tr_next(c, wd, e, k); words += tr_ws(c, e); ev |= e;
#define TR_STEP4(wd) TR_STEP(wd, 0) TR_STEP(wd, 1) TR_STEP(wd, 2) TR_STEP(wd, 3)

// 16 transitions, no branches: e = state entry, words += word starts, ev |= every entry seen
template <class Ctx>
TR_HD void group16(const Ctx &c, const Quad &q, uint32_t &e, uint32_t &words, uint32_t &ev) {
    TR_STEP4(q.a) TR_STEP4(q.b) TR_STEP4(q.c) TR_STEP4(q.d)
}

#define TR_STEPF(wd, k)                                                                     \
    e = tr_next(c, wd, e, k); words += tr_ws(c, e);                                         \
    if (e & tr_evmask(c)) apply_event(c, tr_evid(c, e), msg);
#define TR_STEPF4(wd) TR_STEPF(wd, 0) TR_STEPF(wd, 1) TR_STEPF(wd, 2) TR_STEPF(wd, 3)

// same 16 transitions with the (rare) events applied to message `msg`
template <class Ctx>
TR_HD void group16_events(const Ctx &c, const Quad &q, uint32_t &e, uint32_t &words, uint32_t msg) {
    TR_STEPF4(q.a) TR_STEPF4(q.b) TR_STEPF4(q.c) TR_STEPF4(q.d)
}

template <class Ctx>
TR_HD void half8_events(const Ctx &c, uint32_t wa, uint32_t wb, uint32_t &e, uint32_t &words, uint32_t msg) {
    TR_STEPF4(wa) TR_STEPF4(wb)
}

// counted group: fast pass; a half whose transitions carried an event is replayed with event handling (the word
// counts of the fast pass stand, the replay only applies the events)
template <class Ctx>
TR_HD void group16_counted(const Ctx &c, const Quad &q, uint32_t &e, uint32_t &words, uint32_t msg) {
    if (InlineEvents<Ctx>::value) { group16_events(c, q, e, words, msg); return; }
    const uint32_t e0 = e;
    uint32_t ev = 0, ev2 = 0;
    TR_STEP4(q.a) TR_STEP4(q.b)
    const uint32_t e8 = e;
    { uint32_t &ev = ev2; TR_STEP4(q.c) TR_STEP4(q.d) }
    if ((ev | ev2) & tr_evmask(c)) {
        uint32_t wdummy = 0;
        if (ev & tr_evmask(c)) { uint32_t er = e0; half8_events(c, q.a, q.b, er, wdummy, msg); }
        if (ev2 & tr_evmask(c)) { uint32_t er = e8; half8_events(c, q.c, q.d, er, wdummy, msg); }
    }
}

// same fast pass, but a half whose transitions carried an event is handed to `sink` (e.g. a per-warp queue that replays
// many halves at once with all lanes, instead of one lane re-walking 8 bytes while 31 wait)
template <class Ctx, class Sink>
TR_HD void group16_sink(const Ctx &c, const Quad &q, uint32_t &e, uint32_t &words, uint32_t msg, Sink &sink) {
    const uint32_t e0 = e;
    uint32_t ev = 0, ev2 = 0;
    TR_STEP4(q.a) TR_STEP4(q.b)
    const uint32_t e8 = e;
    { uint32_t &ev = ev2; TR_STEP4(q.c) TR_STEP4(q.d) }
    if ((ev | ev2) & tr_evmask(c)) {
        if (ev & tr_evmask(c)) sink.push(c, e0, q.a, q.b, msg);
        if (ev2 & tr_evmask(c)) sink.push(c, e8, q.c, q.d, msg);
    }
}

TR_HD uint32_t low_mask(int32_t nbytes) { return nbytes <= 0 ? 0u : (nbytes >= 4 ? 0xFFFFFFFFu : ((1u << (8 * nbytes)) - 1u)); }
// bytes [0, h) of the group <- pad
TR_HD Quad mask_head(Quad q, int32_t h, uint32_t pad4) {
    uint32_t m0 = low_mask(h), m1 = low_mask(h - 4), m2 = low_mask(h - 8), m3 = low_mask(h - 12);
    q.a = (q.a & ~m0) | (pad4 & m0); q.b = (q.b & ~m1) | (pad4 & m1);
    q.c = (q.c & ~m2) | (pad4 & m2); q.d = (q.d & ~m3) | (pad4 & m3);
    return q;
}
// bytes [t, 16) of the group <- pad
TR_HD Quad mask_tail(Quad q, int32_t t, uint32_t pad4) {
    uint32_t m0 = ~low_mask(t), m1 = ~low_mask(t - 4), m2 = ~low_mask(t - 8), m3 = ~low_mask(t - 12);
    q.a = (q.a & ~m0) | (pad4 & m0); q.b = (q.b & ~m1) | (pad4 & m1);
    q.c = (q.c & ~m2) | (pad4 & m2); q.d = (q.d & ~m3) | (pad4 & m3);
    return q;
}
template <class Ctx>
TR_HD Quad load_quad(const Ctx &c, int32_t p) { return *reinterpret_cast<const Quad *>(c.stage + p + c.warm); }

// ------------------------------------------------------------------------------------------- fast lane ---
// Walks the lane's chunk [chunk, chunk_end) (tile-relative, chunk 16-aligned).  Every finished segment is flushed with
// atomics except the partial word count of the message at the chunk start, which is returned in (*head_msg,
// *head_words) so the caller can combine neighbouring lanes first.  k = list index of the message containing `chunk`.
TR_HD void lane_scan_fast(const LaneCtx &c, int32_t chunk, int32_t chunk_end, uint32_t k, uint32_t *head_msg, uint32_t *head_words) {
    int32_t mstart = list_at(c, k), mend = list_at(c, k + 1);
    uint32_t msg = c.msg_lo + k;
    // warm-up: state becomes exact at `chunk`
    const int32_t ws = chunk - c.warm;
    uint32_t e = mstart >= ws ? c.start_state : c.neutral_state;
    {
        uint32_t wdummy = 0, evdummy = 0;
        for (int32_t g = ws; g < chunk; g += 16) {
            if (mstart >= g + 16) continue;   // the whole group precedes the message
            Quad q = load_quad(c, g);
            if (mstart > g) q = mask_head(q, mstart - g, c.head_pad4);
            group16(c, q, e, wdummy, evdummy);
        }
    }
    uint32_t words = 0;
    bool first = true;
    *head_msg = msg;
    *head_words = 0;
    bool open = true;   // a message is in progress (its words not flushed yet)
    for (int32_t g = chunk; g < chunk_end; g += 16) {
        const int32_t gend = g + 16;
        Quad q0 = load_quad(c, g);
        if (mend > gend) {   // common case: the message runs through this whole group
            group16_counted(c, q0, e, words, msg);
            continue;
        }
        // one or more messages end inside (g, gend]
        int32_t cur = g;
        for (;;) {
            const bool ends = mend <= gend;
            const int32_t t = ends ? mend : gend;
            Quad q = q0;
            if (cur > g) q = mask_head(q, cur - g, c.head_pad4);
            if (t < gend) q = mask_tail(q, t - g, c.tail_pad4);   // the first pad byte emits the end-of-message output
            group16_counted(c, q, e, words, msg);
            if (!ends) break;   // the (new) message continues into the next group
            if (t == gend) {    // ended exactly on the group boundary: no pad byte was read, take the output from the table
                uint32_t eo = c.eom[e & 0xFFu];
                words += eo >> 7;
                if (eo & 0x7Fu) apply_event(c, eo & 0x7Fu, msg);
            }
            if (c.stage[t - 1 + c.warm] == '?') tr_or(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_Q, 1u);   // HasSuffix "?"
            if (first) { *head_words = words; first = false; }
            else if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
            words = 0;
            open = false;
            if (t >= chunk_end) break;
            do { k++; } while (list_at(c, k + 1) == list_at(c, k));   // next non-empty message
            msg = c.msg_lo + k;
            mend = list_at(c, k + 1);
            e = c.start_state;
            open = true;
            cur = t;
            if (t == gend) break;   // it starts exactly at the next group
        }
    }
    if (open) {   // the chunk ended inside a message: flush the partial count
        if (first) *head_words = words;
        else if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
    }
}

// ----------------------------------------------------------------------------------------- padded lane ---
// Padded arena layout (the engine's preferred wire format): every message starts on a 16-byte boundary and is followed
// by 1..16 copies of the dictionary's pad byte up to the next message.  The pad byte is a RESET symbol of the automaton
// (from every state it emits that state's end-of-message output and lands in START), so a lane never needs to know where
// messages end to keep its state right: boundaries only matter for attributing counts, and they always fall between
// groups.  No masks, no end-of-message table, no second pass.  k = list index of the message containing `chunk`;
// list holds the (padded) message starts.  at_arena_start: nothing precedes `chunk` (first lane of the first tile).
template <class Ctx>
TR_HD void lane_scan_padded(const Ctx &c, int32_t chunk, int32_t chunk_end, uint32_t k, bool at_arena_start,
                            uint32_t *head_msg, uint32_t *head_words) {
    uint32_t e = c.start_state;
    if (!at_arena_start) {
        e = c.neutral_state;
        uint32_t wdummy = 0, evdummy = 0;
        for (int32_t g = chunk - c.warm; g < chunk; g += 16) group16(c, load_quad(c, g), e, wdummy, evdummy);
    }
    int32_t mend = list_at(c, k + 1);
    uint32_t msg = c.msg_lo + k;
    uint32_t words = 0;
    bool first = true;
    *head_msg = msg;
    *head_words = 0;
    for (int32_t g = chunk; g < chunk_end; g += 16) {
        if (g >= mend) {   // the previous message (and its pads) ended exactly here
            if (first) { *head_words = words; first = false; }
            else if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
            words = 0;
            k++;
            msg++;
            mend = list_at(c, k + 1);
        }
        group16_counted(c, load_quad(c, g), e, words, msg);
    }
    if (first) *head_words = words;
    else if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
}

// ------------------------------------------------------------------------------------------- long lanes ---
// Padded arena, long per-lane chunks taken a window at a time: the lane keeps its automaton state, word count and
// current message in registers from window to window, so the warm-up is paid once per chunk (16 bytes per 320) instead
// of once per window.  scan_rows_kernel (kernels.cu) is built from long_begin / long_boundary / group16_sink / long_end
// with its windows staged by the TMA, and reads the arena's partial last row with long_stream_n<1>; long_window and the
// N-chain forms are the same walk in plain loops, kept for the host checks (tests/csrc) and measured as kernel variants
// (several chains per lane bring nothing, the load/store unit is the limit).
struct LongLane {
    uint32_t e, words, k, msg, head_msg, head_words;
    int32_t mend;
    bool first;
};

// warm_bytes: the c.warm bytes that precede the chunk (16-byte aligned), unused when at_arena_start
template <class Ctx>
TR_HD void long_begin(const Ctx &c, LongLane &L, uint32_t k, bool at_arena_start, const uint8_t *warm_bytes) {
    L.e = c.start_state;
    if (!at_arena_start) {
        L.e = c.neutral_state;
        uint32_t wdummy = 0, evdummy = 0;
        for (int32_t g = 0; g < c.warm; g += 16) group16(c, *reinterpret_cast<const Quad *>(warm_bytes + g), L.e, wdummy, evdummy);
    }
    L.k = k;
    L.mend = list_at(c, k + 1);
    L.msg = c.msg_lo + k;
    L.words = 0;
    L.first = true;
    L.head_msg = L.msg;
    L.head_words = 0;
}

// the message (and its pads) that was current ended exactly at `pos`: flush its count, step to the next one.  Messages
// take at least 16 bytes of the padded arena, so one check per 16-byte group is enough.
template <class Ctx>
TR_HD void long_boundary(const Ctx &c, LongLane &L, int32_t pos) {
    if (pos >= L.mend) {
        if (L.first) { L.head_words = L.words; L.first = false; }
        else if (L.words) tr_add(c.counters + (uint64_t)L.msg * CNT_STRIDE + CNT_WORDS, L.words);
        L.words = 0;
        L.k++;
        L.msg++;
        L.mend = list_at(c, L.k + 1);
    }
}

// win: this lane's bytes of the window (16-byte aligned), pos: their position relative to the list origin
template <class Ctx>
TR_HD void long_window(const Ctx &c, LongLane &L, const uint8_t *win, int32_t pos, int32_t nbytes) {
    for (int32_t g = 0; g < nbytes; g += 16) {
        long_boundary(c, L, pos + g);
        group16_counted(c, *reinterpret_cast<const Quad *>(win + g), L.e, L.words, L.msg);
    }
}

// Re-walks 8 bytes from entry e and applies the events met on the way (the counts of the fast pass stand).  Kept out of
// line and rolled: with several chains per lane an inlined copy per chain and half pushes the loop out of the
// instruction cache.
template <class Ctx>
TR_COLD void replay_half(const Ctx &c, uint32_t wa, uint32_t wb, uint32_t e, uint32_t msg) {
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int i = 0; i < 8; i++) {
        const uint32_t wd = i < 4 ? wa : wb;
        e = tr_next(c, wd >> (8 * (i & 3)), e, 0);
        if (e & tr_evmask(c)) apply_event(c, tr_evid(c, e), msg);
    }
}

// N independent chains per lane, interleaved transition by transition: the walk of one chain is a dependent sequence
// (table read -> index arithmetic -> table read), so N chains give the scheduler N reads in flight per lane (N = 1 in
// the shipped kernels).
#define TR_STEPN(fld, k, EV)                                                                   \
    _Pragma("unroll") for (int ch = 0; ch < N; ch++) {                                         \
        L[ch].e = tr_next(c, q[ch].fld, L[ch].e, k); L[ch].words += tr_ws(c, L[ch].e); EV[ch] |= L[ch].e; }
#define TR_STEPN4(fld, EV) TR_STEPN(fld, 0, EV) TR_STEPN(fld, 1, EV) TR_STEPN(fld, 2, EV) TR_STEPN(fld, 3, EV)

#define TR_STEPNF(fld, k)                                                                      \
    _Pragma("unroll") for (int ch = 0; ch < N; ch++) {                                         \
        L[ch].e = tr_next(c, q[ch].fld, L[ch].e, k); L[ch].words += tr_ws(c, L[ch].e);         \
        if (L[ch].e & tr_evmask(c)) apply_event(c, tr_evid(c, L[ch].e), L[ch].msg); }
#define TR_STEPNF4(fld) TR_STEPNF(fld, 0) TR_STEPNF(fld, 1) TR_STEPNF(fld, 2) TR_STEPNF(fld, 3)

template <int N, class Ctx>
TR_HD void group16_counted_n(const Ctx &c, const Quad (&q)[N], LongLane (&L)[N]) {
    if (InlineEvents<Ctx>::value) {
        TR_STEPNF4(a) TR_STEPNF4(b) TR_STEPNF4(c) TR_STEPNF4(d)
        return;
    }
    uint32_t e0[N], e8[N], eva[N], evb[N], any = 0;
#pragma unroll
    for (int ch = 0; ch < N; ch++) { e0[ch] = L[ch].e; eva[ch] = 0; evb[ch] = 0; }
    TR_STEPN4(a, eva) TR_STEPN4(b, eva)
#pragma unroll
    for (int ch = 0; ch < N; ch++) e8[ch] = L[ch].e;
    TR_STEPN4(c, evb) TR_STEPN4(d, evb)
#pragma unroll
    for (int ch = 0; ch < N; ch++) any |= eva[ch] | evb[ch];
    if (any & tr_evmask(c)) {
#pragma unroll
        for (int ch = 0; ch < N; ch++) {
            if (eva[ch] & tr_evmask(c)) replay_half(c, q[ch].a, q[ch].b, e0[ch], L[ch].msg);
            if (evb[ch] & tr_evmask(c)) replay_half(c, q[ch].c, q[ch].d, e8[ch], L[ch].msg);
        }
    }
}

TR_HD Quad stream_quad(const uint8_t *p) {
#ifdef __CUDA_ARCH__
    uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
    Quad q; q.a = v.x; q.b = v.y; q.c = v.z; q.d = v.w;
    return q;
#else
    return *reinterpret_cast<const Quad *>(p);
#endif
}
TR_HD void stream_prefetch(const uint8_t *p) {
#ifdef __CUDA_ARCH__
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}

// Streams nbytes (a multiple of 16) of N chains straight from global memory: src[ch] is the chunk of chain ch, pos[ch]
// its position relative to the list origin.  The next 16 bytes of every chain are loaded into registers while the
// current ones are walked; `limit` is the end of the arena (prefetch hints stay below it).
template <int N, class Ctx>
TR_HD void long_stream_n(const Ctx &c, LongLane (&L)[N], const uint8_t *const (&src)[N], const int32_t (&pos)[N], int32_t nbytes,
                         const uint8_t *limit) {
    Quad cur[N], nxt[N];
#pragma unroll
    for (int ch = 0; ch < N; ch++) { cur[ch] = stream_quad(src[ch]); nxt[ch] = cur[ch]; }
#pragma unroll 2
    for (int32_t g = 0; g < nbytes; g += 16) {
        if (g + 16 < nbytes) {
#pragma unroll
            for (int ch = 0; ch < N; ch++) nxt[ch] = stream_quad(src[ch] + g + 16);
        }
        if ((g & 127) == 0) {
#pragma unroll
            for (int ch = 0; ch < N; ch++)
                if (src[ch] + g + 384 < limit) stream_prefetch(src[ch] + g + 384);
        }
#pragma unroll
        for (int ch = 0; ch < N; ch++) long_boundary(c, L[ch], pos[ch] + g);
        group16_counted_n<N>(c, cur, L);
#pragma unroll
        for (int ch = 0; ch < N; ch++) cur[ch] = nxt[ch];
    }
}

template <class Ctx>
TR_HD void long_end(const Ctx &c, LongLane &L) {
    if (L.first) L.head_words = L.words;
    else if (L.words) tr_add(c.counters + (uint64_t)L.msg * CNT_STRIDE + CNT_WORDS, L.words);
}

// ------------------------------------------------------------------------------------------- slow lane ---
// Fallback for dictionaries without usable pad bytes: same walk, byte at a time around boundaries.
TR_HD void step_slow(const LaneCtx &c, uint32_t byte, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    e = tr_entry(c.tbl, (e & 0xFFu) * kRowBytes + 2u * byte);
    if (count) {
        words += e >> 15;
        uint32_t id = (e >> 8) & 0x7Fu;
        if (id) apply_event(c, id, msg);
    }
}
TR_HD void run_slow(const LaneCtx &c, int32_t a, int32_t b, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    for (int32_t p = a; p < b; p++) step_slow(c, c.stage[p + c.warm], e, words, msg, count);
}
TR_HD void run_range(const LaneCtx &c, int32_t a, int32_t b, uint32_t &e, uint32_t &words, uint32_t msg, bool count) {
    int32_t p = a;
    int32_t head = (16 - ((p + c.warm) & 15)) & 15;   // stage is 16-byte aligned at index 0
    if (head > b - p) head = b - p;
    if (head) { run_slow(c, p, p + head, e, words, msg, count); p += head; }
    for (; p + 16 <= b; p += 16) {
        Quad q = load_quad(c, p);
        if (count) group16_counted(c, q, e, words, msg);
        else { uint32_t w = 0, ev = 0; group16(c, q, e, w, ev); }
    }
    if (p < b) run_slow(c, p, b, e, words, msg, count);
}
TR_HD void lane_scan_slow(const LaneCtx &c, int32_t chunk, int32_t chunk_end, uint32_t k, uint32_t *head_msg, uint32_t *head_words) {
    int32_t mstart = list_at(c, k), mend = list_at(c, k + 1);
    uint32_t msg = c.msg_lo + k;
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
            if (c.stage[pos - 1 + c.warm] == '?') tr_or(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_Q, 1u);
        }
        if (first) { *head_words = words; first = false; }
        else if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
        words = 0;
        if (pos == mend && pos < chunk_end) {
            do { k++; } while (list_at(c, k + 1) == list_at(c, k));
            msg = c.msg_lo + k;
            mend = list_at(c, k + 1);
            e = c.start_state;
        }
    }
}

}  // namespace triage
