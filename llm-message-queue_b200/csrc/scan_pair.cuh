// scan_pair.cuh — the two-bytes-per-lookup lane of the padded-arena scan, written once for device and host.
//
// The one-lookup-per-byte lane (scan_core.cuh) is bound by the rate at which an SM accepts shared-memory loads.  This
// lane halves the lookups and takes the word count out of the automaton:
//   * words: strings.Fields word starts of a 16-byte group by SIMD-in-register compares (ASCII whitespace 9..13, 32;
//     a byte >= 0x80 counts as non-space), no lookups;
//   * events (keywords, question words, sentiment words): a FILTER automaton over 4-bit byte classes ((b & 31) * mul >>
//     shift & 15, arithmetic only) that consumes two bytes per lookup.  It is walked in its Moore form: an entry is one
//     byte, the next state, at state * 260 + class pair of a table at a fixed shared-memory address (one multiply-add,
//     the base is an immediate of the load), and a state >= ffired says "an occurrence of some pattern's class string
//     may have ended in the two bytes that led here".  It over-approximates: bytes that share a class are not told apart;
//   * every group the filter flags, and every group that holds one of the lead bytes C2 C4 C5 E0-E3 (multi-byte
//     whitespace, the fold runes U+017F / U+212A / U+0130) or follows such a group, is handed to `sink`, which walks the
//     group through the EXACT one-byte automaton (its column-merged copy; 16 warm-up bytes + the 16 bytes).  A group the
//     filter alone flagged is walked for its events, its word starts stay the SIMD count; a group with, or right after,
//     one of those lead bytes takes its word starts from the walk too.  Flagged groups are rare on text (under 1 % on the
//     benchmark set).
// A lane takes its bytes a WINDOW (NG groups) at a time: the NG groups are scanned without a branch, and one test per
// window - nothing flagged, no byte >= 0x80, no message boundary inside - decides between adding up the window's word
// starts and the per-group path that pushes flagged groups and steps over message boundaries.
// Exactness: an event that completes in group g depends on at most 17 bytes, so every byte it depends on lies in g-1 or
// g; the filter's class automaton knows every way to complete an event that uses none of those lead bytes (literals of
// any script included) and flags g when it sees one, otherwise g-1 or g holds such a lead byte and g is walked exactly
// anyway.  A group that is not walked exactly has no multi-byte whitespace in or just before it, so its word starts are
// what the SIMD compares count: a byte >= 0x80 outside those runes is never whitespace to Go.
#pragma once
#include <stdint.h>

#include "scan_core.cuh"

namespace triage {

// The device keeps the filter table at this shared-memory address, so that the address of an entry is one multiply-add
// plus an immediate: state * kPairRowBytes + class pair.  Rows are 260 bytes apart: row r starts at bank r mod 32, lanes in
// different states that read the same class pair do not collide.
constexpr uint32_t kPairTableShared = 0x10000u;
constexpr uint32_t kPairRowBytes = 260u;

struct PairCtx {
    const uint8_t *ftbl;        // filter table, Moore form: 256 next states per state (host; the device reads kPairTableShared)
    uint32_t fstart, ffired;    // filter root; states >= ffired were entered by a firing entry
    uint32_t mul, shift;        // byte -> class: ((b & 31) * mul >> shift) & 15
    const uint16_t *xtbl;       // exact automaton, column-merged: entry(state, xcls[byte]), xn per state (shared memory on the device)
    const uint8_t *xcls;
    uint32_t xn;
    uint32_t xtbl_s, xcls_s;    // device: their shared-memory addresses
    const EventPacked *events;  // global memory on the device: read when an event is applied
    uint32_t neutral_state;
    uint32_t pad4;              // reset pad byte x 4
    const int32_t *list;        // message starts of the tile (same meaning as LaneCtx)
    const uint64_t *list_abs;
    int64_t tile_start;
    uint32_t list_n, msg_lo;
    uint32_t *counters;
    const uint8_t *arena;       // device: the arena (a queued group is read back from it when it is walked)
};

// f: a filter state (upper bytes zero); k: which byte of `pairs` holds the class pair (0 or 2)
TR_HD uint32_t pr_fnext(const PairCtx &c, uint32_t pairs, uint32_t f, uint32_t k) {
#if defined(__CUDA_ARCH__)
    uint32_t v;
    asm("ld.shared.u8 %0, [%1+65536];" : "=r"(v) : "r"(f * kPairRowBytes + __byte_perm(pairs, 0u, 0x4440u | k)));
    return v;
#else
    return c.ftbl[(f << 8) | ((pairs >> (8 * k)) & 0xFFu)];
#endif
}
TR_HD uint32_t pr_xnext(const PairCtx &c, uint32_t e, uint32_t byte) {
#if defined(__CUDA_ARCH__)
    uint32_t cls;
    uint16_t v;
    asm("ld.shared.u8 %0, [%1];" : "=r"(cls) : "r"(c.xcls_s + byte));
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(c.xtbl_s + 2u * ((e & 0xFFu) * c.xn + cls)));
    return v;
#else
    return c.xtbl[(e & 0xFFu) * c.xn + c.xcls[byte]];
#endif
}

// bit 7 of every byte: the byte is not ASCII whitespace.  The other bits are junk: users mask.  Only valid for words
// whose bytes are all < 0x80 (the per-byte sums of a byte >= 0x80 carry over): windows with such bytes are recounted
// with pr_nonspace_any.
TR_HD uint32_t pr_nonspace(uint32_t w) {
    const uint32_t ne20 = (w ^ 0x20202020u) + 0x7F7F7F7Fu;    // bit 7: byte != 0x20
    const uint32_t gt13 = w + 0x72727272u;                    // bit 7: byte > 13
    const uint32_t ge9 = w + 0x77777777u;                     // bit 7: byte >= 9
    return ne20 & (gt13 | ~ge9);
}
// the same for any bytes: a byte >= 0x80 is not ASCII whitespace
TR_HD uint32_t pr_nonspace_any(uint32_t w) { return pr_nonspace(w & 0x7F7F7F7Fu) | w; }
// Does the word hold a byte that can start a rune the exact automaton treats specially?  Multi-byte whitespace starts
// with C2, E1, E2 or E3; the fold runes U+017F, U+212A, U+0130 with C5, E2, C4.  (E0 is let in with E1..E3: one test.)
// Every other byte >= 0x80 - accented Latin, Cyrillic, CJK ideographs, emoji, stray continuation bytes - is to the path
// what an ordinary letter is: not whitespace, part of no fold.  Bit 7 of every such byte.
TR_HD uint32_t pr_special(uint32_t w) {
    const uint32_t m = w & 0x7F7F7F7Fu;
    const uint32_t n42 = (m ^ 0x42424242u) + 0x7F7F7F7Fu;     // bit 7: low seven bits != 0x42
    const uint32_t n44 = (m ^ 0x44444444u) + 0x7F7F7F7Fu;
    const uint32_t n45 = (m ^ 0x45454545u) + 0x7F7F7F7Fu;
    const uint32_t n6x = ((w & 0x7C7C7C7Cu) ^ 0x60606060u) + 0x7F7F7F7Fu;   // bit 7: low seven bits not in 0x60..0x63
    return ~(n42 & n44 & n45 & n6x) & w & 0x80808080u;
}
TR_HD uint32_t pr_special(const Quad &q) {
    if (!((q.a | q.b | q.c | q.d) & 0x80808080u)) return 0;
    return pr_special(q.a) | pr_special(q.b) | pr_special(q.c) | pr_special(q.d);
}
TR_HD uint32_t pr_funnel8(uint32_t lo, uint32_t hi) {          // (hi << 8) | (lo >> 24)
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, 8);
#else
    return (hi << 8) | (lo >> 24);
#endif
}
TR_HD uint32_t pr_popc(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)__popc(v);
#else
    return (uint32_t)__builtin_popcount(v);
#endif
}
TR_HD uint32_t pr_max(uint32_t a, uint32_t b) { return a > b ? a : b; }
// class pairs of a word: byte 0 = c0 | c1 << 4, byte 2 = c2 | c3 << 4
template <bool SIMPLE>
TR_HD uint32_t pr_pairs(const PairCtx &c, uint32_t w) {
    uint32_t h;
    if (SIMPLE) h = w & 0x0F0F0F0Fu;                           // mul 1, shift 0
    else h = (((w & 0x1F1F1F1Fu) * c.mul) >> c.shift) & 0x0F0F0F0Fu;
    return h + (h >> 4);
}

struct PairLane {
    LongLane L;            // message tracking and the word count of the current message (L.e is unused)
    uint32_t f;            // filter state
    uint32_t prev_ns;      // pr_nonspace of the previous word (bit 31: the byte before the group)
    uint32_t prev_rare;    // the previous group held one of the special lead bytes
    Quad prevq;            // the 16 bytes before the window (warm-up of an exact walk of its first group)
};

// The exact walk of one flagged group: state from the 16 bytes before it, then words and events of its 16 bytes.
// load(i): word i of the 32 bytes (0..3 the bytes before, 4..7 the group).  Rolled by word: it runs for one group in a
// hundred, its code should not push the scan loop out of the instruction cache.
// count_words: the group's word starts come from this walk too (groups with, or right after, the special lead bytes);
// otherwise the SIMD count of the group stands and only the events are taken.
template <class Ctx, class Load>
TR_HD void pair_replay_words(const Ctx &c, Load load, uint32_t msg, bool count_words) {
    uint32_t e = c.neutral_state, words = 0;
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int i = 0; i < 4; i++) {
        const uint32_t w = load(i);
        e = pr_xnext(c, e, w & 0xFFu);
        e = pr_xnext(c, e, (w >> 8) & 0xFFu);
        e = pr_xnext(c, e, (w >> 16) & 0xFFu);
        e = pr_xnext(c, e, w >> 24);
    }
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int i = 4; i < 8; i++) {
        const uint32_t w = load(i);
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
        for (int k = 0; k < 4; k++) {
            e = pr_xnext(c, e, (w >> (8 * k)) & 0xFFu);
            words += e >> 15;
            if (e & 0x7F00u) apply_event_raw(c.events, c.counters, (e >> 8) & 0x7Fu, msg);
        }
    }
    if (count_words && words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
}
template <class Ctx>
TR_COLD void pair_replay(const Ctx &c, Quad warm, Quad q, uint32_t msg, bool count_words) {
    const uint32_t v[8] = {warm.a, warm.b, warm.c, warm.d, q.a, q.b, q.c, q.d};
    pair_replay_words(c, [&](int i) { return v[i]; }, msg, count_words);
}

// warm: the 16 bytes before the lane's chunk (the reset pad byte x 16 when nothing precedes it: the exact walk then
// starts in the start state, and the byte before the chunk counts as whitespace)
template <bool SIMPLE, class Ctx>
TR_HD void pair_begin(const Ctx &c, PairLane &P, uint32_t k, const Quad &warm) {
    P.L.e = 0;
    P.L.k = k;
    P.L.mend = list_at(c, k + 1);
    P.L.msg = c.msg_lo + k;
    P.L.words = 0;
    P.L.first = true;
    P.L.head_msg = P.L.msg;
    P.L.head_words = 0;
    uint32_t f = c.fstart;
    const uint32_t ww[4] = {warm.a, warm.b, warm.c, warm.d};
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int j = 0; j < 4; j++) {
        const uint32_t p = pr_pairs<SIMPLE>(c, ww[j]);
        f = pr_fnext(c, p, f, 0);
        f = pr_fnext(c, p, f, 2);
    }
    P.f = f;
    P.prev_ns = pr_nonspace_any(warm.d);
    P.prev_rare = pr_special(warm);
    P.prevq = warm;
}

// Word starts of 16 bytes as a bit set: bit 7 - j of byte i <=> byte i of word j starts a word.  pns: pr_nonspace of the
// word before (in, out).  ANY: exact for bytes >= 0x80 too.
template <bool ANY>
TR_HD uint32_t pair_starts16(const Quad &q, uint32_t &pns) {
    const uint32_t qw[4] = {q.a, q.b, q.c, q.d};
    uint32_t u = 0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int j = 0; j < 4; j++) {
        const uint32_t ns = ANY ? pr_nonspace_any(qw[j]) : pr_nonspace(qw[j]);
        u += ((ns & ~pr_funnel8(pns, ns)) & 0x80808080u) >> j;
        pns = ns;
    }
    return u;
}

// 16 bytes without a branch: the filter walks on (f); returns the highest filter state entered
template <bool SIMPLE>
TR_HD uint32_t pair_filter16(const PairCtx &c, const Quad &q, uint32_t &f) {
    const uint32_t qw[4] = {q.a, q.b, q.c, q.d};
    uint32_t m = 0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int j = 0; j < 4; j++) {
        const uint32_t p = pr_pairs<SIMPLE>(c, qw[j]);
        f = pr_fnext(c, p, f, 0); m = pr_max(m, f);
        f = pr_fnext(c, p, f, 2); m = pr_max(m, f);
    }
    return m;
}

// byte k of v
TR_HD uint32_t pr_byte(uint32_t v, uint32_t k) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(v, 0u, 0x4440u | k);
#else
    return (v >> (8 * k)) & 0xFFu;
#endif
}

// NG (<= 5) 16-byte groups at tile-relative position pos.  P.prevq holds the 16 bytes before win[0] (they may no longer
// be readable where they came from).  Returns how many groups went to the sink.
// The groups are scanned without a branch; what the window leaves behind is small: `flags` (bit g: the filter fired in
// group g), the running count of word starts after each group (four of them packed into `cum`, the last one `total`) and
// `hi`.  Windows where nothing fired, no byte is >= 0x80 and no message ends add `total` and are done; the others visit
// only the groups that need a look (message boundary, flagged), unless bytes >= 0x80 make every group one.
template <bool SIMPLE, int NG, class Ctx, class Sink>
TR_HD uint32_t pair_window(const Ctx &c, PairLane &P, const Quad *win, int32_t pos, Sink &sink) {
    static_assert(NG >= 1 && NG <= 5, "the running counts of the first NG - 1 groups share one register");
    uint32_t hi = 0, f = P.f, pns = P.prev_ns, flags = 0, cum = 0, total = 0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int g = 0; g < NG; g++) {
        const Quad q = win[g];
        total += pr_popc(pair_starts16<false>(q, pns));
        if (g < NG - 1) cum |= total << (8 * g);
        if (pair_filter16<SIMPLE>(c, q, f) >= c.ffired) flags |= 1u << g;
        hi |= q.a | q.b | q.c | q.d;
    }
    P.f = f;
    hi &= 0x80808080u;
    const Quad last = win[NG - 1];
    uint32_t pushed = 0;
    if (!hi && !P.prev_rare && pos + 16 * (NG - 1) < P.L.mend) {
        // one message, no byte >= 0x80 in or right before the window: the SIMD counts stand; what the filter flagged is
        // queued for its events only
        P.L.words += total;
        P.prev_ns = pns;
        while (flags) {
#if defined(__CUDA_ARCH__)
            const uint32_t g = (uint32_t)__ffs((int)flags) - 1u;
#else
            const uint32_t g = (uint32_t)__builtin_ctz(flags);
#endif
            flags &= flags - 1u;
            sink.push(c, g ? win[g - 1] : P.prevq, win[g], P.L.msg, pos + 16 * (int32_t)g, false);
            pushed++;
        }
        P.prevq = last;
        return pushed;
    }
    if (hi) {   // bytes >= 0x80 in the window: every group is looked at; those that hold such bytes, or follow one that
                // does, get their word starts from the compares that are exact for them
        uint32_t done = 0;
        bool prev_high = false;   // (P.prev_ns, the word before the window, is always exact)
        pns = P.prev_ns;
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
        for (int g = 0; g < NG; g++) {
            long_boundary(c, P.L, pos + 16 * g);
            const Quad q = win[g];
            const uint32_t upto = g + 1 < NG ? pr_byte(cum, (uint32_t)g) : total;
            const bool high = ((q.a | q.b | q.c | q.d) & 0x80808080u) != 0;
            uint32_t n = upto - done, rare = 0;
            done = upto;
            if (high || prev_high) {
                if (g) pns = pr_nonspace_any(win[g - 1].d);
                n = pr_popc(pair_starts16<true>(q, pns));
                if (high) rare = pr_special(q);
            }
            prev_high = high;
            const bool exact_words = rare || P.prev_rare;        // multi-byte whitespace may be in play: words from the walk
            if (exact_words || ((flags >> g) & 1u)) {
                sink.push(c, g ? win[g - 1] : P.prevq, q, P.L.msg, pos + 16 * g, exact_words);
                pushed++;
            }
            if (!exact_words) P.L.words += n;
            P.prev_rare = rare;
        }
        pns = pr_nonspace_any(last.d);
    } else {
        uint32_t need = flags | (P.prev_rare ? 1u : 0u), done = 0;
        {
            const int32_t gb = (P.L.mend - pos) >> 4;              // first group of the next message, if it is in this window
            if (gb < NG) need |= 1u << (gb < 0 ? 0 : gb);
        }
        while (need) {
#if defined(__CUDA_ARCH__)
            const uint32_t g = (uint32_t)__ffs((int)need) - 1u;
#else
            const uint32_t g = (uint32_t)__builtin_ctz(need);
#endif
            need &= need - 1u;
            const uint32_t before = g ? pr_byte(cum, g - 1) : 0u;   // word starts of the groups before g
            P.L.words += before - done;
            done = before;
            if (pos + 16 * (int32_t)g >= P.L.mend) {
                long_boundary(c, P.L, pos + 16 * (int32_t)g);
                const int32_t gb = (P.L.mend - pos) >> 4;          // short messages: the next one may end in this window too
                if (gb < NG) need |= 1u << gb;
            }
            const bool exact_words = g == 0 && P.prev_rare;
            if (exact_words || ((flags >> g) & 1u)) {
                sink.push(c, g ? win[g - 1] : P.prevq, win[g], P.L.msg, pos + 16 * (int32_t)g, exact_words);
                pushed++;
                if (exact_words) done = g + 1u < (uint32_t)NG ? pr_byte(cum, g) : total;   // its word starts come from the exact walk
            }
        }
        P.L.words += total - done;
        P.prev_rare = 0;
    }
    P.prev_ns = pns;
    P.prevq = last;
    return pushed;
}

// one group (the arena's partial last row, host checks)
template <bool SIMPLE, class Ctx, class Sink>
TR_HD bool pair_group(const Ctx &c, PairLane &P, const Quad &q, int32_t pos, Sink &sink) {
    return pair_window<SIMPLE, 1>(c, P, &q, pos, sink) != 0;
}

template <class Ctx>
TR_HD void pair_end(const Ctx &c, PairLane &P) { long_end(c, P.L); }

// sink that walks a flagged group on the spot (host checks, the partial last row)
struct PairSinkNow {
    template <class Ctx>
    TR_HD void push(const Ctx &c, const Quad &warm, const Quad &q, uint32_t msg, int32_t, bool count_words) { pair_replay(c, warm, q, msg, count_words); }
};

}  // namespace triage
