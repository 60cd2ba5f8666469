// scan_pair.cuh — the two-bytes-per-lookup lane of the padded-arena scan, written once for device and host.
//
// The one-lookup-per-byte lane (scan_core.cuh) is bound by the rate at which an SM accepts shared-memory loads.  This
// lane halves the lookups and takes the word count out of the automaton:
//   * words: strings.Fields word starts of a 16-byte group by SIMD-in-register compares (ASCII whitespace 9..13, 32),
//     no lookups;
//   * events (keywords, question words, sentiment words): a FILTER automaton over 4-bit byte classes ((b & 31) * mul >>
//     shift & 15, arithmetic only) that consumes two bytes per lookup - entry(state, class pair) = next state | "an
//     occurrence of some pattern's class string may have ended in these two bytes".  It over-approximates: bytes that
//     share a class are not told apart;
//   * every group the filter flags, and every group that holds one of the lead bytes C2 C4 C5 E0-E3 (multi-byte
//     whitespace, the fold runes U+017F / U+212A / U+0130) or follows such a group, is handed to `sink`, which walks the group
//     through the EXACT one-byte automaton (its column-merged copy; 16 warm-up bytes + the 16 bytes) and takes words and events from that walk
//     instead.  Flagged groups are rare on text (about 0.5 % on the benchmark set), so the exact walk is off the hot path.
// Exactness: an event that completes in group g depends on at most 17 bytes, so every byte it depends on lies in g-1 or
// g; the filter's class automaton knows every way to complete an event that uses none of those lead bytes (literals of
// any script included) and flags g when it sees one, otherwise g-1 or g holds such a lead byte and g is walked exactly
// anyway.  A group that is not walked exactly has no multi-byte whitespace in or just before it, so its word starts are
// what the SIMD compares count: a byte >= 0x80 outside those runes is never whitespace to Go.
#pragma once
#include <stdint.h>

#include "scan_core.cuh"

namespace triage {

struct PairCtx {
    const uint8_t *ftbl;        // filter table: kRowBytes per state, 256 16-bit entries: next state | fired << 15
    uint32_t ftbl_s;            // device: its shared-memory address
    uint32_t fstart;            // filter root
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
};

TR_HD uint32_t pr_fnext(const PairCtx &c, uint32_t pairs, uint32_t f, uint32_t k) {
#if defined(__CUDA_ARCH__)
    uint32_t addr = __dp4a(__byte_perm(pairs, f, 0x4440u | k), kDpMul, c.ftbl_s);
    uint16_t v;
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
#else
    return *reinterpret_cast<const uint16_t *>(c.ftbl + (f & 0xFFu) * kRowBytes + 2u * ((pairs >> (8 * k)) & 0xFFu));
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

// bit 7 of every byte: the byte is not ASCII whitespace.  Exact for words whose bytes are all < 0x80 (the per-byte sums
// of a word with a byte >= 0x80 may carry over): groups with such bytes recompute it with pr_nonspace_any.
TR_HD uint32_t pr_nonspace(uint32_t w) {
    const uint32_t ne20 = (w ^ 0x20202020u) + 0x7F7F7F7Fu;    // bit 7: byte != 0x20
    const uint32_t gt13 = w + 0x72727272u;                    // bit 7: byte > 13
    const uint32_t ge9 = w + 0x77777777u;                     // bit 7: byte >= 9
    return ne20 & (gt13 | ~ge9) & 0x80808080u;
}
// the same for any bytes: a byte >= 0x80 is not ASCII whitespace
TR_HD uint32_t pr_nonspace_any(uint32_t w) { return (pr_nonspace(w & 0x7F7F7F7Fu) | w) & 0x80808080u; }
// bit 7 of every byte that equals v (exact)
TR_HD uint32_t pr_eq(uint32_t w, uint32_t v4) {
    const uint32_t t = w ^ v4;
    return ~(((t & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t) & 0x80808080u;
}
// Does the word hold a byte that can start a rune the exact automaton treats specially?  Multi-byte whitespace starts
// with C2, E1, E2 or E3; the fold runes U+017F, U+212A, U+0130 with C5, E2, C4.  (E0 is let in with E1..E3: one test.)
// Every other byte >= 0x80 - accented Latin, Cyrillic, CJK ideographs, emoji, stray continuation bytes - is to the path
// what an ordinary letter is: not whitespace, part of no fold.
TR_HD uint32_t pr_special(uint32_t w) {
    return pr_eq(w, 0xC2C2C2C2u) | pr_eq(w, 0xC4C4C4C4u) | pr_eq(w, 0xC5C5C5C5u) | pr_eq(w & 0xFCFCFCFCu, 0xE0E0E0E0u);
}
TR_HD uint32_t pr_funnel8(uint32_t lo, uint32_t hi) {          // (hi << 8) | (lo >> 24)
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, 8);
#else
    return (hi << 8) | (lo >> 24);
#endif
}
TR_HD uint32_t pr_bytesum(uint32_t v) {                        // sum of the four bytes
#if defined(__CUDA_ARCH__)
    return __vsadu4(v, 0u);
#else
    return (v & 0xFFu) + ((v >> 8) & 0xFFu) + ((v >> 16) & 0xFFu) + (v >> 24);
#endif
}
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
    uint32_t f;            // filter state (entry: low byte)
    uint32_t prev_ns;      // non-space bits of the previous word (bit 31: the byte before the group)
    uint32_t prev_rare;    // the previous group held a byte >= 0x80
    Quad prevq;            // the previous 16 bytes (warm-up of an exact walk)
};

// The exact walk of one flagged group: state from the 16 bytes before it, then words and events of its 16 bytes.
template <class Ctx>
TR_COLD void pair_replay(const Ctx &c, Quad warm, Quad q, uint32_t msg) {   // quads by value: they stay in registers at the call
    uint32_t e = c.neutral_state, words = 0;
    const uint32_t ww[4] = {warm.a, warm.b, warm.c, warm.d}, qw[4] = {q.a, q.b, q.c, q.d};
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int i = 0; i < 16; i++) e = pr_xnext(c, e, (ww[i >> 2] >> (8 * (i & 3))) & 0xFFu);
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int i = 0; i < 16; i++) {
        e = pr_xnext(c, e, (qw[i >> 2] >> (8 * (i & 3))) & 0xFFu);
        words += e >> 15;
        if (e & 0x7F00u) apply_event_raw(c.events, c.counters, (e >> 8) & 0x7Fu, msg);
    }
    if (words) tr_add(c.counters + (uint64_t)msg * CNT_STRIDE + CNT_WORDS, words);
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
    P.prev_rare = pr_special(warm.a) | pr_special(warm.b) | pr_special(warm.c) | pr_special(warm.d);
    P.prevq = warm;
}

// one 16-byte group at tile-relative position pos.  prevp: where the 16 bytes before the group can still be read (the
// staging buffer, inside a window); nullptr: they are P.prevq, which the caller then keeps up to date (KEEP: this function
// does it after every group).  Keeping them out of registers inside a window saves a reload of the group per iteration.
template <bool SIMPLE, bool KEEP = true, class Ctx, class Sink>
TR_HD bool pair_group(const Ctx &c, PairLane &P, const Quad &q, int32_t pos, Sink &sink, const Quad *prevp = nullptr) {
    long_boundary(c, P.L, pos);
    const uint32_t high = (q.a | q.b | q.c | q.d) & 0x80808080u;
    const uint32_t qw[4] = {q.a, q.b, q.c, q.d};
    const uint32_t pns0 = P.prev_ns;
    uint32_t f = P.f, fired = 0, starts = 0, pns = pns0, rare = 0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int j = 0; j < 4; j++) {
        const uint32_t w = qw[j];
        const uint32_t ns = pr_nonspace(w);
        const uint32_t before = pr_funnel8(pns, ns);           // non-space bit of the byte before each byte
        starts += pr_bytesum(ns & ~before);                    // 128 per word start
        pns = ns;
        const uint32_t p = pr_pairs<SIMPLE>(c, w);
        f = pr_fnext(c, p, f, 0); fired |= f;
        f = pr_fnext(c, p, f, 2); fired |= f;
    }
    if (high) {   // bytes >= 0x80 in the group: is one of the few lead bytes that matter among them?
        rare = pr_special(q.a) | pr_special(q.b) | pr_special(q.c) | pr_special(q.d);
        if (!rare) {   // other scripts: the group is counted here after all, with space bits that are exact for any byte
            starts = 0; pns = pns0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int j = 0; j < 4; j++) {
                const uint32_t ns = pr_nonspace_any(qw[j]);
                starts += pr_bytesum(ns & ~pr_funnel8(pns, ns));
                pns = ns;
            }
        }
        // (rare: this group and the next are walked exactly; the "byte before" bit handed on is exact whenever the group
        // after those uses it, because that needs the last word of the next group to be free of such bytes)
    }
    P.f = f;
    P.prev_ns = pns;
    const bool flagged = (rare | P.prev_rare | (fired & 0x8000u)) != 0;
    if (flagged) sink.push(c, prevp ? *prevp : P.prevq, q, P.L.msg);
    else P.L.words += starts >> 7;
    P.prev_rare = rare;
    if (KEEP) P.prevq = q;
    return flagged;
}

template <class Ctx>
TR_HD void pair_end(const Ctx &c, PairLane &P) { long_end(c, P.L); }

// sink that walks a flagged group on the spot (host checks, the partial last row, message indices beyond the queue's range)
struct PairSinkNow {
    template <class Ctx>
    TR_HD void push(const Ctx &c, const Quad &warm, const Quad &q, uint32_t msg) { pair_replay(c, warm, q, msg); }
};

}  // namespace triage
