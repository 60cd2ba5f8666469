// regex_walk.cuh — the walk of one compiled pattern (regex_vm.h) over one message, written once for device and host.
// The device runs it in regex_count_kernel (one message per thread); tests/csrc compiles the same function with g++.
//
// Ordered-thread simulation: the thread list is kept in priority order (leftmost start first, then the order the
// pattern prefers: left alternative, greedy more / lazy fewer).  A thread that reaches MATCH records the end and cuts
// off every thread behind it; threads ahead of it keep running and may replace the match.  When the list runs dry the
// recorded match is final: count it and resume the search at its end (FindAllString's non-overlap rule).
#pragma once
#include <stdint.h>

#include "regex_vm.h"
#include "scan_core.cuh"

namespace triage {

struct RegexView {
    const RegexProgram *programs;
    const RegexInst *insts;
    const RegexClass *classes;
    const uint32_t *ranges;
    uint32_t n_programs;
};

// unicode/utf8.DecodeRune: invalid input -> (U+FFFD, width 1)
TR_HD uint32_t rx_decode(const uint8_t *p, uint64_t n, uint32_t *width) {
    const uint32_t b0 = p[0];
    *width = 1;
    if (b0 < 0x80) return b0;
    if (b0 < 0xC2 || b0 > 0xF4) return 0xFFFDu;
    if (b0 < 0xE0) {
        if (n < 2 || (p[1] & 0xC0) != 0x80) return 0xFFFDu;
        *width = 2;
        return ((b0 & 0x1Fu) << 6) | (p[1] & 0x3Fu);
    }
    if (b0 < 0xF0) {
        const uint32_t lo = b0 == 0xE0 ? 0xA0u : 0x80u, hi = b0 == 0xED ? 0x9Fu : 0xBFu;
        if (n < 3 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80) return 0xFFFDu;
        *width = 3;
        return ((b0 & 0x0Fu) << 12) | ((uint32_t)(p[1] & 0x3Fu) << 6) | (p[2] & 0x3Fu);
    }
    const uint32_t lo = b0 == 0xF0 ? 0x90u : 0x80u, hi = b0 == 0xF4 ? 0x8Fu : 0xBFu;
    if (n < 4 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80 || (p[3] & 0xC0) != 0x80) return 0xFFFDu;
    *width = 4;
    return ((b0 & 0x07u) << 18) | ((uint32_t)(p[1] & 0x3Fu) << 12) | ((uint32_t)(p[2] & 0x3Fu) << 6) | (p[3] & 0x3Fu);
}

TR_HD bool rx_class_match(const RegexView &v, uint32_t cls, uint32_t r) {
    const RegexClass &c = v.classes[cls];
    bool in;
    if (r < 128) {
        in = (c.ascii[r >> 5] >> (r & 31)) & 1u;
    } else {
        in = false;
        for (uint32_t k = 0; k < c.n_ranges; k++) {
            const uint32_t lo = v.ranges[2 * (c.first_range + k)], hi = v.ranges[2 * (c.first_range + k) + 1];
            if (r >= lo && r <= hi) { in = true; break; }
        }
    }
    return in != (c.negate != 0);
}

struct RxList { uint8_t pc[kRegexMaxInsts]; uint32_t n; };

// appends pc and everything reachable from it through SPLIT / JMP, in priority order, skipping what `seen` already has
TR_HD void rx_add(const RegexInst *prog, RxList &l, uint64_t &seen, uint32_t pc0) {
    uint8_t stack[kRegexMaxInsts];
    uint32_t sp = 0;
    stack[sp++] = (uint8_t)pc0;
    while (sp) {
        const uint32_t pc = stack[--sp];
        if ((seen >> pc) & 1ull) continue;
        seen |= 1ull << pc;
        const RegexInst in = prog[pc];
        if (in.op == ROP_JMP) { stack[sp++] = (uint8_t)in.x; }
        else if (in.op == ROP_SPLIT) { stack[sp++] = (uint8_t)in.y; stack[sp++] = (uint8_t)in.x; }   // x is popped first
        else l.pc[l.n++] = (uint8_t)pc;
    }
}

// len(re.FindAllString(s, -1)) for program `p` (patterns never match the empty string: the compiler refuses those)
TR_HD uint32_t rx_count(const RegexView &v, uint32_t p, const uint8_t *s, uint64_t n) {
    const RegexInst *prog = v.insts + v.programs[p].first_inst;
    uint32_t count = 0;
    uint64_t pos = 0;
    RxList a, b;
    while (pos < n) {
        RxList *cur = &a, *nxt = &b;
        cur->n = 0;
        int64_t matched = -1;
        uint64_t q = pos;
        for (;;) {
            if (matched < 0 && q < n) {   // a new start has the lowest priority: appended last
                uint64_t seen = 0;
                for (uint32_t i = 0; i < cur->n; i++) seen |= 1ull << cur->pc[i];
                rx_add(prog, *cur, seen, 0);
            }
            if (cur->n == 0) break;
            uint32_t w = 1, r = 0;
            if (q < n) r = rx_decode(s + q, n - q, &w);
            nxt->n = 0;
            uint64_t seen = 0;
            for (uint32_t i = 0; i < cur->n; i++) {
                const RegexInst in = prog[cur->pc[i]];
                if (in.op == ROP_MATCH) { matched = (int64_t)q; break; }   // threads behind this one lose
                if (q < n && rx_class_match(v, in.cls, r)) rx_add(prog, *nxt, seen, (uint32_t)cur->pc[i] + 1);
            }
            if (q >= n) break;
            q += w;
            RxList *t = cur; cur = nxt; nxt = t;
        }
        if (matched < 0) break;   // nothing matches from pos on
        count++;
        pos = (uint64_t)matched;
    }
    return count;
}

}  // namespace triage
