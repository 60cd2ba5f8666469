// scan_wide.cuh — the wide table format ("W") for dictionaries whose automaton has more than 256 states (e.g. 10 K
// keywords): bytes map to classes through a 256-byte table in shared memory, the transition table
// (n_states x n_classes 32-bit entries: next state 20 bits | event id 11 bits | word-start bit) lives in global memory
// and is served by L1/L2 (29 MB for 10 K keywords: L2-resident on B200).  Same lane walk as the dense format
// (scan_core.cuh templates), only the accessors differ.
#pragma once
#include "scan_core.cuh"

namespace triage {

struct WideCtx {
    const uint32_t *wtbl;       // global memory
    const uint8_t *cls;         // byte -> class (shared memory on device)
    uint32_t ncls;
    const EventPacked *events;  // global memory
    const uint8_t *stage;
    const int32_t *list;
    const uint64_t *list_abs;
    int64_t tile_start;
    uint32_t list_n, msg_lo;
    uint32_t *counters;
    int32_t warm;
    uint32_t start_state, neutral_state;
};

TR_HD uint32_t tr_next(const WideCtx &c, uint32_t word, uint32_t entry, uint32_t k) {
    const uint32_t byte = (word >> (8 * k)) & 0xFFu;
    const uint32_t idx = (entry & 0xFFFFFu) * c.ncls + c.cls[byte];
#if defined(__CUDA_ARCH__)
    return __ldg(c.wtbl + idx);
#else
    return c.wtbl[idx];
#endif
}
TR_HD uint32_t tr_ws(const WideCtx &, uint32_t e) { return e >> 31; }
TR_HD uint32_t tr_evmask(const WideCtx &) { return 0x7FF00000u; }
TR_HD uint32_t tr_evid(const WideCtx &, uint32_t e) { return (e >> 20) & 0x7FFu; }
TR_HD void apply_event(const WideCtx &c, uint32_t id, uint32_t msg) { apply_event_raw(c.events, c.counters, id, msg); }
template <> struct InlineEvents<WideCtx> { static constexpr bool value = true; };

}  // namespace triage
