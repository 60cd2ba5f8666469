// scan_host_check.cc — TEST HARNESS ONLY (never linked into libtriage_b200.so).
// Compiles the product's dictionary compiler (csrc/dfa_build.cc) and the per-lane scan body
// (csrc/scan_core.cuh) with g++ and walks every lane of every tile sequentially, so the
// automaton and the tile/chunk/boundary logic can be checked against the oracle on a CPU box.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../llm-message-queue_b200/csrc/dfa_build.h"
#include "../../llm-message-queue_b200/csrc/scan_core.cuh"

using namespace triage;

extern "C" {

// dict: n_pat entries, literals concatenated in lits with lens[]; returns 0 ok, 4 on build error
int hostcheck_scan(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                   const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs,
                   uint32_t nt, uint32_t chunk, int use_abs_list,
                   uint32_t *counters /* n_msgs * 8 */, uint32_t *info /* states, unmin, warm, events, head_pad, tail_pad */,
                   char *err, uint32_t errcap)
{
    // use_abs_list bit 0: global-offset message list; bit 1: force the byte-at-a-time fallback lane
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty()) { std::strncpy(err, e.c_str(), errcap - 1); err[errcap - 1] = 0; return 4; }
    if (dfa.wide) { std::strncpy(err, "automaton needs the wide table format", errcap - 1); return 6; }
    info[0] = dfa.n_states; info[1] = dfa.n_states_unminimised; info[2] = dfa.warm_bytes; info[3] = (uint32_t)dfa.events.size();
    info[4] = (uint32_t)dfa.head_pad; info[5] = (uint32_t)dfa.tail_pad;
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) {
        pack_event(dfa.events[i], &ev[i].x);
    }
    // padded rows, exactly as the engine uploads them
    std::vector<uint8_t> tbl((size_t)dfa.n_states * kRowBytes, 0);
    for (uint32_t st = 0; st < dfa.n_states; st++)
        std::memcpy(tbl.data() + (size_t)st * kRowBytes, dfa.table.data() + (size_t)st * 256, 512);
    const bool fast = !(use_abs_list & 2) && dfa.head_pad >= 0 && dfa.tail_pad >= 0;
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    uint64_t total = offsets[n_msgs];
    uint64_t tile = uint64_t(nt) * chunk;
    int32_t warm = (int32_t)dfa.warm_bytes;
    std::vector<uint8_t> stage_raw(tile + warm + 64);
    uint8_t *stage = stage_raw.data();
    while (reinterpret_cast<uintptr_t>(stage) & 15) stage++;
    uint64_t ntiles = (total + tile - 1) / tile;
    for (uint64_t t = 0; t < ntiles; t++) {
        uint64_t ts = t * tile, te = std::min(total, ts + tile);
        std::memset(stage, 0xEE, tile + warm + 16);
        uint64_t cs = ts >= (uint64_t)warm ? ts - warm : 0;
        std::memcpy(stage + (cs - (ts - warm)), bytes + cs, te - cs);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (offsets[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (offsets[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) {
            int64_t r = (int64_t)offsets[m] - (int64_t)ts;
            const int64_t lim = 1 << 30;
            list.push_back((int32_t)(r < -lim ? -lim : (r > lim ? lim : r)));
        }
        LaneCtx c;
        c.tbl = tbl.data(); c.events = ev.data(); c.eom = dfa.eom.data(); c.stage = stage;
        c.list = (use_abs_list & 1) ? nullptr : list.data(); c.list_abs = offsets; c.tile_start = (int64_t)ts;
        c.list_n = (uint32_t)list.size(); c.msg_lo = mlo; c.counters = counters; c.warm = warm;
        c.start_state = dfa.start_state; c.neutral_state = dfa.neutral_state;
        c.tbl_s = 0;
        c.head_pad4 = 0x01010101u * (uint32_t)(dfa.head_pad & 0xFF); c.tail_pad4 = 0x01010101u * (uint32_t)(dfa.tail_pad & 0xFF);
        int32_t tile_len = (int32_t)(te - ts);
        uint32_t khint = 0;
        for (uint32_t l = 0; l < nt; l++) {
            int32_t ch = (int32_t)(l * chunk);
            if (ch >= tile_len) break;
            int32_t ce = std::min<int32_t>(ch + (int32_t)chunk, tile_len);
            uint32_t hm = 0, hw = 0;
            if ((l & 31) == 0) khint = find_msg(c, ch);           // the warp's first lane searches, the others walk forward
            uint32_t k = find_msg_from(c, ch, khint);
            if (fast) lane_scan_fast(c, ch, ce, k, &hm, &hw);
            else lane_scan_slow(c, ch, ce, k, &hm, &hw);
            if (hw) counters[(uint64_t)hm * 8 + CNT_WORDS] += hw;
        }
    }
    return 0;
}

int hostcheck_parse(const char *s, uint32_t len, int *ci, char *lit, uint32_t cap, uint32_t *lit_len)
{
    bool c; std::string l, e;
    if (!parse_regex_literal(s, len, &c, &l, &e)) return 4;
    if (l.size() > cap) return 5;
    *ci = c; std::memcpy(lit, l.data(), l.size()); *lit_len = (uint32_t)l.size();
    return 0;
}

int hostcheck_bordered(const uint8_t *lit, uint32_t len, int ci)
{
    return literal_is_bordered(std::string(reinterpret_cast<const char *>(lit), len), ci != 0) ? 1 : 0;
}
}

// exports the compiled D8 table (for offline layout / bank-conflict studies)
extern "C" int hostcheck_table(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                               uint16_t *table /* 256*256 */, uint32_t *info /* states, start, neutral, warm */)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty() || dfa.wide) return 4;
    std::memcpy(table, dfa.table.data(), dfa.table.size() * 2);
    info[0] = dfa.n_states; info[1] = dfa.start_state; info[2] = dfa.neutral_state; info[3] = dfa.warm_bytes;
    return 0;
}

// padded-arena variant: re-packs the input the way the engine's packer does (16-byte aligned starts, 1..16 reset pads
// after every message), then walks warp mini-tiles of 32*chunk bytes with lane_scan_padded
extern "C" int hostcheck_scan_padded(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                                     const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t chunk, int use_abs_list,
                                     uint32_t *counters, uint32_t *info, char *err, uint32_t errcap)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty()) { std::strncpy(err, e.c_str(), errcap - 1); err[errcap - 1] = 0; return 4; }
    if (dfa.wide) { std::strncpy(err, "automaton needs the wide table format", errcap - 1); return 6; }
    info[0] = dfa.n_states; info[1] = (uint32_t)dfa.reset_pad;
    if (dfa.reset_pad < 0) { std::strncpy(err, "no reset pad", errcap - 1); return 5; }
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) {
        pack_event(dfa.events[i], &ev[i].x);
    }
    std::vector<uint8_t> tbl((size_t)dfa.n_states * kRowBytes, 0);
    for (uint32_t st = 0; st < dfa.n_states; st++)
        std::memcpy(tbl.data() + (size_t)st * kRowBytes, dfa.table.data() + (size_t)st * 256, 512);
    // pack
    std::vector<uint64_t> poff(n_msgs + 1, 0);
    for (uint32_t i = 0; i < n_msgs; i++) poff[i + 1] = poff[i] + ((offsets[i + 1] - offsets[i]) / 16 + 1) * 16;
    std::vector<uint8_t> arena(poff[n_msgs] + 32, (uint8_t)dfa.reset_pad);
    for (uint32_t i = 0; i < n_msgs; i++) std::memcpy(arena.data() + poff[i], bytes + offsets[i], offsets[i + 1] - offsets[i]);
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    const uint64_t total = poff[n_msgs];
    const uint64_t mt = 32ull * chunk;
    const int32_t warm = (int32_t)dfa.warm_bytes;
    std::vector<uint8_t> stage_raw(mt + warm + 64);
    uint8_t *stage = stage_raw.data();
    while (reinterpret_cast<uintptr_t>(stage) & 15) stage++;
    for (uint64_t ts = 0; ts < total; ts += mt) {
        uint64_t te = std::min(total, ts + mt);
        std::memset(stage, 0xEE, mt + warm + 16);
        uint64_t cs = ts >= (uint64_t)warm ? ts - warm : 0;
        std::memcpy(stage + (cs - (ts - warm)), arena.data() + cs, te - cs);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (poff[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (poff[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) list.push_back((int32_t)((int64_t)poff[m] - (int64_t)ts));
        LaneCtx c;
        c.tbl = tbl.data(); c.events = ev.data(); c.eom = dfa.eom.data(); c.stage = stage;
        c.list = use_abs_list ? nullptr : list.data(); c.list_abs = poff.data(); c.tile_start = (int64_t)ts;
        c.list_n = (uint32_t)list.size(); c.msg_lo = mlo; c.counters = counters; c.warm = warm;
        c.start_state = dfa.start_state; c.neutral_state = dfa.neutral_state; c.tbl_s = 0; c.head_pad4 = c.tail_pad4 = 0;
        int32_t len = (int32_t)(te - ts);
        for (uint32_t l = 0; l < 32; l++) {
            int32_t ch = (int32_t)(l * chunk);
            if (ch >= len) break;
            int32_t ce = std::min<int32_t>(ch + (int32_t)chunk, len);
            uint32_t hm = 0, hw = 0;
            lane_scan_padded(c, ch, ce, find_msg(c, ch), ts == 0 && l == 0, &hm, &hw);
            if (hw) counters[(uint64_t)hm * 8 + CNT_WORDS] += hw;
        }
    }
    // HasSuffix "?" is applied by the finalize kernel in the padded pipeline
    for (uint32_t i = 0; i < n_msgs; i++)
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') counters[(uint64_t)i * 8 + CNT_Q] |= 1;
    return 0;
}


// sequential walk of every message through the EXACT (non-overlapping) automaton: what the device fix-up kernel does
extern "C" int hostcheck_exact(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                               const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t *counters, uint32_t *info)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa x;
    if (!build_dfa(dict, &x, true).empty() || x.wide) return 4;
    info[0] = x.n_states; info[1] = dict_has_bordered(dict) ? 1 : 0;
    for (uint32_t i = 0; i < n_msgs; i++) {
        uint32_t *acc = counters + (size_t)i * 8;
        std::memset(acc, 0, 32);
        uint32_t st = x.start_state;
        auto event = [&](uint32_t id) {
            const EventVec &v = x.events[id];
            for (int k = 0; k < 4; k++) acc[CNT_KW + k] += v.kw[k];
            acc[CNT_POS] += v.pos; acc[CNT_NEG] += v.neg; if (v.question) acc[CNT_Q] |= 1;
        };
        for (uint64_t p = offsets[i]; p < offsets[i + 1]; p++) {
            uint16_t e = x.table[(size_t)st * 256 + bytes[p]];
            st = e & 0xFF; acc[CNT_WORDS] += e >> 15;
            if (e & 0x7F00) event((e >> 8) & 0x7F);
        }
        uint8_t eo = x.eom[st];
        acc[CNT_WORDS] += eo >> 7;
        if (eo & 0x7F) event(eo & 0x7F);
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') acc[CNT_Q] |= 1;
    }
    return 0;
}

// wide-format check: padded arena + lane_scan_padded<WideCtx> over warp mini-tiles (what scan_wide_kernel runs), and,
// when the dictionary has self-overlapping literals, the sequential exact recount of flagged messages
#include "../../llm-message-queue_b200/csrc/scan_wide.cuh"
// exact recount of flagged messages (self-overlapping literals), what fixup_kernel does with the wide format
static int wide_exact_recount(const std::vector<DictEntry> &dict, const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs,
                              uint32_t *counters, char *err, uint32_t errcap)
{
    if (!dict_has_bordered(dict)) return 0;
    Dfa x;
    std::string e2 = build_dfa(dict, &x, true);
    if (!e2.empty() || !x.wide) { std::strncpy(err, "exact automaton", errcap - 1); return 7; }
    for (uint32_t i = 0; i < n_msgs; i++) {
        uint32_t *acc = counters + (size_t)i * 8;
        if (!(acc[CNT_Q] & 2)) continue;
        std::memset(acc, 0, 32);
        uint32_t st = x.start_state_w;
        auto event = [&](uint32_t id) {
            const EventVec &v = x.events[id];
            for (int k = 0; k < 4; k++) acc[CNT_KW + k] += v.kw[k];
            acc[CNT_POS] += v.pos; acc[CNT_NEG] += v.neg; if (v.question) acc[CNT_Q] |= 1;
        };
        for (uint64_t p = offsets[i]; p < offsets[i + 1]; p++) {
            uint32_t t = x.wtable[(size_t)st * x.n_classes + x.cls_of[bytes[p]]];
            st = t & 0xFFFFF; acc[CNT_WORDS] += t >> 31;
            if (t & 0x7FF00000u) event((t >> 20) & 0x7FF);
        }
        uint32_t eo = x.eom_w[st];
        acc[CNT_WORDS] += eo >> 31;
        if (eo & 0x7FFFFFFF) event(eo & 0x7FFFFFFF);
    }
    return 0;
}
extern "C" int hostcheck_scan_wide(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                                   const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t chunk,
                                   uint32_t *counters, uint32_t *info, char *err, uint32_t errcap)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty()) { std::strncpy(err, e.c_str(), errcap - 1); err[errcap - 1] = 0; return 4; }
    info[0] = dfa.n_states; info[1] = (uint32_t)dfa.reset_pad; info[2] = dfa.wide; info[3] = dfa.n_classes; info[4] = dfa.warm_bytes;
    info[5] = (uint32_t)dfa.events.size();
    if (!dfa.wide || dfa.reset_pad < 0) { std::strncpy(err, "not a wide automaton / no reset pad", errcap - 1); return 5; }
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) pack_event(dfa.events[i], &ev[i].x);
    std::vector<uint64_t> poff(n_msgs + 1, 0);
    for (uint32_t i = 0; i < n_msgs; i++) poff[i + 1] = poff[i] + ((offsets[i + 1] - offsets[i]) / 16 + 1) * 16;
    std::vector<uint8_t> arena(poff[n_msgs] + 32, (uint8_t)dfa.reset_pad);
    for (uint32_t i = 0; i < n_msgs; i++) std::memcpy(arena.data() + poff[i], bytes + offsets[i], offsets[i + 1] - offsets[i]);
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    const uint64_t total = poff[n_msgs];
    const uint64_t mt = 32ull * chunk;
    const int32_t warm = (int32_t)dfa.warm_bytes;
    std::vector<uint8_t> stage_raw(mt + warm + 64);
    uint8_t *stage = stage_raw.data();
    while (reinterpret_cast<uintptr_t>(stage) & 15) stage++;
    for (uint64_t ts = 0; ts < total; ts += mt) {
        uint64_t te = std::min(total, ts + mt);
        uint64_t cs = ts >= (uint64_t)warm ? ts - warm : 0;
        std::memcpy(stage + (cs - (ts - warm)), arena.data() + cs, te - cs);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (poff[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (poff[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) list.push_back((int32_t)((int64_t)poff[m] - (int64_t)ts));
        WideCtx c;
        c.wtbl = dfa.wtable.data(); c.cls = dfa.cls_of; c.ncls = dfa.n_classes; c.events = ev.data(); c.stage = stage;
        c.list = list.data(); c.list_abs = poff.data(); c.tile_start = (int64_t)ts; c.list_n = (uint32_t)list.size(); c.msg_lo = mlo;
        c.counters = counters; c.warm = warm; c.start_state = dfa.start_state_w; c.neutral_state = dfa.neutral_state_w;
        int32_t len = (int32_t)(te - ts);
        for (uint32_t l = 0; l < 32; l++) {
            int32_t ch = (int32_t)(l * chunk);
            if (ch >= len) break;
            int32_t ce = std::min<int32_t>(ch + (int32_t)chunk, len);
            uint32_t hm = 0, hw = 0;
            lane_scan_padded(c, ch, ce, find_msg(c, ch), ts == 0 && l == 0, &hm, &hw);
            if (hw) counters[(uint64_t)hm * 8 + CNT_WORDS] += hw;
        }
    }
    if (int rc = wide_exact_recount(dict, offsets, bytes, n_msgs, counters, err, errcap)) return rc;
    for (uint32_t i = 0; i < n_msgs; i++) {
        counters[(uint64_t)i * 8 + CNT_Q] &= 1;
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') counters[(uint64_t)i * 8 + CNT_Q] |= 1;
    }
    return 0;
}

// long-lane variant: 32 lanes x chunk_long bytes per warp iteration, streamed in windows of `window` bytes
extern "C" int hostcheck_scan_long(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                                   const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t chunk_long, uint32_t window,
                                   uint32_t *counters, char *err, uint32_t errcap)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty() || dfa.wide || dfa.reset_pad < 0) { std::strncpy(err, "unsupported", errcap - 1); return 4; }
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) pack_event(dfa.events[i], &ev[i].x);
    std::vector<uint8_t> tbl((size_t)dfa.n_states * kRowBytes, 0);
    for (uint32_t st = 0; st < dfa.n_states; st++)
        std::memcpy(tbl.data() + (size_t)st * kRowBytes, dfa.table.data() + (size_t)st * 256, 512);
    std::vector<uint64_t> poff(n_msgs + 1, 0);
    for (uint32_t i = 0; i < n_msgs; i++) poff[i + 1] = poff[i] + ((offsets[i + 1] - offsets[i]) / 16 + 1) * 16;
    std::vector<uint8_t> arena_raw(poff[n_msgs] + 64, (uint8_t)dfa.reset_pad);
    uint8_t *arena = arena_raw.data();
    while (reinterpret_cast<uintptr_t>(arena) & 15) arena++;
    for (uint32_t i = 0; i < n_msgs; i++) std::memcpy(arena + poff[i], bytes + offsets[i], offsets[i + 1] - offsets[i]);
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    const uint64_t total = poff[n_msgs];
    const uint64_t st_bytes = 32ull * chunk_long;
    const int32_t warm = (int32_t)dfa.warm_bytes;
    for (uint64_t ts = 0; ts < total; ts += st_bytes) {
        uint64_t te = std::min(total, ts + st_bytes);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (poff[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (poff[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) list.push_back((int32_t)((int64_t)poff[m] - (int64_t)ts));
        LaneCtx c;
        c.tbl = tbl.data(); c.events = ev.data(); c.eom = nullptr; c.stage = nullptr;
        c.list = list.data(); c.list_abs = poff.data(); c.tile_start = (int64_t)ts; c.list_n = (uint32_t)list.size(); c.msg_lo = mlo;
        c.counters = counters; c.warm = warm; c.start_state = dfa.start_state; c.neutral_state = dfa.neutral_state; c.tbl_s = 0;
        c.head_pad4 = c.tail_pad4 = 0;
        for (uint32_t l = 0; l < 32; l++) {
            uint64_t cs = ts + (uint64_t)l * chunk_long;
            if (cs >= te) break;
            uint64_t ce = std::min<uint64_t>(cs + chunk_long, te);
            LongLane L;
            long_begin(c, L, find_msg(c, (int32_t)(cs - ts)), cs == 0, cs ? arena + cs - warm : arena);
            for (uint64_t w = cs; w < ce; w += window)
                long_window(c, L, arena + w, (int32_t)(w - ts), (int32_t)std::min<uint64_t>(window, ce - w));
            long_end(c, L);
            if (L.head_words) counters[(uint64_t)L.head_msg * 8 + CNT_WORDS] += L.head_words;
        }
    }
    for (uint32_t i = 0; i < n_msgs; i++) {
        counters[(uint64_t)i * 8 + CNT_Q] &= 1;
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') counters[(uint64_t)i * 8 + CNT_Q] |= 1;
    }
    return 0;
}

// streaming variant: a lane owns `nch` chunks of chunk_long bytes (chunk v = ch * 32 + lane of the super-tile) and walks
// them interleaved with long_stream_n; a chunk cut short by the end of the arena is walked alone
template <int N, class Ctx>
static void stream_lane(const Ctx &c, const uint8_t *arena, uint64_t ts, uint64_t te, uint32_t l, uint32_t chunk_long, int32_t warm,
                        uint32_t *counters)
{
    LongLane L[N];
    const uint8_t *src[N];
    int32_t pos[N], nb[N];
    bool all_full = true;
    for (int ch = 0; ch < N; ch++) {
        uint64_t cs = ts + (uint64_t)(ch * 32 + l) * chunk_long;
        nb[ch] = cs >= te ? 0 : (int32_t)std::min<uint64_t>(chunk_long, te - cs);
        all_full = all_full && nb[ch] == (int32_t)chunk_long;
        src[ch] = arena + cs;
        pos[ch] = (int32_t)(cs - ts);
        if (nb[ch]) long_begin(c, L[ch], find_msg(c, pos[ch]), cs == 0, cs ? arena + cs - warm : arena);
    }
    if (all_full) {
        long_stream_n<N>(c, L, src, pos, (int32_t)chunk_long, arena + te);
    } else {
        for (int ch = 0; ch < N; ch++) {
            if (!nb[ch]) continue;
            LongLane L1[1] = {L[ch]};
            const uint8_t *s1[1] = {src[ch]};
            int32_t p1[1] = {pos[ch]};
            long_stream_n<1>(c, L1, s1, p1, nb[ch], arena + te);
            L[ch] = L1[0];
        }
    }
    for (int ch = 0; ch < N; ch++) {
        if (!nb[ch]) continue;
        long_end(c, L[ch]);
        if (L[ch].head_words) counters[(uint64_t)L[ch].head_msg * 8 + CNT_WORDS] += L[ch].head_words;
    }
}

extern "C" int hostcheck_scan_stream(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                                     const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t chunk_long, uint32_t nch,
                                     uint32_t *counters, char *err, uint32_t errcap)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty() || dfa.wide || dfa.reset_pad < 0 || nch < 1 || nch > 4) { std::strncpy(err, "unsupported", errcap - 1); return 4; }
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) pack_event(dfa.events[i], &ev[i].x);
    std::vector<uint8_t> tbl((size_t)dfa.n_states * kRowBytes, 0);
    for (uint32_t st = 0; st < dfa.n_states; st++)
        std::memcpy(tbl.data() + (size_t)st * kRowBytes, dfa.table.data() + (size_t)st * 256, 512);
    std::vector<uint64_t> poff(n_msgs + 1, 0);
    for (uint32_t i = 0; i < n_msgs; i++) poff[i + 1] = poff[i] + ((offsets[i + 1] - offsets[i]) / 16 + 1) * 16;
    std::vector<uint8_t> arena_raw(poff[n_msgs] + 64, (uint8_t)dfa.reset_pad);
    uint8_t *arena = arena_raw.data();
    while (reinterpret_cast<uintptr_t>(arena) & 15) arena++;
    for (uint32_t i = 0; i < n_msgs; i++) std::memcpy(arena + poff[i], bytes + offsets[i], offsets[i + 1] - offsets[i]);
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    const uint64_t total = poff[n_msgs];
    const uint64_t st_bytes = 32ull * chunk_long * nch;
    const int32_t warm = (int32_t)dfa.warm_bytes;
    for (uint64_t ts = 0; ts < total; ts += st_bytes) {
        uint64_t te = std::min(total, ts + st_bytes);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (poff[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (poff[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) list.push_back((int32_t)((int64_t)poff[m] - (int64_t)ts));
        LaneCtx c;
        c.tbl = tbl.data(); c.events = ev.data(); c.eom = nullptr; c.stage = nullptr;
        c.list = list.data(); c.list_abs = poff.data(); c.tile_start = (int64_t)ts; c.list_n = (uint32_t)list.size(); c.msg_lo = mlo;
        c.counters = counters; c.warm = warm; c.start_state = dfa.start_state; c.neutral_state = dfa.neutral_state; c.tbl_s = 0;
        c.head_pad4 = c.tail_pad4 = 0;
        for (uint32_t l = 0; l < 32; l++) {
            switch (nch) {
                case 1: stream_lane<1>(c, arena, ts, te, l, chunk_long, warm, counters); break;
                case 2: stream_lane<2>(c, arena, ts, te, l, chunk_long, warm, counters); break;
                case 3: stream_lane<3>(c, arena, ts, te, l, chunk_long, warm, counters); break;
                default: stream_lane<4>(c, arena, ts, te, l, chunk_long, warm, counters); break;
            }
        }
    }
    for (uint32_t i = 0; i < n_msgs; i++) {
        counters[(uint64_t)i * 8 + CNT_Q] &= 1;
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') counters[(uint64_t)i * 8 + CNT_Q] |= 1;
    }
    return 0;
}


// wide format, rows streamed from global memory (scan_stream_wide_kernel): lanes own `nch` rows of chunk_long bytes of a
// super-tile and walk them interleaved, events applied where they are met (InlineEvents<WideCtx>)
extern "C" int hostcheck_scan_stream_wide(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                                          const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t chunk_long, uint32_t nch,
                                          uint32_t *counters, char *err, uint32_t errcap)
{
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty() || !dfa.wide || dfa.reset_pad < 0 || nch < 1 || nch > 4 || (chunk_long & 15)) { std::strncpy(err, "unsupported", errcap - 1); return 4; }
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) pack_event(dfa.events[i], &ev[i].x);
    std::vector<uint64_t> poff(n_msgs + 1, 0);
    for (uint32_t i = 0; i < n_msgs; i++) poff[i + 1] = poff[i] + ((offsets[i + 1] - offsets[i]) / 16 + 1) * 16;
    std::vector<uint8_t> arena_raw(poff[n_msgs] + 64, (uint8_t)dfa.reset_pad);
    uint8_t *arena = arena_raw.data();
    while (reinterpret_cast<uintptr_t>(arena) & 15) arena++;
    for (uint32_t i = 0; i < n_msgs; i++) std::memcpy(arena + poff[i], bytes + offsets[i], offsets[i + 1] - offsets[i]);
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    const uint64_t total = poff[n_msgs];
    const uint64_t st_bytes = 32ull * chunk_long * nch;
    const int32_t warm = (int32_t)dfa.warm_bytes;
    for (uint64_t ts = 0; ts < total; ts += st_bytes) {
        uint64_t te = std::min(total, ts + st_bytes);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (poff[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (poff[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) list.push_back((int32_t)((int64_t)poff[m] - (int64_t)ts));
        WideCtx c;
        c.wtbl = dfa.wtable.data(); c.cls = dfa.cls_of; c.ncls = dfa.n_classes; c.events = ev.data(); c.stage = nullptr;
        c.list = list.data(); c.list_abs = poff.data(); c.tile_start = (int64_t)ts; c.list_n = (uint32_t)list.size(); c.msg_lo = mlo;
        c.counters = counters; c.warm = warm; c.start_state = dfa.start_state_w; c.neutral_state = dfa.neutral_state_w;
        for (uint32_t l = 0; l < 32; l++) {
            switch (nch) {
                case 1: stream_lane<1>(c, arena, ts, te, l, chunk_long, warm, counters); break;
                case 2: stream_lane<2>(c, arena, ts, te, l, chunk_long, warm, counters); break;
                case 3: stream_lane<3>(c, arena, ts, te, l, chunk_long, warm, counters); break;
                default: stream_lane<4>(c, arena, ts, te, l, chunk_long, warm, counters); break;
            }
        }
    }
    if (int rc = wide_exact_recount(dict, offsets, bytes, n_msgs, counters, err, errcap)) return rc;
    for (uint32_t i = 0; i < n_msgs; i++) {
        counters[(uint64_t)i * 8 + CNT_Q] &= 1;
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') counters[(uint64_t)i * 8 + CNT_Q] |= 1;
    }
    return 0;
}

// ---- AddKeywordPattern regex subset: compile + the host build of the device walk (regex_walk.cuh) ----
#include "../../llm-message-queue_b200/csrc/regex_vm.h"
// returns 0 and the count, or 4 and the reason the pattern is outside the subset
extern "C" int hostcheck_regex_count(const char *pattern, uint32_t plen, const uint8_t *s, uint64_t n, uint32_t *count, char *err, uint32_t err_cap)
{
    triage::RegexSet set;
    std::string e = triage::regex_compile(pattern, plen, 1, &set);
    if (!e.empty()) { snprintf(err, err_cap, "%s", e.c_str()); return 4; }
    *count = triage::regex_count_host(set, 0, s, n);
    return 0;
}

// ---- two-bytes-per-lookup lane (scan_pair.cuh): rows of chunk_long bytes walked group by group, flagged groups through
//      the exact automaton on the spot; info[0] = filter states, [1] = mul, [2] = shift, [3] = flagged groups, [4] = groups ----
#include "../../llm-message-queue_b200/csrc/scan_pair.cuh"
namespace {
struct CountingSink {
    uint64_t flagged = 0;
    void push(const triage::PairCtx &c, const triage::Quad &warm, const triage::Quad &q, uint32_t msg, int32_t, bool count_words) { flagged++; triage::pair_replay(c, warm, q, msg, count_words); }
};
}
extern "C" int hostcheck_scan_pair(const uint8_t *prios, const uint8_t *cis, const uint32_t *lens, const uint8_t *lits, uint32_t n_pat,
                                   const uint64_t *offsets, const uint8_t *bytes, uint32_t n_msgs, uint32_t chunk_long,
                                   uint32_t *counters, uint64_t *info, char *err, uint32_t errcap)
{
    using namespace triage;
    std::vector<DictEntry> dict;
    size_t o = 0;
    for (uint32_t i = 0; i < n_pat; i++) {
        dict.push_back({prios[i], cis[i] != 0, std::string(reinterpret_cast<const char *>(lits + o), lens[i])});
        o += lens[i];
    }
    Dfa dfa;
    std::string e = build_dfa(dict, &dfa);
    if (!e.empty() || !dfa.filter.ok) { std::strncpy(err, e.empty() ? "no pair filter for this dictionary" : e.c_str(), errcap - 1); return 4; }
    info[0] = dfa.filter.m_states | ((uint64_t)dfa.n_cmp << 32); info[1] = dfa.filter.mul; info[2] = dfa.filter.shift; info[3] = info[4] = 0;
    if (n_msgs == 0) return 0;
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) pack_event(dfa.events[i], &ev[i].x);
    std::vector<uint8_t> xtbl((size_t)dfa.n_states * kRowBytes, 0);
    for (uint32_t st = 0; st < dfa.n_states; st++) std::memcpy(xtbl.data() + (size_t)st * kRowBytes, dfa.table.data() + (size_t)st * 256, 512);
    std::vector<uint64_t> poff(n_msgs + 1, 0);
    for (uint32_t i = 0; i < n_msgs; i++) poff[i + 1] = poff[i] + ((offsets[i + 1] - offsets[i]) / 16 + 1) * 16;
    std::vector<uint8_t> arena_raw(poff[n_msgs] + 64, (uint8_t)dfa.reset_pad);
    uint8_t *arena = arena_raw.data();
    while (reinterpret_cast<uintptr_t>(arena) & 15) arena++;
    for (uint32_t i = 0; i < n_msgs; i++) std::memcpy(arena + poff[i], bytes + offsets[i], offsets[i + 1] - offsets[i]);
    std::memset(counters, 0, sizeof(uint32_t) * 8 * n_msgs);
    const uint64_t total = poff[n_msgs];
    const uint64_t st_bytes = 32ull * chunk_long;
    CountingSink sink;
    const bool simple = dfa.filter.mul == 1 && dfa.filter.shift == 0;
    Quad padq;
    padq.a = padq.b = padq.c = padq.d = 0x01010101u * (uint32_t)(dfa.reset_pad & 0xFF);
    for (uint64_t ts = 0; ts < total; ts += st_bytes) {
        uint64_t te = std::min(total, ts + st_bytes);
        uint32_t lo = 0, hi = n_msgs + 1;
        while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (poff[mid] > ts) hi = mid; else lo = mid + 1; }
        uint32_t mlo = lo - 1, mhi = mlo;
        while (poff[mhi + 1] < te) mhi++;
        std::vector<int32_t> list;
        for (uint32_t m = mlo; m <= mhi + 1; m++) list.push_back((int32_t)((int64_t)poff[m] - (int64_t)ts));
        PairCtx c;
        c.ftbl = dfa.filter.mtable.data(); c.fstart = dfa.filter.m_start; c.ffired = dfa.filter.m_fired_from; c.mul = dfa.filter.mul; c.shift = dfa.filter.shift;
        c.xtbl = dfa.ctable.data(); c.xcls = dfa.ccls; c.xn = dfa.n_cmp; c.events = ev.data(); c.neutral_state = dfa.neutral_state; c.pad4 = padq.a;
        c.list = list.data(); c.list_abs = poff.data(); c.tile_start = (int64_t)ts; c.list_n = (uint32_t)list.size(); c.msg_lo = mlo;
        c.counters = counters; c.arena = nullptr;
        for (uint32_t l = 0; l < 32; l++) {
            uint64_t cs = ts + (uint64_t)l * chunk_long;
            if (cs >= te) break;
            uint64_t ce = std::min<uint64_t>(cs + chunk_long, te);
            PairLane P;
            Quad warm = cs ? *reinterpret_cast<const Quad *>(arena + cs - 16) : padq;
            if (simple) pair_begin<true>(c, P, find_msg(c, (int32_t)(cs - ts)), warm); else pair_begin<false>(c, P, find_msg(c, (int32_t)(cs - ts)), warm);
            // whole 80-byte windows the way the kernel takes them, the rest group by group
            uint64_t g = cs;
            for (; g + 80 <= ce; g += 80) {
                const Quad *win = reinterpret_cast<const Quad *>(arena + g);
                if (simple) pair_window<true, 5>(c, P, win, (int32_t)(g - ts), sink); else pair_window<false, 5>(c, P, win, (int32_t)(g - ts), sink);
                info[4] += 5;
            }
            for (; g < ce; g += 16) {
                const Quad q = *reinterpret_cast<const Quad *>(arena + g);
                if (simple) pair_group<true>(c, P, q, (int32_t)(g - ts), sink); else pair_group<false>(c, P, q, (int32_t)(g - ts), sink);
                info[4]++;
            }
            pair_end(c, P);
            if (P.L.head_words) counters[(uint64_t)P.L.head_msg * 8 + CNT_WORDS] += P.L.head_words;
        }
    }
    info[3] = sink.flagged;
    for (uint32_t i = 0; i < n_msgs; i++) {
        counters[(uint64_t)i * 8 + CNT_Q] &= 1;
        if (offsets[i + 1] > offsets[i] && bytes[offsets[i + 1] - 1] == '?') counters[(uint64_t)i * 8 + CNT_Q] |= 1;
    }
    return 0;
}

// ---- JSON ingest / analysis emit (ingest.cc), engine-free ----
#include "../../llm-message-queue_b200/csrc/ingest.h"
extern "C" int hostcheck_ingest(const uint8_t *bodies, const uint64_t *body_offsets, uint32_t n, int pad, const char *vip_user, int vip_prio,
                                uint8_t *arena, uint64_t cap, uint64_t *offsets, uint32_t *lengths, triage_meta *meta, uint8_t *status,
                                uint64_t *total, char *err, uint32_t errcap)
{
    struct Ctx { std::string user; int prio; } ctx{vip_user ? vip_user : "", vip_prio};
    auto lookup = [](const void *c, const char *id, size_t len) -> int {
        const Ctx *x = static_cast<const Ctx *>(c);
        return (!x->user.empty() && x->user == std::string(id, len)) ? x->prio : 0;
    };
    triage::IngestOut out{arena, cap, offsets, lengths, meta, status};
    std::string e = triage::ingest_json(bodies, body_offsets, n, pad, lookup, &ctx, out, total);
    if (!e.empty()) { snprintf(err, errcap, "%s", e.c_str()); return 5; }
    return 0;
}
extern "C" uint64_t hostcheck_emit(const triage_result *res, uint32_t n, char *out, uint64_t cap, uint64_t *offsets)
{
    return triage::emit_analysis_json(res, n, out, cap, offsets);
}
