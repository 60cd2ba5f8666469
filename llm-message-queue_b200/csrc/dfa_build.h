// dfa_build.h — host-side dictionary compiler: keyword literals + the reference's fixed
// sentiment / question word lists + Go whitespace -> one byte-level Mealy DFA.
//
// What it encodes (reference: internal/preprocessor/preprocessor.go):
//   * keyword patterns, substring semantics, "(?i)" folding with regexp's SimpleFold orbits
//     (s ~ U+017F, k ~ U+212A)                                              :28-43, :144-149
//   * strings.Fields word starts over unicode.IsSpace                       :204-205
//   * whole-field equality with the 5+5 sentiment words after strings.ToLower (:211-223),
//     where U+0130 lowers to 'i' and U+212A to 'k'
//   * question words + " " as substrings of the lower-cased content         :235-243
// Every equivalence is at the byte level; lead bytes are never consumed as continuation
// bytes by Go's UTF-8 decoder, which is what makes that exact ("byte-level exactness").
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace triage {

struct DictEntry {
    uint8_t prio;       // 1..4
    bool ci;            // "(?i)" literal
    std::string bytes;  // valid UTF-8 literal
};

// counters one transition can bump (word starts are a separate bit)
struct EventVec {
    uint16_t kw[4];  // realtime, high, normal, low keyword matches ending on this byte
    uint8_t pos, neg, question;
    uint8_t risk;    // a self-overlapping pattern matched while another of its threads was alive: occurrences may overlap,
                     // the message needs the sequential exact recount (FindAllString counts without overlap)
    bool operator==(const EventVec &o) const;
    bool empty() const;
};

// 16-byte device form of an EventVec (scan_core.cuh EventPacked): x = kw_rt | kw_hi << 16, y = kw_norm | kw_low << 16,
// z = pos | neg << 8 | question << 16 | risk << 24
inline void pack_event(const EventVec &v, uint32_t out[4]) {
    out[0] = uint32_t(v.kw[0]) | uint32_t(v.kw[1]) << 16;
    out[1] = uint32_t(v.kw[2]) | uint32_t(v.kw[3]) << 16;
    out[2] = uint32_t(v.pos) | uint32_t(v.neg) << 8 | uint32_t(v.question) << 16 | uint32_t(v.risk) << 24;
    out[3] = 0;
}

// Dense table format "D8": entry(state, byte) is 16 bits:
//   bits 0..7  next state      bits 8..14 event id (0 = none)      bit 15 word start
constexpr uint32_t kMaxStatesD8 = 256;
constexpr uint32_t kMaxEventsD8 = 127;

// Wide table format "W" (dictionaries beyond 256 states, e.g. 10 K keywords): bytes are first mapped to classes
// (bytes no pattern, whitespace rune or fixed word mentions share class 0), entry(state, class) is 32 bits:
//   bits 0..19 next state      bits 20..30 event id      bit 31 word start
constexpr uint32_t kMaxStatesW = 1u << 20;
constexpr uint32_t kMaxEventsW = 2047;

// Filter automaton of the two-bytes-per-lookup scan lane (scan_pair.cuh): bytes are mapped to 4-bit classes by
// arithmetic, ((b & 31) * mul >> shift) & 15, the automaton recognises the class strings of every all-ASCII way to
// complete an event (keyword literal, question word + " ", whitespace + sentiment word + whitespace) and consumes two
// classes per table entry: entry(state, c0 | c1 << 4) = next state | fired << 15.  It over-approximates (bytes sharing
// a class are not told apart); what it flags is re-walked through the exact automaton.
// The device walks its Moore form - "fired" moved into the state: states >= m_fired_from are entered by a firing entry -
// so that an entry is one byte and its address is state << 8 | class pair, a single byte permute.
struct PairFilter {
    bool ok = false;                  // false: this dictionary keeps to the one-lookup-per-byte lanes
    uint32_t n_states = 0, start = 0;
    uint32_t mul = 1, shift = 0;
    std::vector<uint16_t> table;      // n_states * 256
    std::vector<uint8_t> mtable;      // Moore form: m_states * 256 next states
    uint32_t m_states = 0, m_start = 0, m_fired_from = 0;
    double sample_fire_rate = 0;      // flagged 16-byte groups on the compiler's sample text (diagnostics)
};

struct Dfa {
    PairFilter filter;
    // D8 only: the same automaton with identical byte columns merged - entry(state, ccls[byte]), n_cmp entries per state.
    // Small enough to sit in shared memory next to the filter table; the pair lanes walk flagged groups through it.
    std::vector<uint16_t> ctable;
    uint8_t ccls[256] = {0};
    uint32_t n_cmp = 0;
    uint32_t n_states = 0;
    bool wide = false;
    std::vector<uint16_t> table;      // D8: n_states * 256
    std::vector<uint32_t> wtable;     // W: n_states * n_classes
    uint8_t cls_of[256] = {0};        // W: byte -> class
    uint32_t n_classes = 0;
    uint32_t start_state_w = 0, neutral_state_w = 0;   // W: state ids do not fit the 8-bit fields below
    std::vector<uint8_t> eom;         // per state: bit 7 word start, bits 0..6 event id, applied at end of message (D8)
    std::vector<uint32_t> eom_w;      // same, event id | word start << 31 (both formats)
    std::vector<EventVec> events;     // id -> counters (events[0] is all-zero)
    uint8_t start_state = 0;          // at a message start (field start, nothing pending)
    uint8_t neutral_state = 0;        // mid-field, no partial match: where a warm-up begins
    uint32_t span = 1;                // longest byte distance an event can depend on
    uint32_t warm_bytes = 16;         // span-1 rounded up to 16
    uint32_t n_states_unminimised = 0;
    // pad bytes for the scan's in-register boundary masking (-1: none exists for this dictionary)
    int head_pad = -1;   // START --head_pad--> START with no output
    int tail_pad = -1;   // from every state: reading it emits exactly that state's end-of-message output, then nothing
    int reset_pad = -1;  // tail_pad that additionally lands in START from every state: the padded arena's fill byte
};

// returns "" on success, else a message (too many states / events for D8, bordered pattern, ...)
std::string validate_literal(const std::string &bytes, bool ci);
bool literal_is_bordered(const std::string &bytes, bool ci);
// exact_nonoverlap = false: the definite automaton the chunked scan uses (counts every occurrence, raises `risk`).
// exact_nonoverlap = true: matching a self-overlapping pattern kills its other threads — FindAllString's leftmost
// non-overlapping count, only valid when a message is walked sequentially from its start (not definite).
std::string build_dfa(const std::vector<DictEntry> &dict, Dfa *out, bool exact_nonoverlap = false);
bool dict_has_bordered(const std::vector<DictEntry> &dict);

// "(?i)literal" / "literal" with backslash escapes -> DictEntry fields; false if not a literal
bool parse_regex_literal(const char *s, uint32_t len, bool *ci, std::string *lit, std::string *err);

}  // namespace triage
