// regex_vm.cc — see regex_vm.h.  Host only: recursive-descent parser for the RE2 subset, emission of the rune-level
// program (Thompson construction with ordered SPLITs), and the host build of the walk for tests.
#include "regex_vm.h"

#include <algorithm>
#include <cstring>
#include <memory>

#define TR_HD inline
#include "regex_walk.cuh"

namespace triage {
namespace {

struct RuneSet {                       // sorted, disjoint, inclusive ranges over runes
    std::vector<std::pair<uint32_t, uint32_t>> r;
    void add(uint32_t lo, uint32_t hi) { if (lo <= hi) r.emplace_back(lo, hi); }
    void normalise() {
        std::sort(r.begin(), r.end());
        std::vector<std::pair<uint32_t, uint32_t>> o;
        for (auto &x : r) {
            if (!o.empty() && x.first <= o.back().second + 1) o.back().second = std::max(o.back().second, x.second);
            else o.push_back(x);
        }
        r.swap(o);
    }
    bool has(uint32_t c) const { for (auto &x : r) if (c >= x.first && c <= x.second) return true; return false; }
    bool has_above_ascii() const { for (auto &x : r) if (x.second >= 0x80) return true; return false; }
};

struct Node {
    enum Kind { RUNE, CAT, ALT, REP } kind = RUNE;
    RuneSet set; bool negate = false;          // RUNE
    std::vector<std::unique_ptr<Node>> kids;   // CAT, ALT, REP (one kid)
    uint32_t lo = 1, hi = 1; bool unbounded = false, lazy = false;   // REP: {lo,hi} or {lo,}
};
using NodeP = std::unique_ptr<Node>;

struct Parser {
    const char *s; uint32_t n; uint32_t i = 0;
    bool fold = false, dotall = false;
    std::string err;

    bool more() const { return i < n && err.empty(); }
    uint8_t peek() const { return static_cast<uint8_t>(s[i]); }
    bool fail(const std::string &m) { if (err.empty()) err = m; return false; }

    // one UTF-8 rune of the pattern source
    bool rune(uint32_t *r) {
        uint32_t w = 1;
        *r = rx_decode(reinterpret_cast<const uint8_t *>(s) + i, n - i, &w);
        if (*r == 0xFFFDu) return fail(w == 1 && static_cast<uint8_t>(s[i]) != 0xEF ? "pattern is not valid UTF-8" : "U+FFFD in a pattern is not supported");
        i += w;
        return true;
    }

    // regexp (?i): unicode.SimpleFold orbits that touch ASCII: A-Z <-> a-z, k ~ U+212A, s ~ U+017F
    bool add_folded(RuneSet *set, uint32_t lo, uint32_t hi) {
        set->add(lo, hi);
        if (!fold) return true;
        if (hi >= 0x80) {
            // a range reaching beyond ASCII under (?i) would need the full Unicode fold tables
            return fail("non-ASCII rune in a (?i) pattern is not supported");
        }
        for (uint32_t c = lo; c <= hi; c++) {
            if (c >= 'a' && c <= 'z') set->add(c - 32, c - 32);
            if (c >= 'A' && c <= 'Z') set->add(c + 32, c + 32);
            if (c == 'k' || c == 'K') set->add(0x212A, 0x212A);
            if (c == 's' || c == 'S') set->add(0x17F, 0x17F);
        }
        return true;
    }

    void perl_class(char c, RuneSet *set) {   // RE2: ASCII only, never folded further
        switch (c) {
            case 'd': set->add('0', '9'); break;
            case 'w': set->add('0', '9'); set->add('A', 'Z'); set->add('a', 'z'); set->add('_', '_'); break;
            case 's': set->add('\t', '\n'); set->add('\f', '\r'); set->add(' ', ' '); break;
        }
    }
    void complement(RuneSet *set) {
        set->normalise();
        RuneSet o;
        uint32_t next = 0;
        for (auto &x : set->r) { if (x.first > next) o.add(next, x.first - 1); next = x.second + 1; }
        if (next <= 0x10FFFF) o.add(next, 0x10FFFF);
        *set = o;
    }

    // after a backslash; outside a class `set` gets one item; returns false on error
    bool escape(RuneSet *set, bool in_class) {
        if (i >= n) return fail("trailing backslash");
        const char c = s[i];
        if (c == 'd' || c == 'w' || c == 's') { i++; perl_class(c, set); return true; }
        if (c == 'D' || c == 'W' || c == 'S') {
            i++;
            RuneSet t; perl_class(static_cast<char>(c + 32), &t); complement(&t);
            for (auto &x : t.r) set->add(x.first, x.second);
            return true;
        }
        uint32_t r = 0;
        if (c == 'n') { r = '\n'; i++; }
        else if (c == 't') { r = '\t'; i++; }
        else if (c == 'r') { r = '\r'; i++; }
        else if (c == 'f') { r = '\f'; i++; }
        else if (c == 'v') { r = '\v'; i++; }
        else if (c == 'x') {
            i++;
            if (i + 2 > n || !isxdigit(static_cast<uint8_t>(s[i])) || !isxdigit(static_cast<uint8_t>(s[i + 1]))) return fail("\\x needs two hex digits (\\x{...} is not supported)");
            r = static_cast<uint32_t>(std::stoul(std::string(s + i, 2), nullptr, 16));
            i += 2;
        } else if (std::strchr("\\.+*?()|[]{}^$-/\"'#&~,:;<=>@!%` ", c) != nullptr && c != '\0') { r = static_cast<uint8_t>(c); i++; }
        else return fail(std::string("escape \\") + c + " is not supported (anchors, word boundaries and back-references are outside the subset)");
        (void)in_class;
        return add_folded(set, r, r);
    }

    NodeP parse_class() {   // after '['
        auto nd = std::make_unique<Node>();
        if (more() && peek() == '^') { nd->negate = true; i++; }
        bool first = true;
        for (;;) {
            if (i >= n) { fail("missing ]"); return nullptr; }
            if (peek() == ']' && !first) { i++; break; }
            first = false;
            if (peek() == '[' && i + 1 < n && s[i + 1] == ':') { fail("POSIX classes [[:name:]] are not supported"); return nullptr; }
            uint32_t lo = 0;
            bool single = false;
            if (peek() == '\\') {
                i++;
                const char c = i < n ? s[i] : '\0';
                if (std::strchr("dwsDWS", c) != nullptr && c != '\0') { if (!escape(&nd->set, true)) return nullptr; continue; }
                RuneSet t;
                const bool was = fold; fold = false;
                const bool ok = escape(&t, true);
                fold = was;
                if (!ok) return nullptr;
                lo = t.r[0].first; single = true;
            } else {
                if (!rune(&lo)) return nullptr;
                single = true;
            }
            (void)single;
            uint32_t hi = lo;
            if (i + 1 < n && peek() == '-' && s[i + 1] != ']') {   // range
                i++;
                if (peek() == '\\') {
                    i++;
                    RuneSet t;
                    const bool was = fold; fold = false;
                    const bool ok = escape(&t, true);
                    fold = was;
                    if (!ok) return nullptr;
                    if (t.r.size() != 1 || t.r[0].first != t.r[0].second) { fail("bad range end"); return nullptr; }
                    hi = t.r[0].first;
                } else if (!rune(&hi)) return nullptr;
                if (hi < lo) { fail("invalid character class range"); return nullptr; }
            }
            if (!add_folded(&nd->set, lo, hi)) return nullptr;
        }
        return nd;
    }

    NodeP parse_atom() {
        const uint8_t c = peek();
        if (c == '(') {
            i++;
            if (more() && peek() == '?') {
                if (i + 1 < n && s[i + 1] == ':') i += 2;
                else { fail("only (?: ) groups and a leading (?i)/(?s) are supported"); return nullptr; }
            }
            NodeP a = parse_alt();
            if (!a) return nullptr;
            if (i >= n || s[i] != ')') { fail("missing )"); return nullptr; }
            i++;
            return a;
        }
        if (c == '[') { i++; return parse_class(); }
        auto nd = std::make_unique<Node>();
        if (c == '.') {
            i++;
            if (dotall) nd->set.add(0, 0x10FFFF);
            else { nd->set.add(0, '\n' - 1); nd->set.add('\n' + 1, 0x10FFFF); }
            return nd;
        }
        if (c == '\\') { i++; if (!escape(&nd->set, false)) return nullptr; return nd; }
        if (c == '^' || c == '$') { fail("anchors are not supported"); return nullptr; }
        if (c == '*' || c == '+' || c == '?' ) { fail("missing argument to repetition operator"); return nullptr; }
        if (c == ')' ) { fail("unexpected )"); return nullptr; }
        // '{' that does not start a valid repetition, ']' and '}' are literals in RE2 as well
        uint32_t r = 0;
        if (!rune(&r)) return nullptr;
        if (!add_folded(&nd->set, r, r)) return nullptr;
        return nd;
    }

    bool parse_counted(uint32_t *lo, uint32_t *hi, bool *unbounded) {   // at '{'; leaves i unchanged when it is not a repetition
        uint32_t j = i + 1, a = 0, b = 0;
        bool have_a = false, have_b = false, comma = false;
        while (j < n && isdigit(static_cast<uint8_t>(s[j]))) { a = a * 10 + (s[j] - '0'); have_a = true; j++; if (a > 100000) break; }
        if (j < n && s[j] == ',') { comma = true; j++; while (j < n && isdigit(static_cast<uint8_t>(s[j]))) { b = b * 10 + (s[j] - '0'); have_b = true; j++; if (b > 100000) break; } }
        if (!have_a || j >= n || s[j] != '}') return false;
        i = j + 1;
        *lo = a; *hi = comma ? (have_b ? b : a) : a; *unbounded = comma && !have_b;
        return true;
    }

    NodeP parse_repeat() {
        NodeP a = parse_atom();
        if (!a) return nullptr;
        while (more()) {
            const uint8_t c = peek();
            uint32_t lo = 0, hi = 0; bool unb = false;
            if (c == '*') { lo = 0; unb = true; i++; }
            else if (c == '+') { lo = 1; unb = true; i++; }
            else if (c == '?') { lo = 0; hi = 1; i++; }
            else if (c == '{') { if (!parse_counted(&lo, &hi, &unb)) break; }
            else break;
            if (!unb && hi < lo) { fail("bad repetition {n,m}"); return nullptr; }
            if (lo > kRegexMaxRepeat || (!unb && hi > kRegexMaxRepeat)) { fail("repetition count too large"); return nullptr; }
            auto rep = std::make_unique<Node>();
            rep->kind = Node::REP; rep->lo = lo; rep->hi = hi; rep->unbounded = unb;
            if (more() && peek() == '?') { rep->lazy = true; i++; }
            if (more() && (peek() == '*' || peek() == '+' || peek() == '?') ) { fail("bad repetition operator"); return nullptr; }   // RE2: a** is an error
            rep->kids.push_back(std::move(a));
            a = std::move(rep);
        }
        return a;
    }

    NodeP parse_concat() {
        auto cat = std::make_unique<Node>();
        cat->kind = Node::CAT;
        while (more() && peek() != '|' && peek() != ')') {
            NodeP r = parse_repeat();
            if (!r) return nullptr;
            cat->kids.push_back(std::move(r));
        }
        if (!err.empty()) return nullptr;
        return cat;
    }

    NodeP parse_alt() {
        auto alt = std::make_unique<Node>();
        alt->kind = Node::ALT;
        for (;;) {
            NodeP c = parse_concat();
            if (!c) return nullptr;
            alt->kids.push_back(std::move(c));
            if (more() && peek() == '|') { i++; continue; }
            break;
        }
        if (alt->kids.size() == 1) return std::move(alt->kids[0]);
        return alt;
    }
};

bool nullable(const Node &nd) {
    switch (nd.kind) {
        case Node::RUNE: return false;
        case Node::CAT: for (auto &k : nd.kids) if (!nullable(*k)) return false; return true;
        case Node::ALT: for (auto &k : nd.kids) if (nullable(*k)) return true; return false;
        case Node::REP: return nd.lo == 0 || nullable(*nd.kids[0]);
    }
    return false;
}
// a repeated part that can be empty would loop without consuming input
bool bad_repeat(const Node &nd) {
    if (nd.kind == Node::REP && (nd.unbounded || nd.hi > 1) && nullable(*nd.kids[0])) return true;
    for (auto &k : nd.kids) if (bad_repeat(*k)) return true;
    return false;
}

struct Emitter {
    RegexSet *set;
    std::vector<RegexInst> code;
    bool overflow = false;

    uint32_t emit(RegexInst in) {
        if (code.size() >= kRegexMaxInsts) { overflow = true; return 0; }
        code.push_back(in);
        return static_cast<uint32_t>(code.size() - 1);
    }
    uint16_t add_class(const Node &nd) {
        RuneSet rs = nd.set;
        rs.normalise();
        RegexClass c{};
        c.negate = nd.negate ? 1u : 0u;
        c.first_range = static_cast<uint32_t>(set->ranges.size() / 2);
        for (auto &x : rs.r) {
            for (uint32_t ch = x.first; ch <= std::min<uint32_t>(x.second, 127); ch++) c.ascii[ch >> 5] |= 1u << (ch & 31);
            if (x.second >= 0x80) { set->ranges.push_back(std::max<uint32_t>(x.first, 0x80)); set->ranges.push_back(x.second); c.n_ranges++; }
        }
        set->classes.push_back(c);
        return static_cast<uint16_t>(set->classes.size() - 1);
    }
    void gen(const Node &nd) {
        if (overflow) return;
        switch (nd.kind) {
            case Node::RUNE: emit(RegexInst{ROP_RUNE, 0, add_class(nd), 0, 0}); break;
            case Node::CAT: for (auto &k : nd.kids) gen(*k); break;
            case Node::ALT: {
                std::vector<uint32_t> jumps;
                for (size_t k = 0; k < nd.kids.size(); k++) {
                    uint32_t split = 0;
                    const bool last = k + 1 == nd.kids.size();
                    if (!last) split = emit(RegexInst{ROP_SPLIT, 0, 0, 0, 0});
                    if (!last && !overflow) code[split].x = static_cast<uint16_t>(code.size());
                    gen(*nd.kids[k]);
                    if (!last) { jumps.push_back(emit(RegexInst{ROP_JMP, 0, 0, 0, 0})); if (!overflow) code[split].y = static_cast<uint16_t>(code.size()); }
                }
                for (uint32_t j : jumps) if (!overflow) code[j].x = static_cast<uint16_t>(code.size());
                break;
            }
            case Node::REP: {
                const Node &body = *nd.kids[0];
                for (uint32_t k = 0; k < nd.lo; k++) gen(body);
                if (nd.unbounded) {           // L1: split L2, L3 ; L2: body ; jmp L1 ; L3:
                    const uint32_t l1 = emit(RegexInst{ROP_SPLIT, 0, 0, 0, 0});
                    const uint32_t l2 = static_cast<uint32_t>(code.size());
                    gen(body);
                    emit(RegexInst{ROP_JMP, 0, 0, static_cast<uint16_t>(l1), 0});
                    const uint32_t l3 = static_cast<uint32_t>(code.size());
                    if (!overflow) { code[l1].x = static_cast<uint16_t>(nd.lazy ? l3 : l2); code[l1].y = static_cast<uint16_t>(nd.lazy ? l2 : l3); }
                } else {                      // (body (body (...)?)?)? for the optional copies
                    std::vector<uint32_t> splits;
                    for (uint32_t k = nd.lo; k < nd.hi; k++) {
                        const uint32_t sp = emit(RegexInst{ROP_SPLIT, 0, 0, 0, 0});
                        splits.push_back(sp);
                        const uint32_t b = static_cast<uint32_t>(code.size());
                        if (!overflow) { if (nd.lazy) code[sp].y = static_cast<uint16_t>(b); else code[sp].x = static_cast<uint16_t>(b); }
                        gen(body);
                    }
                    const uint32_t end = static_cast<uint32_t>(code.size());
                    for (uint32_t sp : splits) if (!overflow) { if (nd.lazy) code[sp].x = static_cast<uint16_t>(end); else code[sp].y = static_cast<uint16_t>(end); }
                }
                break;
            }
        }
    }
};

}  // namespace

std::string regex_compile(const char *pattern, uint32_t len, uint32_t prio, RegexSet *set) {
    Parser p{pattern, len};
    if (len >= 2 && pattern[0] == '(' && pattern[1] == '?' ) {   // leading flag group
        uint32_t j = 2;
        bool any = false, ok = true;
        while (j < len && pattern[j] != ')' && pattern[j] != ':') {
            if (pattern[j] == 'i') p.fold = true;
            else if (pattern[j] == 's') p.dotall = true;
            else ok = false;
            any = true; j++;
        }
        if (j < len && pattern[j] == ')' && any) {
            if (!ok) return "only the flags i and s are supported";
            p.i = j + 1;
        } else { p.fold = p.dotall = false; }
    }
    NodeP root = p.parse_alt();
    if (!root || !p.err.empty()) return p.err.empty() ? "syntax error" : p.err;
    if (p.i < len) return p.i < len && pattern[p.i] == ')' ? "unexpected )" : "syntax error";
    if (nullable(*root)) return "a pattern that can match the empty string is not supported";
    if (bad_repeat(*root)) return "a repeated part that can match the empty string is not supported";
    const size_t classes_before = set->classes.size(), ranges_before = set->ranges.size();
    Emitter em{set};
    em.gen(*root);
    em.emit(RegexInst{ROP_MATCH, 0, 0, 0, 0});
    if (em.overflow || set->classes.size() > 65000) {
        set->classes.resize(classes_before); set->ranges.resize(ranges_before);
        return "pattern too large (more than 64 instructions after expanding repetitions)";
    }
    RegexProgram pr{};
    pr.first_inst = static_cast<uint32_t>(set->insts.size());
    pr.n_inst = static_cast<uint32_t>(em.code.size());
    pr.prio = prio;
    set->insts.insert(set->insts.end(), em.code.begin(), em.code.end());
    set->programs.push_back(pr);
    set->sources.emplace_back(pattern, len);
    return "";
}

uint32_t regex_count_host(const RegexSet &set, uint32_t program, const uint8_t *s, uint64_t n) {
    RegexView v{set.programs.data(), set.insts.data(), set.classes.data(), set.ranges.data(), static_cast<uint32_t>(set.programs.size())};
    return rx_count(v, program, s, n);
}

}  // namespace triage
