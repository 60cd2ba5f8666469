// dfa_build.cc — see dfa_build.h.  Host only; runs once per triage_dict_commit.
#include "dfa_build.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>

namespace triage {

bool EventVec::operator==(const EventVec &o) const { return std::memcmp(this, &o, sizeof(EventVec)) == 0; }
bool EventVec::empty() const {
    EventVec z;
    std::memset(&z, 0, sizeof(z));
    return *this == z;
}

namespace {

// ---- UTF-8 (same validity rules as Go's unicode/utf8) -------------------------------------
bool decode_rune(const std::string &s, size_t i, uint32_t *r, uint32_t *w) {
    auto b = [&](size_t k) { return static_cast<uint8_t>(s[k]); };
    uint8_t b0 = b(i);
    size_t n = s.size() - i;
    if (b0 < 0x80) { *r = b0; *w = 1; return true; }
    if (b0 < 0xC2 || b0 > 0xF4) return false;
    if (b0 < 0xE0) {
        if (n < 2 || (b(i + 1) & 0xC0) != 0x80) return false;
        *r = ((b0 & 0x1Fu) << 6) | (b(i + 1) & 0x3Fu); *w = 2; return true;
    }
    if (b0 < 0xF0) {
        uint8_t lo = b0 == 0xE0 ? 0xA0 : 0x80, hi = b0 == 0xED ? 0x9F : 0xBF;
        if (n < 3 || b(i + 1) < lo || b(i + 1) > hi || (b(i + 2) & 0xC0) != 0x80) return false;
        *r = ((b0 & 0x0Fu) << 12) | ((b(i + 1) & 0x3Fu) << 6) | (b(i + 2) & 0x3Fu); *w = 3; return true;
    }
    uint8_t lo = b0 == 0xF0 ? 0x90 : 0x80, hi = b0 == 0xF4 ? 0x8F : 0xBF;
    if (n < 4 || b(i + 1) < lo || b(i + 1) > hi || (b(i + 2) & 0xC0) != 0x80 || (b(i + 3) & 0xC0) != 0x80) return false;
    *r = ((b0 & 0x07u) << 18) | ((b(i + 1) & 0x3Fu) << 12) | ((b(i + 2) & 0x3Fu) << 6) | (b(i + 3) & 0x3Fu);
    *w = 4;
    return true;
}

bool is_ascii_letter(uint32_t r) { return (r >= 'a' && r <= 'z') || (r >= 'A' && r <= 'Z'); }
uint32_t ascii_lower(uint32_t r) { return (r >= 'A' && r <= 'Z') ? r + 32 : r; }

// ---- NFA for the substring patterns (keywords + question words) ---------------------------
struct NfaNode {
    std::vector<std::pair<uint8_t, uint32_t>> edges;  // (byte, target)
    int accept_event = -1;                            // index into pattern events when a pattern ends here
    int pattern = -1;                                 // which pattern's chain this node belongs to
};

struct Nfa {
    std::vector<NfaNode> nodes;  // node 0 = root (always active)
    std::vector<EventVec> accept;  // per accepting node event
    std::vector<uint8_t> bordered; // per pattern: a proper prefix equals a suffix (after folding)
    uint32_t max_bytes = 1;

    Nfa() { nodes.emplace_back(); }
    void finish() { for (auto &n : nodes) std::sort(n.edges.begin(), n.edges.end()); }
    uint32_t add_node() { nodes.emplace_back(); return static_cast<uint32_t>(nodes.size() - 1); }

    // one symbol = a set of alternative byte strings leading from `from` to a fresh node
    // children of shareable (non self-overlapping) patterns, keyed by (from node, alternatives): common prefixes
    // become one trie path, so the live-thread sets stay as small as the pattern depth even for 10 K keywords
    std::map<std::pair<uint32_t, std::string>, uint32_t> shared;

    uint32_t add_symbol(uint32_t from, const std::vector<std::string> &alts, int pattern) {
        std::string key;
        if (pattern < 0) {
            for (const std::string &a : alts) { key += a; key.push_back('\0'); }
            auto it = shared.find({from, key});
            if (it != shared.end()) return it->second;
        }
        uint32_t to = add_node();
        if (pattern < 0) shared[{from, key}] = to;
        nodes[to].pattern = pattern;
        for (const std::string &a : alts) {
            uint32_t cur = from;
            for (size_t k = 0; k < a.size(); k++) {
                uint32_t nxt = (k + 1 == a.size()) ? to : add_node();
                nodes[nxt].pattern = pattern;
                nodes[cur].edges.emplace_back(static_cast<uint8_t>(a[k]), nxt);
                cur = nxt;
            }
        }
        return to;
    }
};

enum FoldKind { FOLD_NONE, FOLD_REGEXP, FOLD_TOLOWER };

// alternatives a content rune may take to match pattern rune r
std::vector<std::string> symbol_alts(uint32_t r, const std::string &raw, FoldKind fold) {
    std::vector<std::string> alts;
    if (fold != FOLD_NONE && is_ascii_letter(r)) {
        uint32_t lo = ascii_lower(r);
        alts.push_back(std::string(1, static_cast<char>(lo)));
        alts.push_back(std::string(1, static_cast<char>(lo - 32)));
        if (lo == 'k') alts.push_back("\xE2\x84\xAA");                       // U+212A, both folds
        if (lo == 's' && fold == FOLD_REGEXP) alts.push_back("\xC5\xBF");    // U+017F, SimpleFold orbit only
        if (lo == 'i' && fold == FOLD_TOLOWER) alts.push_back("\xC4\xB0");   // U+0130, unicode.ToLower only
    } else {
        alts.push_back(raw);
    }
    return alts;
}

void add_event(EventVec *a, const EventVec &b);

uint32_t add_pattern(Nfa *nfa, const std::string &lit, FoldKind fold, const EventVec &ev, bool bordered) {
    uint32_t cur = 0, bytes = 0;
    const int pattern = static_cast<int>(nfa->bordered.size());
    nfa->bordered.push_back(bordered ? 1 : 0);
    for (size_t i = 0; i < lit.size();) {
        uint32_t r = 0, w = 1;
        decode_rune(lit, i, &r, &w);
        std::vector<std::string> alts = symbol_alts(r, lit.substr(i, w), fold);
        size_t longest = 0;
        for (auto &a : alts) longest = std::max(longest, a.size());
        bytes += static_cast<uint32_t>(longest);
        cur = nfa->add_symbol(cur, alts, bordered ? pattern : -1);   // only self-overlapping patterns need private chains
        i += w;
    }
    if (nfa->nodes[cur].accept_event >= 0) {   // a shared end node: another pattern with the same text ends here too
        add_event(&nfa->accept[nfa->nodes[cur].accept_event], ev);
    } else {
        nfa->nodes[cur].accept_event = static_cast<int>(nfa->accept.size());
        nfa->accept.push_back(ev);
    }
    nfa->max_bytes = std::max(nfa->max_bytes, bytes);
    return cur;
}

// ---- sentiment trie (whole-field equality after strings.ToLower) --------------------------
struct SentTrie {
    struct Node { std::map<uint8_t, uint16_t> next; uint8_t pos = 0, neg = 0; };
    std::vector<Node> nodes;  // 0 = root
    uint32_t max_bytes = 0;
    SentTrie() { nodes.emplace_back(); }
    uint16_t child(uint16_t n, uint8_t b) {
        auto it = nodes[n].next.find(b);
        if (it != nodes[n].next.end()) return it->second;
        uint16_t id = static_cast<uint16_t>(nodes.size());
        nodes.emplace_back();
        nodes[n].next[b] = id;
        return id;
    }
    void add(const char *word, bool positive) {
        // byte-level trie; a letter's alternatives rejoin at one node, so walk sets of nodes
        std::vector<uint16_t> cur{0};
        uint32_t bytes = 0;
        for (const char *p = word; *p; p++) {
            std::vector<std::string> alts = symbol_alts(static_cast<uint8_t>(*p), std::string(1, *p), FOLD_TOLOWER);
            size_t longest = 0;
            for (auto &a : alts) longest = std::max(longest, a.size());
            bytes += static_cast<uint32_t>(longest);
            // all alternatives of one symbol must land on the same logical node: use the lower-case
            // single-byte child as the canonical target and alias the others to it.
            std::vector<uint16_t> nxt;
            for (uint16_t n : cur) {
                uint16_t canon = child(n, static_cast<uint8_t>(alts[0][0]));
                for (size_t a = 1; a < alts.size(); a++) {
                    uint16_t c = n;
                    for (size_t k = 0; k + 1 < alts[a].size(); k++) c = child(c, static_cast<uint8_t>(alts[a][k]));
                    nodes[c].next[static_cast<uint8_t>(alts[a].back())] = canon;
                }
                nxt.push_back(canon);
            }
            std::sort(nxt.begin(), nxt.end());
            nxt.erase(std::unique(nxt.begin(), nxt.end()), nxt.end());
            cur = nxt;
        }
        for (uint16_t n : cur) { if (positive) nodes[n].pos++; else nodes[n].neg++; }
        max_bytes = std::max(max_bytes, bytes);
    }
    static constexpr uint16_t DEAD = 0xFFFF;
    uint16_t step(uint16_t n, uint8_t b) const {
        if (n == DEAD) return DEAD;
        auto it = nodes[n].next.find(b);
        return it == nodes[n].next.end() ? DEAD : it->second;
    }
};

// ---- multi-byte whitespace recogniser (unicode.IsSpace beyond ASCII) -----------------------
enum WsProg : uint8_t { WS_NONE = 0, WS_C2, WS_E1, WS_E1_9A, WS_E2, WS_E2_80, WS_E2_81, WS_E3, WS_E3_80 };
enum WsStep { WS_PENDING, WS_COMPLETE, WS_FAIL };

bool is_ascii_space(uint8_t b) { return (b >= 0x09 && b <= 0x0D) || b == 0x20; }

WsProg ws_lead(uint8_t b) {
    switch (b) {
        case 0xC2: return WS_C2;
        case 0xE1: return WS_E1;
        case 0xE2: return WS_E2;
        case 0xE3: return WS_E3;
        default: return WS_NONE;
    }
}

WsStep ws_continue(WsProg p, uint8_t b, WsProg *next) {
    switch (p) {
        case WS_C2: return (b == 0x85 || b == 0xA0) ? WS_COMPLETE : WS_FAIL;            // U+0085, U+00A0
        case WS_E1: if (b == 0x9A) { *next = WS_E1_9A; return WS_PENDING; } return WS_FAIL;
        case WS_E1_9A: return b == 0x80 ? WS_COMPLETE : WS_FAIL;                         // U+1680
        case WS_E2:
            if (b == 0x80) { *next = WS_E2_80; return WS_PENDING; }
            if (b == 0x81) { *next = WS_E2_81; return WS_PENDING; }
            return WS_FAIL;
        case WS_E2_80:                                                                   // U+2000-200A, 2028, 2029, 202F
            return ((b >= 0x80 && b <= 0x8A) || b == 0xA8 || b == 0xA9 || b == 0xAF) ? WS_COMPLETE : WS_FAIL;
        case WS_E2_81: return b == 0x9F ? WS_COMPLETE : WS_FAIL;                         // U+205F
        case WS_E3: if (b == 0x80) { *next = WS_E3_80; return WS_PENDING; } return WS_FAIL;
        case WS_E3_80: return b == 0x80 ? WS_COMPLETE : WS_FAIL;                         // U+3000
        default: return WS_FAIL;
    }
}

// ---- product configuration ------------------------------------------------------------------
struct Config {
    uint8_t in_field = 0;  // 0: last completed rune was whitespace / message start; 1: inside a field
    uint8_t ws = WS_NONE;  // bytes of a possible multi-byte whitespace rune seen so far
    uint16_t sent = 0;     // sentiment trie node (frozen while ws is pending), DEAD when the field cannot match
    std::vector<uint32_t> kw;  // active NFA nodes, sorted
    std::string key() const {
        std::string k;
        k.push_back(static_cast<char>(in_field));
        k.push_back(static_cast<char>(ws));
        k.append(reinterpret_cast<const char *>(&sent), 2);
        k.append(reinterpret_cast<const char *>(kw.data()), kw.size() * 4);
        return k;
    }
};

struct Out {
    bool word_start = false;
    EventVec ev;
};

void add_event(EventVec *a, const EventVec &b) {
    for (int i = 0; i < 4; i++) a->kw[i] = static_cast<uint16_t>(a->kw[i] + b.kw[i]);
    a->pos = static_cast<uint8_t>(a->pos + b.pos);
    a->neg = static_cast<uint8_t>(a->neg + b.neg);
    a->question = (a->question || b.question) ? 1 : 0;
    a->risk = (a->risk || b.risk) ? 1 : 0;
}

struct Machine {
    Nfa nfa;
    SentTrie sent;
    bool exact_nonoverlap = false;

    void ws_complete(Config *c, Out *o) const {
        if (c->in_field && c->sent != SentTrie::DEAD) {
            const auto &n = sent.nodes[c->sent];
            o->ev.pos = static_cast<uint8_t>(o->ev.pos + n.pos);
            o->ev.neg = static_cast<uint8_t>(o->ev.neg + n.neg);
        }
        c->in_field = 0;
        c->sent = 0;
        c->ws = WS_NONE;
    }

    void fresh_byte(Config *c, uint8_t b, Out *o) const {
        if (is_ascii_space(b)) { ws_complete(c, o); return; }
        WsProg lead = ws_lead(b);
        if (lead != WS_NONE) { c->ws = lead; return; }   // undecided: context and sentiment node stay frozen
        if (!c->in_field) {
            o->word_start = true;
            c->in_field = 1;
            c->sent = sent.step(0, b);
        } else {
            c->sent = sent.step(c->sent, b);
        }
    }

    Config step(const Config &in, uint8_t b, Out *o) const {
        Config c = in;
        // keyword / question NFA: consumes every byte, root is always active
        std::vector<uint32_t> nk;
        std::vector<int> matched;   // self-overlapping patterns that completed on this byte
        auto follow = [&](uint32_t node) {
            const auto &edges = nfa.nodes[node].edges;   // sorted by byte (Nfa::finish)
            auto lo = std::lower_bound(edges.begin(), edges.end(), std::make_pair(b, uint32_t(0)));
            for (auto it = lo; it != edges.end() && it->first == b; ++it) {
                const auto &e = *it;
                const NfaNode &t = nfa.nodes[e.second];
                if (t.accept_event >= 0) {
                    add_event(&o->ev, nfa.accept[t.accept_event]);
                    if (t.pattern >= 0 && nfa.bordered[t.pattern]) matched.push_back(t.pattern);
                }
                if (!t.edges.empty()) nk.push_back(e.second);
            }
        };
        follow(0);
        for (uint32_t n : in.kw) follow(n);
        std::sort(nk.begin(), nk.end());
        // NOTE: duplicates are kept out — two NFA paths reaching the same node are the same thread
        nk.erase(std::unique(nk.begin(), nk.end()), nk.end());
        for (int p : matched) {
            // every other live thread of p started inside the occurrence that just ended
            auto inside = [&](uint32_t n) { return nfa.nodes[n].pattern == p; };
            if (exact_nonoverlap) nk.erase(std::remove_if(nk.begin(), nk.end(), inside), nk.end());   // search resumes at the match end
            else if (std::any_of(nk.begin(), nk.end(), inside)) o->ev.risk = 1;                        // occurrences may overlap
        }
        c.kw = nk;
        // whitespace / field context
        if (c.ws != WS_NONE) {
            WsProg nxt = WS_NONE;
            WsStep r = ws_continue(static_cast<WsProg>(c.ws), b, &nxt);
            if (r == WS_PENDING) { c.ws = nxt; return c; }
            if (r == WS_COMPLETE) { ws_complete(&c, o); return c; }
            // the pending bytes were not whitespace: they are field content that matches no sentiment word
            if (!c.in_field) { o->word_start = true; c.in_field = 1; }
            c.sent = SentTrie::DEAD;
            c.ws = WS_NONE;
        }
        fresh_byte(&c, b, o);
        return c;
    }

    Out end_of_message(const Config &c) const {
        Out o;
        std::memset(&o.ev, 0, sizeof(o.ev));
        if (c.ws != WS_NONE) {
            if (!c.in_field) o.word_start = true;   // a truncated sequence is an invalid, non-space rune
        } else if (c.in_field && c.sent != SentTrie::DEAD) {
            o.ev.pos = sent.nodes[c.sent].pos;
            o.ev.neg = sent.nodes[c.sent].neg;
        }
        return o;
    }
};

std::vector<uint32_t> fold_runes(const std::string &s, bool ci) {
    std::vector<uint32_t> out;
    for (size_t i = 0; i < s.size();) {
        uint32_t r = 0, w = 1;
        decode_rune(s, i, &r, &w);
        out.push_back(ci ? ascii_lower(r) : r);
        i += w;
    }
    return out;
}

}  // namespace

std::string validate_literal(const std::string &bytes, bool ci) {
    if (bytes.empty()) return "empty pattern";
    for (size_t i = 0; i < bytes.size();) {
        uint32_t r = 0, w = 1;
        if (!decode_rune(bytes, i, &r, &w)) return "pattern is not valid UTF-8";
        if (r == 0xFFFD) return "U+FFFD in a pattern is not supported";
        if (ci && r >= 0x80) return "non-ASCII rune in a (?i) pattern is not supported";
        i += w;
    }
    return "";
}

bool literal_is_bordered(const std::string &bytes, bool ci) {
    std::vector<uint32_t> r = fold_runes(bytes, ci);
    // prefix function: a non-zero value at the end means a proper prefix equals a suffix
    std::vector<size_t> pi(r.size(), 0);
    for (size_t i = 1; i < r.size(); i++) {
        size_t k = pi[i - 1];
        while (k > 0 && r[i] != r[k]) k = pi[k - 1];
        if (r[i] == r[k]) k++;
        pi[i] = k;
    }
    return !r.empty() && pi.back() > 0;
}

bool parse_regex_literal(const char *s, uint32_t len, bool *ci, std::string *lit, std::string *err) {
    uint32_t i = 0;
    *ci = false;
    lit->clear();
    if (len >= 4 && std::memcmp(s, "(?i)", 4) == 0) { *ci = true; i = 4; }
    static const char meta[] = "\\.+*?()|[]{}^$";
    for (; i < len; i++) {
        char c = s[i];
        if (c == '\\') {
            if (i + 1 >= len) { *err = "trailing backslash"; return false; }
            char n = s[i + 1];
            if (std::strchr(meta, n) == nullptr || n == '\0') { *err = "escape class is not a literal"; return false; }
            lit->push_back(n);
            i++;
        } else if (c != '\0' && std::strchr(meta, c) != nullptr) {
            *err = "regexp metacharacter: only (?i)-prefixed literals are supported";
            return false;
        } else {
            lit->push_back(c);
        }
    }
    if (lit->empty()) { *err = "empty pattern"; return false; }
    return true;
}

bool dict_has_bordered(const std::vector<DictEntry> &dict) {
    for (const DictEntry &d : dict) if (literal_is_bordered(d.bytes, d.ci)) return true;
    return false;
}

// ---- shared-memory placement of the dense table ---------------------------------------------------------------
// The scan reads entry (state, byte) at byte offset state * 516 + 2 * byte of shared memory: bank = (state + byte / 2)
// mod 32.  32 lanes in different states reading different bytes collide on banks; how often depends only on
// state id mod 32 of the states the lanes are usually in.  place_states() renumbers the (minimised) states so that the
// expected number of shared-memory wavefronts per 32-lane lookup is small for ordinary text: it walks the automaton over
// a deterministic sample of English-like text (common words, the dictionary's own patterns, some capitals, digits and
// punctuation), draws 32-lane groups from that walk and does coordinate descent on "most distinct words in one bank"
// over the rotations of the hottest states.  Results never depend on the numbering; a workload unlike the sample only
// sees the same conflict rate an arbitrary numbering would give.
static const char kPriorWords[] =
    "the of and to a in is it you that he was for on are with as i his they be at one have this from or had by hot but some "
    "there we can out other were all your up use word said an each she which do their time if will way about many then them "
    "would write like so these her long make thing see him two has look more day could go come did my sound no most number "
    "over know water than call first people may down side been now find any new work part take get place made live after "
    "back little only round man year came show every me give our under name very through just form much think say help low "
    "line before turn cause same mean differ move right boy old too does tell sentence set three want air well also play "
    "small end put home read hand port large spell add even land here must big high such follow act ask men change went "
    "light kind off need house picture try us again animal point mother world near build self earth father head stand own "
    "page should country found answer school grow study still learn plant cover food sun four thought let keep eye never "
    "last door between city tree cross since hard start might story saw far sea draw left late run while press close night "
    "real life few stop open seem together next white children begin got walk example ease paper often always music those "
    "both mark book letter until mile river car feet care second group carry took rain eat room friend began idea fish "
    "mountain north once base hear horse cut sure watch color face wood main enough plain girl usual young ready above ever "
    "red list though feel talk bird body dog family direct pose leave song measure state product black short please thanks "
    "hello team customer order account issue problem update server error request message service support system data";

// deterministic sample of English-like text: common words, the dictionary's own patterns (4 %), some capitals, digits, punctuation
static std::string prior_sample_text(const std::vector<std::string> &patterns, uint64_t &rng);

static std::vector<uint32_t> place_states(const std::vector<uint8_t> &next /* n * 256 */, uint32_t n, uint32_t start,
                                          const std::vector<std::string> &patterns) {
    std::vector<uint32_t> perm(n);
    for (uint32_t i = 0; i < n; i++) perm[i] = i;
    if (n <= 2) return perm;
    uint64_t rng = 0x9E3779B97F4A7C15ull;
    const std::string text = prior_sample_text(patterns, rng);
    auto rnd = [&]() { rng = rng * 6364136223846793005ull + 1442695040888963407ull; return static_cast<uint32_t>(rng >> 33); };
    // walk: (state, word of the row) per text byte
    std::vector<uint16_t> key(text.size());
    std::vector<uint32_t> heat(n, 0);
    uint32_t st = start;
    for (size_t i = 0; i < text.size(); i++) {
        uint8_t b = static_cast<uint8_t>(text[i]);
        key[i] = static_cast<uint16_t>(st << 7 | b >> 1);
        heat[st]++;
        st = next[static_cast<size_t>(st) * 256 + b];
    }
    // 32-lane samples, de-duplicated (equal keys are one broadcast read)
    constexpr uint32_t kSamples = 1024, kHot = 32, kSweeps = 2;
    std::vector<std::vector<uint16_t>> samples(kSamples);
    std::vector<std::vector<uint32_t>> uses(n);   // state -> samples containing it
    for (uint32_t t = 0; t < kSamples; t++) {
        std::vector<uint16_t> &v = samples[t];
        for (int l = 0; l < 32; l++) v.push_back(key[rnd() % key.size()]);
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
        uint32_t last = ~0u;
        for (uint16_t k : v) if (static_cast<uint32_t>(k >> 7) != last) { last = k >> 7; uses[last].push_back(t); }
    }
    std::vector<uint32_t> order(n);
    for (uint32_t i = 0; i < n; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return heat[a] > heat[b]; });
    std::vector<uint8_t> rot(n);
    for (uint32_t i = 0; i < n; i++) rot[i] = static_cast<uint8_t>(i & 31);
    auto sample_cost = [&](uint32_t t) {
        uint8_t cnt[32] = {0};
        uint32_t mx = 0;
        for (uint16_t k : samples[t]) {
            uint32_t b = (rot[k >> 7] + (k & 127u)) & 31u;
            if (++cnt[b] > mx) mx = cnt[b];
        }
        return mx;
    };
    const uint32_t nhot = std::min<uint32_t>(kHot, n);
    for (uint32_t sweep = 0; sweep < kSweeps; sweep++) {
        for (uint32_t h = 0; h < nhot; h++) {
            const uint32_t s = order[h];
            if (heat[s] == 0) break;
            uint32_t best_cost = ~0u, best_rot = rot[s];
            const uint8_t keep = rot[s];
            for (uint32_t r = 0; r < 32; r++) {
                rot[s] = static_cast<uint8_t>((keep + r) & 31);   // the current rotation first: ties keep it
                uint32_t c = 0;
                for (uint32_t t : uses[s]) c += sample_cost(t);
                if (c < best_cost) { best_cost = c; best_rot = rot[s]; }
            }
            rot[s] = static_cast<uint8_t>(best_rot);
        }
    }
    // rotations -> ids: hot states take a free id with their residue (nearest residue if that one is used up)
    std::vector<bool> taken(n, false);
    std::vector<bool> placed(n, false);
    for (uint32_t h = 0; h < nhot; h++) {
        const uint32_t s = order[h];
        if (heat[s] == 0) break;
        for (uint32_t d = 0; d < 32 && !placed[s]; d++) {
            for (uint32_t id = (rot[s] + d) & 31u; id < n; id += 32) {
                if (!taken[id]) { taken[id] = true; perm[s] = id; placed[s] = true; break; }
            }
        }
    }
    uint32_t free_id = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t s = order[i];
        if (placed[s]) continue;
        while (taken[free_id]) free_id++;
        taken[free_id] = true;
        perm[s] = free_id;
    }
    return perm;
}

static std::string prior_sample_text(const std::vector<std::string> &patterns, uint64_t &rng) {
    auto rnd = [&]() { rng = rng * 6364136223846793005ull + 1442695040888963407ull; return static_cast<uint32_t>(rng >> 33); };
    std::vector<std::string> common;
    for (const char *p = kPriorWords; *p;) {
        const char *q = p;
        while (*q && *q != ' ') q++;
        common.emplace_back(p, q);
        p = *q ? q + 1 : q;
    }
    std::string text;
    while (text.size() < 48 * 1024) {
        std::string w = (!patterns.empty() && rnd() % 100 < 4) ? patterns[rnd() % patterns.size()] : common[rnd() % common.size()];
        uint32_t r = rnd() % 100;
        if (r < 6) { if (!w.empty() && w[0] >= 'a' && w[0] <= 'z') w[0] = static_cast<char>(w[0] - 32); }
        else if (r < 8) { for (char &ch : w) if (ch >= 'a' && ch <= 'z') ch = static_cast<char>(ch - 32); }
        else if (r < 10) w = std::to_string(rnd() % 1000);
        r = rnd() % 100;
        if (r < 6) w.push_back(",.?!:"[rnd() % 5]);
        text += w;
        text.push_back(' ');
    }
    return text;
}

// ---- filter automaton of the two-bytes-per-lookup lane (scan_pair.cuh) --------------------------------------------
namespace {

struct ClassPattern { std::vector<uint16_t> sym; };   // per symbol: the set of 4-bit classes (bit mask) that continue it

inline uint32_t class4(uint32_t b, uint32_t mul, uint32_t shift) { return (((b & 31u) * mul) >> shift) & 15u; }

// all-ASCII ways to complete an event, as class strings
std::vector<ClassPattern> class_patterns(const std::vector<DictEntry> &dict, uint32_t mul, uint32_t shift) {
    std::vector<ClassPattern> out;
    auto letter = [&](uint32_t ch, bool fold) {
        uint16_t m = static_cast<uint16_t>(1u << class4(ch, mul, shift));
        if (fold && is_ascii_letter(ch)) m |= static_cast<uint16_t>(1u << class4(ch ^ 0x20u, mul, shift));
        return m;
    };
    uint16_t ws = 0;
    for (uint32_t b : {0x09u, 0x0Au, 0x0Bu, 0x0Cu, 0x0Du, 0x20u}) ws |= static_cast<uint16_t>(1u << class4(b, mul, shift));
    for (const DictEntry &d : dict) {
        // literals of any script: their bytes map to classes like any others (only the lead bytes C2 C4 C5 E0-E3 force an
        // exact walk on the device, and an occurrence that holds one of those is found there)
        ClassPattern p;
        for (unsigned char ch : d.bytes) p.sym.push_back(letter(ch, d.ci && ch < 0x80));
        if (!p.sym.empty()) out.push_back(p);
    }
    for (const char *q : {"what", "how", "why", "when", "where", "who"}) {
        ClassPattern p;
        for (const char *c = q; *c; c++) p.sym.push_back(letter(static_cast<unsigned char>(*c), true));
        p.sym.push_back(static_cast<uint16_t>(1u << class4(' ', mul, shift)));
        out.push_back(p);
    }
    for (const char *w : {"good", "great", "excellent", "happy", "satisfied", "bad", "terrible", "awful", "angry", "frustrated"}) {
        ClassPattern p;
        p.sym.push_back(ws);                               // field start: ASCII whitespace (or a reset pad) before the word
        for (const char *c = w; *c; c++) p.sym.push_back(letter(static_cast<unsigned char>(*c), true));
        p.sym.push_back(ws);                               // the event is emitted on the whitespace (or pad) that ends the field
        out.push_back(p);
    }
    return out;
}

// subset construction over the 16 classes, composition into class pairs, minimisation.  false: too many states
bool determinise_pairs(const std::vector<ClassPattern> &pats, PairFilter *out) {
    std::map<std::vector<uint32_t>, uint32_t> ids;
    std::vector<std::vector<uint32_t>> states;
    struct T { uint32_t next; bool fired; };
    std::vector<T> single;   // state-major, 16 per state
    auto intern = [&](const std::vector<uint32_t> &v) {
        auto it = ids.find(v);
        if (it != ids.end()) return it->second;
        const uint32_t id = static_cast<uint32_t>(states.size());
        ids.emplace(v, id);
        states.push_back(v);
        return id;
    };
    intern({});
    for (uint32_t s = 0; s < states.size(); s++) {
        if (states.size() > 4096) return false;
        const std::vector<uint32_t> cur = states[s];
        for (uint32_t c = 0; c < 16; c++) {
            std::vector<uint32_t> nx;
            bool fired = false;
            auto advance = [&](uint32_t p, uint32_t idx) {
                if (!((pats[p].sym[idx] >> c) & 1u)) return;
                if (idx + 1 == pats[p].sym.size()) fired = true;
                else nx.push_back(p << 8 | (idx + 1));
            };
            for (uint32_t pos : cur) advance(pos >> 8, pos & 0xFFu);
            for (uint32_t p = 0; p < pats.size(); p++) advance(p, 0);   // an occurrence may start at every byte
            std::sort(nx.begin(), nx.end());
            nx.erase(std::unique(nx.begin(), nx.end()), nx.end());
            single.push_back(T{intern(nx), fired});
        }
    }
    const uint32_t n1 = static_cast<uint32_t>(states.size());
    // pairs, restricted to what is reachable from the root in steps of two
    std::vector<T> pair(static_cast<size_t>(n1) * 256);
    for (uint32_t s = 0; s < n1; s++)
        for (uint32_t c0 = 0; c0 < 16; c0++) {
            const T a = single[s * 16 + c0];
            for (uint32_t c1 = 0; c1 < 16; c1++) {
                const T b = single[a.next * 16 + c1];
                pair[static_cast<size_t>(s) * 256 + (c0 | c1 << 4)] = T{b.next, a.fired || b.fired};
            }
        }
    std::vector<int> reach(n1, -1);
    std::vector<uint32_t> order{0};
    reach[0] = 0;
    for (size_t k = 0; k < order.size(); k++)
        for (uint32_t x = 0; x < 256; x++) {
            const uint32_t t = pair[static_cast<size_t>(order[k]) * 256 + x].next;
            if (reach[t] < 0) { reach[t] = static_cast<int>(order.size()); order.push_back(t); }
        }
    const uint32_t n = static_cast<uint32_t>(order.size());
    // Moore refinement
    std::vector<uint32_t> cls(n, 0);
    {
        std::map<std::string, uint32_t> sig;
        for (uint32_t i = 0; i < n; i++) {
            std::string k(256, '0');
            for (uint32_t x = 0; x < 256; x++) if (pair[static_cast<size_t>(order[i]) * 256 + x].fired) k[x] = '1';
            cls[i] = sig.emplace(k, static_cast<uint32_t>(sig.size())).first->second;
        }
    }
    for (;;) {
        std::map<std::string, uint32_t> sig;
        std::vector<uint32_t> ncl(n);
        for (uint32_t i = 0; i < n; i++) {
            std::string k(reinterpret_cast<const char *>(&cls[i]), 4);
            for (uint32_t x = 0; x < 256; x++) {
                const uint32_t t = cls[reach[pair[static_cast<size_t>(order[i]) * 256 + x].next]];
                k.append(reinterpret_cast<const char *>(&t), 4);
            }
            ncl[i] = sig.emplace(k, static_cast<uint32_t>(sig.size())).first->second;
        }
        const uint32_t before = *std::max_element(cls.begin(), cls.end()) + 1, after = static_cast<uint32_t>(sig.size());
        cls.swap(ncl);
        if (after == before) break;
    }
    const uint32_t nmin = *std::max_element(cls.begin(), cls.end()) + 1;
    if (nmin > 256) return false;
    out->n_states = nmin;
    out->start = cls[0];
    out->table.assign(static_cast<size_t>(nmin) * 256, 0);
    for (uint32_t i = 0; i < n; i++)
        for (uint32_t x = 0; x < 256; x++) {
            const T t = pair[static_cast<size_t>(order[i]) * 256 + x];
            out->table[static_cast<size_t>(cls[i]) * 256 + x] = static_cast<uint16_t>(cls[reach[t.next]] | (t.fired ? 0x8000u : 0u));
        }
    return true;
}

// Moore form of the minimised pair automaton: (state, the entry that led here fired), reachable part, quiet states first.
// false: more than 256 of them
bool moore_pairs(PairFilter *f) {
    std::map<uint32_t, uint32_t> ids;          // state << 1 | fired -> index into `order`
    std::vector<uint32_t> order;
    auto intern = [&](uint32_t key) {
        auto it = ids.find(key);
        if (it != ids.end()) return it->second;
        const uint32_t id = static_cast<uint32_t>(order.size());
        ids.emplace(key, id);
        order.push_back(key);
        return id;
    };
    intern(f->start << 1);
    std::vector<uint32_t> next;                // order-major, 256 per state
    for (size_t k = 0; k < order.size(); k++) {
        if (order.size() > 256) return false;
        const uint32_t s = order[k] >> 1;
        for (uint32_t x = 0; x < 256; x++) {
            const uint16_t e = f->table[static_cast<size_t>(s) * 256 + x];
            next.push_back(intern((e & 0xFFu) << 1 | (e >> 15)));
        }
    }
    const uint32_t n = static_cast<uint32_t>(order.size());
    if (n > 256) return false;
    std::vector<uint32_t> id(n);
    uint32_t quiet = 0;
    for (uint32_t i = 0; i < n; i++) if (!(order[i] & 1u)) id[i] = quiet++;
    uint32_t loud = quiet;
    for (uint32_t i = 0; i < n; i++) if (order[i] & 1u) id[i] = loud++;
    f->m_states = n;
    f->m_start = id[0];
    f->m_fired_from = quiet;
    f->mtable.assign(static_cast<size_t>(n) * 256, 0);
    for (uint32_t i = 0; i < n; i++)
        for (uint32_t x = 0; x < 256; x++) f->mtable[static_cast<size_t>(id[i]) * 256 + x] = static_cast<uint8_t>(id[next[static_cast<size_t>(i) * 256 + x]]);
    return true;
}

double pair_fire_rate(const PairFilter &f, const std::string &text) {
    uint32_t st = f.start, groups = 0, fired_groups = 0;
    for (size_t g = 0; g + 16 <= text.size(); g += 16) {
        bool fired = false;
        for (size_t j = 0; j < 16; j += 2) {
            const uint32_t x = class4(static_cast<unsigned char>(text[g + j]), f.mul, f.shift) |
                               class4(static_cast<unsigned char>(text[g + j + 1]), f.mul, f.shift) << 4;
            const uint16_t e = f.table[static_cast<size_t>(st) * 256 + x];
            st = e & 0xFFu;
            fired = fired || (e & 0x8000u);
        }
        groups++;
        fired_groups += fired ? 1 : 0;
    }
    return groups ? static_cast<double>(fired_groups) / groups : 0.0;
}

void build_pair_filter(const std::vector<DictEntry> &dict, const Dfa &exact, PairFilter *out) {
    *out = PairFilter();
    // the lane relies on a 16-byte warm-up, on reset pads that are ASCII whitespace, and on the dense exact table
    if (exact.wide || exact.warm_bytes != 16 || exact.span > 17) return;
    if (!(exact.reset_pad == 0x20 || (exact.reset_pad >= 0x09 && exact.reset_pad <= 0x0D))) return;
    std::vector<std::string> pats;
    for (const DictEntry &d : dict) pats.push_back(d.bytes);
    uint64_t rng = 0xD1B54A32D192ED03ull;
    const std::string text = prior_sample_text(pats, rng);
    static const uint32_t cand[][2] = {{1, 0}, {5, 2}, {3, 2}, {5, 1}, {3, 0}, {7, 1}};
    PairFilter best;
    for (const auto &mk : cand) {
        PairFilter f;
        f.mul = mk[0]; f.shift = mk[1];
        if (!determinise_pairs(class_patterns(dict, f.mul, f.shift), &f) || !moore_pairs(&f)) continue;
        f.sample_fire_rate = pair_fire_rate(f, text);
        f.ok = true;
        // the first candidate needs no multiply on the device: keep it unless another one is clearly more selective
        if (!best.ok || f.sample_fire_rate * 1.5 + 0.002 < best.sample_fire_rate) best = f;
    }
    // the sample carries the dictionary's own patterns in 4 % of its words (about one group in eight fires by design);
    // a filter that flags much more than that flags ordinary text and is no filter
    if (best.ok && best.sample_fire_rate < 0.30) *out = best;
}

}  // namespace

std::string build_dfa(const std::vector<DictEntry> &dict, Dfa *out, bool exact_nonoverlap) {
    Machine m;
    m.exact_nonoverlap = exact_nonoverlap;
    for (const DictEntry &d : dict) {
        std::string e = validate_literal(d.bytes, d.ci);
        if (!e.empty()) return e;
        if (d.prio < 1 || d.prio > 4) return "priority out of range";
        EventVec ev;
        std::memset(&ev, 0, sizeof(ev));
        ev.kw[d.prio - 1] = 1;
        add_pattern(&m.nfa, d.bytes, d.ci ? FOLD_REGEXP : FOLD_NONE, ev, literal_is_bordered(d.bytes, d.ci));
    }
    // preprocessor.go:51 — question words, matched as keyword + " " in strings.ToLower(content)
    for (const char *q : {"what", "how", "why", "when", "where", "who"}) {
        EventVec ev;
        std::memset(&ev, 0, sizeof(ev));
        ev.question = 1;
        add_pattern(&m.nfa, std::string(q) + " ", FOLD_TOLOWER, ev, false);
    }
    // preprocessor.go:49-50
    for (const char *w : {"good", "great", "excellent", "happy", "satisfied"}) m.sent.add(w, true);
    for (const char *w : {"bad", "terrible", "awful", "angry", "frustrated"}) m.sent.add(w, false);

    m.nfa.finish();
    // ---- byte classes: bytes that nothing mentions behave identically (non-space, match nothing) ----
    bool relevant[256] = {false};
    for (const NfaNode &nd : m.nfa.nodes) for (auto &e : nd.edges) relevant[e.first] = true;
    for (const auto &nd : m.sent.nodes) for (auto &kv : nd.next) relevant[kv.first] = true;
    for (uint32_t b : {0x09u, 0x0Au, 0x0Bu, 0x0Cu, 0x0Du, 0x20u, 0xC2u, 0x85u, 0xA0u, 0xE1u, 0x9Au, 0xE2u, 0x81u, 0x9Fu, 0xE3u, 0xA8u, 0xA9u, 0xAFu}) relevant[b] = true;
    for (uint32_t b = 0x80; b <= 0x8A; b++) relevant[b] = true;
    uint8_t cls_of[256];
    std::vector<uint8_t> rep;   // representative byte per class; class 0 = "other"
    {
        int other = -1;
        for (int b = 0; b < 256; b++) if (!relevant[b]) { other = b; break; }
        rep.push_back(static_cast<uint8_t>(other < 0 ? 0 : other));
        for (int b = 0; b < 256; b++) {
            if (relevant[b] || other < 0) { cls_of[b] = static_cast<uint8_t>(rep.size() < 256 ? rep.size() : 0); if (rep.size() < 256) rep.push_back(static_cast<uint8_t>(b)); }
            else cls_of[b] = 0;
        }
        if (other < 0) {   // every byte value is mentioned somewhere: no shared class (class 0 is byte 0's own)
            rep.clear();
            for (int b = 0; b < 256; b++) { cls_of[b] = static_cast<uint8_t>(b); rep.push_back(static_cast<uint8_t>(b)); }
        }
    }
    const uint32_t ncls = static_cast<uint32_t>(rep.size());

    // ---- reachable configurations (BFS) ----
    std::unordered_map<std::string, uint32_t> ids;
    std::vector<Config> cfgs;
    auto intern = [&](const Config &c) {
        std::string k = c.key();
        auto it = ids.find(k);
        if (it != ids.end()) return it->second;
        uint32_t id = static_cast<uint32_t>(cfgs.size());
        ids.emplace(std::move(k), id);
        cfgs.push_back(c);
        return id;
    };
    Config start;   // field start, nothing pending
    Config neutral; // mid-field, no partial matches
    neutral.in_field = 1;
    neutral.sent = SentTrie::DEAD;
    uint32_t start_id = intern(start), neutral_id = intern(neutral);

    std::vector<EventVec> events(1);
    std::memset(&events[0], 0, sizeof(EventVec));
    auto event_id = [&](const EventVec &e) -> uint32_t {
        for (uint32_t i = 0; i < events.size(); i++) if (events[i] == e) return i;
        events.push_back(e);
        return static_cast<uint32_t>(events.size() - 1);
    };

    struct Trans { uint32_t next; uint16_t ev; uint8_t ws; };
    std::vector<Trans> trans;  // cfg-major, ncls per configuration
    std::vector<uint32_t> eom_ev;
    std::vector<uint8_t> eom_ws;
    for (uint32_t s = 0; s < cfgs.size(); s++) {
        if (cfgs.size() > kMaxStatesW) return "dictionary automaton exceeds 2^20 states";
        Config cur = cfgs[s];
        for (uint32_t c = 0; c < ncls; c++) {
            Out o;
            std::memset(&o.ev, 0, sizeof(o.ev));
            Config nx = m.step(cur, rep[c], &o);
            uint32_t eid = event_id(o.ev);
            if (eid > kMaxEventsW) return "dictionary automaton has too many distinct event vectors";
            trans.push_back({intern(nx), static_cast<uint16_t>(eid), static_cast<uint8_t>(o.word_start ? 1 : 0)});
        }
        Out e = m.end_of_message(cur);
        eom_ev.push_back(event_id(e.ev));
        eom_ws.push_back(e.word_start ? 1 : 0);
    }
    uint32_t n = static_cast<uint32_t>(cfgs.size());
    out->n_states_unminimised = n;

    // ---- minimise (Moore partition refinement on outputs + successor classes); skipped for very large automata ----
    std::vector<uint32_t> cls(n, 0);
    if (n <= 8192) {
        {
            std::map<std::string, uint32_t> sig;
            for (uint32_t s = 0; s < n; s++) {
                std::string k;
                k.append(reinterpret_cast<const char *>(&eom_ev[s]), 4);
                k.push_back(static_cast<char>(eom_ws[s]));
                for (uint32_t c = 0; c < ncls; c++) {
                    const Trans &t = trans[(size_t)s * ncls + c];
                    k.append(reinterpret_cast<const char *>(&t.ev), 2);
                    k.push_back(static_cast<char>(t.ws));
                }
                auto it = sig.emplace(k, static_cast<uint32_t>(sig.size())).first;
                cls[s] = it->second;
            }
        }
        for (;;) {
            std::map<std::string, uint32_t> sig;
            std::vector<uint32_t> ncl(n);
            for (uint32_t s = 0; s < n; s++) {
                std::string k;
                k.append(reinterpret_cast<const char *>(&cls[s]), 4);
                for (uint32_t c = 0; c < ncls; c++) k.append(reinterpret_cast<const char *>(&cls[trans[(size_t)s * ncls + c].next]), 4);
                auto it = sig.emplace(k, static_cast<uint32_t>(sig.size())).first;
                ncl[s] = it->second;
            }
            uint32_t before = *std::max_element(cls.begin(), cls.end()) + 1;
            uint32_t after = static_cast<uint32_t>(sig.size());
            cls.swap(ncl);
            if (after == before) break;
        }
    } else {
        for (uint32_t s = 0; s < n; s++) cls[s] = s;
    }
    uint32_t nmin = *std::max_element(cls.begin(), cls.end()) + 1;
    const bool wide = nmin > kMaxStatesD8 || events.size() - 1 > kMaxEventsD8;
    if (!wide) {   // renumber the states for few shared-memory bank conflicts (place_states)
        std::vector<uint8_t> mnext(static_cast<size_t>(nmin) * 256, 0);
        for (uint32_t s = 0; s < n; s++)
            for (uint32_t b = 0; b < 256; b++)
                mnext[static_cast<size_t>(cls[s]) * 256 + b] = static_cast<uint8_t>(cls[trans[(size_t)s * ncls + cls_of[b]].next]);
        std::vector<std::string> pats;
        for (const DictEntry &d : dict) pats.push_back(d.bytes);
        for (const char *w : {"good", "great", "excellent", "happy", "satisfied", "bad", "terrible", "awful", "angry", "frustrated",
                              "what", "how", "why", "when", "where", "who"}) pats.push_back(w);
        std::vector<uint32_t> perm = place_states(mnext, nmin, cls[start_id], pats);
        for (uint32_t s = 0; s < n; s++) cls[s] = perm[cls[s]];
    }

    out->n_states = nmin;
    out->wide = wide;
    out->n_classes = ncls;
    std::memcpy(out->cls_of, cls_of, 256);
    out->eom.assign(nmin, 0);
    out->table.clear();
    out->wtable.clear();
    if (!wide) out->table.assign(static_cast<size_t>(nmin) * 256, 0);
    else out->wtable.assign(static_cast<size_t>(nmin) * ncls, 0);
    std::vector<bool> done(nmin, false);
    std::vector<uint32_t> eom_full(nmin, 0);   // event id | ws << 31 (wide eom ids may exceed 7 bits)
    for (uint32_t s = 0; s < n; s++) {
        uint32_t c = cls[s];
        if (done[c]) continue;
        done[c] = true;
        if (!wide) {
            for (uint32_t b = 0; b < 256; b++) {
                const Trans &t = trans[(size_t)s * ncls + cls_of[b]];
                out->table[static_cast<size_t>(c) * 256 + b] = static_cast<uint16_t>(cls[t.next] | (uint32_t(t.ev) << 8) | (t.ws ? 0x8000u : 0u));
            }
            out->eom[c] = static_cast<uint8_t>(eom_ev[s] | (eom_ws[s] ? 0x80u : 0u));
        } else {
            for (uint32_t k = 0; k < ncls; k++) {
                const Trans &t = trans[(size_t)s * ncls + k];
                out->wtable[static_cast<size_t>(c) * ncls + k] = cls[t.next] | (uint32_t(t.ev) << 20) | (t.ws ? 0x80000000u : 0u);
            }
        }
        eom_full[c] = eom_ev[s] | (eom_ws[s] ? 0x80000000u : 0u);
    }
    out->eom_w = eom_full;
    out->events = events;
    out->start_state = static_cast<uint8_t>(cls[start_id]);
    out->neutral_state = static_cast<uint8_t>(cls[neutral_id]);
    out->start_state_w = cls[start_id];
    out->neutral_state_w = cls[neutral_id];
    // longest byte distance an event can depend on: a keyword's own bytes, or for a sentiment word
    // the tail of a multi-byte whitespace before it (2), the word, and the whitespace after it (3)
    out->span = std::max<uint32_t>(m.nfa.max_bytes, m.sent.max_bytes + 5);
    out->warm_bytes = ((out->span - 1 + 15) / 16) * 16;
    if (out->warm_bytes == 0) out->warm_bytes = 16;
    // pad bytes (see scan_core.cuh): searched over all 256 values, whitespace first
    struct TView { uint32_t next; uint32_t ev; bool ws; };
    auto tv = [&](uint32_t st, uint32_t b) -> TView {
        if (!wide) { uint16_t t = out->table[static_cast<size_t>(st) * 256 + b]; return {uint32_t(t & 0xFF), uint32_t((t >> 8) & 0x7F), (t & 0x8000) != 0}; }
        uint32_t t = out->wtable[static_cast<size_t>(st) * ncls + cls_of[b]];
        return {t & 0xFFFFFu, (t >> 20) & 0x7FFu, (t >> 31) != 0};
    };
    const uint32_t st0 = cls[start_id];
    std::vector<uint32_t> cand = {0x09, 0x0B, 0x0C, 0x0A, 0x0D, 0x20};
    for (uint32_t b = 0; b < 256; b++) cand.push_back(b);
    out->head_pad = out->tail_pad = out->reset_pad = -1;
    for (uint32_t b : cand) {
        if (out->reset_pad < 0) {
            bool ok = eom_full[st0] == 0;
            for (uint32_t st = 0; st < nmin && ok; st++) {
                TView t = tv(st, b);
                uint32_t emitted = t.ev | (t.ws ? 0x80000000u : 0u);
                if (emitted != eom_full[st] || t.next != st0) ok = false;
            }
            if (ok) out->reset_pad = static_cast<int>(b);
        }
        if (out->head_pad < 0) { TView t = tv(st0, b); if (t.next == st0 && t.ev == 0 && !t.ws) out->head_pad = static_cast<int>(b); }
        if (out->tail_pad < 0) {
            bool ok = true;
            for (uint32_t st = 0; st < nmin && ok; st++) {
                TView t = tv(st, b);
                uint32_t emitted = t.ev | (t.ws ? 0x80000000u : 0u);
                if (emitted != eom_full[st]) { ok = false; break; }
                TView t2 = tv(t.next, b);
                if (t2.next != t.next || t2.ev != 0 || t2.ws) ok = false;   // further pads: same state, no output
            }
            if (ok) out->tail_pad = static_cast<int>(b);
        }
    }
    out->ctable.clear();
    out->n_cmp = 0;
    if (!wide) {   // column-merged copy of the dense table
        std::map<std::vector<uint16_t>, uint32_t> cols;
        std::vector<uint32_t> rep;
        for (uint32_t b = 0; b < 256; b++) {
            std::vector<uint16_t> col(nmin);
            for (uint32_t st = 0; st < nmin; st++) col[st] = out->table[static_cast<size_t>(st) * 256 + b];
            auto it = cols.find(col);
            if (it == cols.end()) { it = cols.emplace(col, static_cast<uint32_t>(rep.size())).first; rep.push_back(b); }
            out->ccls[b] = static_cast<uint8_t>(it->second);
        }
        out->n_cmp = static_cast<uint32_t>(rep.size());
        out->ctable.resize(static_cast<size_t>(nmin) * out->n_cmp);
        for (uint32_t st = 0; st < nmin; st++)
            for (uint32_t c = 0; c < out->n_cmp; c++) out->ctable[static_cast<size_t>(st) * out->n_cmp + c] = out->table[static_cast<size_t>(st) * 256 + rep[c]];
    }
    if (!exact_nonoverlap) build_pair_filter(dict, *out, &out->filter);
    return "";
}

}  // namespace triage
