// ingest.cc — the wire format on either side of the path, in batches and without a per-message heap detour (host only).
//
//   triage_ingest_json        api/handlers.go:160-176: c.ShouldBindJSON(&message) for a batch of request bodies.  Reads the
//                             fields of models.Message (pkg/models/message.go:58-74) the path uses straight into the padded
//                             arena + triage_meta records; every other field is only type-checked, because a type error in
//                             ANY known field makes the reference answer 400 and drop the message.
//   triage_emit_analysis_json api/handlers.go:180-190: json.Marshal(AnalyzeMessageContent(content)) -> Metadata["analysis"],
//                             for a batch of results.
//
// encoding/json semantics restated here (Go 1.21): object keys match the struct's json names exactly or under
// strings.EqualFold; unknown keys are skipped; null leaves a field as it is; a later duplicate key overrides; strings
// are unquoted with invalid UTF-8 bytes and lone surrogate escapes replaced by U+FFFD; integers must be written without
// fraction or exponent and fit int64; time.Time fields take RFC 3339 strings; Decoder.Decode reads ONE value and ignores
// what follows it.  Metadata is a map[string]interface{}: its keys are case-sensitive.
#include <algorithm>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/triage.h"
#include "ingest.h"

namespace triage {
namespace {

struct Cur {
    const uint8_t *p, *end;
    bool ok = true;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool eat(char c) { ws(); if (p < end && *p == (uint8_t)c) { p++; return true; } return false; }
    int peek() { ws(); return p < end ? *p : -1; }
};

void put_rune(std::string *out, uint32_t r) {
    if (r < 0x80) out->push_back((char)r);
    else if (r < 0x800) { out->push_back((char)(0xC0 | (r >> 6))); out->push_back((char)(0x80 | (r & 0x3F))); }
    else if (r < 0x10000) { out->push_back((char)(0xE0 | (r >> 12))); out->push_back((char)(0x80 | ((r >> 6) & 0x3F))); out->push_back((char)(0x80 | (r & 0x3F))); }
    else { out->push_back((char)(0xF0 | (r >> 18))); out->push_back((char)(0x80 | ((r >> 12) & 0x3F))); out->push_back((char)(0x80 | ((r >> 6) & 0x3F))); out->push_back((char)(0x80 | (r & 0x3F))); }
}

// width of the valid UTF-8 sequence at p (unicode/utf8 rules), 0 if invalid
uint32_t utf8_width(const uint8_t *p, size_t n) {
    const uint8_t b0 = p[0];
    if (b0 < 0x80) return 1;
    if (b0 < 0xC2 || b0 > 0xF4) return 0;
    if (b0 < 0xE0) return (n >= 2 && (p[1] & 0xC0) == 0x80) ? 2 : 0;
    if (b0 < 0xF0) {
        const uint8_t lo = b0 == 0xE0 ? 0xA0 : 0x80, hi = b0 == 0xED ? 0x9F : 0xBF;
        return (n >= 3 && p[1] >= lo && p[1] <= hi && (p[2] & 0xC0) == 0x80) ? 3 : 0;
    }
    const uint8_t lo = b0 == 0xF0 ? 0x90 : 0x80, hi = b0 == 0xF4 ? 0x8F : 0xBF;
    return (n >= 4 && p[1] >= lo && p[1] <= hi && (p[2] & 0xC0) == 0x80 && (p[3] & 0xC0) == 0x80) ? 4 : 0;
}

int hex4(const uint8_t *p) {
    int v = 0;
    for (int i = 0; i < 4; i++) {
        const int c = p[i];
        int d = c >= '0' && c <= '9' ? c - '0' : c >= 'a' && c <= 'f' ? c - 'a' + 10 : c >= 'A' && c <= 'F' ? c - 'A' + 10 : -1;
        if (d < 0) return -1;
        v = v * 16 + d;
    }
    return v;
}

// JSON string at the cursor (after ws); out == nullptr only validates.  encoding/json unquote.
bool parse_string(Cur &c, std::string *out) {
    if (!c.eat('"')) return c.ok = false;
    if (out) out->clear();
    while (c.p < c.end) {
        {   // a run of plain bytes (printable ASCII other than '"' and '\\'), eight at a time: copied as it is
            const uint8_t *s = c.p;
#if defined(__SSE2__)
            while (c.p + 16 <= c.end) {   // 16 at a time where the host has SSE2 (signed compares: bytes >= 0x80 are negative)
                const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i *>(c.p));
                const __m128i stop = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(v, _mm_set1_epi8('"')), _mm_cmpeq_epi8(v, _mm_set1_epi8('\\'))),
                                                  _mm_cmplt_epi8(v, _mm_set1_epi8(0x20)));
                const int mask = _mm_movemask_epi8(stop);
                if (mask) { c.p += __builtin_ctz((unsigned)mask); break; }
                c.p += 16;
            }
#endif
            while (c.p + 8 <= c.end) {
                uint64_t w;
                std::memcpy(&w, c.p, 8);
                const uint64_t k01 = 0x0101010101010101ull, k80 = 0x8080808080808080ull;
                const uint64_t q = w ^ (k01 * 0x22), bs = w ^ (k01 * 0x5C);
                const uint64_t stop = (w | ((w - k01 * 0x20) & ~w) | ((q - k01) & ~q) | ((bs - k01) & ~bs)) & k80;
                if (stop) break;
                c.p += 8;
            }
            while (c.p < c.end && *c.p >= 0x20 && *c.p < 0x80 && *c.p != '"' && *c.p != '\\') c.p++;
            if (out && c.p != s) out->append(reinterpret_cast<const char *>(s), (size_t)(c.p - s));
            if (c.p >= c.end) break;
        }
        const uint8_t b = *c.p;
        if (b == '"') { c.p++; return true; }
        if (b < 0x20) return c.ok = false;                       // raw control character: syntax error
        if (b == '\\') {
            if (c.p + 1 >= c.end) return c.ok = false;
            const uint8_t e = c.p[1];
            c.p += 2;
            uint32_t r;
            switch (e) {
                case '"': r = '"'; break; case '\\': r = '\\'; break; case '/': r = '/'; break;
                case 'b': r = '\b'; break; case 'f': r = '\f'; break; case 'n': r = '\n'; break;
                case 'r': r = '\r'; break; case 't': r = '\t'; break;
                case 'u': {
                    if (c.p + 4 > c.end) return c.ok = false;
                    int v = hex4(c.p);
                    if (v < 0) return c.ok = false;
                    c.p += 4;
                    r = (uint32_t)v;
                    if (r >= 0xD800 && r < 0xDC00) {              // high surrogate: needs \uDC00..DFFF right behind it
                        int lo = -1;
                        if (c.p + 6 <= c.end && c.p[0] == '\\' && c.p[1] == 'u') lo = hex4(c.p + 2);
                        if (lo >= 0xDC00 && lo < 0xE000) { r = 0x10000 + ((r - 0xD800) << 10) + (uint32_t)(lo - 0xDC00); c.p += 6; }
                        else r = 0xFFFD;
                    } else if (r >= 0xDC00 && r < 0xE000) r = 0xFFFD;
                    break;
                }
                default: return c.ok = false;
            }
            if (out) put_rune(out, r);
            continue;
        }
        const uint32_t w = utf8_width(c.p, (size_t)(c.end - c.p));
        if (w == 0) { if (out) put_rune(out, 0xFFFD); c.p++; continue; }   // invalid byte -> U+FFFD
        if (out) out->append(reinterpret_cast<const char *>(c.p), w);
        c.p += w;
    }
    return c.ok = false;
}

// JSON number; *is_int: written without fraction / exponent; value saturates (overflow flagged)
bool parse_number(Cur &c, bool *is_int, int64_t *val, bool *overflow) {
    c.ws();
    const uint8_t *s = c.p;
    bool neg = false;
    if (c.p < c.end && *c.p == '-') { neg = true; c.p++; }
    if (c.p >= c.end) return c.ok = false;
    uint64_t v = 0;
    bool ovf = false;
    if (*c.p == '0') c.p++;
    else if (*c.p >= '1' && *c.p <= '9') {
        while (c.p < c.end && *c.p >= '0' && *c.p <= '9') {
            if (v > (UINT64_MAX - 9) / 10) ovf = true; else v = v * 10 + (*c.p - '0');
            c.p++;
        }
    } else return c.ok = false;
    *is_int = true;
    if (c.p < c.end && *c.p == '.') {
        *is_int = false;
        c.p++;
        if (c.p >= c.end || *c.p < '0' || *c.p > '9') return c.ok = false;
        while (c.p < c.end && *c.p >= '0' && *c.p <= '9') c.p++;
    }
    if (c.p < c.end && (*c.p == 'e' || *c.p == 'E')) {
        *is_int = false;
        c.p++;
        if (c.p < c.end && (*c.p == '+' || *c.p == '-')) c.p++;
        if (c.p >= c.end || *c.p < '0' || *c.p > '9') return c.ok = false;
        while (c.p < c.end && *c.p >= '0' && *c.p <= '9') c.p++;
    }
    (void)s;
    if (neg) { if (v > (uint64_t)INT64_MAX + 1) ovf = true; *val = ovf ? INT64_MIN : (int64_t)(0 - v); }
    else { if (v > (uint64_t)INT64_MAX) ovf = true; *val = ovf ? INT64_MAX : (int64_t)v; }
    *overflow = ovf;
    return true;
}

bool lit(Cur &c, const char *w) {
    const size_t n = std::strlen(w);
    if ((size_t)(c.end - c.p) < n || std::memcmp(c.p, w, n) != 0) return c.ok = false;
    c.p += n;
    return true;
}

// any JSON value, validated and skipped.  kind: 's' string, 'n' number, 'o' object, 'a' array, 't' true/false, 'z' null
bool skip_value(Cur &c, char *kind, int depth = 0) {
    if (depth > 512) return c.ok = false;
    const int ch = c.peek();
    char k = 0;
    if (ch == '"') { k = 's'; if (!parse_string(c, nullptr)) return false; }
    else if (ch == '{') {
        k = 'o'; c.p++;
        if (!c.eat('}')) {
            for (;;) {
                if (!parse_string(c, nullptr)) return false;
                if (!c.eat(':')) return c.ok = false;
                char kk;
                if (!skip_value(c, &kk, depth + 1)) return false;
                if (c.eat(',')) continue;
                if (c.eat('}')) break;
                return c.ok = false;
            }
        }
    } else if (ch == '[') {
        k = 'a'; c.p++;
        if (!c.eat(']')) {
            for (;;) {
                char kk;
                if (!skip_value(c, &kk, depth + 1)) return false;
                if (c.eat(',')) continue;
                if (c.eat(']')) break;
                return c.ok = false;
            }
        }
    } else if (ch == 't') { k = 't'; if (!lit(c, "true")) return false; }
    else if (ch == 'f') { k = 't'; if (!lit(c, "false")) return false; }
    else if (ch == 'n') { k = 'z'; if (!lit(c, "null")) return false; }
    else if (ch == '-' || (ch >= '0' && ch <= '9')) { k = 'n'; bool i, o; int64_t v; if (!parse_number(c, &i, &v, &o)) return false; }
    else return c.ok = false;
    if (kind) *kind = k;
    return true;
}


// strings.EqualFold(key, name) for an ASCII lower-case `name`: ASCII case, plus the two non-ASCII runes whose simple
// folding lands on an ASCII letter (U+017F -> s, U+212A -> k)
bool key_is(const std::string &key, const char *name) {
    size_t i = 0;
    for (const char *n = name; *n; n++) {
        if (i >= key.size()) return false;
        const uint8_t b = (uint8_t)key[i];
        if (b < 0x80) { const char lo = (b >= 'A' && b <= 'Z') ? (char)(b + 32) : (char)b; if (lo != *n) return false; i++; }
        else if (*n == 's' && b == 0xC5 && i + 1 < key.size() && (uint8_t)key[i + 1] == 0xBF) i += 2;
        else if (*n == 'k' && b == 0xE2 && i + 2 < key.size() && (uint8_t)key[i + 1] == 0x84 && (uint8_t)key[i + 2] == 0xAA) i += 3;
        else return false;
    }
    return i == key.size();
}

// time.Time.UnmarshalJSON: RFC 3339, "2006-01-02T15:04:05[.frac](Z|+hh:mm|-hh:mm)"
bool rfc3339(const std::string &t) {
    auto dig = [&](size_t i) { return i < t.size() && t[i] >= '0' && t[i] <= '9'; };
    auto num = [&](size_t i) { return (t[i] - '0') * 10 + (t[i + 1] - '0'); };
    if (t.size() < 20) return false;
    for (size_t i : {0u, 1u, 2u, 3u, 5u, 6u, 8u, 9u, 11u, 12u, 14u, 15u, 17u, 18u}) if (!dig(i)) return false;
    if (t[4] != '-' || t[7] != '-' || (t[10] != 'T' && t[10] != 't') || t[13] != ':' || t[16] != ':') return false;
    const int year = num(0) * 100 + num(2), mon = num(5), day = num(8), hh = num(11), mm = num(14), ss = num(17);
    static const int mdays[] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
    const bool leap = (year % 4 == 0 && year % 100 != 0) || year % 400 == 0;
    if (mon < 1 || mon > 12 || day < 1 || day > mdays[mon - 1] + (mon == 2 && leap ? 1 : 0) || hh > 23 || mm > 59 || ss > 59) return false;
    size_t i = 19;
    if (t[i] == '.' || t[i] == ',') { i++; if (!dig(i)) return false; while (dig(i)) i++; }
    if (i >= t.size()) return false;
    if (t[i] == 'Z' || t[i] == 'z') return i + 1 == t.size();
    if (t[i] != '+' && t[i] != '-') return false;
    if (i + 6 != t.size() || !dig(i + 1) || !dig(i + 2) || t[i + 3] != ':' || !dig(i + 4) || !dig(i + 5)) return false;
    return num(i + 1) <= 23 && num(i + 4) <= 59;
}

// strings.ToLower as far as the comparison with realtime/high/normal/low can tell: ASCII, U+0130 -> i, U+212A -> k
uint8_t priority_name_code(const std::string &s) {
    std::string lo;
    for (size_t i = 0; i < s.size();) {
        const uint8_t b = (uint8_t)s[i];
        if (b < 0x80) { lo.push_back((b >= 'A' && b <= 'Z') ? (char)(b + 32) : (char)b); i++; }
        else if (b == 0xC4 && i + 1 < s.size() && (uint8_t)s[i + 1] == 0xB0) { lo.push_back('i'); i += 2; }
        else if (b == 0xE2 && i + 2 < s.size() && (uint8_t)s[i + 1] == 0x84 && (uint8_t)s[i + 2] == 0xAA) { lo.push_back('k'); i += 3; }
        else { lo.push_back((char)b); i++; }
    }
    if (lo == "realtime") return TRIAGE_PRIO_REALTIME;
    if (lo == "high") return TRIAGE_PRIO_HIGH;
    if (lo == "normal") return TRIAGE_PRIO_NORMAL;
    if (lo == "low") return TRIAGE_PRIO_LOW;
    return 0;
}

struct Decoded {
    std::string content, user_id;
    int64_t priority = 0;
    bool has_meta = false;            // Metadata != nil
    uint8_t user_priority = TRIAGE_UP_ABSENT, meta_priority = 0;
    bool vip = false;
    void reset() {                    // (keeps the strings' capacity from message to message)
        content.clear(); user_id.clear();
        priority = 0; has_meta = false; user_priority = TRIAGE_UP_ABSENT; meta_priority = 0; vip = false;
    }
};

// Metadata map[string]interface{}: only three keys matter; a later duplicate overrides (the map entry is replaced)
bool decode_metadata(Cur &c, Decoded *d) {
    c.p++;   // '{'
    d->has_meta = true;
    if (c.eat('}')) return true;
    std::string key, sval;
    for (;;) {
        if (!parse_string(c, &key)) return false;
        if (!c.eat(':')) return c.ok = false;
        const bool up = key == "user_priority", mp = key == "priority", vip = key == "vip_user";
        if ((up || mp) && c.peek() == '"') {
            if (!parse_string(c, &sval)) return false;
            const uint8_t code = priority_name_code(sval);
            if (up) d->user_priority = code ? code : (uint8_t)TRIAGE_UP_UNRECOGNISED;   // a string that names no priority still takes the branch (:68-82)
            else d->meta_priority = code;
        } else {
            char k;
            if (!skip_value(c, &k)) return false;
            if (up) d->user_priority = TRIAGE_UP_ABSENT;        // not a string: the .(string) assertion fails
            if (mp) d->meta_priority = 0;
        }
        if (vip) d->vip = true;                                  // presence of the key is the rule's condition (queue_factory.go:217)
        if (c.eat(',')) continue;
        if (c.eat('}')) return true;
        return c.ok = false;
    }
}

// one request body -> Decoded; false: the handler would answer 400
bool decode_message(const uint8_t *p, const uint8_t *end, Decoded *d) {
    Cur c{p, end};
    const int first = c.peek();
    if (first == 'n') return lit(c, "null");                    // null into a struct: nothing happens
    if (first != '{') return false;                               // any other JSON value: UnmarshalTypeError (or a syntax error)
    c.p++;
    bool type_error = false;
    if (c.eat('}')) return true;
    std::string key, sval;
    for (;;) {
        if (!parse_string(c, &key)) return false;
        if (!c.eat(':')) return false;
        const int vk = c.peek();
        enum { STR, INT, TIME, TIMEPTR, META, OTHER } want = OTHER;
        int field = -1;
        static const struct { const char *name; int kind; } kFields[] = {
            {"id", STR}, {"conversation_id", STR}, {"user_id", STR}, {"content", STR}, {"priority", INT}, {"status", STR},
            {"queue_name", STR}, {"retry_count", INT}, {"max_retries", INT}, {"timeout", INT}, {"created_at", TIME},
            {"updated_at", TIME}, {"scheduled_at", TIMEPTR}, {"completed_at", TIMEPTR}, {"metadata", META}};
        for (int f = 0; f < 15 && field < 0; f++) if (key == kFields[f].name) field = f;          // exact name first
        for (int f = 0; f < 15 && field < 0; f++) if (key_is(key, kFields[f].name)) field = f;    // then strings.EqualFold
        if (field >= 0) want = (decltype(want))kFields[field].kind;
        if (want == OTHER) { char k; if (!skip_value(c, &k)) return false; }
        else if (vk == 'n') { if (!lit(c, "null")) return false; if (want == META) { d->has_meta = false; d->user_priority = 0; d->meta_priority = 0; d->vip = false; } }
        else if (want == STR) {
            if (vk != '"') { char k; if (!skip_value(c, &k)) return false; type_error = true; }
            else {
                if (!parse_string(c, field == 3 ? &d->content : field == 2 ? &d->user_id : nullptr)) return false;   // (a later duplicate overrides)
            }
        } else if (want == INT) {
            if (vk == '-' || (vk >= '0' && vk <= '9')) {
                bool is_int, ovf; int64_t v;
                if (!parse_number(c, &is_int, &v, &ovf)) return false;
                if (!is_int || ovf) type_error = true;
                else if (field == 4) d->priority = v;
            } else { char k; if (!skip_value(c, &k)) return false; type_error = true; }
        } else if (want == TIME || want == TIMEPTR) {
            if (vk != '"') { char k; if (!skip_value(c, &k)) return false; type_error = true; }
            else { if (!parse_string(c, &sval)) return false; if (!rfc3339(sval)) type_error = true; }
        } else {   // META
            if (vk != '{') { char k; if (!skip_value(c, &k)) return false; type_error = true; }
            else if (!decode_metadata(c, d)) return false;
        }
        if (c.eat(',')) continue;
        if (c.eat('}')) break;
        return false;
    }
    return !type_error;   // the decoder reads the whole object and reports the first type error at the end
}

}  // namespace

std::string ingest_json(const uint8_t *bodies, const uint64_t *body_offsets, uint32_t n, int pad, UserLookup lookup, const void *ctx,
                        const IngestOut &out, uint64_t *total) {
    if ((n && (!bodies || !body_offsets)) || !out.offsets || !out.lengths || !out.meta || !out.status || !total) return "null argument";
    if (pad < 0 || pad > 255) return "this dictionary has no pad byte for the padded arena";
    // Messages are independent until their place in the arena is fixed, so the bodies are decoded a block at a time on a
    // few threads (TRIAGE_INGEST_THREADS, default: up to 16), then laid out in order and copied into the arena by the same
    // threads.  The lookup callback is called from those threads: it only reads the engine's table.
    unsigned nthr = 1;
    if (n >= 2048) {
        nthr = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
        if (const char *v = getenv("TRIAGE_INGEST_THREADS")) { const int x = atoi(v); if (x >= 1 && x <= 64) nthr = (unsigned)x; }
    }
    constexpr uint32_t kBlock = 8192;
    std::vector<Decoded> dec(std::min<uint32_t>(n, kBlock));
    std::string err;
    uint64_t o = 0;
    auto parallel = [&](uint32_t count, auto &&fn) {
        if (nthr <= 1 || count < 512) { fn(0u, count); return; }
        std::vector<std::thread> th;
        const uint32_t per = (count + nthr - 1) / nthr;
        for (unsigned t = 0; t < nthr; t++) {
            const uint32_t lo = std::min(count, t * per), hi = std::min(count, lo + per);
            if (lo < hi) th.emplace_back([&fn, lo, hi] { fn(lo, hi); });
        }
        for (auto &x : th) x.join();
    };
    for (uint32_t base = 0; base < n; base += kBlock) {
        const uint32_t cnt = std::min<uint32_t>(kBlock, n - base);
        parallel(cnt, [&](uint32_t lo, uint32_t hi) {
            for (uint32_t j = lo; j < hi; j++) {
                const uint32_t i = base + j;
                Decoded &d = dec[j];
                d.reset();
                const bool ok = decode_message(bodies + body_offsets[i], bodies + body_offsets[i + 1], &d);
                if (!ok) d.reset();                              // rejected: an empty slot keeps the indices aligned
                out.status[i] = ok ? 0 : 1;
                triage_meta m;
                std::memset(&m, 0, sizeof(m));
                m.preset = (uint8_t)((d.priority < 0 || d.priority > 254) ? 255 : d.priority);   // any other explicit priority: left untouched (:63-65)
                m.user_priority = d.user_priority;
                m.meta_priority = d.meta_priority;
                m.flags = d.vip ? TRIAGE_META_VIP : 0;
                if (ok && lookup) { const int up = lookup(ctx, d.user_id.data(), d.user_id.size()); m.user_default = (uint8_t)((up < 0 || up > 254) ? 0 : up); }
                out.meta[i] = m;
            }
        });
        for (uint32_t j = 0; j < cnt; j++) {                     // layout: 1..16 pad bytes after every message
            const uint64_t len = dec[j].content.size();
            if (len > 0xFFFFFFFFull) return "message longer than 4 GiB";
            out.offsets[base + j] = o;
            out.lengths[base + j] = (uint32_t)len;
            o += (len / 16 + 1) * 16;
        }
        if (out.arena) {
            if (o > out.cap) return "arena too small";
            out.offsets[base + cnt] = o;
            parallel(cnt, [&](uint32_t lo, uint32_t hi) {
                for (uint32_t j = lo; j < hi; j++) {
                    const uint64_t at = out.offsets[base + j], len = dec[j].content.size(), slot = out.offsets[base + j + 1] - at;
                    if (len) std::memcpy(out.arena + at, dec[j].content.data(), len);
                    std::memset(out.arena + at + len, pad, slot - len);
                }
            });
        }
    }
    out.offsets[n] = o;
    *total = o;
    return err;
}

uint64_t emit_analysis_json(const triage_result *res, uint32_t n, char *out, uint64_t cap, uint64_t *offsets) {
    // json.Marshal of map[string]interface{}{"word_count": int, "sentiment": string, "is_question": bool}: keys sorted.
    // Written by hand (no printf): the lengths first, then the bytes, on a few threads for large batches.
    static const char *const kHead[2] = {"{\"is_question\":false,\"sentiment\":\"", "{\"is_question\":true,\"sentiment\":\""};
    static const char *const kSent[3] = {"negative", "neutral", "positive"};
    static const char kMid[] = "\",\"word_count\":";
    static const uint32_t kHeadLen[2] = {(uint32_t)std::strlen(kHead[0]), (uint32_t)std::strlen(kHead[1])};
    static const uint32_t kSentLen[3] = {8, 7, 8};
    const uint32_t kMidLen = (uint32_t)sizeof(kMid) - 1;
    auto digits = [](uint32_t v) { uint32_t d = 1; while (v >= 10) { v /= 10; d++; } return d; };
    auto sent_of = [](const triage_result &r) { return r.sentiment > 0 ? 2 : (r.sentiment < 0 ? 0 : 1); };
    std::vector<uint64_t> own;
    uint64_t *offs = offsets;
    if (!offs && out) { own.resize((size_t)n + 1); offs = own.data(); }
    uint64_t o = 0;
    for (uint32_t i = 0; i < n; i++) {
        if (offs) offs[i] = o;
        o += kHeadLen[res[i].question ? 1 : 0] + kSentLen[sent_of(res[i])] + kMidLen + digits(res[i].words) + 1;
    }
    if (offs) offs[n] = o;
    if (!out) return o;
    auto fill = [&](uint32_t lo, uint32_t hi) {
        for (uint32_t i = lo; i < hi; i++) {
            if (offs[i + 1] > cap) break;                       // nothing is written beyond cap
            char *p = out + offs[i];
            const int q = res[i].question ? 1 : 0, sn = sent_of(res[i]);
            std::memcpy(p, kHead[q], kHeadLen[q]); p += kHeadLen[q];
            std::memcpy(p, kSent[sn], kSentLen[sn]); p += kSentLen[sn];
            std::memcpy(p, kMid, kMidLen); p += kMidLen;
            uint32_t v = res[i].words;
            const uint32_t d = digits(v);
            for (uint32_t k = d; k-- > 0;) { p[k] = (char)('0' + v % 10); v /= 10; }
            p[d] = '}';
        }
    };
    unsigned nthr = n >= 65536 ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1;
    if (nthr <= 1) fill(0, n);
    else {
        std::vector<std::thread> th;
        const uint32_t per = (n + nthr - 1) / nthr;
        for (unsigned t = 0; t < nthr; t++) {
            const uint32_t lo = std::min(n, t * per), hi = std::min(n, lo + per);
            if (lo < hi) th.emplace_back([&fill, lo, hi] { fill(lo, hi); });
        }
        for (auto &x : th) x.join();
    }
    return o;
}

}  // namespace triage
