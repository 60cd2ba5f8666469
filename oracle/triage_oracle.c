/*
 * triage_oracle.c — TEST INFRASTRUCTURE (see triage_oracle.h for scope, pins and
 * the reference lines this restates).  Plain C, integer/byte arithmetic only.
 *
 * The work structure deliberately mirrors the Go code (one full pass per keyword
 * pattern, materialised strings.Fields slice, per-field lower-cased copies, two
 * lower-cased content copies per question word, one mutex-free but strictly
 * sequential container/heap per level queue) so that timing it gives a CPU
 * baseline of the same shape as the reference ("port", not "reference").
 */
#define _GNU_SOURCE
#include "triage_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ---------------------------------------------------------------- utf8 ---- */
/* unicode/utf8.DecodeRune: returns rune and width; invalid -> (0xFFFD, 1).   */
#define RUNE_ERROR 0xFFFDu
static uint32_t decode_rune(const uint8_t *p, uint64_t n, uint32_t *width)
{
    uint8_t b0 = p[0];
    if (b0 < 0x80) { *width = 1; return b0; }
    if (b0 < 0xC2 || b0 > 0xF4) { *width = 1; return RUNE_ERROR; }
    if (b0 < 0xE0) {
        if (n < 2 || (p[1] & 0xC0) != 0x80) { *width = 1; return RUNE_ERROR; }
        *width = 2;
        return ((uint32_t)(b0 & 0x1F) << 6) | (p[1] & 0x3F);
    }
    if (b0 < 0xF0) {
        uint8_t lo = 0x80, hi = 0xBF;
        if (b0 == 0xE0) lo = 0xA0;
        if (b0 == 0xED) hi = 0x9F;
        if (n < 3 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80) { *width = 1; return RUNE_ERROR; }
        *width = 3;
        return ((uint32_t)(b0 & 0x0F) << 12) | ((uint32_t)(p[1] & 0x3F) << 6) | (p[2] & 0x3F);
    }
    {
        uint8_t lo = 0x80, hi = 0xBF;
        if (b0 == 0xF0) lo = 0x90;
        if (b0 == 0xF4) hi = 0x8F;
        if (n < 4 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80 || (p[3] & 0xC0) != 0x80) {
            *width = 1; return RUNE_ERROR;
        }
        *width = 4;
        return ((uint32_t)(b0 & 0x07) << 18) | ((uint32_t)(p[1] & 0x3F) << 12) |
               ((uint32_t)(p[2] & 0x3F) << 6) | (p[3] & 0x3F);
    }
}

/* unicode.IsSpace (White_Space property) */
static int is_space_rune(uint32_t r)
{
    if (r <= 0xFF) return (r >= 0x09 && r <= 0x0D) || r == 0x20 || r == 0x85 || r == 0xA0;
    if (r == 0x1680) return 1;
    if (r >= 0x2000 && r <= 0x200A) return 1;
    return r == 0x2028 || r == 0x2029 || r == 0x202F || r == 0x205F || r == 0x3000;
}

/* ------------------------------------------------------------ patterns ---- */
typedef struct {
    int prio, ci;
    uint32_t nrunes;
    uint32_t *runes;
} pat_t;

/* "fast" mode (orc_dict_set_fast): same scores as the per-pattern passes, computed in one pass for the patterns where
 * that is provably the same number.  A case-sensitive literal without a border (no proper prefix equal to a suffix)
 * cannot overlap itself, so its leftmost non-overlapping count is its occurrence count; those literals go into one
 * byte trie that is walked from every rune start.  Every other pattern ((?i) literals, self-overlapping literals)
 * keeps its own count_nonoverlap pass.  Used for the full-size goldens of the 10 000-keyword config, where 10 000
 * passes per message are out of reach; tests/test_oracle.py checks it against the per-pattern mode.              */
typedef struct { uint32_t child_base; uint32_t add[5]; uint8_t has_add; } tnode_t;
struct orc_dict {
    pat_t *pats;
    uint32_t n, cap;
    int fast;
    /* fast mode: open-addressed (node, byte) -> node map + per node score increments */
    uint64_t *tkeys; uint32_t *tvals; uint32_t tmask;
    tnode_t *tnodes; uint32_t tn;
    uint32_t *slow_idx; uint32_t nslow;
    /* AddKeywordPattern with a regular expression (regex_oracle.c) */
    orc_regex **rx; int *rx_prio; uint32_t nrx, rxcap;
};

/* regexp (?i) on a literal rune: unicode.SimpleFold orbit membership.
 * Orbits touching ASCII: {A..Z}<->{a..z}; {K,k,U+212A}; {S,s,U+017F}.       */
static int rune_match(uint32_t pr, uint32_t cr, int ci)
{
    if (pr == cr) return 1;
    if (!ci) return 0;
    if (pr >= 'A' && pr <= 'Z') pr += 32;
    if (pr >= 'a' && pr <= 'z') {
        if (cr == pr || cr == pr - 32) return 1;
        if (pr == 's' && cr == 0x17F) return 1;
        if (pr == 'k' && cr == 0x212A) return 1;
    }
    return 0;
}

orc_dict *orc_dict_new(int with_defaults)
{
    orc_dict *d = (orc_dict *)calloc(1, sizeof(*d));
    if (with_defaults) {
        /* preprocessor.go:28-29, each compiled as "(?i)"+pattern (:32,:39) */
        static const char *rt[] = {"immediate", "emergency", "asap", "right now"};
        static const char *hi[] = {"urgent", "important", "priority", "critical", "soon"};
        for (int i = 0; i < 4; i++) orc_dict_add(d, ORC_PRIO_REALTIME, (const uint8_t *)rt[i], (uint32_t)strlen(rt[i]), 1);
        for (int i = 0; i < 5; i++) orc_dict_add(d, ORC_PRIO_HIGH, (const uint8_t *)hi[i], (uint32_t)strlen(hi[i]), 1);
    }
    return d;
}

int orc_dict_add(orc_dict *d, int prio, const uint8_t *lit, uint32_t len, int ci)
{
    if (len == 0 || prio < 1 || prio > 4) return 4;
    uint32_t *runes = (uint32_t *)malloc(sizeof(uint32_t) * len);
    uint32_t nr = 0;
    for (uint32_t i = 0; i < len;) {
        uint32_t w, r = decode_rune(lit + i, len - i, &w);
        if (r == RUNE_ERROR) { free(runes); return 4; }       /* invalid UTF-8 / U+FFFD literal */
        if (ci && r >= 0x80) { free(runes); return 4; }        /* non-ASCII fold orbits unsupported */
        runes[nr++] = r;
        i += w;
    }
    if (d->n == d->cap) {
        d->cap = d->cap ? d->cap * 2 : 16;
        d->pats = (pat_t *)realloc(d->pats, sizeof(pat_t) * d->cap);
    }
    d->fast = 0;   /* the trie of the fast mode is stale now: orc_dict_set_fast again after the last add */
    d->pats[d->n].prio = prio;
    d->pats[d->n].ci = ci;
    d->pats[d->n].nrunes = nr;
    d->pats[d->n].runes = runes;
    d->n++;
    return 0;
}

void orc_dict_free(orc_dict *d)
{
    if (!d) return;
    for (uint32_t i = 0; i < d->n; i++) free(d->pats[i].runes);
    free(d->pats);
    free(d->tkeys); free(d->tvals); free(d->tnodes); free(d->slow_idx);
    for (uint32_t i = 0; i < d->nrx; i++) orc_regex_free(d->rx[i]);
    free(d->rx); free(d->rx_prio);
    free(d);
}

uint32_t orc_dict_size(const orc_dict *d) { return d->n; }

/* len(pattern.FindAllString(content, -1)) for a literal: successive leftmost
 * non-overlapping matches, search resumes at the previous match end.         */
static uint32_t count_nonoverlap(const pat_t *p, const uint8_t *s, uint64_t n)
{
    uint32_t count = 0;
    uint64_t i = 0;
    while (i < n) {
        uint64_t j = i;
        uint32_t k = 0;
        while (k < p->nrunes && j < n) {
            uint32_t w, r = decode_rune(s + j, n - j, &w);
            if (!rune_match(p->runes[k], r, p->ci)) break;
            j += w;
            k++;
        }
        if (k == p->nrunes) {
            count++;
            i = j;
        } else {
            uint32_t w;
            (void)decode_rune(s + i, n - i, &w);
            i += w;
        }
    }
    return count;
}

/* ------------------------------------------------------ fast mode -------- */
static int pat_bordered(const pat_t *p)
{
    /* prefix function over (folded) runes: non-zero at the end = a proper prefix equals a suffix */
    uint32_t m = p->nrunes;
    if (m < 2) return 0;
    uint32_t *pi = (uint32_t *)calloc(m, sizeof(uint32_t));
    for (uint32_t i = 1; i < m; i++) {
        uint32_t k = pi[i - 1];
        while (k > 0 && p->runes[i] != p->runes[k]) k = pi[k - 1];
        if (p->runes[i] == p->runes[k]) k++;
        pi[i] = k;
    }
    int b = pi[m - 1] > 0;
    free(pi);
    return b;
}
static uint32_t encode_rune(uint32_t r, uint8_t out[4])
{
    if (r < 0x80) { out[0] = (uint8_t)r; return 1; }
    if (r < 0x800) { out[0] = (uint8_t)(0xC0 | (r >> 6)); out[1] = (uint8_t)(0x80 | (r & 0x3F)); return 2; }
    if (r < 0x10000) { out[0] = (uint8_t)(0xE0 | (r >> 12)); out[1] = (uint8_t)(0x80 | ((r >> 6) & 0x3F)); out[2] = (uint8_t)(0x80 | (r & 0x3F)); return 3; }
    out[0] = (uint8_t)(0xF0 | (r >> 18)); out[1] = (uint8_t)(0x80 | ((r >> 12) & 0x3F)); out[2] = (uint8_t)(0x80 | ((r >> 6) & 0x3F));
    out[3] = (uint8_t)(0x80 | (r & 0x3F));
    return 4;
}
static uint32_t tmap_get(const orc_dict *d, uint32_t node, uint8_t b)
{
    uint64_t key = ((uint64_t)node << 8 | b) + 1;
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    for (uint32_t i = (uint32_t)(h >> 40) & d->tmask;; i = (i + 1) & d->tmask) {
        if (d->tkeys[i] == key) return d->tvals[i];
        if (d->tkeys[i] == 0) return 0;
    }
}
static void tmap_put(orc_dict *d, uint32_t node, uint8_t b, uint32_t val)
{
    uint64_t key = ((uint64_t)node << 8 | b) + 1;
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    uint32_t i = (uint32_t)(h >> 40) & d->tmask;
    while (d->tkeys[i] != 0) i = (i + 1) & d->tmask;
    d->tkeys[i] = key;
    d->tvals[i] = val;
}
void orc_dict_set_fast(orc_dict *d, int on)
{
    free(d->tkeys); free(d->tvals); free(d->tnodes); free(d->slow_idx);
    d->tkeys = NULL; d->tvals = NULL; d->tnodes = NULL; d->slow_idx = NULL;
    d->fast = 0; d->tn = 0; d->nslow = 0;
    if (!on) return;
    uint64_t total_bytes = 1;
    for (uint32_t i = 0; i < d->n; i++) total_bytes += 4ull * d->pats[i].nrunes;
    uint32_t cap = 16;
    while (cap < total_bytes * 2) cap <<= 1;
    d->tmask = cap - 1;
    d->tkeys = (uint64_t *)calloc(cap, sizeof(uint64_t));
    d->tvals = (uint32_t *)calloc(cap, sizeof(uint32_t));
    d->tnodes = (tnode_t *)calloc(total_bytes + 1, sizeof(tnode_t));
    d->slow_idx = (uint32_t *)malloc(sizeof(uint32_t) * (d->n ? d->n : 1));
    d->tn = 1;   /* node 0 = root; child id 0 = absent */
    for (uint32_t i = 0; i < d->n; i++) {
        const pat_t *p = &d->pats[i];
        if (p->ci || pat_bordered(p)) { d->slow_idx[d->nslow++] = i; continue; }
        uint32_t node = 0;
        for (uint32_t k = 0; k < p->nrunes; k++) {
            uint8_t enc[4];
            uint32_t w = encode_rune(p->runes[k], enc);
            for (uint32_t j = 0; j < w; j++) {
                uint32_t nx = tmap_get(d, node, enc[j]);
                if (!nx) { nx = d->tn++; tmap_put(d, node, enc[j], nx); }
                node = nx;
            }
        }
        d->tnodes[node].add[p->prio]++;
        d->tnodes[node].has_add = 1;
    }
    d->fast = 1;
}
static void scores_fast(const orc_dict *d, const uint8_t *s, uint64_t n, uint32_t scores[5])
{
    for (uint32_t k = 0; k < d->nslow; k++) {
        const pat_t *p = &d->pats[d->slow_idx[k]];
        if (p->ci) { scores[p->prio] += count_nonoverlap(p, s, n); continue; }
        /* self-overlapping case-sensitive literal: leftmost non-overlapping occurrences by byte search.  A byte-level
         * occurrence of a valid UTF-8 literal always starts and ends on rune boundaries of the text (its first byte
         * is ASCII or a lead byte, never a continuation byte), so this is the rune-level count.                    */
        uint8_t enc[64 * 4];
        uint32_t m = 0;
        if (p->nrunes > 64) { scores[p->prio] += count_nonoverlap(p, s, n); continue; }
        for (uint32_t r = 0; r < p->nrunes; r++) m += encode_rune(p->runes[r], enc + m);
        const uint8_t *q = s, *end = s + n;
        while ((q = (const uint8_t *)memmem(q, (size_t)(end - q), enc, m)) != NULL) { scores[p->prio]++; q += m; }
    }
    if (d->tn <= 1) return;
    for (uint64_t i = 0; i < n;) {
        uint32_t node = 0;
        for (uint64_t j = i; j < n; j++) {
            node = tmap_get(d, node, s[j]);
            if (!node) break;
            if (d->tnodes[node].has_add) for (int p = 1; p <= 4; p++) scores[p] += d->tnodes[node].add[p];
        }
        uint32_t w;
        (void)decode_rune(s + i, n - i, &w);   /* matches start at rune starts only (regexp works on runes) */
        i += w;
    }
}

/* preprocessor.go:141-149 */
void orc_scores(const orc_dict *d, const uint8_t *s, uint64_t n, uint32_t scores[5])
{
    for (int p = 0; p < 5; p++) scores[p] = 0;
    for (uint32_t i = 0; i < d->nrx; i++) scores[d->rx_prio[i]] += orc_regex_count(d->rx[i], s, n);
    if (d->fast) { scores_fast(d, s, n, scores); return; }
    for (uint32_t i = 0; i < d->n; i++) scores[d->pats[i].prio] += count_nonoverlap(&d->pats[i], s, n);
}

/* preprocessor.go:176-184 with a pattern that is not a plain literal; 0 ok, 4 outside the restated subset */
int orc_dict_add_regex(orc_dict *d, int prio, const uint8_t *pattern, uint32_t len)
{
    if (prio < 1 || prio > 4) return 4;
    orc_regex *re = orc_regex_compile(pattern, len);
    if (!re) return 4;
    if (d->nrx == d->rxcap) {
        d->rxcap = d->rxcap ? d->rxcap * 2 : 8;
        d->rx = (orc_regex **)realloc(d->rx, sizeof(orc_regex *) * d->rxcap);
        d->rx_prio = (int *)realloc(d->rx_prio, sizeof(int) * d->rxcap);
    }
    d->rx[d->nrx] = re; d->rx_prio[d->nrx] = prio; d->nrx++;
    return 0;
}

/* ----------------------------------------------------------- analysis ---- */
/* strings.ToLower restricted to what can change an equality / substring test
 * against an all-ASCII needle: ASCII A-Z, U+0130 -> 'i', U+212A -> 'k'.  Other
 * non-ASCII runes stay non-ASCII under unicode.ToLower, so their bytes are
 * copied through; an invalid byte becomes EF BF BD as strings.Map does.       */
static uint64_t go_lower(const uint8_t *s, uint64_t n, uint8_t *out)
{
    uint64_t o = 0;
    for (uint64_t i = 0; i < n;) {
        uint8_t b = s[i];
        if (b < 0x80) {
            out[o++] = (b >= 'A' && b <= 'Z') ? (uint8_t)(b + 32) : b;
            i++;
            continue;
        }
        uint32_t w, r = decode_rune(s + i, n - i, &w);
        if (r == RUNE_ERROR && w == 1) { out[o++] = 0xEF; out[o++] = 0xBF; out[o++] = 0xBD; }
        else if (r == 0x130) out[o++] = 'i';
        else if (r == 0x212A) out[o++] = 'k';
        else { memcpy(out + o, s + i, w); o += w; }
        i += w;
    }
    return o;
}

static const char *POSITIVE[5] = {"good", "great", "excellent", "happy", "satisfied"};      /* :49 */
static const char *NEGATIVE[5] = {"bad", "terrible", "awful", "angry", "frustrated"};       /* :50 */
static const char *QUESTION[6] = {"what", "how", "why", "when", "where", "who"};            /* :51 */

typedef struct { uint64_t off, len; } field_t;

/* preprocessor.go:204-243 (and the twin at :257-296) */
void orc_analyze(const uint8_t *s, uint64_t n, uint32_t *words, int32_t *pos, int32_t *neg, int *question)
{
    /* strings.Fields: materialise the slice like the reference does */
    uint32_t nf = 0, cap = 64;
    field_t *f = (field_t *)malloc(sizeof(field_t) * cap);
    int in = 0;
    uint64_t start = 0;
    for (uint64_t i = 0; i < n;) {
        uint32_t w, r = decode_rune(s + i, n - i, &w);
        if (is_space_rune(r)) {
            if (in) {
                if (nf == cap) { cap *= 2; f = (field_t *)realloc(f, sizeof(field_t) * cap); }
                f[nf].off = start; f[nf].len = i - start; nf++;
                in = 0;
            }
        } else if (!in) { in = 1; start = i; }
        i += w;
    }
    if (in) {
        if (nf == cap) { cap *= 2; f = (field_t *)realloc(f, sizeof(field_t) * cap); }
        f[nf].off = start; f[nf].len = n - start; nf++;
    }
    *words = nf;

    int32_t p = 0, q = 0;
    uint8_t tmp[64];
    for (uint32_t i = 0; i < nf; i++) {
        /* a field longer than any sentiment word (10) cannot equal one even after
         * lowering shrinks it: İ (2 bytes) -> i, so bound by 3 bytes per rune. */
        if (f[i].len > 20) continue;
        uint64_t m = go_lower(s + f[i].off, f[i].len, tmp);
        for (int k = 0; k < 5; k++) {
            if (m == strlen(POSITIVE[k]) && memcmp(tmp, POSITIVE[k], m) == 0) p++;
            if (m == strlen(NEGATIVE[k]) && memcmp(tmp, NEGATIVE[k], m) == 0) q++;
        }
    }
    free(f);
    *pos = p;
    *neg = q;

    int isq = 0;
    if (n > 0 && s[n - 1] == '?') isq = 1;                      /* HasSuffix(content, "?") */
    else if (n > 0) {
        uint8_t *low = (uint8_t *)malloc(n * 3 + 1);
        for (int k = 0; k < 6 && !isq; k++) {
            char needle[8];
            size_t nl = strlen(QUESTION[k]);
            memcpy(needle, QUESTION[k], nl);
            needle[nl++] = ' ';
            uint64_t m = go_lower(s, n, low);                   /* strings.ToLower(content) */
            if (memmem(low, m, needle, nl)) { isq = 1; break; }
            m = go_lower(s, n, low);                            /* second ToLower for HasPrefix */
            if (m >= nl && memcmp(low, needle, nl) == 0) { isq = 1; break; }
        }
        free(low);
    }
    *question = isq;
}

/* ------------------------------------------------------ ProcessMessage ---- */
static void process_one(const orc_dict *d, int default_prio, const uint8_t *s, uint64_t n,
                        const orc_meta *m, orc_result *r)
{
    memset(r, 0, sizeof(*r));
    int preset = m->preset;
    /* :63-65 respect an explicit non-default priority: no analysis at all */
    if (preset != ORC_PRIO_NORMAL && preset != 0) {
        r->label = (uint8_t)preset;
        r->queue_level = r->label;
        return;
    }
    int label = preset, reason = ORC_REASON_NONE, tie = 0;
    uint32_t sc[5] = {0, 0, 0, 0, 0};
    if (m->user_priority != ORC_UP_ABSENT) {                    /* :68-82 */
        if (m->user_priority >= 1 && m->user_priority <= 4) { label = m->user_priority; reason = ORC_REASON_USER_OVERRIDE; }
    } else if (m->user_default != 0) {                          /* :83-86 */
        label = m->user_default;
        reason = ORC_REASON_USER_DEFAULT;
    } else {                                                    /* :88-93 */
        if (n == 0) {                                           /* :120-138 */
            label = (m->meta_priority >= 1 && m->meta_priority <= 4) ? m->meta_priority : default_prio;
        } else {
            orc_scores(d, s, n, sc);
            /* :152-165.  Go ranges over a map in random order with strict '>', so a tie at
             * the maximum is genuinely nondeterministic in the reference.  Convention:
             * the smallest (most urgent) priority wins and the message is flagged TIE.   */
            uint32_t best = 0;
            int bp = default_prio;
            for (int p = 1; p <= 4; p++) {
                if (sc[p] > best) { best = sc[p]; bp = p; }
            }
            if (best > 0) for (int p = bp + 1; p <= 4; p++) if (sc[p] == best) tie = 1;
            label = bp;
        }
        if (label != preset) reason = ORC_REASON_CONTENT_KEYWORDS;
    }
    r->label = (uint8_t)label;
    r->queue_level = r->label;
    r->reason = (uint8_t)reason;
    r->flags = ORC_RF_ANALYZED | (tie ? ORC_RF_TIE : 0);
    r->score_rt = sc[1] > 0xFFFF ? 0xFFFF : (uint16_t)sc[1];
    r->score_hi = sc[2] > 0xFFFF ? 0xFFFF : (uint16_t)sc[2];
    if (n > 0) {                                                /* :199-201 */
        uint32_t words; int32_t pos, neg; int q;
        orc_analyze(s, n, &words, &pos, &neg, &q);
        r->words = words;
        r->sentiment = (int8_t)(pos > neg ? 1 : (neg > pos ? -1 : 0));
        r->question = (uint8_t)q;
        r->flags |= ORC_RF_HAS_ANALYSIS;
    }
}

typedef struct {
    const orc_dict *d; int default_prio;
    const uint64_t *offsets; const uint8_t *bytes; const orc_meta *meta;
    uint32_t lo, hi; orc_result *out;
} job_t;

static void *worker(void *arg)
{
    job_t *j = (job_t *)arg;
    for (uint32_t i = j->lo; i < j->hi; i++)
        process_one(j->d, j->default_prio, j->bytes + j->offsets[i], j->offsets[i + 1] - j->offsets[i], &j->meta[i], &j->out[i]);
    return NULL;
}

void orc_process_batch(const orc_dict *d, int default_prio, const uint64_t *offsets,
                       const uint8_t *bytes, const orc_meta *meta, uint32_t n,
                       orc_result *out, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    if ((uint32_t)nthreads > n) nthreads = n ? (int)n : 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * nthreads);
    job_t *jobs = (job_t *)malloc(sizeof(job_t) * nthreads);
    /* shard by bytes so threads finish together */
    uint64_t total = offsets[n];
    uint32_t lo = 0;
    for (int t = 0; t < nthreads; t++) {
        uint64_t target = total / nthreads * (t + 1);
        uint32_t hi = lo;
        if (t == nthreads - 1) hi = n;
        else while (hi < n && offsets[hi] < target) hi++;
        jobs[t] = (job_t){d, default_prio, offsets, bytes, meta, lo, hi, out};
        lo = hi;
    }
    for (int t = 1; t < nthreads; t++) pthread_create(&th[t], NULL, worker, &jobs[t]);
    worker(&jobs[0]);
    for (int t = 1; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

/* --------------------------------------------------- container/heap ------ */
typedef struct { int prio; uint64_t ts; uint32_t id; } item_t;
struct orc_heap { item_t *a; uint32_t n, cap; };

/* queue.go:22-27 */
static int heap_less(const item_t *x, const item_t *y)
{
    if (x->prio == y->prio) return x->ts < y->ts;
    return x->prio < y->prio;
}
static void heap_swap(orc_heap *h, uint32_t i, uint32_t j) { item_t t = h->a[i]; h->a[i] = h->a[j]; h->a[j] = t; }
static void heap_up(orc_heap *h, uint32_t j)
{
    for (;;) {
        if (j == 0) break;
        uint32_t i = (j - 1) / 2;
        if (i == j || !heap_less(&h->a[j], &h->a[i])) break;
        heap_swap(h, i, j);
        j = i;
    }
}
static void heap_down(orc_heap *h, uint32_t i0, uint32_t n)
{
    uint32_t i = i0;
    for (;;) {
        uint64_t j1 = 2ull * i + 1;
        if (j1 >= n) break;
        uint32_t j = (uint32_t)j1;
        if (j1 + 1 < n && heap_less(&h->a[j1 + 1], &h->a[j1])) j = (uint32_t)j1 + 1;
        if (!heap_less(&h->a[j], &h->a[i])) break;
        heap_swap(h, i, j);
        i = j;
    }
}
orc_heap *orc_heap_new(void) { return (orc_heap *)calloc(1, sizeof(orc_heap)); }
void orc_heap_free(orc_heap *h) { if (h) { free(h->a); free(h); } }
uint32_t orc_heap_len(const orc_heap *h) { return h->n; }
void orc_heap_push(orc_heap *h, int prio, uint64_t ts, uint32_t id)
{
    if (h->n == h->cap) { h->cap = h->cap ? h->cap * 2 : 1024; h->a = (item_t *)realloc(h->a, sizeof(item_t) * h->cap); }
    h->a[h->n] = (item_t){prio, ts, id};
    h->n++;
    heap_up(h, h->n - 1);                                       /* heap.Push */
}
int orc_heap_pop(orc_heap *h, uint32_t *id)
{
    if (h->n == 0) return 0;
    uint32_t n = h->n - 1;                                      /* heap.Pop */
    heap_swap(h, 0, n);
    heap_down(h, 0, n);
    *id = h->a[n].id;
    h->n = n;
    return 1;
}

/* ------------------------------------------- rules, admission, order ----- */
uint32_t orc_admit_and_order(orc_result *res, const orc_meta *meta, const uint64_t *offsets,
                             const uint64_t *ts, uint32_t n, int rules_on, int64_t max_size,
                             uint32_t *order_out)
{
    orc_heap *q[5];
    for (int p = 0; p < 5; p++) q[p] = orc_heap_new();
    uint32_t admitted = 0;
    for (uint32_t i = 0; i < n; i++) {
        /* caller picks the queue from the pre-rule priority: PushMessage(fmt.Sprint(msg.Priority), msg)
         * (api/handlers.go:202); only the four level queues exist (cmd/queue-manager/main.go:58-63) */
        int level = res[i].label;
        res[i].queue_level = (uint8_t)level;
        if (level < 1 || level > 4) { res[i].accepted = 0; res[i].admit = ORC_ADMIT_NOT_FOUND; continue; }
        if (rules_on) {                                         /* queue_manager.go:451-466, first match wins */
            if (meta[i].flags & ORC_FLAG_VIP) res[i].label = ORC_PRIO_HIGH;                 /* queue_factory.go:214-221 */
            else if (offsets[i + 1] - offsets[i] > 10000) res[i].label = ORC_PRIO_LOW;      /* :223-230 */
        }
        if (max_size > 0 && (int64_t)orc_heap_len(q[level]) >= max_size) {                  /* queue.go:101-103 */
            res[i].accepted = 0; res[i].admit = ORC_ADMIT_FULL; continue;
        }
        res[i].accepted = 1;
        res[i].admit = ORC_ADMIT_OK;
        orc_heap_push(q[level], res[i].label, ts[i], i);        /* queue.go:105-111 */
        admitted++;
    }
    uint32_t k = 0;
    for (int p = 1; p <= 4; p++) {                              /* cmd/queue-manager/main.go:112 */
        uint32_t id;
        while (orc_heap_pop(q[p], &id)) order_out[k++] = id;
    }
    for (int p = 0; p < 5; p++) orc_heap_free(q[p]);
    return admitted;
}

uint32_t orc_pipeline(const orc_dict *d, int default_prio, const uint64_t *offsets,
                      const uint8_t *bytes, const orc_meta *meta, const uint64_t *ts, uint32_t n,
                      int rules_on, int64_t max_size, orc_result *out, uint32_t *order_out,
                      int nthreads)
{
    orc_process_batch(d, default_prio, offsets, bytes, meta, n, out, nthreads);
    return orc_admit_and_order(out, meta, offsets, ts, n, rules_on, max_size, order_out);
}

/* ExponentialBackoff.NextBackoff, internal/priorityqueue/worker.go:274-290: float64 product, cap, truncation */
int64_t orc_next_backoff_ns(int64_t initial_ns, int64_t max_ns, double factor, int retry_count)
{
    if (retry_count <= 0) return initial_ns;                    /* :275-277 */
    double backoff = (double)initial_ns;                        /* :280 */
    for (int i = 1; i < retry_count; i++) backoff *= factor;    /* :281-283 */
    if (backoff > (double)max_ns) return max_ns;                /* :285-287 */
    return (int64_t)backoff;                                    /* :289 time.Duration(backoff) */
}
