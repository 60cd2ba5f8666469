/*
 * gen.c — TEST INFRASTRUCTURE: deterministic synthetic message sets for the five
 * BASELINE.json configs.  Integer arithmetic only, so the
 * same seed gives the same bytes on every machine.  No reference code involved.
 */
#include "triage_oracle.h"
#include <stdlib.h>
#include <string.h>

/* xoshiro256** seeded by splitmix64 */
typedef struct { uint64_t s[4]; } rng_t;
static uint64_t splitmix64(uint64_t *x)
{
    uint64_t z = (*x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t rng_next(rng_t *r)
{
    uint64_t *s = r->s;
    uint64_t result = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
    s[2] ^= t; s[3] = rotl(s[3], 45);
    return result;
}
static void rng_seed(rng_t *r, uint64_t seed) { for (int i = 0; i < 4; i++) r->s[i] = splitmix64(&seed); }
static uint32_t rng_below(rng_t *r, uint32_t n) { return (uint32_t)(((rng_next(r) >> 32) * (uint64_t)n) >> 32); }
static int rng_chance(rng_t *r, uint32_t per10000) { return rng_below(r, 10000) < per10000; }

/* ------------------------------------------------------------ vocabulary -- */
#define VOCAB 2048
static const char *BASE_WORDS[] = {
    "the","of","and","to","in","is","that","for","it","as","with","on","be","at","by","this","have","from","or",
    "one","had","not","but","all","were","we","there","can","an","your","which","their","said","if","do","will",
    "each","about","up","out","them","then","she","many","some","so","these","would","other","into","has","more",
    "her","two","like","him","see","time","could","no","make","than","first","been","its","now","people","my",
    "made","over","did","down","only","way","find","use","may","water","long","little","very","after","words",
    "called","just","most","know","get","through","back","much","go","new","write","our","me","man","too","any",
    "day","same","look","think","also","around","another","came","come","work","three","must","because","does",
    "part","even","place","well","such","here","take","things","help","put","years","different","away","again",
    "off","went","old","number","great1","tell","men","say","small","every","found","still","between","name",
    "should","home","big","give","air","line","set","own","under","read","last","never","us","left","end","along",
    "while","might","next","sound","below","saw","something","thought","both","few","those","always","looked",
    "show","large","often","together","asked","house","world","going","want","school","until","form","food",
    "keep","children","feet","land","side","without","boy","once","animals","life","enough","took","sometimes",
    "four","head","above","kind","began","almost","live","page","got","earth","need","far","hand","high1","year",
    "mother","light","parts","country","father","let","night","following","picture","being","study","second",
    "eyes","paper","group","run","story","since","white","days","hard","open","example","begin","river","carry",
    "state","once","book","hear","stop","without","later","miss","idea","eat","face","watch","far","real","almost",
    "let","girl","mountains","cut","young","talk","list","song","leave","family","body","music","color","stand",
    "sun","questions","fish","area","mark","dog","horse","birds","problem","complete","room","knew","ever","piece",
    "told","usually","friends","easy","heard","order","red","door","sure","become","top","ship","across","today",
    "during","short","better","best","however","low","hours","black","products","happened","whole","measure",
    "remember","early","waves","reached","listen","wind","rock","space","covered","fast","several","hold","himself",
    "toward","five","step","morning","passed","vowel","true","hundred","against","pattern","numeral","table","north",
    "slowly","money","map","farm","pulled","draw","voice","seen","cold","cried","plan","notice","south","sing","war",
    "ground","fall","king","town","unit","figure","certain","field","travel","wood","fire","upon",
    /* near-misses for the keyword / sentiment / question dictionaries */
    "urge","urgency1","imminent","immense","emerge","emerged","critic","critics","prior","priory","import","imports",
    "son","soot","sap","rig","rights","nowhere1","goods","badge","badly","greater","happen","angr","angle","awe",
    "satisfy","terribly","excel","frustrate","whatever1","whom","whose","hows","whyever","whence","wherever1",
    "message","server","request","queue","model","token","latency","customer","account","billing","invoice",
    "ticket","support","update","release","deploy","cluster","session","context","prompt","answer","reply",
};
/* entries ending in a digit are dropped: they exist only so the list above can show which
 * near-misses were considered and rejected because they CONTAIN a dictionary word. */

static const char *ONSET[] = {"b","c","d","f","g","h","j","k","l","m","n","p","r","s","t","v","w","z","br","cl","dr","fl","gl","pl","st","tr","sk","sp","th","ch","sh",""};
static const char *NUCLEUS[] = {"a","e","i","o","u","ai","ea","ee","io","ou","oa","ie"};
static const char *CODA[] = {"","n","r","s","t","l","m","d","ck","ng","st","nt","rd","x","p","ll"};

static const char *BANNED[] = {
    "immediate","emergency","asap","right now","urgent","important","priority","critical","soon",
    "what","how","why","when","where","who",
};
static const char *BANNED_EQ[] = {"good","great","excellent","happy","satisfied","bad","terrible","awful","angry","frustrated"};

static char g_vocab[VOCAB][16];
static uint8_t g_vlen[VOCAB];
static uint64_t g_cdf[VOCAB];
static int g_vocab_ready = 0;

static int word_ok(const char *w)
{
    size_t n = strlen(w);
    if (n == 0 || n > 14) return 0;
    if (w[n - 1] >= '0' && w[n - 1] <= '9') return 0;
    for (size_t i = 0; i < sizeof(BANNED) / sizeof(BANNED[0]); i++) if (strstr(w, BANNED[i])) return 0;
    for (size_t i = 0; i < sizeof(BANNED_EQ) / sizeof(BANNED_EQ[0]); i++) if (strcmp(w, BANNED_EQ[i]) == 0) return 0;
    return 1;
}

static void build_vocab(void)
{
    if (g_vocab_ready) return;
    uint32_t n = 0;
    size_t nb = sizeof(BASE_WORDS) / sizeof(BASE_WORDS[0]);
    for (size_t i = 0; i < nb && n < VOCAB; i++) {
        if (!word_ok(BASE_WORDS[i])) continue;
        int dup = 0;
        for (uint32_t j = 0; j < n; j++) if (strcmp(g_vocab[j], BASE_WORDS[i]) == 0) { dup = 1; break; }
        if (dup) continue;
        strcpy(g_vocab[n++], BASE_WORDS[i]);
    }
    rng_t r;
    rng_seed(&r, 0x5EEDF00Dull);
    while (n < VOCAB) {
        char w[32] = "";
        uint32_t syl = 1 + rng_below(&r, 3);
        for (uint32_t s = 0; s < syl; s++) {
            strcat(w, ONSET[rng_below(&r, 32)]);
            strcat(w, NUCLEUS[rng_below(&r, 12)]);
            strcat(w, CODA[rng_below(&r, 16)]);
        }
        if (!word_ok(w) || strlen(w) < 2) continue;
        int dup = 0;
        for (uint32_t j = 0; j < n; j++) if (strcmp(g_vocab[j], w) == 0) { dup = 1; break; }
        if (dup) continue;
        strcpy(g_vocab[n++], w);
    }
    uint64_t acc = 0;
    for (uint32_t i = 0; i < VOCAB; i++) {          /* Zipf s = 1 */
        g_vlen[i] = (uint8_t)strlen(g_vocab[i]);
        acc += (1ull << 40) / (i + 1);
        g_cdf[i] = acc;
    }
    g_vocab_ready = 1;
}

static uint32_t zipf_word(rng_t *r)
{
    uint64_t x = rng_next(r) % g_cdf[VOCAB - 1];
    uint32_t lo = 0, hi = VOCAB - 1;
    while (lo < hi) {
        uint32_t mid = (lo + hi) / 2;
        if (g_cdf[mid] > x) hi = mid; else lo = mid + 1;
    }
    return lo;
}

/* ---------------------------------------------------------------- emit ---- */
typedef struct { uint8_t *p; uint64_t n, lim; int dry; } buf_t;
static void put(buf_t *b, const void *s, uint64_t n)
{
    /* never write past the current message's end: the next bytes belong to another message
     * (possibly another thread's block) */
    if (!b->dry && b->n < b->lim) memcpy(b->p + b->n, s, n < b->lim - b->n ? n : b->lim - b->n);
    b->n += n;
}
static void put_cased(buf_t *b, const char *w, uint32_t n, int mode, rng_t *r)
{
    /* mode 0 as-is, 1 Title, 2 UPPER, 3 random per-letter case */
    char t[48];
    for (uint32_t i = 0; i < n; i++) {
        char c = w[i];
        int up = mode == 2 || (mode == 1 && i == 0) || (mode == 3 && (rng_next(r) & 1));
        if (up && c >= 'a' && c <= 'z') c = (char)(c - 32);
        t[i] = c;
    }
    put(b, t, n);
}

static const char *RT_KW[] = {"immediate","emergency","asap","right now"};
static const char *HI_KW[] = {"urgent","important","priority","critical","soon"};
static const char *SENT[] = {"good","great","excellent","happy","satisfied","bad","terrible","awful","angry","frustrated"};
static const char *QW[] = {"what","how","why","when","where","who"};
static const char *EDGE_WS[] = {"\xC2\xA0","\xE3\x80\x80","\xE2\x80\x83","\xC2\x85","\xE1\x9A\x80","\xE2\x80\xA8","\xE2\x81\x9F","\t","\n","\r\n","\v","\f"};
static const char *EDGE_TOK[] = {
    "a\xC5\xBF" "ap", "\xC5\xBFoon", "ris\xE2\x84\xAA", "cr\xC4\xB0tical", "sat\xC4\xB0sfied", "terr\xC4\xB0" "ble", "\xE2\x84\xAA",
    "\xE6\xB6\x88\xE6\x81\xAF", "\xE7\xB4\xA7\xE6\x80\xA5", "\xF0\x9F\x98\x80", "\xFF", "\xC3", "\xE2\x80", "\xC0\xA0", "\xED\xA0\x80", "\x80\x80",
    "URGENT", "AsAp", "Right Now", "right\xC2\xA0now", "right  now", "happy!", "good", "BAD", "somehow", "WHAT", "\x00", "\x1C", "\x1F",
    "i\xCC\x87mmediate", "\xC4\xB1mportant", "\xE2\x80\x8Bsoon", "good\xE3\x80\x80", "\xEF\xBF\xBD",
};

/* one message body of (about) target bytes */
static void gen_message(rng_t *r, buf_t *b, uint32_t target, int n_rt, int n_hi, int n_sent, int n_q,
                        int end_q, int edgy)
{
    uint64_t start = b->n;
    b->lim = start + target;
    /* spread the insertions over the message */
    uint32_t approx_tokens = target / 6 + 1;
    int ins_total = n_rt + n_hi + n_sent + n_q;
    uint32_t tok = 0;
    while (b->n - start < target) {
        int did = 0;
        if (ins_total > 0 && rng_below(r, approx_tokens) < (uint32_t)ins_total * 2 + 1) {
            int pick = (int)rng_below(r, (uint32_t)ins_total);
            const char *w;
            int kind;
            if (pick < n_rt) { w = RT_KW[rng_below(r, 4)]; n_rt--; kind = 0; }
            else if (pick < n_rt + n_hi) { w = HI_KW[rng_below(r, 5)]; n_hi--; kind = 0; }
            else if (pick < n_rt + n_hi + n_sent) { w = SENT[rng_below(r, 10)]; n_sent--; kind = 1; }
            else { w = QW[rng_below(r, 6)]; n_q--; kind = 2; }
            ins_total--;
            int mode = (int)rng_below(r, 4);
            if (kind == 0 && rng_chance(r, 2000)) {           /* embedded as a substring */
                uint32_t v = zipf_word(r);
                put(b, g_vocab[v], g_vlen[v]);
                put_cased(b, w, (uint32_t)strlen(w), mode, r);
                put(b, "x", 1);
            } else {
                put_cased(b, w, (uint32_t)strlen(w), mode, r);
                if (kind == 1 && rng_chance(r, 1500)) put(b, "!", 1);   /* "happy!" is not a sentiment field */
            }
            did = 1;
        }
        if (!did) {
            if (edgy && rng_chance(r, 1500)) {
                const char *t = EDGE_TOK[rng_below(r, sizeof(EDGE_TOK) / sizeof(EDGE_TOK[0]))];
                uint32_t tl = (uint32_t)strlen(t);
                if (tl == 0) { put(b, "\0", 1); } else put(b, t, tl);
            } else {
                uint32_t v = zipf_word(r);
                uint32_t c = rng_below(r, 100);
                int mode = c < 10 ? 1 : (c < 12 ? 2 : 0);
                put_cased(b, g_vocab[v], g_vlen[v], mode, r);
                if (rng_chance(r, 300)) put(b, &",.!"[rng_below(r, 3)], 1);
            }
        }
        tok++;
        if (edgy && rng_chance(r, 2500)) {
            const char *ws = EDGE_WS[rng_below(r, sizeof(EDGE_WS) / sizeof(EDGE_WS[0]))];
            put(b, ws, strlen(ws));
        } else if (rng_chance(r, 100)) put(b, "\n", 1);
        else if (rng_chance(r, 50)) put(b, "  ", 2);
        else put(b, " ", 1);
    }
    /* exact length: truncate (may cut a token; for edgy messages this also makes ragged UTF-8 tails) */
    b->n = start + target;
    if (end_q && target > 0) { if (!b->dry) b->p[b->n - 1] = '?'; }
}

static uint32_t pick_len(rng_t *r, const orc_gen_cfg *c)
{
    if (c->mix == 1) {                    /* C3: 70% U[256,1024] / 25% U[1024,4096] / 5% U[4096,16384] */
        uint32_t x = rng_below(r, 100);
        if (x < 70) return 256 + rng_below(r, 1024 - 256 + 1);
        if (x < 95) return 1024 + rng_below(r, 4096 - 1024 + 1);
        return 4096 + rng_below(r, 16384 - 4096 + 1);
    }
    return c->len_lo + rng_below(r, c->len_hi - c->len_lo + 1);
}

#define GEN_BLOCK 1024u

/* messages are generated in independent blocks of GEN_BLOCK, each with its own stream seeded
 * from (seed, block), so any number of threads produces the same bytes */
static uint64_t gen_block(const orc_gen_cfg *cfg, uint32_t blk, uint64_t base, uint64_t *offsets,
                          uint8_t *bytes, orc_meta *meta, uint64_t *ts)
{
    rng_t r;
    uint64_t sd = cfg->seed + 0x632BE59BD9B4E019ull * (uint64_t)(blk + 1);
    rng_seed(&r, sd);
    buf_t b = {bytes, base, 0, bytes == NULL};
    uint32_t edge_rate = cfg->plain ? 0 : (cfg->edge_permille_x10 ? cfg->edge_permille_x10 : 50);
    uint32_t burst_left = 0;
    uint32_t lo = blk * GEN_BLOCK, hi = lo + GEN_BLOCK;
    if (hi > cfg->n) hi = cfg->n;
    for (uint32_t i = lo; i < hi; i++) {
        orc_meta m;
        memset(&m, 0, sizeof(m));
        m.preset = ORC_PRIO_NORMAL;
        uint32_t len = pick_len(&r, cfg);
        int edgy = rng_chance(&r, edge_rate);
        int n_rt = 0, n_hi = 0, n_sent = 0, n_q = 0, end_q = 0;
        if (cfg->mix == 2) {              /* C5-style skew: realtime bursts, rest mostly quiet */
            if (burst_left == 0 && rng_below(&r, 20000) == 0) burst_left = 64 + rng_below(&r, 512);
            if (burst_left > 0) { burst_left--; n_rt = 1 + (int)rng_below(&r, 3); }
            else if (rng_chance(&r, 300)) n_hi = 1 + (int)rng_below(&r, 2);
        } else {
            if (rng_chance(&r, 200)) n_rt = 1 + (int)rng_below(&r, 3);
            if (rng_chance(&r, 800)) n_hi = 1 + (int)rng_below(&r, 3);
        }
        if (rng_chance(&r, 500)) n_sent = 1 + (int)rng_below(&r, 3);
        if (rng_chance(&r, 300)) n_q = 1;
        if (rng_chance(&r, 100)) end_q = 1;
        if (!cfg->plain) {
            uint32_t x = rng_below(&r, 1000);
            if (x < 50) { static const uint8_t ps[4] = {1, 2, 4, 0}; m.preset = ps[rng_below(&r, 4)]; }
            else if (x < 80) { m.user_priority = (uint8_t)(1 + rng_below(&r, 5)); }      /* 5 = unrecognised */
            else if (x < 100) { m.user_default = (uint8_t)(1 + rng_below(&r, 4)); }
            if (rng_chance(&r, 50)) m.flags |= ORC_FLAG_VIP;
            if (rng_chance(&r, 100)) m.meta_priority = (uint8_t)(1 + rng_below(&r, 4));
            if (rng_chance(&r, 10)) len = 0;                                             /* empty content */
            else if (edgy && rng_chance(&r, 1000)) len = 10000 + rng_below(&r, 2);       /* rule-2 boundary */
            else if (edgy && rng_chance(&r, 500)) len = 1 + rng_below(&r, 24);           /* tiny */
        }
        if (offsets) offsets[i] = b.n;
        if (len) gen_message(&r, &b, len, n_rt, n_hi, n_sent, n_q, end_q, edgy);
        if (meta) meta[i] = m;
        if (ts) ts[i] = 1700000000000000000ull + (uint64_t)i * 100ull;
    }
    return b.n - base;
}

typedef struct {
    const orc_gen_cfg *cfg; uint32_t nblk, tid, nthr; uint64_t *blk_size; const uint64_t *blk_base;
    uint64_t *offsets; uint8_t *bytes; orc_meta *meta; uint64_t *ts; int phase;
} gjob_t;

static void *gen_worker(void *arg)
{
    gjob_t *j = (gjob_t *)arg;
    for (uint32_t b = j->tid; b < j->nblk; b += j->nthr) {
        if (j->phase == 0) j->blk_size[b] = gen_block(j->cfg, b, 0, NULL, NULL, NULL, NULL);
        else gen_block(j->cfg, b, j->blk_base[b], j->offsets, j->bytes, j->meta, j->ts);
    }
    return NULL;
}

#include <pthread.h>
#include <unistd.h>

uint64_t orc_generate(const orc_gen_cfg *cfg, uint64_t *offsets, uint8_t *bytes, orc_meta *meta, uint64_t *ts)
{
    build_vocab();
    uint32_t nblk = (cfg->n + GEN_BLOCK - 1) / GEN_BLOCK;
    if (nblk == 0) { if (offsets) offsets[0] = 0; return 0; }
    long nc = sysconf(_SC_NPROCESSORS_ONLN);
    uint32_t nthr = nc < 1 ? 1 : (nc > 64 ? 64 : (uint32_t)nc);
    if (nthr > nblk) nthr = nblk;
    uint64_t *size = (uint64_t *)calloc(nblk, sizeof(uint64_t));
    uint64_t *base = (uint64_t *)calloc(nblk + 1, sizeof(uint64_t));
    pthread_t th[64];
    gjob_t jobs[64];
    for (int phase = 0; phase < 2; phase++) {
        if (phase == 1 && bytes == NULL) break;
        for (uint32_t t = 0; t < nthr; t++) {
            jobs[t] = (gjob_t){cfg, nblk, t, nthr, size, base, offsets, bytes, meta, ts, phase};
            if (t) pthread_create(&th[t], NULL, gen_worker, &jobs[t]);
        }
        gen_worker(&jobs[0]);
        for (uint32_t t = 1; t < nthr; t++) pthread_join(th[t], NULL);
        if (phase == 0) for (uint32_t b = 0; b < nblk; b++) base[b + 1] = base[b] + size[b];
    }
    uint64_t total = base[nblk];
    if (offsets) offsets[cfg->n] = total;
    free(size);
    free(base);
    return total;
}
