/*
 * regex_oracle.c — TEST INFRASTRUCTURE (see triage_oracle.h).  AddKeywordPattern with a real regular expression
 * (internal/preprocessor/preprocessor.go:176-184 compiles the string with regexp.Compile, :146 counts
 * len(pattern.FindAllString(content, -1))).  Go's regexp cannot run here, so this restates the published semantics of
 * the syntax subset the tests use: leftmost-first matching (the match a backtracking search finds first: left
 * alternative before right, greedy repetition tries more first, lazy tries fewer first), successive non-overlapping
 * matches, everything on runes with invalid UTF-8 bytes as U+FFFD of width 1.
 *
 * Method: the pattern is parsed into position-independent code fragments (sets, ordered splits, jumps) and matched by
 * a depth-first backtracking search with an explicit stack, anchored at every rune start in turn - the definition of
 * leftmost-first, and a different algorithm from the product's ordered-thread simulation.  A second, independent check
 * is oracle/py_oracle.py, which hands the same patterns to Python's `re` with explicit character classes.
 */
#define _GNU_SOURCE
#include "triage_oracle.h"
#include <ctype.h>
#include <stdlib.h>
#include <string.h>

enum { I_SET, I_SPLIT, I_JMP, I_MATCH };
typedef struct { uint32_t lo, hi; } rr_t;
typedef struct { rr_t *r; int n, cap, negate; } rset_t;
typedef struct { int op; int a, b; rset_t set; } inst_t;      /* a, b: jump targets RELATIVE to this instruction */
typedef struct { inst_t *v; int n; } frag_t;
struct orc_regex { inst_t *code; int n; };

typedef struct { const uint8_t *s; size_t n, i; int fold, dotall, err; } ps_t;

static uint32_t dec(const uint8_t *p, size_t n, uint32_t *w)
{
    uint8_t b0 = p[0];
    *w = 1;
    if (b0 < 0x80) return b0;
    if (b0 < 0xC2 || b0 > 0xF4) return 0xFFFD;
    if (b0 < 0xE0) { if (n < 2 || (p[1] & 0xC0) != 0x80) return 0xFFFD; *w = 2; return ((b0 & 0x1Fu) << 6) | (p[1] & 0x3Fu); }
    if (b0 < 0xF0) {
        uint8_t lo = b0 == 0xE0 ? 0xA0 : 0x80, hi = b0 == 0xED ? 0x9F : 0xBF;
        if (n < 3 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80) return 0xFFFD;
        *w = 3;
        return ((b0 & 0x0Fu) << 12) | ((uint32_t)(p[1] & 0x3F) << 6) | (p[2] & 0x3Fu);
    }
    {
        uint8_t lo = b0 == 0xF0 ? 0x90 : 0x80, hi = b0 == 0xF4 ? 0x8F : 0xBF;
        if (n < 4 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80 || (p[3] & 0xC0) != 0x80) return 0xFFFD;
        *w = 4;
        return ((b0 & 0x07u) << 18) | ((uint32_t)(p[1] & 0x3F) << 12) | ((uint32_t)(p[2] & 0x3F) << 6) | (p[3] & 0x3Fu);
    }
}

static void set_add(rset_t *s, uint32_t lo, uint32_t hi)
{
    if (s->n == s->cap) { s->cap = s->cap ? s->cap * 2 : 8; s->r = (rr_t *)realloc(s->r, sizeof(rr_t) * s->cap); }
    s->r[s->n].lo = lo; s->r[s->n].hi = hi; s->n++;
}
/* (?i): unicode.SimpleFold orbits that touch ASCII */
static void set_add_fold(ps_t *p, rset_t *s, uint32_t lo, uint32_t hi)
{
    set_add(s, lo, hi);
    if (!p->fold) return;
    if (hi >= 0x80) { p->err = 1; return; }
    for (uint32_t c = lo; c <= hi; c++) {
        if (c >= 'a' && c <= 'z') set_add(s, c - 32, c - 32);
        if (c >= 'A' && c <= 'Z') set_add(s, c + 32, c + 32);
        if (c == 'k' || c == 'K') set_add(s, 0x212A, 0x212A);
        if (c == 's' || c == 'S') set_add(s, 0x17F, 0x17F);
    }
}
static int set_has(const rset_t *s, uint32_t c)
{
    int in = 0;
    for (int i = 0; i < s->n; i++) if (c >= s->r[i].lo && c <= s->r[i].hi) { in = 1; break; }
    return in != s->negate;
}
static void perl(rset_t *s, int c, int neg)
{
    rset_t t = {0};
    if (c == 'd') set_add(&t, '0', '9');
    if (c == 'w') { set_add(&t, '0', '9'); set_add(&t, 'A', 'Z'); set_add(&t, 'a', 'z'); set_add(&t, '_', '_'); }
    if (c == 's') { set_add(&t, '\t', '\n'); set_add(&t, '\f', '\r'); set_add(&t, ' ', ' '); }
    if (!neg) { for (int i = 0; i < t.n; i++) set_add(s, t.r[i].lo, t.r[i].hi); }
    else for (uint32_t ch = 0; ch <= 0x10FFFF; ) {      /* complement, range by range */
        int in = 0; uint32_t upto = 0;
        for (int i = 0; i < t.n; i++) if (ch >= t.r[i].lo && ch <= t.r[i].hi) { in = 1; upto = t.r[i].hi; }
        if (in) { ch = upto + 1; continue; }
        uint32_t next = 0x110000;
        for (int i = 0; i < t.n; i++) if (t.r[i].lo > ch && t.r[i].lo < next) next = t.r[i].lo;
        set_add(s, ch, next - 1);
        ch = next;
    }
    free(t.r);
}

static frag_t f_empty(void) { frag_t f = {NULL, 0}; return f; }
static frag_t f_one(inst_t in) { frag_t f; f.v = (inst_t *)malloc(sizeof(inst_t)); f.v[0] = in; f.n = 1; return f; }
static rset_t set_copy(const rset_t *s)
{
    rset_t c = *s;
    c.r = (rr_t *)malloc(sizeof(rr_t) * (s->n ? s->n : 1));
    memcpy(c.r, s->r, sizeof(rr_t) * s->n);
    c.cap = s->n ? s->n : 1;
    return c;
}
static frag_t f_copy(frag_t a)
{
    frag_t f; f.n = a.n; f.v = (inst_t *)malloc(sizeof(inst_t) * (a.n ? a.n : 1));
    for (int i = 0; i < a.n; i++) { f.v[i] = a.v[i]; if (a.v[i].op == I_SET) f.v[i].set = set_copy(&a.v[i].set); }
    return f;
}
static void f_free(frag_t a) { for (int i = 0; i < a.n; i++) if (a.v[i].op == I_SET) free(a.v[i].set.r); free(a.v); }
static frag_t f_cat(frag_t a, frag_t b)      /* consumes both */
{
    frag_t f; f.n = a.n + b.n; f.v = (inst_t *)malloc(sizeof(inst_t) * (f.n ? f.n : 1));
    memcpy(f.v, a.v, sizeof(inst_t) * a.n); memcpy(f.v + a.n, b.v, sizeof(inst_t) * b.n);
    free(a.v); free(b.v);
    return f;
}
static inst_t mk(int op, int a, int b) { inst_t in; memset(&in, 0, sizeof(in)); in.op = op; in.a = a; in.b = b; return in; }
/* a | b, a preferred:  split(+1, +len(a)+2) a jmp(+len(b)+1) b */
static frag_t f_alt(frag_t a, frag_t b)
{
    int la = a.n, lb = b.n;
    frag_t f = f_cat(f_one(mk(I_SPLIT, 1, la + 2)), a);
    f = f_cat(f, f_one(mk(I_JMP, lb + 1, 0)));
    return f_cat(f, b);
}
/* a? : split(+1, +len+1) a   (lazy: targets swapped) */
static frag_t f_quest(frag_t a, int lazy)
{
    int la = a.n;
    return f_cat(f_one(lazy ? mk(I_SPLIT, la + 1, 1) : mk(I_SPLIT, 1, la + 1)), a);
}
/* a* : L1: split(+1, +len+2) a jmp(-len-1) */
static frag_t f_star(frag_t a, int lazy)
{
    int la = a.n;
    frag_t f = f_cat(f_one(lazy ? mk(I_SPLIT, la + 2, 1) : mk(I_SPLIT, 1, la + 2)), a);
    return f_cat(f, f_one(mk(I_JMP, -(la + 1), 0)));
}

static int nullable_code(const inst_t *c, int n);
static frag_t p_alt(ps_t *p);

static int p_escape(ps_t *p, rset_t *set, int in_class, uint32_t *single)
{
    if (p->i >= p->n) { p->err = 1; return 0; }
    int c = p->s[p->i];
    if (c == 'd' || c == 'w' || c == 's') { p->i++; perl(set, c, 0); return 1; }
    if (c == 'D' || c == 'W' || c == 'S') { p->i++; perl(set, c + 32, 1); return 1; }
    uint32_t r;
    if (c == 'n') r = '\n'; else if (c == 't') r = '\t'; else if (c == 'r') r = '\r'; else if (c == 'f') r = '\f'; else if (c == 'v') r = '\v';
    else if (c == 'x') {
        if (p->i + 3 > p->n || !isxdigit(p->s[p->i + 1]) || !isxdigit(p->s[p->i + 2])) { p->err = 1; return 0; }
        char h[3] = {(char)p->s[p->i + 1], (char)p->s[p->i + 2], 0};
        r = (uint32_t)strtoul(h, NULL, 16);
        p->i += 2;
    } else if (ispunct(c) || c == ' ') r = (uint32_t)c;
    else { p->err = 1; return 0; }      /* \b \B \A \z \1 ...: outside the subset */
    p->i++;
    (void)in_class;
    if (single) { *single = r; return 2; }
    set_add_fold(p, set, r, r);
    return 1;
}

static frag_t p_class(ps_t *p)
{
    inst_t in = mk(I_SET, 0, 0);
    if (p->i < p->n && p->s[p->i] == '^') { in.set.negate = 1; p->i++; }
    int first = 1;
    for (;;) {
        if (p->i >= p->n) { p->err = 1; break; }
        if (p->s[p->i] == ']' && !first) { p->i++; break; }
        first = 0;
        if (p->s[p->i] == '[' && p->i + 1 < p->n && p->s[p->i + 1] == ':') { p->err = 1; break; }
        uint32_t lo, hi, w;
        if (p->s[p->i] == '\\') {
            p->i++;
            int c = p->i < p->n ? p->s[p->i] : 0;
            if (c && strchr("dwsDWS", c)) { p_escape(p, &in.set, 1, NULL); continue; }
            if (p_escape(p, &in.set, 1, &lo) != 2) break;
        } else { lo = dec(p->s + p->i, p->n - p->i, &w); if (lo == 0xFFFD) { p->err = 1; break; } p->i += w; }
        hi = lo;
        if (p->i + 1 < p->n && p->s[p->i] == '-' && p->s[p->i + 1] != ']') {
            p->i++;
            if (p->s[p->i] == '\\') { p->i++; if (p_escape(p, &in.set, 1, &hi) != 2) { p->err = 1; break; } }
            else { hi = dec(p->s + p->i, p->n - p->i, &w); if (hi == 0xFFFD) { p->err = 1; break; } p->i += w; }
            if (hi < lo) { p->err = 1; break; }
        }
        set_add_fold(p, &in.set, lo, hi);
    }
    return f_one(in);
}

static frag_t p_atom(ps_t *p)
{
    int c = p->s[p->i];
    if (c == '(') {
        p->i++;
        if (p->i < p->n && p->s[p->i] == '?') {
            if (p->i + 1 < p->n && p->s[p->i + 1] == ':') p->i += 2; else { p->err = 1; return f_empty(); }
        }
        frag_t a = p_alt(p);
        if (p->i >= p->n || p->s[p->i] != ')') { p->err = 1; return a; }
        p->i++;
        return a;
    }
    if (c == '[') { p->i++; return p_class(p); }
    inst_t in = mk(I_SET, 0, 0);
    if (c == '.') {
        p->i++;
        if (p->dotall) set_add(&in.set, 0, 0x10FFFF); else { set_add(&in.set, 0, 9); set_add(&in.set, 11, 0x10FFFF); }
        return f_one(in);
    }
    if (c == '\\') { p->i++; p_escape(p, &in.set, 0, NULL); return f_one(in); }
    if (c == '^' || c == '$' || c == '*' || c == '+' || c == '?' || c == ')') { p->err = 1; return f_one(in); }
    uint32_t w, r = dec(p->s + p->i, p->n - p->i, &w);
    if (r == 0xFFFD) { p->err = 1; return f_one(in); }
    p->i += w;
    set_add_fold(p, &in.set, r, r);
    return f_one(in);
}

static int p_counted(ps_t *p, int *lo, int *hi, int *unb)
{
    size_t j = p->i + 1; long a = 0, b = 0; int ha = 0, hb = 0, comma = 0;
    while (j < p->n && isdigit(p->s[j]) && a < 100000) { a = a * 10 + (p->s[j] - '0'); ha = 1; j++; }
    if (j < p->n && p->s[j] == ',') { comma = 1; j++; while (j < p->n && isdigit(p->s[j]) && b < 100000) { b = b * 10 + (p->s[j] - '0'); hb = 1; j++; } }
    if (!ha || j >= p->n || p->s[j] != '}') return 0;
    p->i = j + 1;
    *lo = (int)a; *hi = comma ? (hb ? (int)b : (int)a) : (int)a; *unb = comma && !hb;
    return 1;
}

static frag_t p_repeat(ps_t *p)
{
    frag_t a = p_atom(p);
    while (!p->err && p->i < p->n) {
        int c = p->s[p->i], lo = 0, hi = 0, unb = 0;
        if (c == '*') { unb = 1; p->i++; }
        else if (c == '+') { lo = 1; unb = 1; p->i++; }
        else if (c == '?') { hi = 1; p->i++; }
        else if (c == '{') { if (!p_counted(p, &lo, &hi, &unb)) break; }
        else break;
        if ((!unb && hi < lo) || lo > 100 || hi > 100) { p->err = 1; break; }
        int lazy = 0;
        if (p->i < p->n && p->s[p->i] == '?') { lazy = 1; p->i++; }
        if (p->i < p->n && (p->s[p->i] == '*' || p->s[p->i] == '+' || p->s[p->i] == '?')) { p->err = 1; break; }
        if ((unb || hi > 1) && nullable_code(a.v, a.n)) { p->err = 1; break; }    /* empty-width loops: outside the subset */
        frag_t out = f_empty();
        for (int k = 0; k < lo; k++) out = f_cat(out, f_copy(a));
        if (unb) out = f_cat(out, f_star(f_copy(a), lazy));
        else {                                        /* x{lo,hi}: (x(x(...)?)?)? nested from the inside */
            frag_t opt = f_empty();
            for (int k = lo; k < hi; k++) opt = f_quest(f_cat(f_copy(a), opt), lazy);
            out = f_cat(out, opt);
        }
        f_free(a);
        a = out;
        if (a.n > 4096) { p->err = 1; break; }
    }
    return a;
}

static frag_t p_concat(ps_t *p)
{
    frag_t f = f_empty();
    while (!p->err && p->i < p->n && p->s[p->i] != '|' && p->s[p->i] != ')') f = f_cat(f, p_repeat(p));
    return f;
}
static frag_t p_alt(ps_t *p)
{
    frag_t f = p_concat(p);
    if (!p->err && p->i < p->n && p->s[p->i] == '|') { p->i++; frag_t rest = p_alt(p); f = f_alt(f, rest); }   /* a|b|c = a|(b|c): same preference order */
    return f;
}

/* can the code reach its end without consuming a rune? (small fixed-point over the relative-jump code) */
static int nullable_code(const inst_t *c, int n)
{
    if (n == 0) return 1;
    char *reach = (char *)calloc((size_t)n + 1, 1);
    int *stack = (int *)malloc(sizeof(int) * (size_t)(2 * n + 4)), sp = 0, res = 0;
    stack[sp++] = 0;
    while (sp) {
        int pc = stack[--sp];
        if (pc >= n) { res = 1; break; }
        if (reach[pc]) continue;
        reach[pc] = 1;
        if (c[pc].op == I_JMP) stack[sp++] = pc + c[pc].a;
        else if (c[pc].op == I_SPLIT) { stack[sp++] = pc + c[pc].a; stack[sp++] = pc + c[pc].b; }
    }
    free(reach); free(stack);
    return res;
}

orc_regex *orc_regex_compile(const uint8_t *pattern, uint32_t len)
{
    ps_t p; memset(&p, 0, sizeof(p));
    p.s = pattern; p.n = len;
    if (len >= 2 && pattern[0] == '(' && pattern[1] == '?') {      /* leading flag group only */
        size_t j = 2; int any = 0, ok = 1, fold = 0, dotall = 0;
        while (j < len && pattern[j] != ')' && pattern[j] != ':') { if (pattern[j] == 'i') fold = 1; else if (pattern[j] == 's') dotall = 1; else ok = 0; any = 1; j++; }
        if (j < len && pattern[j] == ')' && any) { if (!ok) return NULL; p.fold = fold; p.dotall = dotall; p.i = j + 1; }
    }
    frag_t f = p_alt(&p);
    if (p.err || p.i < p.n || nullable_code(f.v, f.n)) { f_free(f); return NULL; }
    f = f_cat(f, f_one(mk(I_MATCH, 0, 0)));
    orc_regex *re = (orc_regex *)malloc(sizeof(*re));
    re->code = f.v; re->n = f.n;
    return re;
}

void orc_regex_free(orc_regex *re)
{
    if (!re) return;
    frag_t f = {re->code, re->n};
    f_free(f);
    free(re);
}

typedef struct { int pc; uint64_t pos; } bt_t;

/* leftmost-first match anchored at `start`: depth-first, preferred branch first; returns the end or -1 */
static int64_t match_at(const orc_regex *re, const uint8_t *s, uint64_t n, uint64_t start, bt_t **stk, size_t *cap)
{
    size_t sp = 0;
    (*stk)[sp].pc = 0; (*stk)[sp].pos = start; sp++;
    while (sp) {
        sp--;
        int pc = (*stk)[sp].pc; uint64_t pos = (*stk)[sp].pos;
        for (;;) {
            const inst_t *in = &re->code[pc];
            if (in->op == I_MATCH) return (int64_t)pos;
            if (in->op == I_JMP) { pc += in->a; continue; }
            if (in->op == I_SPLIT) {
                if (sp + 1 >= *cap) { *cap *= 2; *stk = (bt_t *)realloc(*stk, sizeof(bt_t) * *cap); }
                (*stk)[sp].pc = pc + in->b; (*stk)[sp].pos = pos; sp++;       /* the other branch, tried after this one fails */
                pc += in->a;
                continue;
            }
            if (pos >= n) break;
            uint32_t w, r = dec(s + pos, n - pos, &w);
            if (!set_has(&in->set, r)) break;
            pos += w; pc++;
        }
    }
    return -1;
}

uint32_t orc_regex_count(const orc_regex *re, const uint8_t *s, uint64_t n)
{
    uint32_t count = 0;
    size_t cap = 64;
    bt_t *stk = (bt_t *)malloc(sizeof(bt_t) * cap);
    for (uint64_t i = 0; i < n;) {
        int64_t e = match_at(re, s, n, i, &stk, &cap);
        if (e >= 0) { count++; i = (uint64_t)e; }      /* never empty: nullable patterns are refused */
        else { uint32_t w; (void)dec(s + i, n - i, &w); i += w; }
    }
    free(stk);
    return count;
}
