/*
 * sse_inspect_oracle.c -- brute-force restatement of the content-inspection matcher.
 *
 * TEST INFRASTRUCTURE ONLY (see sse_oracle.h).
 *
 * PARITY STATUS.  The reference never runs inspection (src/sse-adapter/sse_handler.go:193-208 is
 * a comment block); its pattern sets are snippets:
 *   Set A  docs/security-inspection-patterns.md:205-209 (regexes, Redact) and :229-234 (literals, Alert)
 *   Set B  src/spin-functions/nats-subscriber/src/lib.rs:59-64, :79-84, PINNED by the three Rust
 *          known-answer tests at lib.rs:108-124 (tests/golden/hand_vectors.py SET_B_KNOWN).
 * Everything else (offsets, streaming across tokens, Set A) is PARITY UNPINNED: the semantics are
 * this project's definition (DESIGN.md section "Inspection").
 *
 * Deliberately a different algorithm from the product (which compiles the patterns to DFAs): every
 * pattern is tried at every start position with a backtracking matcher over the raw item list.
 *
 * Semantics restated here:
 *   - text = decoded token bytes; \d [0-9], \w [0-9A-Za-z_], \b between a word and a non-word byte,
 *     bytes >= 0x80 and the text edges are non-word; literals fold ASCII case only;
 *   - a match (b,e) becomes DECIDABLE at byte index d = e if the pattern ends in \b (the byte after
 *     the match is needed), else d = e-1; end of text counts as one more, non-word, position;
 *   - per_token: the lowest-numbered pattern that matches anywhere wins; end = its smallest e;
 *     begin = smallest b among its matches ending there;
 *   - streaming (text = all tokens of the conversation so far, no separator; the record owns
 *     [s,t), plus the end-of-text position if it is the done token): among the matches decidable
 *     inside the record the smallest (d, e, id) wins; offsets are relative to s; begin is only
 *     reported for fixed-length patterns (a variable-length match may have started in an earlier
 *     token the matcher no longer holds): verdict bit 16 + begin = INT32_MIN otherwise.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "sse_oracle.h"

#define MAX_ITEMS 256
#define UNBOUNDED 0xFFFFFFFFu

typedef struct {
    int is_assert;
    uint8_t set[32];
    uint32_t min, max;
} item_t;

typedef struct {
    uint32_t kind, action, n_items, fixed_len; /* fixed_len 0 = variable */
    int ends_in_assert;
    item_t items[MAX_ITEMS];
} pat_t;

struct sseo_inspector {
    uint32_t n;
    pat_t *p;
};

static void set_add(uint8_t *s, unsigned b) { s[b >> 3] |= (uint8_t)(1u << (b & 7)); }
static int set_has(const uint8_t *s, unsigned b) { return (s[b >> 3] >> (b & 7)) & 1; }
static int wordb(int c) { return c >= 0 && ((c >= '0' && c <= '9') || (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z') || c == '_'); }

static int esc_class(char c, uint8_t *s)
{
    unsigned b;
    switch (c) {
    case 'd': for (b = '0'; b <= '9'; b++) set_add(s, b); return 1;
    case 'w': for (b = 0; b < 128; b++) if (wordb((int)b)) set_add(s, b); return 1;
    case 's': set_add(s, ' '); set_add(s, '\t'); set_add(s, '\n'); set_add(s, '\r'); set_add(s, '\f'); set_add(s, '\v'); return 1;
    case 'n': set_add(s, '\n'); return 1;
    case 't': set_add(s, '\t'); return 1;
    case 'r': set_add(s, '\r'); return 1;
    default:
        if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9')) return 0;
        set_add(s, (unsigned char)c);
        return 1;
    }
}

static int parse_num(const char *re, size_t *j, uint32_t *v)
{
    if (re[*j] < '0' || re[*j] > '9') return 0;
    *v = 0;
    while (re[*j] >= '0' && re[*j] <= '9') { *v = *v * 10 + (uint32_t)(re[*j] - '0'); if (*v > 1000) return 0; (*j)++; }
    return 1;
}

static int parse_regex(const char *re, pat_t *p)
{
    size_t i = 0, n = strlen(re);
    p->n_items = 0;
    while (i < n) {
        if (p->n_items >= MAX_ITEMS) return 0;
        item_t *it = &p->items[p->n_items];
        memset(it, 0, sizeof *it);
        it->min = it->max = 1;
        char c = re[i];
        if (c == '\\') {
            if (i + 1 >= n) return 0;
            char e = re[i + 1];
            i += 2;
            if (e == 'b') { it->is_assert = 1; p->n_items++; continue; }
            if (!esc_class(e, it->set)) return 0;
        } else if (c == '[') {
            int neg = 0, first = 1;
            i++;
            if (i < n && re[i] == '^') { neg = 1; i++; }
            while (i < n && (re[i] != ']' || first)) {
                uint8_t one[32];
                unsigned lo;
                int single = 1;
                first = 0;
                memset(one, 0, sizeof one);
                if (re[i] == '\\') {
                    if (i + 1 >= n || !esc_class(re[i + 1], one)) return 0;
                    lo = (unsigned char)re[i + 1];
                    single = !(re[i + 1] == 'd' || re[i + 1] == 'w' || re[i + 1] == 's');
                    i += 2;
                } else { lo = (unsigned char)re[i]; set_add(one, lo); i++; }
                if (single && i + 1 < n && re[i] == '-' && re[i + 1] != ']') {
                    unsigned hi = (unsigned char)re[i + 1], b;
                    i += 2;
                    if (hi < lo) return 0;
                    for (b = lo; b <= hi; b++) set_add(it->set, b);
                } else {
                    for (unsigned b = 0; b < 256; b++) if (set_has(one, b)) set_add(it->set, b);
                }
            }
            if (i >= n) return 0;
            i++;
            if (neg) for (int k = 0; k < 32; k++) it->set[k] = (uint8_t)~it->set[k];
        } else if (c == '.') {
            for (unsigned b = 0; b < 256; b++) if (b != '\n') set_add(it->set, b);
            i++;
        } else if (c == '(' || c == ')' || c == '|' || c == '^' || c == '$' || c == '+' || c == '*' || c == '?' || c == '{') {
            return 0;
        } else { set_add(it->set, (unsigned char)c); i++; }
        if (i < n) {
            char q = re[i];
            if (q == '+') { it->min = 1; it->max = UNBOUNDED; i++; }
            else if (q == '*') { it->min = 0; it->max = UNBOUNDED; i++; }
            else if (q == '?') { it->min = 0; it->max = 1; i++; }
            else if (q == '{') {
                size_t j = i + 1;
                uint32_t a, b;
                if (!parse_num(re, &j, &a)) return 0;
                b = a;
                if (re[j] == ',') { j++; if (re[j] == '}') b = UNBOUNDED; else if (!parse_num(re, &j, &b)) return 0; }
                if (re[j] != '}' || (b != UNBOUNDED && b < a)) return 0;
                it->min = a; it->max = b; i = j + 1;
            }
        }
        p->n_items++;
    }
    return 1;
}

sseo_inspector *sseo_inspect_new(const sseo_pattern *pats, uint32_t n)
{
    if (n == 0 || n > 32) return NULL;
    sseo_inspector *I = (sseo_inspector *)calloc(1, sizeof *I);
    I->n = n;
    I->p = (pat_t *)calloc(n, sizeof(pat_t));
    for (uint32_t k = 0; k < n; k++) {
        pat_t *p = &I->p[k];
        p->kind = pats[k].kind; p->action = pats[k].action;
        if (p->kind == 0) {
            size_t L = strlen(pats[k].text);
            if (L == 0 || L > MAX_ITEMS) goto bad;
            for (size_t i = 0; i < L; i++) {
                unsigned char c = (unsigned char)pats[k].text[i];
                item_t *it = &p->items[i];
                it->min = it->max = 1;
                set_add(it->set, c);
                if (c >= 'a' && c <= 'z') set_add(it->set, c - 32u);
                if (c >= 'A' && c <= 'Z') set_add(it->set, c + 32u);
            }
            p->n_items = (uint32_t)L;
        } else if (p->kind == 1) {
            if (!parse_regex(pats[k].text, p)) goto bad;
        } else goto bad;
        uint64_t mn = 0, mx = 0;
        int unb = 0;
        for (uint32_t i = 0; i < p->n_items; i++) if (!p->items[i].is_assert) {
            mn += p->items[i].min;
            if (p->items[i].max == UNBOUNDED) unb = 1; else mx += p->items[i].max;
        }
        if (mn == 0) goto bad;
        p->fixed_len = (!unb && mn == mx) ? (uint32_t)mn : 0;
        p->ends_in_assert = p->n_items && p->items[p->n_items - 1].is_assert;
    }
    return I;
bad:
    free(I->p); free(I);
    return NULL;
}

void sseo_inspect_free(sseo_inspector *I) { if (I) { free(I->p); free(I); } }

/* Backtracking: all end positions of matches of items[i..] starting at pos. eot: text really ends at len
 * (position len is then a known non-word position); otherwise a \b at pos==len cannot be decided yet. */
static void match_from(const pat_t *p, uint32_t i, const uint8_t *T, uint32_t len, int eot, uint32_t pos, uint8_t *ends)
{
    if (i == p->n_items) { ends[pos] = 1; return; }
    const item_t *it = &p->items[i];
    if (it->is_assert) {
        if (pos == len && !eot) return;
        int a = pos > 0 ? wordb(T[pos - 1]) : 0;
        int b = pos < len ? wordb(T[pos]) : 0;
        if (a != b) match_from(p, i + 1, T, len, eot, pos, ends);
        return;
    }
    uint32_t k = 0, q = pos;
    while (k < it->min) { if (q >= len || !set_has(it->set, T[q])) return; q++; k++; }
    for (;;) {
        match_from(p, i + 1, T, len, eot, q, ends);
        if (it->max != UNBOUNDED && k >= it->max) break;
        if (q >= len || !set_has(it->set, T[q])) break;
        q++; k++;
    }
}

#define NO_MATCH 0xFFFFFFFFu

/* per_token mode */
void sseo_inspect_token(const sseo_inspector *I, const uint8_t *T, uint32_t len, uint32_t *verdict, int32_t *begin, int32_t *end)
{
    uint8_t *ends = (uint8_t *)malloc(len + 2);
    *verdict = 0; *begin = 0; *end = 0;
    for (uint32_t k = 0; k < I->n; k++) {
        uint32_t best_e = NO_MATCH, best_b = 0;
        for (uint32_t b = 0; b < len; b++) {
            memset(ends, 0, len + 2);
            match_from(&I->p[k], 0, T, len, 1, b, ends);
            for (uint32_t e = b + 1; e <= len; e++)
                if (ends[e] && (e < best_e)) { best_e = e; best_b = b; }   /* b ascending: first hit at e is the leftmost */
        }
        if (best_e != NO_MATCH) {
            *verdict = I->p[k].action | (k << 8);
            *begin = (int32_t)best_b; *end = (int32_t)best_e;
            break;
        }
    }
    free(ends);
}

/* streaming mode: T[0,t) is everything the conversation has produced, the record owns [s,t). */
void sseo_inspect_stream(const sseo_inspector *I, const uint8_t *T, uint32_t s, uint32_t t, int done,
                         uint32_t *verdict, int32_t *begin, int32_t *end)
{
    uint8_t *ends = (uint8_t *)malloc(t + 2);
    uint32_t best_d = NO_MATCH, best_e = 0, best_k = 0;
    *verdict = 0; *begin = 0; *end = 0;
    for (uint32_t k = 0; k < I->n; k++) {
        const pat_t *p = &I->p[k];
        for (uint32_t b = 0; b < t; b++) {
            memset(ends, 0, t + 2);
            match_from(p, 0, T, t, done, b, ends);
            for (uint32_t e = b + 1; e <= t; e++) {
                if (!ends[e]) continue;
                uint32_t d = p->ends_in_assert ? e : e - 1;
                if (d < s || d > t || (d == t && !done)) continue;
                if (d < best_d || (d == best_d && (e < best_e || (e == best_e && k < best_k)))) { best_d = d; best_e = e; best_k = k; }
            }
        }
    }
    if (best_d != NO_MATCH) {
        const pat_t *p = &I->p[best_k];
        *verdict = p->action | (best_k << 8);
        *end = (int32_t)best_e - (int32_t)s;
        if (p->fixed_len) *begin = (int32_t)best_e - (int32_t)p->fixed_len - (int32_t)s;
        else { *verdict |= 1u << 16; *begin = INT32_MIN; }
    }
    free(ends);
}
