/*
 * sse_oracle.c -- CPU restatement of the reference's per-token hot path, in plain C.
 *
 * TEST INFRASTRUCTURE ONLY (see sse_oracle.h).  PARITY UNPINNED for decode / encode /
 * frame: the algorithm is Go 1.21 encoding/json + fmt (third-party to the reference,
 * not vendored, no Go toolchain here); each function cites the reference call site it
 * stands in for and the stdlib routine whose published behaviour it restates.
 *
 * Structure deliberately mirrors the Go library (validate whole input, then walk and
 * materialise decoded strings, then append-encode) and NOT the CUDA kernels, which
 * stream runes without materialising anything: two differently shaped implementations
 * must agree byte for byte.
 */
#define _POSIX_C_SOURCE 200809L
#include "sse_oracle.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define MAX_NESTING 10000 /* encoding/json scanner.go: maxNestingDepth */

/* ------------------------------------------------------------------ utf8 ---- */
/* unicode/utf8.DecodeRune: returns the encoded size; an invalid or truncated
 * sequence yields U+FFFD with size 1. */
static int utf8_decode(const uint8_t *p, size_t n, uint32_t *r)
{
    uint8_t b0 = p[0];
    if (b0 < 0x80) { *r = b0; return 1; }
    if (b0 < 0xC2 || b0 > 0xF4) goto bad;
    if (b0 < 0xE0) {
        if (n < 2 || (p[1] & 0xC0) != 0x80) goto bad;
        *r = ((uint32_t)(b0 & 0x1F) << 6) | (p[1] & 0x3F);
        return 2;
    }
    if (b0 < 0xF0) {
        uint8_t lo = 0x80, hi = 0xBF;
        if (b0 == 0xE0) lo = 0xA0;
        if (b0 == 0xED) hi = 0x9F;
        if (n < 3 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80) goto bad;
        *r = ((uint32_t)(b0 & 0x0F) << 12) | ((uint32_t)(p[1] & 0x3F) << 6) | (p[2] & 0x3F);
        return 3;
    }
    {
        uint8_t lo = 0x80, hi = 0xBF;
        if (b0 == 0xF0) lo = 0x90;
        if (b0 == 0xF4) hi = 0x8F;
        if (n < 4 || p[1] < lo || p[1] > hi || (p[2] & 0xC0) != 0x80 || (p[3] & 0xC0) != 0x80)
            goto bad;
        *r = ((uint32_t)(b0 & 0x07) << 18) | ((uint32_t)(p[1] & 0x3F) << 12) |
             ((uint32_t)(p[2] & 0x3F) << 6) | (p[3] & 0x3F);
        return 4;
    }
bad:
    *r = 0xFFFD;
    return 1;
}

/* unicode/utf8.EncodeRune for a valid scalar value. */
static int utf8_encode(uint8_t *d, uint32_t r)
{
    if (r < 0x80) { d[0] = (uint8_t)r; return 1; }
    if (r < 0x800) { d[0] = 0xC0 | (r >> 6); d[1] = 0x80 | (r & 0x3F); return 2; }
    if (r < 0x10000) {
        d[0] = 0xE0 | (r >> 12); d[1] = 0x80 | ((r >> 6) & 0x3F); d[2] = 0x80 | (r & 0x3F);
        return 3;
    }
    d[0] = 0xF0 | (r >> 18); d[1] = 0x80 | ((r >> 12) & 0x3F);
    d[2] = 0x80 | ((r >> 6) & 0x3F); d[3] = 0x80 | (r & 0x3F);
    return 4;
}

/* ------------------------------------------------------------- validation ---- */
static int is_ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }
static int is_digit(uint8_t c) { return c >= '0' && c <= '9'; }
static int is_hex(uint8_t c)
{
    return is_digit(c) || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F');
}

/* scanner.go stateInString/stateInStringEsc*: p[i] is the opening quote; returns the
 * index one past the closing quote, or 0 on error. */
static size_t scan_string(const uint8_t *p, size_t n, size_t i)
{
    for (i++; i < n; i++) {
        uint8_t c = p[i];
        if (c == '"') return i + 1;
        if (c < 0x20) return 0;
        if (c == '\\') {
            if (++i >= n) return 0;
            switch (p[i]) {
            case 'b': case 'f': case 'n': case 'r': case 't': case '\\': case '/': case '"':
                break;
            case 'u':
                if (i + 4 >= n) return 0;
                if (!is_hex(p[i + 1]) || !is_hex(p[i + 2]) || !is_hex(p[i + 3]) || !is_hex(p[i + 4]))
                    return 0;
                i += 4;
                break;
            default:
                return 0;
            }
        }
    }
    return 0;
}

/* scanner.go stateNeg/state0/state1/stateDot/stateE...: returns index past the
 * number, or 0 on error. */
static size_t scan_number(const uint8_t *p, size_t n, size_t i)
{
    if (p[i] == '-') { if (++i >= n) return 0; }
    if (p[i] == '0') i++;
    else if (p[i] >= '1' && p[i] <= '9') { while (i < n && is_digit(p[i])) i++; }
    else return 0;
    if (i < n && p[i] == '.') {
        i++;
        if (i >= n || !is_digit(p[i])) return 0;
        while (i < n && is_digit(p[i])) i++;
    }
    if (i < n && (p[i] == 'e' || p[i] == 'E')) {
        i++;
        if (i < n && (p[i] == '+' || p[i] == '-')) i++;
        if (i >= n || !is_digit(p[i])) return 0;
        while (i < n && is_digit(p[i])) i++;
    }
    return i;
}

static size_t scan_literal(const uint8_t *p, size_t n, size_t i, const char *lit)
{
    size_t L = strlen(lit);
    if (i + L > n || memcmp(p + i, lit, L) != 0) return 0;
    return i + L;
}

/* encoding/json checkValid (decode.go) over the scanner state machine (scanner.go). */
static int json_valid(const uint8_t *p, size_t n)
{
    enum { BEGIN_VALUE, BEGIN_VALUE_OR_EMPTY, BEGIN_KEY_OR_EMPTY, BEGIN_KEY, END_VALUE } st = BEGIN_VALUE;
    static __thread uint8_t stack[MAX_NESTING / 8 + 2]; /* 1 = object, 0 = array */
    size_t depth = 0, i = 0;
    for (;;) {
        while (i < n && is_ws(p[i])) i++;
        if (i >= n) return st == END_VALUE && depth == 0; /* scanner.eof() */
        uint8_t c = p[i];
        switch (st) {
        case BEGIN_VALUE_OR_EMPTY:
            if (c == ']') { depth--; i++; st = END_VALUE; break; }
            /* fallthrough */
        case BEGIN_VALUE:
            if (c == '{' || c == '[') {
                if (depth + 1 > MAX_NESTING) return 0; /* "exceeded max depth" */
                if (c == '{') stack[depth >> 3] |= (uint8_t)(1u << (depth & 7));
                else          stack[depth >> 3] &= (uint8_t)~(1u << (depth & 7));
                depth++; i++;
                st = (c == '{') ? BEGIN_KEY_OR_EMPTY : BEGIN_VALUE_OR_EMPTY;
                break;
            }
            if (c == '"') i = scan_string(p, n, i);
            else if (c == '-' || is_digit(c)) i = scan_number(p, n, i);
            else if (c == 't') i = scan_literal(p, n, i, "true");
            else if (c == 'f') i = scan_literal(p, n, i, "false");
            else if (c == 'n') i = scan_literal(p, n, i, "null");
            else return 0;
            if (!i) return 0;
            st = END_VALUE;
            break;
        case BEGIN_KEY_OR_EMPTY:
            if (c == '}') { depth--; i++; st = END_VALUE; break; }
            /* fallthrough */
        case BEGIN_KEY:
            if (c != '"') return 0;
            i = scan_string(p, n, i);
            if (!i) return 0;
            while (i < n && is_ws(p[i])) i++;
            if (i >= n || p[i] != ':') return 0;
            i++;
            st = BEGIN_VALUE;
            break;
        case END_VALUE:
            if (depth == 0) return 0; /* stateEndTop: only white space may follow */
            if (stack[(depth - 1) >> 3] & (1u << ((depth - 1) & 7))) {
                if (c == ',') st = BEGIN_KEY;
                else if (c == '}') depth--;
                else return 0;
            } else {
                if (c == ',') st = BEGIN_VALUE;
                else if (c == ']') depth--;
                else return 0;
            }
            i++;
            break;
        }
    }
}

/* -------------------------------------------------------------- decoding ---- */
static int hex4(const uint8_t *s)
{
    int v = 0;
    for (int k = 0; k < 4; k++) {
        uint8_t c = s[k];
        int d;
        if (c >= '0' && c <= '9') d = c - '0';
        else if (c >= 'a' && c <= 'f') d = c - 'a' + 10;
        else if (c >= 'A' && c <= 'F') d = c - 'A' + 10;
        else return -1;
        v = v * 16 + d;
    }
    return v;
}

/* decode.go getu4 */
static int getu4(const uint8_t *s, size_t n)
{
    if (n < 6 || s[0] != '\\' || s[1] != 'u') return -1;
    return hex4(s + 2);
}

/* decode.go unquoteBytes on the bytes BETWEEN the quotes of an already-valid JSON
 * string.  Returns decoded length (always valid UTF-8). */
static uint32_t unquote(const uint8_t *s, size_t n, uint8_t *dst)
{
    size_t r = 0;
    uint32_t w = 0;
    while (r < n) {
        uint8_t c = s[r];
        if (c == '\\') {
            r++;
            switch (s[r]) {
            case '"': case '\\': case '/': dst[w++] = s[r]; r++; break;
            case 'b': dst[w++] = '\b'; r++; break;
            case 'f': dst[w++] = '\f'; r++; break;
            case 'n': dst[w++] = '\n'; r++; break;
            case 'r': dst[w++] = '\r'; r++; break;
            case 't': dst[w++] = '\t'; r++; break;
            case 'u': {
                r--;
                int rr = getu4(s + r, n - r);
                r += 6;
                if (rr >= 0xD800 && rr < 0xE000) { /* utf16.IsSurrogate */
                    int rr1 = getu4(s + r, n - r);
                    if (rr < 0xDC00 && rr1 >= 0xDC00 && rr1 < 0xE000) { /* utf16.DecodeRune */
                        uint32_t dec = (((uint32_t)rr - 0xD800) << 10 | ((uint32_t)rr1 - 0xDC00)) + 0x10000;
                        r += 6;
                        w += utf8_encode(dst + w, dec);
                        break;
                    }
                    rr = 0xFFFD;
                }
                w += utf8_encode(dst + w, (uint32_t)rr);
                break;
            }
            default: r++; break; /* unreachable after validation */
            }
        } else if (c < 0x80) {
            dst[w++] = c;
            r++;
        } else {
            uint32_t rune;
            r += utf8_decode(s + r, n - r, &rune);
            w += utf8_encode(dst + w, rune);
        }
    }
    return w;
}

/* fold.go foldRune: smallest rune of the simple-fold orbit.  Only the orbits that
 * contain an ASCII letter matter for the five ASCII field names: {K,k,U+212A} and
 * {S,s,U+017F}; every other non-ASCII rune folds to something non-ASCII. */
static uint32_t fold_rune(uint32_t r)
{
    if (r >= 'a' && r <= 'z') return r - 32;
    if (r == 0x212A) return 'K';
    if (r == 0x017F) return 'S';
    return r;
}

static const char *const FIELD_UPPER[5] = { "CONVERSATION_ID", "TOKEN", "SEQUENCE", "DONE", "TIMESTAMP" };

/* decode.go object(): exact name else foldName(key) lookup (== bytes.EqualFold).
 * key is the unquoted key.  Returns field index or -1. */
static int match_field(const uint8_t *key, uint32_t klen)
{
    for (int f = 0; f < 5; f++) {
        const char *name = FIELD_UPPER[f];
        size_t L = strlen(name), k = 0, i = 0;
        int ok = 1;
        while (i < klen) {
            uint32_t r;
            i += utf8_decode(key + i, klen - i, &r);
            if (k >= L || fold_rune(r) != (uint32_t)(uint8_t)name[k]) { ok = 0; break; }
            k++;
        }
        if (ok && k == L) return f;
    }
    return -1;
}

/* Skip one value of an already-valid document; returns index past it. */
static size_t skip_value(const uint8_t *p, size_t i)
{
    uint8_t c = p[i];
    if (c == '"') {
        for (i++; p[i] != '"'; i++) if (p[i] == '\\') i++;
        return i + 1;
    }
    if (c == '{' || c == '[') {
        size_t depth = 0;
        for (;; i++) {
            c = p[i];
            if (c == '"') { for (i++; p[i] != '"'; i++) if (p[i] == '\\') i++; }
            else if (c == '{' || c == '[') depth++;
            else if (c == '}' || c == ']') { if (--depth == 0) return i + 1; }
        }
    }
    if (c == 't' || c == 'n') return i + 4;
    if (c == 'f') return i + 5;
    while (p[i] == '-' || p[i] == '+' || p[i] == '.' || p[i] == 'e' || p[i] == 'E' || is_digit(p[i])) i++;
    return i;
}

/* strconv.ParseInt(literal, 10, 64) on a JSON number literal (decode.go literalStore,
 * reflect.Int64 case): anything but -?digits, or out of int64 range, fails. */
static int parse_int64(const uint8_t *s, size_t n, int64_t *out)
{
    size_t i = 0;
    int neg = 0;
    if (s[0] == '-') { neg = 1; i = 1; }
    if (i >= n) return 0;
    uint64_t v = 0;
    for (; i < n; i++) {
        if (!is_digit(s[i])) return 0;
        uint64_t d = (uint64_t)(s[i] - '0');
        if (v > (UINT64_MAX - d) / 10) return 0;
        v = v * 10 + d;
    }
    if (neg) { if (v > (uint64_t)1 << 63) return 0; *out = (int64_t)(0 - v); }
    else     { if (v > (uint64_t)INT64_MAX) return 0; *out = (int64_t)v; }
    return 1;
}

int sseo_unmarshal(const uint8_t *p, size_t n, sseo_msg *m, uint8_t *scratch)
{
    static const uint8_t empty[1] = { 0 };
    m->cid = empty; m->cid_len = 0; m->tok = empty; m->tok_len = 0;
    m->seq = 0; m->ts = 0; m->done = 0;
    if (!json_valid(p, n)) return (int)SSEO_ERR_SYNTAX;

    uint8_t *cid_buf = scratch, *tok_buf = scratch + 3 * n + 8;
    int type_err = 0;
    size_t i = 0;
    while (is_ws(p[i])) i++;
    if (p[i] == 'n') return (int)SSEO_OK;   /* literalStore: null into a struct is a no-op */
    if (p[i] != '{') return (int)SSEO_ERR_TYPE; /* array/string/number/bool into struct */
    i++;
    for (;;) {
        while (is_ws(p[i])) i++;
        if (p[i] == '}') break;
        size_t ks = i + 1;
        i = skip_value(p, i);
        size_t ke = i - 1;
        /* keys are unquoted before lookup (decode.go object()) */
        uint8_t *kbuf = (uint8_t *)malloc(3 * (ke - ks) + 8);
        uint32_t klen = unquote(p + ks, ke - ks, kbuf);
        int f = match_field(kbuf, klen);
        free(kbuf);
        while (is_ws(p[i])) i++;
        i++; /* ':' */
        while (is_ws(p[i])) i++;
        size_t vs = i;
        i = skip_value(p, i);
        size_t ve = i;
        uint8_t c = p[vs];
        if (f >= 0 && c != 'n') { /* null leaves the field unchanged */
            switch (f) {
            case 0:
                if (c == '"') { m->cid_len = unquote(p + vs + 1, ve - vs - 2, cid_buf); m->cid = cid_buf; }
                else type_err = 1;
                break;
            case 1:
                if (c == '"') { m->tok_len = unquote(p + vs + 1, ve - vs - 2, tok_buf); m->tok = tok_buf; }
                else type_err = 1;
                break;
            case 2:
                if (!((c == '-' || is_digit(c)) && parse_int64(p + vs, ve - vs, &m->seq))) type_err = 1;
                break;
            case 4:
                if (!((c == '-' || is_digit(c)) && parse_int64(p + vs, ve - vs, &m->ts))) type_err = 1;
                break;
            case 3:
                if (c == 't') m->done = 1;
                else if (c == 'f') m->done = 0;
                else type_err = 1;
                break;
            }
        }
        while (is_ws(p[i])) i++;
        if (p[i] == ',') { i++; continue; }
        break; /* '}' */
    }
    /* decode.go: the first UnmarshalTypeError is saved, decoding continues, the error
     * is returned at the end -> the caller drops the message (sse_handler.go:137-140). */
    return type_err ? (int)SSEO_ERR_TYPE : (int)SSEO_OK;
}

/* -------------------------------------------------------------- encoding ---- */
static const char HEXD[] = "0123456789abcdef";

/* encode.go appendString(dst, src, escapeHTML=true). */
static size_t append_string(uint8_t *d, const uint8_t *s, size_t n, int go122)
{
    size_t w = 0, i = 0;
    d[w++] = '"';
    while (i < n) {
        uint8_t b = s[i];
        if (b < 0x80) {
            if (b >= 0x20 && b != '"' && b != '\\' && b != '<' && b != '>' && b != '&') {
                d[w++] = b; i++; continue; /* htmlSafeSet */
            }
            d[w++] = '\\';
            switch (b) {
            case '\\': case '"': d[w++] = b; break;
            case '\n': d[w++] = 'n'; break;
            case '\r': d[w++] = 'r'; break;
            case '\t': d[w++] = 't'; break;
            default:
                if (go122 && b == '\b') { d[w++] = 'b'; break; }
                if (go122 && b == '\f') { d[w++] = 'f'; break; }
                d[w++] = 'u'; d[w++] = '0'; d[w++] = '0';
                d[w++] = (uint8_t)HEXD[b >> 4]; d[w++] = (uint8_t)HEXD[b & 0xF];
            }
            i++;
            continue;
        }
        uint32_t r;
        int size = utf8_decode(s + i, n - i, &r);
        if (r == 0xFFFD && size == 1) { memcpy(d + w, "\\ufffd", 6); w += 6; i++; continue; }
        if (r == 0x2028 || r == 0x2029) {
            memcpy(d + w, "\\u202", 5); w += 5; d[w++] = (uint8_t)HEXD[r & 0xF];
            i += (size_t)size;
            continue;
        }
        memcpy(d + w, s + i, (size_t)size);
        w += (size_t)size; i += (size_t)size;
    }
    d[w++] = '"';
    return w;
}

/* strconv.AppendInt(dst, v, 10) == fmt %d for int64 */
static size_t append_int(uint8_t *d, int64_t v)
{
    char tmp[24];
    int k = 0;
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    do { tmp[k++] = (char)('0' + u % 10); u /= 10; } while (u);
    size_t w = 0;
    if (v < 0) d[w++] = '-';
    while (k) d[w++] = (uint8_t)tmp[--k];
    return w;
}

static size_t put(uint8_t *d, const char *s) { size_t L = strlen(s); memcpy(d, s, L); return L; }

size_t sseo_marshal_bound(const sseo_msg *m) { return 66 + 6 * ((size_t)m->cid_len + m->tok_len) + 20 + 20 + 5; }

size_t sseo_marshal(const sseo_msg *m, uint8_t *d, int go122)
{
    size_t w = 0;
    w += put(d + w, "{\"conversation_id\":");
    w += append_string(d + w, m->cid, m->cid_len, go122);
    w += put(d + w, ",\"token\":");
    w += append_string(d + w, m->tok, m->tok_len, go122);
    w += put(d + w, ",\"sequence\":");
    w += append_int(d + w, m->seq);
    w += put(d + w, ",\"done\":");
    w += put(d + w, m->done ? "true" : "false");
    w += put(d + w, ",\"timestamp\":");
    w += append_int(d + w, m->ts);
    d[w++] = '}';
    return w;
}

/* Producer side: publishToken (llm-stream-proxy/main.go:230-245) and the load generator's producer
 * (demo/load-generator/main.go:215-223) for n messages given as columns.  Record i = [channel "chat.<cid>.tokens"
 * when with_channel (main.go:240)] + json.Marshal(TokenMessage{...}); records are packed back to back, off[i]..off[i+1]. */
size_t sseo_encode_batch(const uint8_t *cid, const uint32_t *cid_off, const uint8_t *tok, const uint32_t *tok_off,
                         const int64_t *seq, const uint8_t *done, const int64_t *ts, uint32_t n, int go122, int with_channel,
                         uint8_t *out, size_t out_cap, uint64_t *off, uint32_t *chan_len)
{
    size_t w = 0;
    for (uint32_t i = 0; i < n; i++) {
        sseo_msg m;
        m.cid = cid + cid_off[i]; m.cid_len = cid_off[i + 1] - cid_off[i];
        m.tok = tok + tok_off[i]; m.tok_len = tok_off[i + 1] - tok_off[i];
        m.seq = seq[i]; m.done = done[i] != 0; m.ts = ts[i];
        if (w + 12 + m.cid_len + sseo_marshal_bound(&m) > out_cap) return (size_t)-1;
        off[i] = w;
        uint32_t cl = 0;
        if (with_channel) {
            w += put(out + w, "chat.");
            memcpy(out + w, m.cid, m.cid_len); w += m.cid_len;
            w += put(out + w, ".tokens");
            cl = 12 + m.cid_len;
        }
        if (chan_len) chan_len[i] = cl;
        w += sseo_marshal(&m, out + w, go122);
    }
    off[n] = w;
    return w;
}

size_t sseo_frame(const sseo_msg *m, uint8_t *d, int go122)
{
    size_t w = 0;
    w += put(d + w, "event: token\n");      /* sse_handler.go:212 */
    w += put(d + w, "id: ");                /* :213 */
    w += append_int(d + w, m->seq);
    d[w++] = '\n';
    w += put(d + w, "data: ");              /* :214 */
    w += sseo_marshal(m, d + w, go122);
    d[w++] = '\n'; d[w++] = '\n';
    return w;
}

/* sse_handler.go:92-96: fmt.Sscanf(lastEventID, "%d", &startSequence); startSequence++.
 * fmt/scan.go scanInt for verb 'd': skip blanks (a newline is an error), optional
 * sign, then a run of [0-9_]; strconv.ParseInt(tok,10,64) rejects '_' and overflow;
 * on any error the target keeps its zero value. */
int64_t sseo_start_sequence(const char *id)
{
    if (!id || !*id) return 0;
    const unsigned char *s = (const unsigned char *)id;
    uint64_t v = 0;
    int ok = 1, neg = 0, nd = 0;
    while (*s == ' ' || *s == '\t' || *s == '\v' || *s == '\f' || (*s == '\r' && s[1] != '\n')) s++;
    if (*s == '+' || *s == '-') { neg = (*s == '-'); s++; }
    for (; is_digit(*s) || *s == '_'; s++, nd++) {
        if (*s == '_') { ok = 0; continue; }
        uint64_t d = (uint64_t)(*s - '0');
        if (v > (UINT64_MAX - d) / 10) { ok = 0; continue; }
        v = v * 10 + d;
    }
    if (!nd) ok = 0;
    if (ok && !neg && v > (uint64_t)INT64_MAX) ok = 0;
    if (ok && neg && v > (uint64_t)1 << 63) ok = 0;
    uint64_t start = ok ? (neg ? 0 - v : v) : 0;
    return (int64_t)(start + 1); /* Go int64 ++ wraps */
}

/* ----------------------------------------------------------------- batch ---- */
int sseo_adapter_batch(const uint8_t *in, const uint32_t *in_off, uint32_t n,
                       const int64_t *start_seq, int go122,
                       uint8_t *out, uint64_t out_cap, uint64_t *out_off, uint32_t *status)
{
    uint64_t w = 0;
    size_t scap = 0;
    uint8_t *scratch = NULL;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *p = in + in_off[i];
        size_t len = in_off[i + 1] - in_off[i];
        if (6 * len + 16 > scap) { scap = 2 * (6 * len + 16); free(scratch); scratch = (uint8_t *)malloc(scap); }
        sseo_msg m;
        out_off[i] = w;
        uint32_t st = (uint32_t)sseo_unmarshal(p, len, &m, scratch);
        /* sse_handler.go:143 */
        if (st == SSEO_OK && start_seq && start_seq[i] > 0 && m.seq <= start_seq[i]) st = SSEO_FILTERED;
        if (st == SSEO_OK) {
            if (w + sseo_marshal_bound(&m) + 48 > out_cap) { free(scratch); return -1; }
            w += sseo_frame(&m, out + w, go122);
            if (m.done) st |= SSEO_DONE; /* sse_handler.go:220 */
        }
        status[i] = st;
    }
    out_off[n] = w;
    free(scratch);
    return 0;
}

int sseo_bridge_batch(const uint8_t *in, const uint32_t *in_off, uint32_t n, int go122,
                      uint8_t *out, uint64_t out_cap, uint64_t *out_off,
                      uint32_t *subj_len, uint32_t *status)
{
    uint64_t w = 0;
    size_t scap = 0;
    uint8_t *scratch = NULL;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *p = in + in_off[i];
        size_t len = in_off[i + 1] - in_off[i];
        if (6 * len + 16 > scap) { scap = 2 * (6 * len + 16); free(scratch); scratch = (uint8_t *)malloc(scap); }
        sseo_msg m;
        out_off[i] = w;
        subj_len[i] = 0;
        uint32_t st = (uint32_t)sseo_unmarshal(p, len, &m, scratch);
        if (st == SSEO_OK) {
            if (w + sseo_marshal_bound(&m) + m.cid_len + 16 > out_cap) { free(scratch); return -1; }
            /* redis-nats-bridge/main.go:160: fmt.Sprintf("chat.%s.tokens", token.ConversationID) */
            uint64_t w0 = w;
            w += put(out + w, "chat.");
            memcpy(out + w, m.cid, m.cid_len); w += m.cid_len;
            w += put(out + w, ".tokens");
            subj_len[i] = (uint32_t)(w - w0);
            w += sseo_marshal(&m, out + w, go122); /* main.go:161 */
            if (m.done) st |= SSEO_DONE;
        }
        status[i] = st;
    }
    out_off[n] = w;
    free(scratch);
    return 0;
}

/* Inverse view for tests and the encoder bench: json.Unmarshal every payload (all must decode) into the columns the
 * producer-side encoder takes.  cid / tok receive the decoded strings back to back; returns 0, or -1 - i for the first
 * payload i that does not decode, or -(n+2) when a capacity is too small. */
int64_t sseo_decode_columns(const uint8_t *in, const uint32_t *in_off, uint32_t n,
                            uint8_t *cid, size_t cid_cap, uint32_t *cid_off, uint8_t *tok, size_t tok_cap, uint32_t *tok_off,
                            int64_t *seq, uint8_t *done, int64_t *ts)
{
    size_t scap = 0, cw = 0, tw = 0;
    uint8_t *scratch = NULL;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *p = in + in_off[i];
        size_t len = in_off[i + 1] - in_off[i];
        if (6 * len + 16 > scap) { scap = 2 * (6 * len + 16); free(scratch); scratch = (uint8_t *)malloc(scap); }
        sseo_msg m;
        if (sseo_unmarshal(p, len, &m, scratch) != SSEO_OK) { free(scratch); return -1 - (int64_t)i; }
        if (cw + m.cid_len > cid_cap || tw + m.tok_len > tok_cap) { free(scratch); return -((int64_t)n + 2); }
        cid_off[i] = (uint32_t)cw; memcpy(cid + cw, m.cid, m.cid_len); cw += m.cid_len;
        tok_off[i] = (uint32_t)tw; memcpy(tok + tw, m.tok, m.tok_len); tw += m.tok_len;
        seq[i] = m.seq; done[i] = m.done ? 1 : 0; ts[i] = m.ts;
    }
    cid_off[n] = (uint32_t)cw; tok_off[n] = (uint32_t)tw;
    free(scratch);
    return 0;
}

/* Bridge records as complete NATS client-protocol messages: what nc.Publish(subject, data) (redis-nats-bridge/main.go:164)
 * puts on the wire.  The reference delegates this to github.com/nats-io/nats.go v1.31.0 (src/redis-nats-bridge/go.mod),
 * which is not vendored; its publish() writes "PUB " subject " " decimal(len(data)) "\r\n" data "\r\n" (the NATS client
 * protocol's PUB <subject> [reply-to] <#bytes>; no reply subject, no headers on this path). */
int sseo_bridge_pub_batch(const uint8_t *in, const uint32_t *in_off, uint32_t n, int go122,
                          uint8_t *out, uint64_t out_cap, uint64_t *out_off,
                          uint32_t *subj_len, uint32_t *status)
{
    uint64_t w = 0;
    size_t scap = 0;
    uint8_t *scratch = NULL, *tmp = NULL;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *p = in + in_off[i];
        size_t len = in_off[i + 1] - in_off[i];
        if (6 * len + 16 > scap) {
            scap = 2 * (6 * len + 16);
            free(scratch); free(tmp);
            scratch = (uint8_t *)malloc(scap); tmp = (uint8_t *)malloc(scap + 128);
        }
        sseo_msg m;
        out_off[i] = w;
        subj_len[i] = 0;
        uint32_t st = (uint32_t)sseo_unmarshal(p, len, &m, scratch);
        if (st == SSEO_OK) {
            if (w + sseo_marshal_bound(&m) + m.cid_len + 48 > out_cap) { free(scratch); free(tmp); return -1; }
            size_t plen = sseo_marshal(&m, tmp, go122);
            w += put(out + w, "PUB ");
            uint64_t w0 = w;
            w += put(out + w, "chat.");
            memcpy(out + w, m.cid, m.cid_len); w += m.cid_len;
            w += put(out + w, ".tokens");
            subj_len[i] = (uint32_t)(w - w0);
            out[w++] = ' ';
            w += append_int(out + w, (int64_t)plen);
            out[w++] = '\r'; out[w++] = '\n';
            memcpy(out + w, tmp, plen); w += plen;
            out[w++] = '\r'; out[w++] = '\n';
            if (m.done) st |= SSEO_DONE;
        }
        status[i] = st;
    }
    out_off[n] = w;
    free(scratch); free(tmp);
    return 0;
}

/* Producer records as the RESP2 command go-redis writes for rdb.Publish(ctx, channel, data) (llm-stream-proxy/main.go:244).
 * github.com/redis/go-redis/v9 v9.3.0 (not vendored): Publish builds NewIntCmd(ctx, "publish", channel, message) and the
 * protocol writer emits an array of bulk strings: *3 CRLF $7 CRLF publish CRLF $<n> CRLF channel CRLF $<m> CRLF data CRLF. */
size_t sseo_encode_resp_batch(const uint8_t *cid, const uint32_t *cid_off, const uint8_t *tok, const uint32_t *tok_off,
                              const int64_t *seq, const uint8_t *done, const int64_t *ts, uint32_t n, int go122,
                              uint8_t *out, size_t out_cap, uint64_t *off, uint32_t *chan_len)
{
    size_t w = 0;
    for (uint32_t i = 0; i < n; i++) {
        sseo_msg m;
        m.cid = cid + cid_off[i]; m.cid_len = cid_off[i + 1] - cid_off[i];
        m.tok = tok + tok_off[i]; m.tok_len = tok_off[i + 1] - tok_off[i];
        m.seq = seq[i]; m.done = done[i] != 0; m.ts = ts[i];
        size_t bound = sseo_marshal_bound(&m);
        if (w + 96 + m.cid_len + bound > out_cap) return (size_t)-1;
        uint8_t *tmp = (uint8_t *)malloc(bound);
        size_t plen = sseo_marshal(&m, tmp, go122);
        off[i] = w;
        w += put(out + w, "*3\r\n$7\r\npublish\r\n$");
        w += append_int(out + w, (int64_t)(12 + m.cid_len));
        w += put(out + w, "\r\nchat.");
        memcpy(out + w, m.cid, m.cid_len); w += m.cid_len;
        w += put(out + w, ".tokens\r\n$");
        w += append_int(out + w, (int64_t)plen);
        w += put(out + w, "\r\n");
        memcpy(out + w, tmp, plen); w += plen;
        w += put(out + w, "\r\n");
        free(tmp);
        if (chan_len) chan_len[i] = 12 + m.cid_len;
    }
    off[n] = w;
    return w;
}

/* ------------------------------------------------------------- baseline ---- */
typedef struct {
    const uint8_t *in; const uint32_t *in_off; uint32_t lo, hi;
    uint8_t *out; uint64_t out_cap; uint64_t frames, bytes;
} bench_job;

static void *bench_worker(void *arg)
{
    bench_job *j = (bench_job *)arg;
    uint8_t scratch[6 * 4096 + 16];
    uint8_t *big = NULL;
    uint64_t w = 0, frames = 0, total = 0;
    for (uint32_t i = j->lo; i < j->hi; i++) {
        const uint8_t *p = j->in + j->in_off[i];
        size_t len = j->in_off[i + 1] - j->in_off[i];
        uint8_t *s = scratch;
        if (len > 4096) { free(big); big = (uint8_t *)malloc(6 * len + 16); s = big; }
        sseo_msg m;
        if (sseo_unmarshal(p, len, &m, s) == (int)SSEO_OK) {
            if (w + sseo_marshal_bound(&m) + 48 > j->out_cap) w = 0; /* ring: timing only */
            size_t fl = sseo_frame(&m, j->out + w, 0);
            w += fl; total += fl;
            frames++;
        }
    }
    free(big);
    j->frames = frames;
    j->bytes = total;
    return NULL;
}

double sseo_bench_adapter(const uint8_t *in, const uint32_t *in_off, uint32_t n,
                          int threads, int reps, uint64_t *frames, uint64_t *out_bytes)
{
    if (threads < 1) threads = 1;
    bench_job *jobs = (bench_job *)calloc((size_t)threads, sizeof(bench_job));
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        uint32_t lo = (uint32_t)((uint64_t)n * (uint64_t)t / (uint64_t)threads);
        uint32_t hi = (uint32_t)((uint64_t)n * (uint64_t)(t + 1) / (uint64_t)threads);
        uint64_t in_bytes = in_off[hi] - in_off[lo];
        jobs[t].in = in; jobs[t].in_off = in_off; jobs[t].lo = lo; jobs[t].hi = hi;
        jobs[t].out_cap = 2 * in_bytes + 64ull * (hi - lo) + (1u << 20);
        jobs[t].out = (uint8_t *)malloc(jobs[t].out_cap);
    }
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        struct timespec t0, t1;
        clock_gettime(CLOCK_MONOTONIC, &t0);
        for (int t = 0; t < threads; t++) pthread_create(&tid[t], NULL, bench_worker, &jobs[t]);
        for (int t = 0; t < threads; t++) pthread_join(tid[t], NULL);
        clock_gettime(CLOCK_MONOTONIC, &t1);
        double dt = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
        if (dt < best) best = dt;
    }
    uint64_t f = 0, b = 0;
    for (int t = 0; t < threads; t++) { f += jobs[t].frames; b += jobs[t].bytes; free(jobs[t].out); }
    if (frames) *frames = f;
    if (out_bytes) *out_bytes = b;
    free(jobs); free(tid);
    return best;
}

/* The same baseline with its buffers prepared once: sseo_bench_new allocates and touches every worker's output buffer,
 * sseo_bench_run times ONE pass of all workers (wall clock), sseo_bench_free releases.  bench.py's reference arm uses
 * it so that a timed step pays no page faults for freshly malloc'ed output (the GPU arm's buffers are allocated once too). */
struct sseo_bench {
    int threads;
    bench_job *jobs;
    pthread_t *tid;
};

sseo_bench *sseo_bench_new(const uint8_t *in, const uint32_t *in_off, uint32_t n, int threads)
{
    if (threads < 1) threads = 1;
    sseo_bench *b = (sseo_bench *)calloc(1, sizeof *b);
    b->threads = threads;
    b->jobs = (bench_job *)calloc((size_t)threads, sizeof(bench_job));
    b->tid = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        uint32_t lo = (uint32_t)((uint64_t)n * (uint64_t)t / (uint64_t)threads);
        uint32_t hi = (uint32_t)((uint64_t)n * (uint64_t)(t + 1) / (uint64_t)threads);
        uint64_t in_bytes = in_off[hi] - in_off[lo];
        b->jobs[t].in = in; b->jobs[t].in_off = in_off; b->jobs[t].lo = lo; b->jobs[t].hi = hi;
        b->jobs[t].out_cap = 2 * in_bytes + 64ull * (hi - lo) + (1u << 20);
        b->jobs[t].out = (uint8_t *)malloc(b->jobs[t].out_cap);
        if (!b->jobs[t].out) { sseo_bench_free(b); return NULL; }
        memset(b->jobs[t].out, 0, b->jobs[t].out_cap);
    }
    return b;
}

double sseo_bench_run(sseo_bench *b, uint64_t *frames, uint64_t *out_bytes)
{
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < b->threads; t++) pthread_create(&b->tid[t], NULL, bench_worker, &b->jobs[t]);
    for (int t = 0; t < b->threads; t++) pthread_join(b->tid[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    uint64_t f = 0, by = 0;
    for (int t = 0; t < b->threads; t++) { f += b->jobs[t].frames; by += b->jobs[t].bytes; }
    if (frames) *frames = f;
    if (out_bytes) *out_bytes = by;
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

void sseo_bench_free(sseo_bench *b)
{
    if (!b) return;
    for (int t = 0; t < b->threads; t++) free(b->jobs[t].out);
    free(b->jobs); free(b->tid); free(b);
}

/* hash/fnv New64a: the conversation -> GPU shard function. */
uint64_t sseo_fnv1a64(const uint8_t *p, size_t n)
{
    uint64_t h = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 0x100000001b3ull; }
    return h;
}

/* ---------------------------------------------------------------- row N3, decode half ----
 * The proxy's read of the vLLM stream: src/llm-stream-proxy/main.go:188-226 with StreamChunk of main.go:45-53.
 * PARITY UNPINNED (the reference has no test for it).  Written the way the Go code reads: TrimSpace, prefix test,
 * "[DONE]", json.Unmarshal into a materialised struct (choices as a heap slice with Go's growth rule, strings
 * unquoted into their own buffers), then the loop over the choices. */
typedef struct { uint8_t *content; uint32_t content_len; int has_finish; } o_choice;
typedef struct { o_choice *v; uint32_t len, cap; int type_err; } o_chunk;

static int go_is_space(uint32_t r)
{
    if (r == ' ' || (r >= 9 && r <= 13) || r == 0x85 || r == 0xA0 || r == 0x1680 || (r >= 0x2000 && r <= 0x200A)) return 1;
    return r == 0x2028 || r == 0x2029 || r == 0x202F || r == 0x205F || r == 0x3000;
}
static void trim_space(const uint8_t *p, size_t *b, size_t *e)
{
    while (*b < *e) {                                           /* TrimLeftFunc: decode runes from the front */
        uint32_t r;
        int n = utf8_decode(p + *b, *e - *b, &r);
        if (!(n >= 1 && go_is_space(r) && !(r == 0xFFFD && n == 1))) break;
        *b += (size_t)n;
    }
    while (*e > *b) {                                           /* TrimRightFunc: utf8.DecodeLastRune */
        size_t s = *e - 1;
        size_t lim = *e - *b >= 4 ? *e - 4 : *b;
        while (s > lim && (p[s] & 0xC0) == 0x80) s--;
        uint32_t r;
        int n = utf8_decode(p + s, *e - s, &r);
        if (s + (size_t)n != *e) { r = p[*e - 1]; s = *e - 1; if (r >= 0x80) break; }   /* not one whole rune at the end: invalid byte, not a space */
        if (!go_is_space(r)) break;
        *e = s;
    }
}
static int key_is(const uint8_t *lit, size_t n, const char *upper)
{
    uint8_t *k = (uint8_t *)malloc(n + 4);
    uint32_t kl = unquote(lit, n, k), i = 0;
    size_t L = strlen(upper), j = 0;
    int ok = 1;
    while (i < kl) {
        uint32_t r;
        i += (uint32_t)utf8_decode(k + i, kl - i, &r);
        if (j >= L || fold_rune(r) != (uint32_t)(uint8_t)upper[j]) { ok = 0; break; }
        j++;
    }
    free(k);
    return ok && j == L;
}
static size_t ws(const uint8_t *p, size_t i) { while (is_ws(p[i])) i++; return i; }
static size_t str_end(const uint8_t *p, size_t i) { for (i++; p[i] != '"'; i++) if (p[i] == '\\') i++; return i + 1; }

static size_t decode_delta(const uint8_t *p, size_t i, o_choice *c, int *type_err)
{
    if (p[i] == 'n') return i + 4;
    if (p[i] != '{') { *type_err = 1; return skip_value(p, i); }
    i = ws(p, i + 1);
    if (p[i] == '}') return i + 1;
    for (;;) {
        size_t ke = str_end(p, i);
        int is_content = key_is(p + i + 1, ke - i - 2, "CONTENT");
        i = ws(p, ws(p, ke) + 1);
        if (is_content && p[i] == '"') {
            size_t se = str_end(p, i);
            free(c->content);
            c->content = (uint8_t *)malloc(3 * (se - i) + 4);
            c->content_len = unquote(p + i + 1, se - i - 2, c->content);
            i = se;
        } else {
            if (is_content && p[i] != 'n') *type_err = 1;
            i = skip_value(p, i);
        }
        i = ws(p, i);
        if (p[i] == ',') { i = ws(p, i + 1); continue; }
        return i + 1;
    }
}
static size_t decode_choice(const uint8_t *p, size_t i, o_choice *c, int *type_err)
{
    if (p[i] == 'n') return i + 4;
    if (p[i] != '{') { *type_err = 1; return skip_value(p, i); }
    i = ws(p, i + 1);
    if (p[i] == '}') return i + 1;
    for (;;) {
        size_t ke = str_end(p, i);
        int is_delta = key_is(p + i + 1, ke - i - 2, "DELTA"), is_fin = key_is(p + i + 1, ke - i - 2, "FINISH_REASON");
        i = ws(p, ws(p, ke) + 1);
        if (is_delta) i = decode_delta(p, i, c, type_err);
        else if (is_fin) {
            if (p[i] == '"') c->has_finish = 1;
            else if (p[i] == 'n') c->has_finish = 0;
            else *type_err = 1;
            i = skip_value(p, i);
        } else i = skip_value(p, i);
        i = ws(p, i);
        if (p[i] == ',') { i = ws(p, i + 1); continue; }
        return i + 1;
    }
}
static void chunk_free(o_chunk *ch) { for (uint32_t k = 0; k < ch->cap; k++) free(ch->v[k].content); free(ch->v); ch->v = NULL; ch->len = ch->cap = 0; }

static void decode_chunk(const uint8_t *p, o_chunk *ch)
{
    size_t i = ws(p, 0);
    if (p[i] == 'n') return;
    if (p[i] != '{') { ch->type_err = 1; return; }
    i = ws(p, i + 1);
    if (p[i] == '}') return;
    for (;;) {
        size_t ke = str_end(p, i);
        int is_choices = key_is(p + i + 1, ke - i - 2, "CHOICES");
        i = ws(p, ws(p, ke) + 1);
        if (is_choices && p[i] == '[') {                         /* decode.go array() into a slice */
            uint32_t n = 0;
            i = ws(p, i + 1);
            if (p[i] == ']') i++;
            else for (;;) {
                if (n >= ch->cap) {                              /* grow: new backing array, the first len elements move over */
                    uint32_t nc = ch->cap + ch->cap / 2;
                    if (nc < 4) nc = 4;
                    o_choice *nv = (o_choice *)calloc(nc, sizeof *nv);
                    for (uint32_t k = 0; k < ch->len; k++) { nv[k] = ch->v[k]; ch->v[k].content = NULL; }
                    for (uint32_t k = 0; k < ch->cap; k++) free(ch->v[k].content);
                    free(ch->v);
                    ch->v = nv; ch->cap = nc;
                }
                if (n >= ch->len) ch->len = n + 1;
                i = decode_choice(p, i, &ch->v[n], &ch->type_err);
                n++;
                i = ws(p, i);
                if (p[i] == ',') { i = ws(p, i + 1); continue; }
                i++;
                break;
            }
            if (n < ch->len) ch->len = n;
            if (n == 0) chunk_free(ch);                          /* v.Set(reflect.MakeSlice(v.Type(), 0, 0)) */
        } else if (is_choices && p[i] == 'n') { chunk_free(ch); i += 4; }
        else {
            if (is_choices) ch->type_err = 1;
            i = skip_value(p, i);
        }
        i = ws(p, i);
        if (p[i] == ',') { i = ws(p, i + 1); continue; }
        return;
    }
}

/* One line.  Appends the messages it publishes to (tok, tok_off, done) starting at message *n_msgs / byte *w, returns
 * the status word (kind | 0x100 ends the stream). */
static uint32_t chunk_line(const uint8_t *line, size_t len, uint8_t *tok, uint32_t *tok_off, uint8_t *done, uint32_t *n_msgs, size_t *w)
{
    size_t b = 0, e = len;
    trim_space(line, &b, &e);
    if (e - b < 6 || memcmp(line + b, "data: ", 6) != 0) return 0;
    b += 6;
    if (e - b == 6 && memcmp(line + b, "[DONE]", 6) == 0) {
        tok_off[*n_msgs] = (uint32_t)*w; done[*n_msgs] = 1; (*n_msgs)++;
        memcpy(tok + *w, "[DONE]", 6); *w += 6;
        return 2 | 0x100;
    }
    if (!json_valid(line + b, e - b)) return 0;
    uint8_t *doc = (uint8_t *)malloc(e - b + 8);                /* NUL-terminated copy: the walkers above run to delimiters */
    memcpy(doc, line + b, e - b);
    memset(doc + (e - b), 0, 8);
    o_chunk ch = {0, 0, 0, 0};
    decode_chunk(doc, &ch);
    uint32_t st = 0;
    if (!ch.type_err) {
        st = 1;
        for (uint32_t k = 0; k < ch.len; k++) {
            if (ch.v[k].content_len) {
                tok_off[*n_msgs] = (uint32_t)*w; done[*n_msgs] = 0; (*n_msgs)++;
                memcpy(tok + *w, ch.v[k].content, ch.v[k].content_len); *w += ch.v[k].content_len;
            }
            if (ch.v[k].has_finish) {
                tok_off[*n_msgs] = (uint32_t)*w; done[*n_msgs] = 1; (*n_msgs)++;
                memcpy(tok + *w, "[DONE]", 6); *w += 6;
                st |= 0x100;
                break;
            }
        }
    }
    chunk_free(&ch);
    free(doc);
    return st;
}

/* tok must hold 3 x the input bytes + 6 per line; tok_off/done/msg_line one entry per possible message (<= bytes/8 + lines) */
uint32_t sseo_chunk_decode(const uint8_t *lines, const uint32_t *line_off, uint32_t n_lines, uint8_t *tok, uint32_t *tok_off,
                           uint8_t *done, uint32_t *msg_line, uint32_t *line_status)
{
    uint32_t n = 0;
    size_t w = 0;
    for (uint32_t i = 0; i < n_lines; i++) {
        const uint32_t before = n;
        line_status[i] = chunk_line(lines + line_off[i], line_off[i + 1] - line_off[i], tok, tok_off, done, &n, &w);
        for (uint32_t k = before; k < n; k++) msg_line[k] = i;
    }
    tok_off[n] = (uint32_t)w;
    return n;
}
