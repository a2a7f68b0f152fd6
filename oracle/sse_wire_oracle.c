/*
 * sse_wire_oracle.c -- CPU restatement of the receiving side of the two wire protocols.
 *
 * TEST INFRASTRUCTURE: only tests/ may link or call this.  PARITY UNPINNED: in the reference the bytes are split by
 * third-party client libraries that are not vendored -- nats.go v1.31.0 (src/sse-adapter/go.mod:7) hands msg.Data to
 * the adapter's callback (src/sse-adapter/sse_handler.go:135, :316), go-redis v9.3.0 (src/redis-nats-bridge/go.mod:7)
 * hands msg.Payload to the bridge loop (src/redis-nats-bridge/main.go:115, :127, :142) -- and the reference has no test
 * or vector for it.  What is restated here are the published protocols: the NATS client protocol (MSG / HMSG / PING /
 * PONG / +OK / -ERR / INFO, as nats.go's byte-at-a-time parser accepts them) and RESP2/RESP3 pub/sub pushes.
 *
 * Shape: NATS is written as the byte-at-a-time state machine a streaming client runs; RESP as a recursive descent
 * over values.  Both are deliberately different from the product's line-oriented walk (csrc/sse_wire.cuh).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "sse_oracle.h"

typedef struct { sseo_wire_msg *v; uint32_t n, cap; } msgvec;
static void push(msgvec *mv, sseo_wire_msg m)
{
    if (mv->n == mv->cap) { mv->cap = mv->cap ? mv->cap * 2 : 16; mv->v = (sseo_wire_msg *)realloc(mv->v, mv->cap * sizeof *mv->v); }
    mv->v[mv->n++] = m;
}

static int is_sp(uint8_t c) { return c == ' ' || c == '\t'; }
static int lower(uint8_t c) { return c | 0x20; }

/* ---- NATS: state machine over bytes, committing after each complete protocol message ---- */
enum { OP_START, IN_LINE, IN_PAYLOAD, MSG_END };

static int parse_u64(const uint8_t *p, size_t n, uint64_t *out)
{
    if (n == 0 || n > 18) return 0;
    uint64_t v = 0;
    for (size_t i = 0; i < n; i++) { if (p[i] < '0' || p[i] > '9') return 0; v = v * 10 + (uint64_t)(p[i] - '0'); }
    *out = v;
    return 1;
}

static void nats_conn(const uint8_t *buf, uint32_t b, uint32_t e, msgvec *mv, sseo_wire_conn *c)
{
    int state = OP_START;
    uint32_t line_start = b, commit = b, need = 0, pay_start = 0;
    sseo_wire_msg cur;
    memset(&cur, 0, sizeof cur);
    memset(c, 0, sizeof *c);
    for (uint32_t i = b; i < e; i++) {
        const uint8_t ch = buf[i];
        switch (state) {
        case OP_START:
            line_start = i;
            state = IN_LINE;
            __attribute__((fallthrough));                 /* the first byte belongs to the line */
        case IN_LINE: {
            if (ch != '\n') break;
            uint32_t le = i;
            if (le > line_start && buf[le - 1] == '\r') le--;
            /* split into fields */
            uint32_t fb[8], fe[8], nf = 0, k = line_start;
            while (k < le && nf < 8) {
                while (k < le && is_sp(buf[k])) k++;
                if (k >= le) break;
                fb[nf] = k;
                while (k < le && !is_sp(buf[k])) k++;
                fe[nf++] = k;
            }
            if (nf == 0) goto violation;
            const uint8_t *v = buf + fb[0];
            const uint32_t vl = fe[0] - fb[0];
            const int msg = vl == 3 && lower(v[0]) == 'm' && lower(v[1]) == 's' && lower(v[2]) == 'g';
            const int hmsg = vl == 4 && lower(v[0]) == 'h' && lower(v[1]) == 'm' && lower(v[2]) == 's' && lower(v[3]) == 'g';
            if (msg || hmsg) {
                const uint32_t args = nf - 1, base = msg ? 3 : 4;
                uint64_t sid = 0, total = 0, hdr = 0;
                if (args != base && args != base + 1) goto violation;
                if (!parse_u64(buf + fb[2], fe[2] - fb[2], &sid)) goto violation;
                if (!parse_u64(buf + fb[nf - 1], fe[nf - 1] - fb[nf - 1], &total)) goto violation;
                if (hmsg && !parse_u64(buf + fb[nf - 2], fe[nf - 2] - fb[nf - 2], &hdr)) goto violation;
                if (hdr > total || total > 0x7FFFFFFFull) goto violation;
                cur.subj_off = fb[1]; cur.subj_len = fe[1] - fb[1]; cur.sid = sid;
                cur.payload_off = i + 1 + (uint32_t)hdr; cur.payload_len = (uint32_t)(total - hdr);
                pay_start = i + 1; need = (uint32_t)total;
                state = need ? IN_PAYLOAD : MSG_END;
                break;
            }
            if (vl == 4 && lower(v[0]) == 'p' && lower(v[1]) == 'i' && lower(v[2]) == 'n' && lower(v[3]) == 'g') c->pings++;
            else if (vl == 4 && lower(v[0]) == 'p' && lower(v[1]) == 'o' && lower(v[2]) == 'n' && lower(v[3]) == 'g') { }
            else if (vl == 3 && v[0] == '+' && lower(v[1]) == 'o' && lower(v[2]) == 'k') { }
            else if (vl >= 4 && v[0] == '-' && lower(v[1]) == 'e' && lower(v[2]) == 'r' && lower(v[3]) == 'r') c->status |= 2;
            else if (vl == 4 && lower(v[0]) == 'i' && lower(v[1]) == 'n' && lower(v[2]) == 'f' && lower(v[3]) == 'o') { }
            else goto violation;
            commit = i + 1;
            state = OP_START;
            break;
        }
        case IN_PAYLOAD:
            if (i + 1 - pay_start == need) state = MSG_END;
            break;
        case MSG_END:
            if (ch == '\n') {
                push(mv, cur);
                c->n_msgs++; c->payload_bytes += cur.payload_len;
                commit = i + 1;
                state = OP_START;
            }
            break;
        }
    }
    c->consumed = commit - b;
    return;
violation:
    c->status |= 1;
    c->consumed = line_start - b;
}

/* ---- RESP: recursive descent.  Returns the position after the value, 0 = incomplete, UINT32_MAX = violation ---- */
typedef struct { uint32_t off, len; int is_bulk; } resp_elem;

static uint32_t resp_crlf_line(const uint8_t *buf, uint32_t p, uint32_t e, uint32_t *tb, uint32_t *te)
{
    const uint8_t *nl = (const uint8_t *)memchr(buf + p, '\n', e - p);
    if (!nl) return 0;
    const uint32_t q = (uint32_t)(nl - buf);
    if (q < p + 2 || buf[q - 1] != '\r') return UINT32_MAX;
    *tb = p + 1; *te = q - 1;
    return q + 1;
}
static int resp_num(const uint8_t *buf, uint32_t tb, uint32_t te, int64_t *out)
{
    int neg = 0;
    uint64_t v;
    if (tb < te && buf[tb] == '-') { neg = 1; tb++; }
    if (!parse_u64(buf + tb, te - tb, &v)) return 0;
    *out = neg ? -(int64_t)v : (int64_t)v;
    return 1;
}
/* elems (may be NULL): receives the direct children when this value is an aggregate (first 5 only, count in *n_elems) */
static uint32_t resp_value(const uint8_t *buf, uint32_t p, uint32_t e, int depth, resp_elem *elems, uint32_t *n_elems, int *is_aggregate, resp_elem *self)
{
    uint32_t tb, te;
    if (p >= e) return 0;
    const uint8_t t = buf[p];
    const uint32_t nx = resp_crlf_line(buf, p, e, &tb, &te);
    if (nx == 0 || nx == UINT32_MAX) return nx;
    if (self) { self->off = 0; self->len = 0; self->is_bulk = 0; }
    if (is_aggregate) *is_aggregate = 0;
    if (t == '*' || t == '>' || t == '~') {
        int64_t n;
        if (!resp_num(buf, tb, te, &n) || n > 0x7FFFFFFF) return UINT32_MAX;
        if (is_aggregate) *is_aggregate = 1;
        if (n_elems) *n_elems = n > 0 ? (uint32_t)n : 0;
        if (n > 0 && depth >= 8) return UINT32_MAX;
        uint32_t cur = nx;
        for (int64_t k = 0; k < n; k++) {
            resp_elem child;
            int child_agg = 0;
            cur = resp_value(buf, cur, e, depth + 1, NULL, NULL, &child_agg, &child);
            if (cur == 0 || cur == UINT32_MAX) return cur;
            if (child_agg) child.is_bulk = 0;
            if (elems && k < 5) elems[k] = child;
        }
        return cur;
    }
    if (t == '$' || t == '=' || t == '!') {
        int64_t n;
        if (!resp_num(buf, tb, te, &n) || n > 0x7FFFFFFF || n < -1) return UINT32_MAX;
        if (n < 0) return nx;
        if ((uint64_t)nx + (uint64_t)n + 2 > e) return 0;
        if (buf[nx + n] != '\r' || buf[nx + n + 1] != '\n') return UINT32_MAX;
        if (self) { self->off = nx; self->len = (uint32_t)n; self->is_bulk = t == '$'; }
        return nx + (uint32_t)n + 2;
    }
    if (t == '+' || t == '-' || t == ':' || t == '_' || t == '#' || t == ',' || t == '(') return nx;
    return UINT32_MAX;
}

static void resp_conn(const uint8_t *buf, uint32_t b, uint32_t e, msgvec *mv, sseo_wire_conn *c)
{
    memset(c, 0, sizeof *c);
    uint32_t p = b;
    while (p < e) {
        resp_elem el[5];
        uint32_t n = 0;
        int agg = 0;
        memset(el, 0, sizeof el);
        const uint32_t nx = resp_value(buf, p, e, 0, el, &n, &agg, NULL);
        if (nx == 0) break;
        if (nx == UINT32_MAX) { c->status |= 1; break; }
        if (agg && (n == 3 || n == 4)) {
            int all_bulk = 1;
            for (uint32_t k = 0; k < n; k++) all_bulk = all_bulk && el[k].is_bulk;
            const char *kind = n == 3 ? "message" : "pmessage";
            if (all_bulk && el[0].len == strlen(kind) && memcmp(buf + el[0].off, kind, el[0].len) == 0) {
                sseo_wire_msg m;
                m.payload_off = el[n - 1].off; m.payload_len = el[n - 1].len;
                m.subj_off = el[n - 2].off; m.subj_len = el[n - 2].len; m.sid = 0;
                push(mv, m);
                c->n_msgs++; c->payload_bytes += m.payload_len;
            }
        }
        p = nx;
    }
    c->consumed = p - b;
}

/* Splits every connection's bytes buf[buf_off[c] .. buf_off[c+1]).  *msgs is malloc'ed (caller frees with free()),
 * deliveries in connection order; conns[n_conn] receives the per-connection summary. */
uint32_t sseo_wire_split(uint32_t proto, const uint8_t *buf, const uint32_t *buf_off, uint32_t n_conn, sseo_wire_msg **msgs,
                         uint32_t *msg_conn_out_cap_unused, sseo_wire_conn *conns)
{
    (void)msg_conn_out_cap_unused;
    msgvec mv = {0, 0, 0};
    for (uint32_t c = 0; c < n_conn; c++) {
        const uint32_t before = mv.n;
        if (proto == 1) resp_conn(buf, buf_off[c], buf_off[c + 1], &mv, &conns[c]);
        else nats_conn(buf, buf_off[c], buf_off[c + 1], &mv, &conns[c]);
        for (uint32_t k = before; k < mv.n; k++) mv.v[k].conn = c;
    }
    *msgs = mv.v;
    return mv.n;
}
void sseo_free(void *p) { free(p); }
