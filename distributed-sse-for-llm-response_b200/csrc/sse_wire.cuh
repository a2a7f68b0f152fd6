// sse_wire.cuh -- receiving side of the two wire protocols the hot path sits between: splitting a
// connection's socket bytes into the messages the services hand to their callbacks.
//
// In the reference this is third-party client code, not vendored: nats.go v1.31.0's parser delivers `msg.Data` to the
// adapter's callback (src/sse-adapter/sse_handler.go:135, :316) and go-redis v9.3.0's reader delivers `msg.Payload` to the
// bridge loop (src/redis-nats-bridge/main.go:115, :127, :142).  The byte layouts are restated from the NATS client
// protocol and from RESP2/RESP3 -- PARITY UNPINNED (no test or vector in the reference covers them).
//
// A byte stream is a length-prefixed chain, so ONE connection is sequential; a box holds thousands of connections,
// and those are independent: one lane walks one connection's buffer.  The walk is run twice (count, then emit at the
// offsets an exclusive scan of the counts gives) so that the messages of a batch come out in connection order, densely.
//
// NATS, server -> client (what this accepts; control lines end in LF with an optional CR before it, fields are separated
// by runs of space/tab, the verb is case-insensitive, as nats.go's parser has it):
//     MSG <subject> <sid> [reply-to] <#bytes>CRLF<payload>CRLF
//     HMSG <subject> <sid> [reply-to] <#header bytes> <#total bytes>CRLF<headers><payload>CRLF      (payload = total - header bytes)
//     PING | PONG | +OK | -ERR <text> | INFO <json>                                               (counted / skipped)
//   after the payload everything up to the next LF is skipped (nats.go: MSG_END).
// RESP2 / RESP3, Redis -> subscriber: a top-level array `*<n>` (or push `><n>`) of bulk strings whose first element is
//     message  <channel> <payload>          or          pmessage <pattern> <channel> <payload>
//   is a delivery; every other complete top-level value (subscribe acknowledgements, pongs, simple strings, integers,
//   errors, nulls, nested arrays to depth 8) is skipped.  Lines end in CRLF exactly.
// A connection's walk stops at the first incomplete message (its bytes stay in the socket buffer: `consumed` says how
// many were used) or at a protocol violation (status WIRE_ST_ERROR, `consumed` = where it happened).
#pragma once
#include <stdint.h>
#include "sse_record.cuh"

namespace sse {

enum : uint32_t { WIRE_NATS = 0, WIRE_RESP = 1 };
enum : uint32_t { WIRE_ST_OK = 0, WIRE_ST_ERROR = 1, WIRE_ST_SERVER_ERR = 2 /* bit: a -ERR line was seen */ };

struct WireMsg {                 // one delivery inside a connection's buffer (absolute offsets into the socket arena)
    uint32_t payload_off, payload_len;
    uint32_t subj_off, subj_len; // NATS subject / Redis channel
    uint64_t sid;                // NATS subscription id (0 for RESP)
};
struct WireConn { uint32_t n_msgs, payload_bytes, consumed, status, pings; };

SSE_HD bool wire_sp(uint32_t c) { return c == ' ' || c == '\t'; }

// decimal digits [b, e) -> value; false on anything else or more than 18 digits
template <class Src> SSE_HD bool wire_uint(const Src& s, uint32_t b, uint32_t e, uint64_t& v) {
    if (b >= e || e - b > 18) return false;
    v = 0;
    for (; b < e; b++) { const uint32_t d = s.ld8(b) - '0'; if (d > 9) return false; v = v * 10 + d; }
    return true;
}

// Walks one connection's bytes [b, e).  emit(k, msg) is called for delivery number k (0-based) in stream order.
template <class Src, class Emit>
SSE_HD WireConn wire_walk_nats(const Src& s, uint32_t b, uint32_t e, Emit&& emit) {
    WireConn r{0, 0, 0, WIRE_ST_OK, 0};
    uint32_t p = b;
    for (;;) {
        r.consumed = p - b;
        const uint32_t q = find_byte2(s, p, e, '\n', '\n');
        if (q >= e) return r;                                       // no complete control line
        uint32_t le = q;                                            // line = [p, le)
        if (le > p && s.ld8(le - 1) == '\r') le--;
        // fields
        uint32_t fb[7], fe[7], nf = 0, i = p;
        while (i < le && nf < 7) {
            while (i < le && wire_sp(s.ld8(i))) i++;
            if (i >= le) break;
            fb[nf] = i;
            i = find_byte2(s, i, le, ' ', '\t');
            fe[nf++] = i;
        }
        if (nf == 0) { r.status |= WIRE_ST_ERROR; return r; }
        const uint32_t vl = fe[0] - fb[0];
        const uint32_t c0 = s.ld8(fb[0]) | 0x20, c1 = vl > 1 ? (s.ld8(fb[0] + 1) | 0x20) : 0, c2 = vl > 2 ? (s.ld8(fb[0] + 2) | 0x20) : 0,
                       c3 = vl > 3 ? (s.ld8(fb[0] + 3) | 0x20) : 0;
        const bool is_msg = vl == 3 && c0 == 'm' && c1 == 's' && c2 == 'g';
        const bool is_hmsg = vl == 4 && c0 == 'h' && c1 == 'm' && c2 == 's' && c3 == 'g';
        if (is_msg || is_hmsg) {
            const uint32_t na = nf - 1, want = is_msg ? 3u : 4u;
            if (na != want && na != want + 1) { r.status |= WIRE_ST_ERROR; return r; }
            uint64_t sid, total, hdr = 0;
            if (!wire_uint(s, fb[2], fe[2], sid) || !wire_uint(s, fb[nf - 1], fe[nf - 1], total) ||
                (is_hmsg && !wire_uint(s, fb[nf - 2], fe[nf - 2], hdr)) || hdr > total || total > 0x7FFFFFFFull) { r.status |= WIRE_ST_ERROR; return r; }
            const uint64_t pend = (uint64_t)q + 1 + total;
            if (pend > e) return r;                                 // payload not complete yet
            const uint32_t t = find_byte2(s, (uint32_t)pend, e, '\n', '\n');   // nats.go MSG_END: skip to the LF that closes the message
            if (t >= e) return r;
            WireMsg m;
            m.payload_off = q + 1 + (uint32_t)hdr; m.payload_len = (uint32_t)(total - hdr);
            m.subj_off = fb[1]; m.subj_len = fe[1] - fb[1]; m.sid = sid;
            emit(r.n_msgs, m);
            r.n_msgs++; r.payload_bytes += m.payload_len;
            p = t + 1;
            continue;
        }
        const bool ping = vl == 4 && c0 == 'p' && c1 == 'i' && c2 == 'n' && c3 == 'g';
        const bool pong = vl == 4 && c0 == 'p' && c1 == 'o' && c2 == 'n' && c3 == 'g';
        const bool ok = vl == 3 && s.ld8(fb[0]) == '+' && c1 == 'o' && c2 == 'k';
        const bool err = vl >= 4 && s.ld8(fb[0]) == '-' && c1 == 'e' && c2 == 'r' && c3 == 'r';
        const bool info = vl == 4 && c0 == 'i' && c1 == 'n' && c2 == 'f' && c3 == 'o';
        if (ping) r.pings++;
        else if (err) r.status |= WIRE_ST_SERVER_ERR;
        else if (!(pong || ok || info)) { r.status |= WIRE_ST_ERROR; return r; }
        p = q + 1;
    }
}

// One RESP line "<type><text>CRLF" starting at p: returns the position after the LF, 0 if incomplete, 0xFFFFFFFF on a
// malformed terminator; [tb, te) is the text.
template <class Src> SSE_HD uint32_t resp_line(const Src& s, uint32_t p, uint32_t e, uint32_t& tb, uint32_t& te) {
    const uint32_t q = find_byte2(s, p + 1, e, '\n', '\n');
    if (q >= e) return 0;
    if (q < p + 2 || s.ld8(q - 1) != '\r') return 0xFFFFFFFFu;
    tb = p + 1; te = q - 1;
    return q + 1;
}
template <class Src> SSE_HD bool resp_int(const Src& s, uint32_t b, uint32_t e, int64_t& v) {
    bool neg = false;
    if (b < e && s.ld8(b) == '-') { neg = true; b++; }
    uint64_t u;
    if (!wire_uint(s, b, e, u)) return false;
    v = neg ? -(int64_t)u : (int64_t)u;
    return true;
}

template <class Src, class Emit>
SSE_HD WireConn wire_walk_resp(const Src& s, uint32_t b, uint32_t e, Emit&& emit) {
    WireConn r{0, 0, 0, WIRE_ST_OK, 0};
    constexpr int MAXD = 8;
    uint32_t p = b;
    for (;;) {                                                      // one top-level value per trip
        r.consumed = p - b;
        if (p >= e) return r;
        int64_t left[MAXD];                                         // elements still to read at each open array level
        int depth = 0;
        // bulk strings of the top-level array, in order (only the first four matter)
        uint32_t eb[4], el[4], ne = 0;
        int64_t top_n = -1;
        bool top_array = false, flat = true;
        uint32_t cur = p;
        for (;;) {                                                  // one value (possibly nested) per trip
            if (cur >= e) return r;                                 // incomplete: leave the whole top-level value in the buffer
            const uint32_t t = s.ld8(cur);
            uint32_t tb, te;
            const uint32_t nx = resp_line(s, cur, e, tb, te);
            if (nx == 0) return r;
            if (nx == 0xFFFFFFFFu) { r.status |= WIRE_ST_ERROR; return r; }
            bool closed_value = true;
            if (t == '*' || t == '>' || t == '~') {                 // array / push / set
                int64_t n;
                if (!resp_int(s, tb, te, n) || n > 0x7FFFFFFF) { r.status |= WIRE_ST_ERROR; return r; }
                if (depth == 0) { top_array = true; top_n = n; } else flat = false;
                if (n > 0) {
                    if (depth >= MAXD) { r.status |= WIRE_ST_ERROR; return r; }
                    left[depth++] = n;
                    closed_value = false;
                }
                cur = nx;
            } else if (t == '$' || t == '=' || t == '!') {          // bulk string / verbatim / bulk error
                int64_t n;
                if (!resp_int(s, tb, te, n) || n > 0x7FFFFFFF || n < -1) { r.status |= WIRE_ST_ERROR; return r; }
                if (n >= 0) {
                    const uint64_t end = (uint64_t)nx + (uint64_t)n + 2;
                    if (end > e) return r;                          // bulk not complete yet
                    if (s.ld8(nx + (uint32_t)n) != '\r' || s.ld8(nx + (uint32_t)n + 1) != '\n') { r.status |= WIRE_ST_ERROR; return r; }
                    if (depth == 1 && t == '$' && ne < 4) { eb[ne] = nx; el[ne] = (uint32_t)n; ne++; }
                    else if (depth == 1) flat = false;
                    cur = (uint32_t)end;
                } else { cur = nx; if (depth == 1) flat = false; }
            } else if (t == '+' || t == '-' || t == ':' || t == '_' || t == '#' || t == ',' || t == '(') {
                if (depth == 1) flat = false;
                cur = nx;                                           // simple string / error / integer / null / bool / double / big number
            } else { r.status |= WIRE_ST_ERROR; return r; }
            if (closed_value) {
                while (depth > 0 && --left[depth - 1] == 0) depth--;
            }
            if (depth == 0) break;
        }
        // a delivery?  go-redis: reply[0] is "message" (3 elements) or "pmessage" (4 elements), exact spelling
        if (top_array && flat && (ne == 3 || ne == 4) && top_n == (int64_t)ne) {
            const char* kind = ne == 3 ? "message" : "pmessage";
            const uint32_t kl = ne == 3 ? 7u : 8u;
            bool same = el[0] == kl;
            for (uint32_t k = 0; same && k < kl; k++) same = s.ld8(eb[0] + k) == (uint32_t)(uint8_t)kind[k];
            if (same) {
                WireMsg m;
                m.payload_off = eb[ne - 1]; m.payload_len = el[ne - 1];
                m.subj_off = eb[ne - 2]; m.subj_len = el[ne - 2]; m.sid = 0;
                emit(r.n_msgs, m);
                r.n_msgs++; r.payload_bytes += m.payload_len;
            }
        }
        p = cur;
    }
}

template <class Src, class Emit>
SSE_HD WireConn wire_walk(uint32_t proto, const Src& s, uint32_t b, uint32_t e, Emit&& emit) {
    return proto == WIRE_RESP ? wire_walk_resp(s, b, e, emit) : wire_walk_nats(s, b, e, emit);
}

}  // namespace sse
