// sse_chunk.cuh -- decode half of row N3: the proxy's read of the vLLM (OpenAI chat-completions) SSE stream.
//
// Reference: src/llm-stream-proxy/main.go:188-226 -- per line of the response body:
//     line = strings.TrimSpace(line)                      skip unless it starts with "data: "
//     data == "[DONE]"                                    -> publish "[DONE]" (done) and stop reading the stream
//     json.Unmarshal(data, &StreamChunk{})  error        -> skip the line
//     for each choice: content != "" -> publish it;  finish_reason != nil -> publish "[DONE]" (done) and stop
// with StreamChunk = { choices []{ delta {content string}, finish_reason *string } } (main.go:45-53).
// Go's decoding rules apply as everywhere on this path: whole-document syntax check first, keys
// matched exactly or under simple case folding, unknown keys skipped, null a no-op for strings / structs and nil for
// the pointer, a wrong-typed value makes Unmarshal return an error (the line is skipped) even though decoding went on,
// duplicate keys decode into the same value again -- for `choices` that means element-wise into the existing slice,
// which is then cut to the new length (decode.go array()).  The choices of one line are kept in a fixed array:
// a line with more than CHUNK_MAX_CHOICES choices gets the status CHUNK_TOO_MANY instead of a silent difference.
//
// One lane decodes one line.  What comes out per line: a status word, the number of messages it publishes (token
// messages, then possibly the "[DONE]" sentinel) and their decoded token bytes -- the columns ssegpu_encode_launch
// takes.  Sequence numbers, conversation ids and timestamps are per-chat state of the host (sequence++ per token,
// sentinel = sequence + 1, time.Now()); lines of a chat that follow a line flagged CHUNK_ENDS are never read by the
// reference, the host drops them.
#pragma once
#include <stdint.h>
#include "sse_record.cuh"
#include "sse_inspect.cuh"      // for_each_decoded_byte

namespace sse {

enum : uint32_t { CHUNK_SKIP = 0, CHUNK_DATA = 1, CHUNK_DONE_LINE = 2, CHUNK_KIND_MASK = 3,
                  CHUNK_ENDS = 0x100,          // the reference stops reading this chat's stream after this line
                  CHUNK_TOO_MANY = 0x200 };    // more than CHUNK_MAX_CHOICES choices in one line: not decoded (status only)
constexpr uint32_t CHUNK_MAX_CHOICES = 16;

struct ChunkChoice { uint32_t cb, ce; uint8_t has_content, finish; };   // content literal range (between the quotes), finish_reason != nil
struct ChunkEvents {
    ChunkChoice ch[CHUNK_MAX_CHOICES];
    uint32_t len, cap;          // the Go slice's length and capacity
    bool too_many, type_err;
};

// unicode.IsSpace for the UTF-8 sequence starting at p (e exclusive): returns its length, 0 if it is not white space
template <class Src> SSE_HD uint32_t go_space_at(const Src& s, uint32_t p, uint32_t e) {
    const uint32_t c = s.ld8(p);
    if (c == ' ' || (c >= 9 && c <= 13)) return 1;
    if (c == 0xC2 && p + 1 < e) { const uint32_t d = s.ld8(p + 1); return (d == 0x85 || d == 0xA0) ? 2u : 0u; }
    if (p + 2 < e) {
        const uint32_t d = s.ld8(p + 1), f = s.ld8(p + 2);
        if (c == 0xE1 && d == 0x9A && f == 0x80) return 3;                                   // U+1680
        if (c == 0xE2 && d == 0x80 && ((f >= 0x80 && f <= 0x8A) || f == 0xA8 || f == 0xA9 || f == 0xAF)) return 3;   // U+2000-200A, 2028, 2029, 202F
        if (c == 0xE2 && d == 0x81 && f == 0x9F) return 3;                                   // U+205F
        if (c == 0xE3 && d == 0x80 && f == 0x80) return 3;                                   // U+3000
    }
    return 0;
}
// strings.TrimSpace over [b, e)
template <class Src> SSE_HD void go_trim_space(const Src& s, uint32_t& b, uint32_t& e) {
    for (;;) { if (b >= e) return; const uint32_t n = go_space_at(s, b, e); if (!n) break; b += n; }
    for (;;) {                   // from the end: a space ends at e, so look at the last 1, 2 or 3 bytes
        if (e <= b) return;
        if (go_space_at(s, e - 1, e) == 1) { e -= 1; continue; }
        if (e - b >= 2 && (s.ld8(e - 2) & 0xC0) != 0x80 && go_space_at(s, e - 2, e) == 2) { e -= 2; continue; }
        if (e - b >= 3 && (s.ld8(e - 3) & 0xC0) != 0x80 && go_space_at(s, e - 3, e) == 3) { e -= 3; continue; }
        return;
    }
}

// ---- walking VALID JSON (the syntax check has run): positions only ----
template <class Src> SSE_HD uint32_t jv_ws(const Src& s, uint32_t p) { while (is_ws(s.ld8(p))) p++; return p; }
template <class Src> SSE_HD uint32_t jv_string_end(const Src& s, uint32_t p) {       // p on the opening quote -> past the closing quote
    for (p++;;) {                                    // (valid JSON: the closing quote exists; a word at a time between escapes)
        p = find_byte2(s, p, 0xFFFFFFF0u, '"', '\\');
        if (s.ld8(p) == '"') return p + 1;
        p += 2;                                       // backslash + the byte it escapes
    }
}
template <class Src> SSE_HD uint32_t jv_skip(const Src& s, uint32_t p) {             // p on the first byte of a value -> past it
    uint32_t depth = 0;
    for (;;) {
        const uint32_t c = s.ld8(p);
        if (c == '"') { p = jv_string_end(s, p); if (!depth) return p; continue; }
        if (c == '{' || c == '[') { depth++; p++; continue; }
        if (c == '}' || c == ']') { p++; if (--depth == 0) return p; continue; }
        if (!depth) {                                                                 // scalar: number / true / false / null
            while (true) { const uint32_t d = s.ld8(p); if (d == ',' || d == '}' || d == ']' || is_ws(d) || d == 0) break; p++; }
            return p;
        }
        p++;
    }
}
// does the key literal [b, e) name `upper` (ASCII upper case, length n) exactly or under simple case folding?
template <class Src> SSE_HD bool jv_key_is(const Src& s, uint32_t b, uint32_t e, const char* upper, uint32_t n) {
    uint32_t k = 0;
    while (b < e) {
        uint32_t r = next_rune(s, b, e);
        if (r >= 'a' && r <= 'z') r -= 32;
        else if (r == 0x212A) r = 'K';
        else if (r == 0x017F) r = 'S';
        if (k >= n || r != (uint32_t)(uint8_t)upper[k]) return false;
        k++;
    }
    return k == n;
}

// One element of `choices` (p on its first byte) decoded into c; returns the position past it.
template <class Src> SSE_HD uint32_t chunk_choice(const Src& s, uint32_t p, ChunkChoice& c, bool& type_err) {
    const uint32_t t = s.ld8(p);
    if (t == 'n') return p + 4;                                                       // null: the element keeps what it has
    if (t != '{') { type_err = true; return jv_skip(s, p); }
    p = jv_ws(s, p + 1);
    if (s.ld8(p) == '}') return p + 1;
    for (;;) {
        const uint32_t kb = p + 1, kend = jv_string_end(s, p);
        p = jv_ws(s, jv_ws(s, kend) + 1);                                             // past the colon
        const uint32_t v = s.ld8(p);
        if (jv_key_is(s, kb, kend - 1, "DELTA", 5)) {
            if (v == '{') {
                uint32_t q = jv_ws(s, p + 1);
                if (s.ld8(q) == '}') p = q + 1;
                else for (;;) {
                    const uint32_t kb2 = q + 1, kend2 = jv_string_end(s, q);
                    q = jv_ws(s, jv_ws(s, kend2) + 1);
                    const uint32_t v2 = s.ld8(q);
                    if (jv_key_is(s, kb2, kend2 - 1, "CONTENT", 7)) {
                        if (v2 == '"') { const uint32_t se = jv_string_end(s, q); c.cb = q + 1; c.ce = se - 1; c.has_content = 1; q = se; }
                        else { if (v2 != 'n') type_err = true; q = jv_skip(s, q); }
                    } else q = jv_skip(s, q);
                    q = jv_ws(s, q);
                    if (s.ld8(q) == ',') { q = jv_ws(s, q + 1); continue; }
                    p = q + 1;                                                        // '}'
                    break;
                }
            } else { if (v != 'n') type_err = true; p = jv_skip(s, p); }
        } else if (jv_key_is(s, kb, kend - 1, "FINISH_REASON", 13)) {
            if (v == '"') { c.finish = 1; p = jv_string_end(s, p); }
            else if (v == 'n') { c.finish = 0; p += 4; }                              // null sets the pointer to nil
            else { type_err = true; p = jv_skip(s, p); }
        } else p = jv_skip(s, p);
        p = jv_ws(s, p);
        if (s.ld8(p) == ',') { p = jv_ws(s, p + 1); continue; }
        return p + 1;                                                                 // '}'
    }
}

// json.Unmarshal(data, &StreamChunk{}) over the VALID document [b, e): the resulting choices.
template <class Src> SSE_HD void chunk_unmarshal(const Src& s, uint32_t b, ChunkEvents& ev) {
    ev.len = 0; ev.cap = 0; ev.too_many = false; ev.type_err = false;
    uint32_t p = jv_ws(s, b);
    const uint32_t t = s.ld8(p);
    if (t == 'n') return;                                                             // null: zero struct
    if (t != '{') { ev.type_err = true; return; }
    p = jv_ws(s, p + 1);
    if (s.ld8(p) == '}') return;
    for (;;) {
        const uint32_t kb = p + 1, kend = jv_string_end(s, p);
        p = jv_ws(s, jv_ws(s, kend) + 1);
        const uint32_t v = s.ld8(p);
        if (jv_key_is(s, kb, kend - 1, "CHOICES", 7)) {
            if (v == '[') {
                // decode.go array() into a slice: element i decodes INTO the existing element when there is one, the
                // slice grows by half (at least to 4) with zeroed new elements, and is cut to the new length at the end
                uint32_t i = 0;
                uint32_t q = jv_ws(s, p + 1);
                if (s.ld8(q) == ']') q++;
                else for (;;) {
                    if (i >= CHUNK_MAX_CHOICES) { ev.too_many = true; bool te = false; ChunkChoice dummy{0, 0, 0, 0}; q = chunk_choice(s, q, dummy, te); }
                    else {
                        if (i >= ev.cap) {                                            // grow: copy the first len elements, zero the rest
                            uint32_t nc = ev.cap + ev.cap / 2;
                            if (nc < 4) nc = 4;
                            if (nc > CHUNK_MAX_CHOICES) nc = CHUNK_MAX_CHOICES;
                            for (uint32_t k = ev.len; k < nc; k++) ev.ch[k] = ChunkChoice{0, 0, 0, 0};
                            ev.cap = nc;
                        }
                        if (i >= ev.len) ev.len = i + 1;                              // (exposes whatever an earlier, longer decode left there)
                        q = chunk_choice(s, q, ev.ch[i], ev.type_err);
                    }
                    i++;
                    q = jv_ws(s, q);
                    if (s.ld8(q) == ',') { q = jv_ws(s, q + 1); continue; }
                    q++;                                                              // ']'
                    break;
                }
                if (i < ev.len) ev.len = i;
                if (i == 0) { ev.len = 0; ev.cap = 0; }                               // "[]": a fresh empty slice replaces the old one
                p = q;
            } else if (v == 'n') { ev.len = 0; ev.cap = 0; p += 4; }                  // null: the slice becomes nil
            else { ev.type_err = true; p = jv_skip(s, p); }
        } else p = jv_skip(s, p);
        p = jv_ws(s, p);
        if (s.ld8(p) == ',') { p = jv_ws(s, p + 1); continue; }
        return;
    }
}

// The whole line [b, e): status word; `ev` holds the choices when the kind is CHUNK_DATA.  `nest` as for parse_general.
template <class Src> SSE_HD uint32_t chunk_line(const Src& s, uint32_t b, uint32_t e, ChunkEvents& ev, uint32_t* nest) {
    ev.len = 0; ev.cap = 0; ev.too_many = false; ev.type_err = false;
    go_trim_space(s, b, e);
    if (e - b < 6) return CHUNK_SKIP;
    const char* pre = "data: ";
    for (uint32_t k = 0; k < 6; k++) if (s.ld8(b + k) != (uint32_t)(uint8_t)pre[k]) return CHUNK_SKIP;
    b += 6;
    if (e - b == 6) {
        const char* dn = "[DONE]";
        bool same = true;
        for (uint32_t k = 0; k < 6; k++) same = same && s.ld8(b + k) == (uint32_t)(uint8_t)dn[k];
        if (same) return CHUNK_DONE_LINE | CHUNK_ENDS;
    }
    Plan unused;
    if (parse_general(s, b, e - b, unused, nest, true) == ST_ERR_SYNTAX) return CHUNK_SKIP;   // whole-document syntax check, as Unmarshal does first
    chunk_unmarshal(s, b, ev);
    if (ev.too_many) return CHUNK_DATA | CHUNK_TOO_MANY;
    if (ev.type_err) { ev.len = 0; return CHUNK_SKIP; }
    uint32_t st = CHUNK_DATA;
    for (uint32_t k = 0; k < ev.len; k++) if (ev.ch[k].finish) { st |= CHUNK_ENDS; ev.len = k + 1; break; }   // the loop returns at the first finish_reason
    return st;
}

// messages the line publishes: token messages (non-empty contents, in order), then the sentinel if the stream ends here
SSE_HD uint32_t chunk_message_count(uint32_t st, const ChunkEvents& ev) {
    if (st & CHUNK_TOO_MANY) return 0;
    uint32_t n = (st & CHUNK_ENDS) ? 1u : 0u;
    if ((st & CHUNK_KIND_MASK) == CHUNK_DATA) for (uint32_t k = 0; k < ev.len; k++) n += (ev.ch[k].has_content && ev.ch[k].ce > ev.ch[k].cb) ? 1u : 0u;
    return n;
}
template <class Src> SSE_HD uint32_t chunk_content_bytes(const Src& s, const ChunkChoice& c) {
    uint32_t n = 0;
    for_each_decoded_byte(s, c.cb, c.ce, [&](uint32_t) { n++; });
    return n;
}

}  // namespace sse
