// sse_record.cuh -- per-record logic of the SSE hot path (decode -> filter -> encode -> frame).
//
// What it replaces in the reference (all Go, CPU, one goroutine hop per token):
//   json.Unmarshal(msg.Data,&token)           src/sse-adapter/sse_handler.go:136-140, :317-321
//                                             src/redis-nats-bridge/main.go:153-157
//   sequence filter                           src/sse-adapter/sse_handler.go:142-145
//   json.Marshal(token) + 3x fmt.Fprintf      src/sse-adapter/sse_handler.go:211-214, :427-430
//   "chat."+cid+".tokens" + json.Marshal      src/redis-nats-bridge/main.go:160-161
//
// Design (B200-first, not a translation of encoding/json):
//   * One thread owns one record; the record bytes sit in shared memory (TMA-staged tile)
//     or, for oversize tiles, in global memory.  Nothing is materialised: the decoder
//     yields a PLAN (source ranges of the winning string literals + the integers) and
//     the encoder streams runes from those ranges straight into the output.
//   * Canonical records -- exactly what Go's own json.Marshal produced, which is what
//     the proxy and the load generator publish -- satisfy Marshal(Unmarshal(x)) == x.
//     parse_fast() proves canonicity with word-wide SWAR tests, and the frame is then
//     header + verbatim payload copy.  Everything else takes parse_general(), a complete
//     single-pass restatement of the scanner+decoder rules.
//   * All functions are __host__ __device__ so the very same source is unit-tested on the
//     CPU build box (tests/hostcheck) before a GPU minute is spent.  That host build is a
//     test double only; the shipped library contains device code alone.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define SSE_HD __host__ __device__ __forceinline__
#define SSE_HD_NOINLINE __host__ __device__ __noinline__
#define SSE_CX __host__ __device__ constexpr
#else
#define SSE_CX constexpr
#define SSE_HD inline
#define SSE_HD_NOINLINE
#endif

namespace sse {

// status word (mirrors include/ssegpu.h)
enum : uint32_t {
    ST_OK = 0, ST_ERR_SYNTAX = 1, ST_ERR_TYPE = 2, ST_FILTERED = 3, ST_INSPECT_DROP = 4, ST_NO_INTEREST = 5,
    ST_PENDING = 0xE,            // internal: left to the tail phase of the tile kernel, never visible to the caller
    ST_CODE_MASK = 0xF, ST_DONE = 0x100, ST_REDACTED = 0x200, ST_HEADER_ONLY = 0x400,
};
// op flags (mirrors include/ssegpu.h)
enum : uint32_t { OP_BRIDGE = 1u, OP_GO122 = 2u, OP_NO_FAST = 4u, OP_ORDERED = 8u, OP_INSPECT = 16u, OP_INSPECT_INLINE = 32u, OP_MATCH = 64u,
                  OP_NATS_PUB = 128u, OP_RESP = 256u, OP_HEADER_ONLY = 512u };

constexpr uint32_t MAX_NESTING = 10000;            // encoding/json scanner.go maxNestingDepth
constexpr uint32_t NEST_WORDS = MAX_NESTING / 32 + 1;

struct Plan {
    uint32_t cid_b, cid_e;   // content range (between the quotes) of the winning conversation_id literal
    uint32_t tok_b, tok_e;   // same for token; b==e means ""
    int64_t  seq, ts;
    uint32_t done;           // 0/1
    uint32_t fast;           // 1: payload is canonical, JSON out == payload bytes
    uint32_t seq_b, seq_n;   // fast only: where the sequence literal sits in the payload (its text IS fmt %d)
    uint32_t cid_esc;        // fast only: the id literal contains a backslash escape (decoded id != raw bytes)
    uint32_t tok_esc;        // fast only: the token literal contains a backslash escape (decoded text != raw bytes)
    uint32_t tok_redacted;   // inline inspection replaced the token text with "[REDACTED]" (sse_handler.go:205-207)
};

// ---------------------------------------------------------------- accessors ----
// A Src gives byte/word reads relative to a 16-byte aligned tile origin; a Dst the
// matching writes.  Generic-pointer versions work on host and device (global or shared);
// the kernels add ld.shared/st.shared specialisations.
struct PtrSrc {
    const uint8_t* base;
    SSE_HD uint32_t ld8(uint32_t i) const { return base[i]; }
    SSE_HD uint32_t ld32a(uint32_t i) const { return *reinterpret_cast<const uint32_t*>(base + i); }
};
struct PtrDst {                 // 64-bit positions: the global output arena
    typedef uint64_t pos_t;
    uint8_t* base;
    SSE_HD void st8(uint64_t i, uint32_t v) const { base[i] = (uint8_t)v; }
    SSE_HD void st32a(uint64_t i, uint32_t v) const { *reinterpret_cast<uint32_t*>(base + i) = v; }
};
struct Ptr32Dst {               // 32-bit positions: a shared-memory staging tile
    typedef uint32_t pos_t;
    uint8_t* base;
    SSE_HD void st8(uint32_t i, uint32_t v) const { base[i] = (uint8_t)v; }
    SSE_HD void st32a(uint32_t i, uint32_t v) const { *reinterpret_cast<uint32_t*>(base + i) = v; }
};

SSE_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {   // sh in {0,8,16,24}
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}

// little-endian 4 bytes at an arbitrary byte position (may read up to 3 bytes past pos+3's
// word: tile buffers are padded)
template <class Src> SSE_HD uint32_t ld32u(const Src& s, uint32_t pos) {
    const uint32_t a = pos & ~3u;
    return funnel_r(s.ld32a(a), s.ld32a(a + 4), (pos & 3u) * 8);   // branch-free: lanes differ in alignment
}

SSE_CX uint32_t w4(const char* s) {
    return (uint32_t)(uint8_t)s[0] | ((uint32_t)(uint8_t)s[1] << 8) | ((uint32_t)(uint8_t)s[2] << 16) |
           ((uint32_t)(uint8_t)s[3] << 24);
}

// ---------------------------------------------------------------------- utf8 ----
// unicode/utf8.DecodeRune over [pos,end): size in the return value, rune in r; invalid -> (U+FFFD,1)
template <class Src> SSE_HD uint32_t utf8_decode(const Src& s, uint32_t pos, uint32_t end, uint32_t& r) {
    uint32_t b0 = s.ld8(pos);
    r = 0xFFFD;
    if (b0 < 0x80) { r = b0; return 1; }
    if (b0 < 0xC2 || b0 > 0xF4) return 1;
    uint32_t need = b0 < 0xE0 ? 1 : (b0 < 0xF0 ? 2 : 3);
    if (pos + need + 1 > end) return 1;                // truncated
    uint32_t b1 = s.ld8(pos + 1);
    uint32_t lo = 0x80, hi = 0xBF;
    if (b0 == 0xE0) lo = 0xA0; else if (b0 == 0xED) hi = 0x9F; else if (b0 == 0xF0) lo = 0x90; else if (b0 == 0xF4) hi = 0x8F;
    if (b1 < lo || b1 > hi) return 1;
    if (need == 1) { r = ((b0 & 0x1F) << 6) | (b1 & 0x3F); return 2; }
    uint32_t b2 = s.ld8(pos + 2);
    if ((b2 & 0xC0) != 0x80) return 1;
    if (need == 2) { r = ((b0 & 0x0F) << 12) | ((b1 & 0x3F) << 6) | (b2 & 0x3F); return 3; }
    uint32_t b3 = s.ld8(pos + 3);
    if ((b3 & 0xC0) != 0x80) return 1;
    r = ((b0 & 0x07) << 18) | ((b1 & 0x3F) << 12) | ((b2 & 0x3F) << 6) | (b3 & 0x3F);
    return 4;
}

SSE_HD int hexval(uint32_t c) {
    if (c >= '0' && c <= '9') return (int)c - '0';
    c |= 0x20;
    if (c >= 'a' && c <= 'f') return (int)c - 'a' + 10;
    return -1;
}

template <class Src> SSE_HD int hex4(const Src& s, uint32_t pos) {
    int a = hexval(s.ld8(pos)), b = hexval(s.ld8(pos + 1)), c = hexval(s.ld8(pos + 2)), d = hexval(s.ld8(pos + 3));
    if ((a | b | c | d) < 0) return -1;
    return (a << 12) | (b << 8) | (c << 4) | d;
}

// Next decoded rune of a VALID JSON string literal's content [pos,end)  (decode.go unquoteBytes).
template <class Src> SSE_HD uint32_t next_rune(const Src& s, uint32_t& pos, uint32_t end) {
    uint32_t c = s.ld8(pos);
    if (c == '\\') {
        uint32_t e = s.ld8(pos + 1);
        if (e != 'u') {
            pos += 2;
            switch (e) {
            case 'b': return 8;  case 'f': return 12; case 'n': return 10;
            case 'r': return 13; case 't': return 9;  default: return e;   // " \ /
            }
        }
        uint32_t rr = (uint32_t)hex4(s, pos + 2);
        pos += 6;
        if (rr >= 0xD800 && rr < 0xE000) {             // utf16.IsSurrogate
            if (rr < 0xDC00 && pos + 6 <= end && s.ld8(pos) == '\\' && s.ld8(pos + 1) == 'u') {
                uint32_t r2 = (uint32_t)hex4(s, pos + 2);
                if (r2 >= 0xDC00 && r2 < 0xE000) {      // utf16.DecodeRune
                    pos += 6;
                    return (((rr - 0xD800) << 10) | (r2 - 0xDC00)) + 0x10000;
                }
            }
            return 0xFFFD;
        }
        return rr;
    }
    if (c < 0x80) { pos += 1; return c; }
    uint32_t r;
    pos += utf8_decode(s, pos, end, r);
    return r;
}

// --------------------------------------------------------------- byte writer ----
// Streams bytes to a Dst at an arbitrary byte position.  Bytes gather in a 64-bit accumulator
// and leave as aligned 32-bit stores; only the stream's first and last word, which it shares
// with the neighbouring records (written by other threads), are written bytewise.  Feeding it
// is branch-light: put_n() is shift/or plus one predicated flush.  EMIT=false only counts.
template <class Dst, bool EMIT> struct Writer {
    typedef typename Dst::pos_t pos_t;
    const Dst& d;
    pos_t wpos;         // aligned position of the word being filled
    uint64_t acc;       // gathered bytes, bit 0 = byte 0 of the word at wpos
    uint32_t sh;        // bits of acc in use (8 * bytes), 0..56
    uint32_t first;     // first byte of the word at wpos that belongs to this stream (0 after the first flush)
    SSE_HD Writer(const Dst& dst, pos_t start)
        : d(dst), wpos(start & ~(pos_t)3), acc(0), sh(((uint32_t)start & 3u) * 8), first((uint32_t)start & 3u) {}
    SSE_HD pos_t pos() const { return wpos + (sh >> 3); }
    SSE_HD void flush_word() {                          // sh >= 32: the word at wpos is complete
        if (EMIT) {                                     // straight-line predicated stores: lanes differ in `first`
            const uint32_t v = (uint32_t)acc;
            if (first == 0) d.st32a(wpos, v);
            if (first != 0 && first <= 1) d.st8(wpos + 1, (v >> 8) & 0xFF);
            if (first != 0 && first <= 2) d.st8(wpos + 2, (v >> 16) & 0xFF);
            if (first != 0) d.st8(wpos + 3, v >> 24);
            first = 0;
        }
        acc >>= 32; sh -= 32; wpos += 4;
    }
    SSE_HD void put_n(uint32_t w, uint32_t n) {         // low n (1..4) bytes of w, little endian
        if (EMIT) acc |= (uint64_t)(w & (0xFFFFFFFFu >> (32 - 8 * n))) << sh;
        sh += 8 * n;
        if (sh >= 32) flush_word();
    }
    SSE_HD void put(uint32_t b) { put_n(b, 1); }
    SSE_HD void put32(uint32_t w) {
        if (EMIT) acc |= (uint64_t)w << sh;
        sh += 32;
        flush_word();
    }
    // put32 for streams whose first partial word is already out (first == 0): no first-word test
    SSE_HD void put32_nf(uint32_t w) {
        if (EMIT) { acc |= (uint64_t)w << sh; d.st32a(wpos, (uint32_t)acc); }
        acc >>= 32; wpos += 4;
    }
    SSE_HD void put_n_nf(uint32_t w, uint32_t n) {
        if (EMIT) acc |= (uint64_t)(w & (0xFFFFFFFFu >> (32 - 8 * n))) << sh;
        sh += 8 * n;
        if (sh >= 32) { if (EMIT) d.st32a(wpos, (uint32_t)acc); acc >>= 32; sh -= 32; wpos += 4; }
    }
    SSE_HD void store_word(uint32_t v) {                // only when sh == 0 and first == 0
        if (EMIT) d.st32a(wpos, v);
        wpos += 4;
    }
    SSE_HD void store_words4(uint32_t v0, uint32_t v1, uint32_t v2, uint32_t v3) {     // same precondition
        if (EMIT) { d.st32a(wpos, v0); d.st32a(wpos + 4, v1); d.st32a(wpos + 8, v2); d.st32a(wpos + 12, v3); }
        wpos += 16;
    }
    SSE_HD void store_words_upto3(uint32_t v0, uint32_t v1, uint32_t v2, uint32_t r) {  // the first r (0..3) of them
        if (EMIT) {
            if (r > 0) d.st32a(wpos, v0);
            if (r > 1) d.st32a(wpos + 4, v1);
            if (r > 2) d.st32a(wpos + 8, v2);
        }
        wpos += 4 * r;
    }
    SSE_HD void skip(uint32_t n) {                      // counting only
        uint32_t t = (sh >> 3) + n;
        wpos += t & ~3u; sh = (t & 3u) * 8;
    }
    SSE_HD void finish() {                              // trailing partial word: bytes [first, sh/8)
        if (EMIT) {
            const uint32_t k1 = sh >> 3, v = (uint32_t)acc;
            if (first == 0 && k1 >= 1) d.st8(wpos, v & 0xFF);
            if (first <= 1 && k1 >= 2) d.st8(wpos + 1, (v >> 8) & 0xFF);
            if (first <= 2 && k1 >= 3) d.st8(wpos + 2, (v >> 16) & 0xFF);
        }
    }
};

template <class W> SSE_HD void put_lit_nf(W& w, const char* s, uint32_t n) {     // after the first flush
    uint32_t i = 0;
    for (; i + 4 <= n; i += 4) w.put32_nf(w4(s + i));
    if (i < n) {
        uint32_t v = 0;
        for (uint32_t k = 0; i + k < n; k++) v |= (uint32_t)(uint8_t)s[i + k] << (8 * k);
        w.put_n_nf(v, n - i);
    }
}
template <class W> SSE_HD void put_lit(W& w, const char* s, uint32_t n) {
    uint32_t i = 0;
    for (; i + 4 <= n; i += 4) w.put32(w4(s + i));
    if (i < n) {
        uint32_t v = 0;
        for (uint32_t k = 0; i + k < n; k++) v |= (uint32_t)(uint8_t)s[i + k] << (8 * k);
        w.put_n(v, n - i);
    }
}
#define SSE_PUT_LIT(w, lit) put_lit(w, lit, (uint32_t)(sizeof(lit) - 1))
#define SSE_PUT_LIT_NF(w, lit) put_lit_nf(w, lit, (uint32_t)(sizeof(lit) - 1))

SSE_HD uint32_t digits_u32(uint32_t x) {               // x < 10^8
    return 1u + (x >= 10u) + (x >= 100u) + (x >= 1000u) + (x >= 10000u) + (x >= 100000u) + (x >= 1000000u) + (x >= 10000000u);
}
SSE_HD uint32_t dec_digits(uint64_t u) {
    if (u < 100000000ull) return digits_u32((uint32_t)u);
    if (u < 10000000000000000ull) return 8u + digits_u32((uint32_t)(u / 100000000ull));
    return 16u + digits_u32((uint32_t)(u / 10000000000000000ull));
}
SSE_HD uint32_t dec_len(int64_t v) {                   // strconv.AppendInt / fmt %d length
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    return dec_digits(u) + (v < 0 ? 1u : 0u);
}
SSE_HD uint32_t ascii4(uint32_t x) {                   // x < 10^4 -> its four digits, most significant in byte 0
    const uint32_t d0 = x / 1000u, r0 = x - d0 * 1000u, d1 = r0 / 100u, r1 = r0 - d1 * 100u, d2 = r1 / 10u, d3 = r1 - d2 * 10u;
    return 0x30303030u + (d0 | d1 << 8 | d2 << 16 | d3 << 24);
}
template <class W> SSE_HD void put_dec8(W& w, uint32_t c, uint32_t k) {   // the low k (1..8) decimal digits of c < 10^8
    const uint32_t hi = c / 10000u, lo = c - hi * 10000u;
    const uint64_t v = ((uint64_t)ascii4(hi) | (uint64_t)ascii4(lo) << 32) >> (8u * (8u - k));
    if (k > 4) { w.put32((uint32_t)v); w.put_n((uint32_t)(v >> 32), k - 4); }
    else w.put_n((uint32_t)v, k);
}
template <class W> SSE_HD void put_int(W& w, int64_t v) {
    const uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    if (v < 0) w.put('-');
    if (u < 100000000ull) { put_dec8(w, (uint32_t)u, digits_u32((uint32_t)u)); return; }
    const uint64_t a = u / 10000000000000000ull, r = u - a * 10000000000000000ull;
    const uint32_t b = (uint32_t)(r / 100000000ull), c = (uint32_t)(r - (uint64_t)b * 100000000ull);
    if (a) { put_dec8(w, (uint32_t)a, digits_u32((uint32_t)a)); put_dec8(w, b, 8); }
    else put_dec8(w, b, digits_u32(b));
    put_dec8(w, c, 8);
}

SSE_HD uint32_t hexl(uint32_t v) { return v < 10 ? '0' + v : 'a' + (v - 10); }

// encode.go appendString(escapeHTML=true) for one decoded rune.
template <class W> SSE_HD void put_rune_json(W& w, uint32_t r, bool go122) {
    if (r < 0x80) {
        if (r >= 0x20 && r != '"' && r != '\\' && r != '<' && r != '>' && r != '&') { w.put(r); return; }
        w.put('\\');
        switch (r) {
        case '\\': case '"': w.put(r); return;
        case '\n': w.put('n'); return;
        case '\r': w.put('r'); return;
        case '\t': w.put('t'); return;
        default: break;
        }
        if (go122 && r == 8)  { w.put('b'); return; }
        if (go122 && r == 12) { w.put('f'); return; }
        w.put('u'); w.put('0'); w.put('0'); w.put(hexl(r >> 4)); w.put(hexl(r & 0xF));
        return;
    }
    if (r == 0x2028 || r == 0x2029) { w.put('\\'); w.put('u'); w.put('2'); w.put('0'); w.put('2'); w.put(hexl(r & 0xF)); return; }
    if (r < 0x800) { w.put(0xC0 | (r >> 6)); w.put(0x80 | (r & 0x3F)); return; }
    if (r < 0x10000) { w.put(0xE0 | (r >> 12)); w.put(0x80 | ((r >> 6) & 0x3F)); w.put(0x80 | (r & 0x3F)); return; }
    w.put(0xF0 | (r >> 18)); w.put(0x80 | ((r >> 12) & 0x3F)); w.put(0x80 | ((r >> 6) & 0x3F)); w.put(0x80 | (r & 0x3F));
}
template <class W> SSE_HD void put_rune_raw(W& w, uint32_t r) {   // utf8.EncodeRune
    if (r < 0x80) { w.put(r); return; }
    if (r < 0x800) { w.put(0xC0 | (r >> 6)); w.put(0x80 | (r & 0x3F)); return; }
    if (r < 0x10000) { w.put(0xE0 | (r >> 12)); w.put(0x80 | ((r >> 6) & 0x3F)); w.put(0x80 | (r & 0x3F)); return; }
    w.put(0xF0 | (r >> 18)); w.put(0x80 | ((r >> 12) & 0x3F)); w.put(0x80 | ((r >> 6) & 0x3F)); w.put(0x80 | (r & 0x3F));
}

template <class Src> SSE_HD bool go_string_is_plain(const Src& s, uint32_t b, uint32_t e);
template <class Src, class W> SSE_HD void put_copy(const Src& s, uint32_t sp, uint32_t n, W& w);

// Re-encode the interior [b, e) of a JSON string literal.  A literal made of plain bytes only (no escape, nothing
// appendString would escape, no byte >= 0x80) decodes to itself and encodes to itself: word-wise test, word-wise copy.
template <class Src, class W> SSE_HD void put_json_string(const Src& s, uint32_t b, uint32_t e, W& w, bool go122) {
    w.put('"');
    if (go_string_is_plain(s, b, e)) put_copy(s, b, e - b, w);
    else while (b < e) put_rune_json(w, next_rune(s, b, e), go122);
    w.put('"');
}

// verbatim copy of n source bytes starting at byte position sp: bytes until the destination is
// word aligned, then one aligned load + funnel shift + aligned store per word (the relative
// alignment of source and destination is constant), then the tail bytes.  NF: the writer's
// first partial word has already been flushed (true whenever a header went out before).
template <bool NF, class Src, class W> SSE_HD void put_copy_t(const Src& s, uint32_t sp, uint32_t n, W& w) {
    uint32_t k = (4u - (w.sh >> 3)) & 3u;
    if (k > n) k = n;
    if (k) { if (NF) w.put_n_nf(ld32u(s, sp), k); else w.put_n(ld32u(s, sp), k); sp += k; n -= k; }
    const uint32_t nw = n >> 2;
    if (nw) {
        // full groups of four words in a plain loop (offsets are immediates, no per-word predicates), then the 0..3
        // remaining words as straight-line predicated code: lanes differ only in the trip count and meet again right
        // after the loop.  Loads past the last word stay inside the padded tile.
        uint32_t a = sp & ~3u;
        const uint32_t shb = (sp & 3u) * 8;
        uint32_t lo = s.ld32a(a);
        uint32_t i = 0;
        for (; i + 8 <= nw; i += 8) {                // eight words per trip: no register shuffling between the halves
            const uint32_t h0 = s.ld32a(a + 4), h1 = s.ld32a(a + 8), h2 = s.ld32a(a + 12), h3 = s.ld32a(a + 16);
            const uint32_t h4 = s.ld32a(a + 20), h5 = s.ld32a(a + 24), h6 = s.ld32a(a + 28), h7 = s.ld32a(a + 32);
            w.store_words4(funnel_r(lo, h0, shb), funnel_r(h0, h1, shb), funnel_r(h1, h2, shb), funnel_r(h2, h3, shb));
            w.store_words4(funnel_r(h3, h4, shb), funnel_r(h4, h5, shb), funnel_r(h5, h6, shb), funnel_r(h6, h7, shb));
            lo = h7; a += 32;
        }
        if (i + 4 <= nw) {
            const uint32_t h0 = s.ld32a(a + 4), h1 = s.ld32a(a + 8), h2 = s.ld32a(a + 12), h3 = s.ld32a(a + 16);
            w.store_words4(funnel_r(lo, h0, shb), funnel_r(h0, h1, shb), funnel_r(h1, h2, shb), funnel_r(h2, h3, shb));
            lo = h3; a += 16; i += 4;
        }
        const uint32_t r = nw - i;
        const uint32_t h0 = s.ld32a(a + 4), h1 = s.ld32a(a + 8), h2 = s.ld32a(a + 12);
        w.store_words_upto3(funnel_r(lo, h0, shb), funnel_r(h0, h1, shb), funnel_r(h1, h2, shb), r);
        sp += nw * 4; n -= nw * 4;
    }
    if (n) { if (NF) w.put_n_nf(ld32u(s, sp), n); else w.put_n(ld32u(s, sp), n); }
}
template <class Src, class W> SSE_HD void put_copy(const Src& s, uint32_t sp, uint32_t n, W& w) { put_copy_t<false>(s, sp, n, w); }
template <class Src> SSE_HD void ld3u(const Src& s, uint32_t pos, uint32_t& w0, uint32_t& w1, uint32_t& w2);
// the same for at most 12 bytes (a canonical sequence literal has 1..9), after the first flush: three unaligned word
// loads and three puts instead of the alignment prologue / word loop / tail of the general copy
template <class Src, class W> SSE_HD void put_copy_small_nf(const Src& s, uint32_t sp, uint32_t n, W& w) {
    uint32_t w0, w1, w2;
    ld3u(s, sp, w0, w1, w2);
    w.put_n_nf(w0, n < 4 ? n : 4);
    if (n > 4) w.put_n_nf(w1, n < 8 ? n - 4 : 4);
    if (n > 8) w.put_n_nf(w2, n - 8);
}

// ------------------------------------------------------------------ emission ----
// Writes the record's output (SSE frame, or bridge subject+JSON) at w.pos.  With a counting
// writer it measures the length.  `o` = record start inside Src.
template <class Src, class W>
SSE_HD void put_marshal(const Src& s, uint32_t o, const Plan& pl, bool go122, W& w) {
    SSE_PUT_LIT(w, "{\"conversation_id\":");
    put_json_string(s, o + pl.cid_b, o + pl.cid_e, w, go122);
    SSE_PUT_LIT(w, ",\"token\":");
    if (pl.tok_redacted) SSE_PUT_LIT(w, "\"[REDACTED]\"");   // RedactedContent of the Spin inspector (lib.rs:73)
    else put_json_string(s, o + pl.tok_b, o + pl.tok_e, w, go122);
    SSE_PUT_LIT(w, ",\"sequence\":");
    put_int(w, pl.seq);
    SSE_PUT_LIT(w, ",\"done\":");
    if (pl.done) SSE_PUT_LIT(w, "true"); else SSE_PUT_LIT(w, "false");
    SSE_PUT_LIT(w, ",\"timestamp\":");
    put_int(w, pl.ts);
    w.put('}');
}
template <class Src, class W>
SSE_HD void emit_record(const Src& s, uint32_t o, uint32_t L, const Plan& pl, uint32_t op, W& w, uint32_t* subj_len) {
    const bool go122 = (op & OP_GO122) != 0;
    const bool pub = (op & (OP_BRIDGE | OP_NATS_PUB)) == (OP_BRIDGE | OP_NATS_PUB);
    if (op & OP_BRIDGE) {
        // NATS client protocol, as nats.go's publish() writes it: PUB <subject> <#bytes>\r\n<payload>\r\n
        if (pub) SSE_PUT_LIT(w, "PUB ");
        // redis-nats-bridge/main.go:160  fmt.Sprintf("chat.%s.tokens", token.ConversationID)
        auto p0 = w.pos();
        SSE_PUT_LIT(w, "chat.");
        for (uint32_t b = o + pl.cid_b, e = o + pl.cid_e; b < e;) put_rune_raw(w, next_rune(s, b, e));
        SSE_PUT_LIT(w, ".tokens");
        if (subj_len) *subj_len = (uint32_t)(w.pos() - p0);
        if (pub) {
            uint32_t plen = L;
            if (!pl.fast) {
                Ptr32Dst nd{nullptr};
                Writer<Ptr32Dst, false> c(nd, 0);
                put_marshal(s, o, pl, go122, c);
                plen = c.pos();
            }
            w.put(' '); put_int(w, (int64_t)plen); w.put('\r'); w.put('\n');
        }
    } else {
        SSE_PUT_LIT(w, "event: token\nid: ");            // sse_handler.go:212-213
        put_int(w, pl.seq);
        SSE_PUT_LIT(w, "\ndata: ");                      // :214
    }
    if (pl.fast) put_copy(s, o, L, w);                   // Marshal(Unmarshal(x)) == x for canonical x
    else put_marshal(s, o, pl, go122, w);
    if (!(op & OP_BRIDGE)) { w.put('\n'); w.put('\n'); }
    if (pub) { w.put('\r'); w.put('\n'); }
    w.finish();
}

// ------------------------------------------------------------- canonical path ----
// A byte is "plain" inside a canonical Go-marshalled string when appendString copies it
// through unchanged and unquoteBytes reads it unchanged: 0x20..0x7F except " \ < > &.
// Returns a mask with bit 7 of byte k set if byte k MAY be special (exact for the lowest
// flagged byte, which is the only one used).
SSE_HD uint32_t special_mask(uint32_t w) {
    const uint32_t H = 0x80808080u, L1 = 0x01010101u;
    uint32_t lt = (w - 0x20202020u) & ~w;                       // < 0x20
    uint32_t a = (w & 0xFBFBFBFBu) ^ 0x22222222u;               // '"' (0x22) or '&' (0x26)
    uint32_t b = (w & 0xFDFDFDFDu) ^ 0x3C3C3C3Cu;               // '<' (0x3C) or '>' (0x3E)
    uint32_t c = w ^ 0x5C5C5C5Cu;                               // '\\'
    uint32_t z = ((a - L1) & ~a) | ((b - L1) & ~b) | ((c - L1) & ~c);
    return (w | lt | z) & H;                                    // w & H: >= 0x80
}
// bit 7 of byte k set iff byte k is not an ASCII digit (exact for every byte)
SSE_HD uint32_t nondigit_mask(uint32_t w) {
    uint32_t x = w ^ 0x30303030u;                               // digits -> 0x00..0x09
    return (((x & 0x7F7F7F7Fu) + 0x76767676u) | x) & 0x80808080u;
}

SSE_HD uint32_t first_byte_idx(uint32_t m) {                    // m != 0, flags at bit 7 of each byte
#if defined(__CUDA_ARCH__)
    return (uint32_t)(__ffs((int)m) - 1) >> 3;
#else
    return (uint32_t)__builtin_ctz(m) >> 3;
#endif
}

// \uXXXX spellings json.Marshal itself produces (so that they survive decode->encode verbatim)
SSE_HD bool canon_u_escape(uint32_t h, bool go122) {           // h = the four hex bytes, little endian
    const uint32_t h3 = h >> 24, h012 = h & 0xFFFFFFu;
    if (h012 == (w4("000\0") & 0xFFFFFFu)) {                    // \u000X: controls without a short form
        if (go122 && (h3 == '8' || h3 == 'c')) return false;
        return (h3 >= '0' && h3 <= '8') || h3 == 'b' || h3 == 'c' || h3 == 'e' || h3 == 'f';
    }
    if (h012 == (w4("001\0") & 0xFFFFFFu)) return (h3 >= '0' && h3 <= '9') || (h3 >= 'a' && h3 <= 'f');
    if (h012 == (w4("003\0") & 0xFFFFFFu)) return h3 == 'c' || h3 == 'e';                  // < >
    if (h012 == (w4("002\0") & 0xFFFFFFu)) return h3 == '6';                               // &
    if (h012 == (w4("202\0") & 0xFFFFFFu)) return h3 == '8' || h3 == '9';                  // U+2028 / U+2029
    return false;
}

// The canonical path is written SINGLE-EXIT: no early returns, failures only clear an `ok` flag and the code
// runs on (positions clamped to the record end, so every read stays inside the padded tile).  With early
// returns the lanes of a warp that leave a variable-length loop at different trips never reconverge (the
// only common post-dominator is the end of the function) and everything after the first string -- literal
// compares, both integers, the emission set-up -- ran at half the lanes (ncu r01: 16.0 active threads).

// Scans a canonical string body starting at pos (just after the opening quote); on success pos is left ON
// the closing quote.  Fails on anything json.Marshal would not have produced verbatim -- the general path
// then decides what it really is.
template <class Src> SSE_HD bool scan_canon_string(const Src& s, uint32_t& pos, uint32_t end, bool go122, bool& has_esc) {
    bool ok = true, fin = false;
    while (!fin) {
        // word loop: one aligned load per 4 bytes, the previous words are carried, the next trip's are in flight
        uint32_t a = pos & ~3u;
        const uint32_t shb = (pos & 3u) * 8;
        uint32_t lo = s.ld32a(a), h1 = s.ld32a(a + 4), h2 = s.ld32a(a + 8), w = 0, m = 0;
        while (pos < end) {                                     // 8 bytes per trip
            const uint32_t n1 = s.ld32a(a + 12), n2 = s.ld32a(a + 16);
            const uint32_t wa = funnel_r(lo, h1, shb), wb = funnel_r(h1, h2, shb);
            const uint32_t ma = special_mask(wa), mb = special_mask(wb);
            if (ma | mb) {
                const bool first = ma != 0;
                w = first ? wa : wb; m = first ? ma : mb; pos += first ? 0u : 4u;
                break;
            }
            lo = h2; h1 = n1; h2 = n2; a += 8; pos += 8;
        }
        if (m == 0) { ok = false; fin = true; }                 // ran off the record without a closing quote
        else {
            const uint32_t k = first_byte_idx(m);
            pos += k;
            const uint32_t c = (w >> (8 * k)) & 0xFF;
            if (c == '"') { ok = pos < end; fin = true; }
            else if (c == '\\') {
                has_esc = true;
                const uint32_t e = s.ld8(pos + 1);
                if (e == '"' || e == '\\' || e == 'n' || e == 'r' || e == 't' || (go122 && (e == 'b' || e == 'f'))) pos += 2;
                else if (e == 'u' && canon_u_escape(ld32u(s, pos + 2), go122)) pos += 6;
                else { ok = false; fin = true; }
            } else if (c >= 0x80) {                             // raw multi-byte UTF-8: decoded == raw
                uint32_t r;
                const uint32_t n = utf8_decode(s, pos, end, r);
                if (n == 1 || r == 0x2028 || r == 0x2029) { ok = false; fin = true; }
                else pos += n;
            } else { ok = false; fin = true; }                  // control byte, or < > & raw
        }
    }
    if (pos > end) pos = end;
    return ok;
}

// four ASCII digits (first digit in the low byte) -> value
SSE_HD uint32_t digits4(uint32_t w) {
    uint32_t t = w & 0x0F0F0F0Fu;
    t = t * 10 + (t >> 8);                                      // byte0 = d0*10+d1, byte2 = d2*10+d3
    return (t & 0xFF) * 100 + ((t >> 16) & 0xFF);
}

// canonical integer literal with a value: -?(0|[1-9][0-9]*) of at most 8 digits (sequence numbers; longer
// ones go to the general path, which parses any int64).  Leaves pos just past the literal.
template <class Src> SSE_HD bool scan_canon_int(const Src& s, uint32_t& pos, uint32_t end, int64_t& out) {
    const uint32_t a = pos & ~3u, shb = (pos & 3u) * 8;
    const uint32_t x0 = s.ld32a(a), x1 = s.ld32a(a + 4), x2 = s.ld32a(a + 8), x3 = s.ld32a(a + 12);
    uint32_t w0 = funnel_r(x0, x1, shb), w1 = funnel_r(x1, x2, shb);
    const uint32_t w2 = funnel_r(x2, x3, shb);
    const uint32_t neg = ((w0 & 0xFF) == '-') ? 1u : 0u;
    if (neg) { w0 = funnel_r(w0, w1, 8); w1 = funnel_r(w1, w2, 8); }             // drop the sign byte
    const uint32_t start = pos + neg;
    const uint32_t m0 = nondigit_mask(w0), m1 = nondigit_mask(w1);
    const uint32_t nd = m0 ? first_byte_idx(m0) : (m1 ? 4 + first_byte_idx(m1) : 9);   // 9 = "more than 8"
    const uint32_t d0 = w0 & 0xFF;
    bool ok = nd >= 1 && nd <= 8 && start + nd <= end;
    ok = ok && !(d0 == '0' && (nd != 1 || neg));                // leading zero, "-0"
    const uint32_t ndc = nd > 8 ? 8 : (nd < 1 ? 1 : nd);
    // right-align the digits inside 4-digit groups: missing leading digits become zeros
    const uint32_t k = ndc > 4 ? ndc - 4 : ndc;                 // digits in the leading group (1..4)
    const uint32_t hi = digits4(w0 << (8 * (4 - k)));
    const uint32_t lo4 = k == 4 ? w1 : funnel_r(w0, w1, 8 * k);
    const uint32_t v = ndc > 4 ? hi * 10000 + digits4(lo4) : hi;
    out = neg ? -(int64_t)v : (int64_t)v;
    pos = start + ndc;
    if (pos > end) pos = end;
    return ok;
}

// The timestamp literal is the record's tail: it occupies [pos, end-1) and '}' sits at end-1.
// Its value is not needed on the canonical path, only that it IS canonical: six independent
// aligned loads, five funnel shifts, digit masks -- straight-line, no search loop.
template <class Src> SSE_HD bool check_canon_tail_int(const Src& s, uint32_t pos, uint32_t end) {
    const uint32_t lim = end - 1;
    const uint32_t a = pos & ~3u, shb = (pos & 3u) * 8;
    const uint32_t x0 = s.ld32a(a), x1 = s.ld32a(a + 4), x2 = s.ld32a(a + 8), x3 = s.ld32a(a + 12), x4 = s.ld32a(a + 16),
                   x5 = s.ld32a(a + 20);
    uint32_t ws[6] = {funnel_r(x0, x1, shb), funnel_r(x1, x2, shb), funnel_r(x2, x3, shb), funnel_r(x3, x4, shb), funnel_r(x4, x5, shb), x5 >> shb};
    const uint32_t neg = ((ws[0] & 0xFF) == '-') ? 1u : 0u;
    if (neg) {                                                  // drop the sign byte (timestamps are positive in practice)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 5; k++) ws[k] = funnel_r(ws[k], ws[k + 1], 8);
    }
    const uint32_t start = pos + neg;
    const uint32_t nd = lim > start ? lim - start : 0;
    const uint32_t d0 = ws[0] & 0xFF;
    bool ok = nd >= 1 && nd <= 19;
    ok = ok && !(d0 == '0' && (nd != 1 || neg));                // leading zero, "-0"
    ok = ok && !(nd == 19 && d0 == '9');                        // might overflow: exact check in the general path
    uint32_t bad = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < 5; k++) {
        const int c = (int)nd - 4 * k;                          // digits expected in word k
        const uint32_t keep = c >= 4 ? 0x80808080u : (c > 0 ? (0x80808080u >> (8 * (4 - c))) : 0u);
        bad |= nondigit_mask(ws[k]) & keep;
    }
    return ok && bad == 0 && s.ld8(lim) == '}';
}

// three unaligned words at pos with four aligned loads
template <class Src> SSE_HD void ld3u(const Src& s, uint32_t pos, uint32_t& w0, uint32_t& w1, uint32_t& w2) {
    const uint32_t a = pos & ~3u, shb = (pos & 3u) * 8;
    uint32_t x0 = s.ld32a(a), x1 = s.ld32a(a + 4), x2 = s.ld32a(a + 8), x3 = s.ld32a(a + 12);
    w0 = funnel_r(x0, x1, shb); w1 = funnel_r(x1, x2, shb); w2 = funnel_r(x2, x3, shb);
}

// Hook called by parse_fast right after the conversation id has been scanned (absolute range [b, e) in s; `plain`: the
// record is canonical so far and the id has no escape).  The interest lookup uses it to hash the id and put its first
// table probe in flight while the rest of the record is still being parsed.
struct NoHook { template <class Src> SSE_HD void operator()(const Src&, uint32_t, uint32_t, bool) const {} };

// True iff the record is byte-for-byte what json.Marshal(TokenMessage) would emit.
template <class Src, class Hook> SSE_HD bool parse_fast(const Src& s, uint32_t o, uint32_t L, uint32_t op, Plan& pl, Hook&& after_cid) {
    const bool go122 = (op & OP_GO122) != 0;
    bool ok = L >= 72;
    const uint32_t end = o + (ok ? L : 0u);                     // a too-short record runs through with an empty range
    uint32_t w0, w1, w2, d;
    // {"conversation_id":"   (20 bytes)
    {
        const uint32_t a = o & ~3u, shb = (o & 3u) * 8;
        uint32_t x0 = s.ld32a(a), x1 = s.ld32a(a + 4), x2 = s.ld32a(a + 8), x3 = s.ld32a(a + 12), x4 = s.ld32a(a + 16), x5 = s.ld32a(a + 20);
        d = (funnel_r(x0, x1, shb) ^ w4("{\"co")) | (funnel_r(x1, x2, shb) ^ w4("nver")) | (funnel_r(x2, x3, shb) ^ w4("sati")) |
            (funnel_r(x3, x4, shb) ^ w4("on_i")) | (funnel_r(x4, x5, shb) ^ w4("d\":\""));
        ok = ok && d == 0;
    }
    uint32_t pos = ok ? o + 20 : end;
    bool cid_esc = false, tok_esc = false;
    pl.cid_b = 20;
    ok = scan_canon_string(s, pos, end, go122, cid_esc) && ok;
    pl.cid_e = pos - o;
    after_cid(s, o + 20, pos, ok && !cid_esc);
    // ","token":"   (11 bytes)
    ld3u(s, pos, w0, w1, w2);
    d = (w0 ^ w4("\",\"t")) | (w1 ^ w4("oken")) | ((w2 ^ w4("\":\"\0")) & 0xFFFFFFu);
    ok = ok && d == 0 && pos + 11 <= end;
    pos = ok ? pos + 11 : end;
    pl.tok_b = pos - o;
    ok = scan_canon_string(s, pos, end, go122, tok_esc) && ok;
    pl.tok_e = pos - o;
    // ","sequence":   (13 bytes)
    ld3u(s, pos, w0, w1, w2);
    d = (w0 ^ w4("\",\"s")) | (w1 ^ w4("eque")) | (w2 ^ w4("nce\"")) | (s.ld8(pos + 12) ^ ':');
    ok = ok && d == 0 && pos + 14 <= end;
    pos = ok ? pos + 13 : end;
    pl.seq_b = pos - o;
    ok = scan_canon_int(s, pos, end, pl.seq) && ok;
    pl.seq_n = pos - o - pl.seq_b;
    // ,"done":   (8 bytes) + true|false
    ld3u(s, pos, w0, w1, w2);
    d = (w0 ^ w4(",\"do")) | (w1 ^ w4("ne\":"));
    const bool is_true = w2 == w4("true");
    const bool is_false = w2 == w4("fals") && s.ld8(pos + 12) == 'e';
    ok = ok && d == 0 && pos + 13 <= end && (is_true || is_false);
    pl.done = is_true ? 1u : 0u;
    pos = ok ? pos + (is_true ? 12u : 13u) : end;
    // ,"timestamp":   (13 bytes)
    ld3u(s, pos, w0, w1, w2);
    d = (w0 ^ w4(",\"ti")) | (w1 ^ w4("mest")) | (w2 ^ w4("amp\"")) | (s.ld8(pos + 12) ^ ':');
    ok = ok && d == 0 && pos + 14 <= end;
    pos = ok ? pos + 13 : end;
    pl.ts = 0;                                                  // never read for canonical records
    ok = check_canon_tail_int(s, pos, ok ? end : pos + 1) && ok;
    pl.fast = ok ? 1u : 0u;
    pl.tok_redacted = 0;
    pl.cid_esc = cid_esc ? 1u : 0u;   // the bridge subject needs the DECODED id
    pl.tok_esc = tok_esc ? 1u : 0u;
    return ok;
}

template <class Src> SSE_HD bool parse_fast(const Src& s, uint32_t o, uint32_t L, uint32_t op, Plan& pl) { return parse_fast(s, o, L, op, pl, NoHook()); }

// First position in [p, e) whose byte is c0 or c1, else e.  Word-wise (aligned 4-byte loads, exact zero-byte test on the
// lowest match): one load per four bytes on the walk's dependent chain instead of one per byte.  Reads up to 3 bytes past e.
template <class Src> SSE_HD uint32_t find_byte2(const Src& s, uint32_t p, uint32_t e, uint32_t c0, uint32_t c1) {
    if (p >= e) return e;
    uint32_t a = p & ~3u;
    uint32_t w = s.ld32a(a);
    const uint32_t before = ~(0xFFFFFFFFu << ((p & 3u) * 8));      // bytes in front of p
    const uint32_t k0 = c0 * 0x01010101u, k1 = c1 * 0x01010101u;
    for (;;) {
        uint32_t x0 = w ^ k0, x1 = w ^ k1;
        if (a < p) { x0 |= before; x1 |= before; }                  // first word only: what lies in front of p never matches
        const uint32_t m = (((x0 - 0x01010101u) & ~x0) | ((x1 - 0x01010101u) & ~x1)) & 0x80808080u;
        if (m) { const uint32_t q = a + first_byte_idx(m); return q < e ? q : e; }
        a += 4;
        if (a >= e) return e;
        w = s.ld32a(a);
    }
}

// --------------------------------------------------------------- general path ----
SSE_HD bool is_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }
SSE_HD bool is_digit(uint32_t c) { return c - '0' <= 9u; }

// scanner.go stateInString*: pos on the opening quote; returns position past the closing
// quote or 0 on error.
template <class Src> SSE_HD uint32_t scan_string(const Src& s, uint32_t pos, uint32_t end) {
    for (pos++; pos < end; pos++) {
        // runs of plain bytes four at a time (special_mask also stops at < > & and bytes >= 0x80, which the byte step
        // below lets through)
        while (pos + 4 <= end) {
            const uint32_t m = special_mask(ld32u(s, pos));
            if (m) { pos += first_byte_idx(m); break; }
            pos += 4;
        }
        if (pos >= end) break;
        uint32_t c = s.ld8(pos);
        if (c == '"') return pos + 1;
        if (c < 0x20) return 0;
        if (c == '\\') {
            if (++pos >= end) return 0;
            uint32_t e = s.ld8(pos);
            if (e == 'u') {
                if (pos + 4 >= end) return 0;
                if (hex4(s, pos + 1) < 0) return 0;
                pos += 4;
            } else if (!(e == 'b' || e == 'f' || e == 'n' || e == 'r' || e == 't' || e == '\\' || e == '/' || e == '"')) return 0;
        }
    }
    return 0;
}

// scanner.go number states; returns position past the number or 0.  *plain_int is set when
// the literal is -?digits (the only shape strconv.ParseInt accepts).
template <class Src> SSE_HD uint32_t scan_number(const Src& s, uint32_t pos, uint32_t end, bool* plain_int) {
    *plain_int = true;
    if (s.ld8(pos) == '-') { if (++pos >= end) return 0; }
    uint32_t c = s.ld8(pos);
    if (c == '0') pos++;
    else if (c >= '1' && c <= '9') { while (pos < end && is_digit(s.ld8(pos))) pos++; }
    else return 0;
    if (pos < end && s.ld8(pos) == '.') {
        *plain_int = false;
        pos++;
        if (pos >= end || !is_digit(s.ld8(pos))) return 0;
        while (pos < end && is_digit(s.ld8(pos))) pos++;
    }
    if (pos < end && (s.ld8(pos) | 0x20) == 'e') {
        *plain_int = false;
        pos++;
        if (pos < end && (s.ld8(pos) == '+' || s.ld8(pos) == '-')) pos++;
        if (pos >= end || !is_digit(s.ld8(pos))) return 0;
        while (pos < end && is_digit(s.ld8(pos))) pos++;
    }
    return pos;
}

// strconv.ParseInt(lit,10,64) for a literal already known to be -?digits
template <class Src> SSE_HD bool parse_int64(const Src& s, uint32_t b, uint32_t e, int64_t& out) {
    bool neg = false;
    if (s.ld8(b) == '-') { neg = true; b++; }
    // a JSON integer literal has no leading zeros, so 20 digits or more are out of range whatever they are, and up to
    // 19 digits (< 10^19 < 2^64) accumulate without overflow: no per-digit range test
    if (e - b > 19) return false;
    uint64_t v = 0;
    for (; b < e; b++) v = v * 10 + (uint64_t)(s.ld8(b) - '0');
    if (neg) { if (v > (1ull << 63)) return false; out = (int64_t)(0 - v); }
    else     { if (v > 0x7FFFFFFFFFFFFFFFull) return false; out = (int64_t)v; }
    return true;
}

// decode.go object(): field whose name equals the unquoted key exactly, else under simple
// case folding (fold.go foldName).  Only {K,k,U+212A} and {S,s,U+017F} fold onto ASCII.
template <class Src> SSE_HD int match_field(const Src& s, uint32_t b, uint32_t e) {
    // the exact names first, compared as words (what every producer that uses the struct tags sends); a name that
    // matches exactly is also the one the folded comparison below would find
    {
        const uint32_t n = e - b, w0 = ld32u(s, b);
        if (n == 5 && w0 == w4("toke") && s.ld8(b + 4) == 'n') return 1;
        if (n == 4 && w0 == w4("done")) return 3;
        if (n == 8 && w0 == w4("sequ") && ld32u(s, b + 4) == w4("ence")) return 2;
        if (n == 9 && w0 == w4("time") && ld32u(s, b + 4) == w4("stam") && s.ld8(b + 8) == 'p') return 4;
        if (n == 15 && w0 == w4("conv") && ld32u(s, b + 4) == w4("ersa") && ld32u(s, b + 8) == w4("tion") && ld32u(s, b + 11) == w4("n_id")) return 0;
    }
    uint32_t alive = 0x1F, k = 0;
    while (b < e && alive) {
        uint32_t r = next_rune(s, b, e);
        if (r >= 'a' && r <= 'z') r -= 32;
        else if (r == 0x212A) r = 'K';
        else if (r == 0x017F) r = 'S';
        const char* n0 = "CONVERSATION_ID"; const char* n1 = "TOKEN"; const char* n2 = "SEQUENCE";
        const char* n3 = "DONE"; const char* n4 = "TIMESTAMP";
        if (k >= 15 || r != (uint32_t)(uint8_t)n0[k]) alive &= ~1u;
        if (k >= 5  || r != (uint32_t)(uint8_t)n1[k]) alive &= ~2u;
        if (k >= 8  || r != (uint32_t)(uint8_t)n2[k]) alive &= ~4u;
        if (k >= 4  || r != (uint32_t)(uint8_t)n3[k]) alive &= ~8u;
        if (k >= 9  || r != (uint32_t)(uint8_t)n4[k]) alive &= ~16u;
        k++;
    }
    if (b < e) return -1;
    if ((alive & 1u) && k == 15) return 0;
    if ((alive & 2u) && k == 5) return 1;
    if ((alive & 4u) && k == 8) return 2;
    if ((alive & 8u) && k == 4) return 3;
    if ((alive & 16u) && k == 9) return 4;
    return -1;
}

// json.Unmarshal(data,&TokenMessage{}) in one pass: whole-document syntax validation
// (scanner.go) with the depth-1 field assignments of decode.go applied on the way.  A syntax
// error anywhere wins over any type error, exactly as checkValid-then-decode does.
// `nest` is a MAX_NESTING-bit stack (1 = object, 0 = array), touched only when nesting.
template <class Src>
SSE_HD_NOINLINE uint32_t parse_general(const Src& s, uint32_t o, uint32_t L, Plan& pl, uint32_t* nest, bool syntax_only = false) {
    pl.cid_b = pl.cid_e = pl.tok_b = pl.tok_e = 0;
    pl.seq = 0; pl.ts = 0; pl.done = 0; pl.fast = 0; pl.seq_b = pl.seq_n = 0; pl.cid_esc = 0; pl.tok_esc = 0; pl.tok_redacted = 0;
    enum { BEGIN_VALUE, BEGIN_VALUE_OR_EMPTY, BEGIN_KEY_OR_EMPTY, BEGIN_KEY, END_VALUE };
    uint32_t st = BEGIN_VALUE, depth = 0, pos = o;
    const uint32_t end = o + L;
    bool type_err = false, top_obj = false;
    int field = -1;
    for (;;) {
        while (pos < end && is_ws(s.ld8(pos))) pos++;
        if (pos >= end) {
            if (!(st == END_VALUE && depth == 0)) return ST_ERR_SYNTAX;
            return type_err ? ST_ERR_TYPE : ST_OK;
        }
        uint32_t c = s.ld8(pos);
        switch (st) {
        case BEGIN_VALUE_OR_EMPTY:
            if (c == ']') { depth--; pos++; st = END_VALUE; break; }
            /* fallthrough */
        case BEGIN_VALUE: {
            const bool member = (depth == 1 && top_obj);          // value of a top-level object member
            const int f = member ? field : -1;
            if (c == '{' || c == '[') {
                if (depth + 1 > MAX_NESTING) return ST_ERR_SYNTAX;
                if (depth == 0) { top_obj = (c == '{'); if (!top_obj) type_err = true; }
                if (f >= 0) type_err = true;                      // object/array into string/int/bool
                if (c == '{') nest[depth >> 5] |= 1u << (depth & 31); else nest[depth >> 5] &= ~(1u << (depth & 31));
                depth++; pos++;
                st = (c == '{') ? BEGIN_KEY_OR_EMPTY : BEGIN_VALUE_OR_EMPTY;
                break;
            }
            if (c == '"') {
                uint32_t e = scan_string(s, pos, end);
                if (!e) return ST_ERR_SYNTAX;
                if (depth == 0) type_err = true;
                if (f == 0) { pl.cid_b = pos + 1 - o; pl.cid_e = e - 1 - o; }
                else if (f == 1) { pl.tok_b = pos + 1 - o; pl.tok_e = e - 1 - o; }
                else if (f >= 2) type_err = true;
                pos = e;
            } else if (c == '-' || is_digit(c)) {
                bool plain;
                uint32_t e = scan_number(s, pos, end, &plain);
                if (!e) return ST_ERR_SYNTAX;
                if (depth == 0) type_err = true;
                if (f == 2) { if (!plain || !parse_int64(s, pos, e, pl.seq)) type_err = true; }
                else if (f == 4) { if (!plain || !parse_int64(s, pos, e, pl.ts)) type_err = true; }
                else if (f >= 0) type_err = true;
                pos = e;
            } else if (c == 't') {
                if (pos + 4 > end || ld32u(s, pos) != w4("true")) return ST_ERR_SYNTAX;
                if (depth == 0) type_err = true;
                if (f == 3) pl.done = 1; else if (f >= 0) type_err = true;
                pos += 4;
            } else if (c == 'f') {
                if (pos + 5 > end || ld32u(s, pos) != w4("fals") || s.ld8(pos + 4) != 'e') return ST_ERR_SYNTAX;
                if (depth == 0) type_err = true;
                if (f == 3) pl.done = 0; else if (f >= 0) type_err = true;
                pos += 5;
            } else if (c == 'n') {
                if (pos + 4 > end || ld32u(s, pos) != w4("null")) return ST_ERR_SYNTAX;
                pos += 4;                                         // null: no-op everywhere, also at top level
            } else return ST_ERR_SYNTAX;
            st = END_VALUE;
            break;
        }
        case BEGIN_KEY_OR_EMPTY:
            if (c == '}') { depth--; pos++; st = END_VALUE; break; }
            /* fallthrough */
        case BEGIN_KEY: {
            if (c != '"') return ST_ERR_SYNTAX;
            uint32_t e = scan_string(s, pos, end);
            if (!e) return ST_ERR_SYNTAX;
            if (depth == 1 && top_obj && !syntax_only) field = match_field(s, pos + 1, e - 1);   // (syntax_only: another struct's document, sse_chunk.cuh)
            pos = e;
            while (pos < end && is_ws(s.ld8(pos))) pos++;
            if (pos >= end || s.ld8(pos) != ':') return ST_ERR_SYNTAX;
            pos++;
            st = BEGIN_VALUE;
            break;
        }
        default: /* END_VALUE */
            if (depth == 0) return ST_ERR_SYNTAX;                 // stateEndTop
            if (nest[(depth - 1) >> 5] & (1u << ((depth - 1) & 31))) {
                if (c == ',') st = BEGIN_KEY; else if (c == '}') depth--; else return ST_ERR_SYNTAX;
            } else {
                if (c == ',') st = BEGIN_VALUE; else if (c == ']') depth--; else return ST_ERR_SYNTAX;
            }
            pos++;
            break;
        }
    }
}

// -------------------------------------------------------------------- record ----
// The kernel keeps the canonical path inline (shared-memory loads/stores, ~1k instructions)
// and reaches everything else through two out-of-line calls, so the hot loop stays small in
// the instruction cache.  tests/hostcheck runs the same dispatch on the CPU.

// May the record be emitted by emit_record_fast()?
SSE_HD bool fast_emit_ok(const Plan& pl, uint32_t op) {
    return pl.fast && !((op & OP_BRIDGE) && pl.cid_esc);
}

// Status word and output length of a canonical record from the few facts its frame depends on (the emit pass of
// ordered mode recomputes them from the plan word the counting pass left, instead of loading them).
SSE_HD void fast_status_len(uint32_t op, uint32_t L, uint32_t cid_len, uint32_t seq_n, bool done, bool filtered, uint32_t& st, uint32_t& out_len) {
    out_len = 0;
    if (filtered) { st = ST_FILTERED; return; }
    out_len = (op & OP_BRIDGE) ? 12 + cid_len + L : 26 + seq_n + L;
    if ((op & (OP_BRIDGE | OP_NATS_PUB)) == (OP_BRIDGE | OP_NATS_PUB)) out_len += 9 + dec_digits(L);   // "PUB " + " " + digits + CRLF, CRLF
    st = done ? (ST_OK | ST_DONE) : ST_OK;
    if (op & OP_HEADER_ONLY) {      // the payload (which IS the JSON) and the trailer stay with the host: only what precedes them is produced
        out_len -= L + ((op & OP_BRIDGE) ? ((op & OP_NATS_PUB) ? 2u : 0u) : 2u);
        st |= ST_HEADER_ONLY;
    }
}

// Pass 1, canonical records only.  false = not canonical (or fast path disabled): call plan_record().
template <class Src, class Hook>
SSE_HD bool plan_record_fast(const Src& s, uint32_t o, uint32_t L, uint32_t op, int64_t start_seq,
                             Plan& pl, uint32_t& st, uint32_t& out_len, Hook&& after_cid) {
    pl.fast = 0;
    if ((op & OP_NO_FAST) || !parse_fast(s, o, L, op, pl, after_cid)) return false;
    if (!fast_emit_ok(pl, op)) return false;
    const bool filtered = !(op & OP_BRIDGE) && start_seq > 0 && pl.seq <= start_seq;                      // sse_handler.go:143
    fast_status_len(op, L, pl.cid_e - pl.cid_b, pl.seq_n, pl.done != 0, filtered, st, out_len);
    return true;
}

template <class Src>
SSE_HD bool plan_record_fast(const Src& s, uint32_t o, uint32_t L, uint32_t op, int64_t start_seq,
                             Plan& pl, uint32_t& st, uint32_t& out_len) {
    return plan_record_fast(s, o, L, op, start_seq, pl, st, out_len, NoHook());
}

// Pass 2 for records accepted by plan_record_fast(): header + verbatim payload (+ trailer).
template <class Src, class W>
SSE_HD void emit_record_fast(const Src& s, uint32_t o, uint32_t L, const Plan& pl, uint32_t op, W& w, uint32_t* subj_len) {
    const bool body = !(op & OP_HEADER_ONLY);
    if (op & OP_BRIDGE) {
        const bool pub = (op & OP_NATS_PUB) != 0;
        if (pub) { w.put32(w4("PUB ")); SSE_PUT_LIT_NF(w, "chat."); }   // nats.go publish(): PUB <subject> <#bytes>CRLF<payload>CRLF
        else { w.put32(w4("chat")); w.put_n_nf('.', 1); }  // redis-nats-bridge/main.go:160 (first word: bytewise if shared)
        put_copy_t<true>(s, o + pl.cid_b, pl.cid_e - pl.cid_b, w);
        SSE_PUT_LIT_NF(w, ".tokens");
        if (subj_len) *subj_len = 12 + (pl.cid_e - pl.cid_b);
        if (pub) { w.put_n_nf(' ', 1); put_int(w, (int64_t)L); w.put_n_nf(w4("\r\n\0\0"), 2); }
        if (body) {
            put_copy_t<true>(s, o, L, w);                 // main.go:161
            if (pub) w.put_n_nf(w4("\r\n\0\0"), 2);
        }
    } else {
        w.put32(w4("even"));                              // sse_handler.go:212-213
        SSE_PUT_LIT_NF(w, "t: token\nid: ");
        put_copy_small_nf(s, o + pl.seq_b, pl.seq_n, w);  // canonical literal == fmt %d (1..9 bytes on this path)
        SSE_PUT_LIT_NF(w, "\ndata: ");                    // :214
        if (body) {
            put_copy_t<true>(s, o, L, w);
            w.put_n_nf(w4("\n\n\0\0"), 2);
        }
    }
    w.finish();
}

// Output length of an already decoded record (after inline inspection changed the plan).
template <class Src>
SSE_HD uint32_t measure_record(const Src& s, uint32_t o, uint32_t L, const Plan& pl, uint32_t op) {
    Ptr32Dst nd{nullptr};
    Writer<Ptr32Dst, false> w(nd, 0);
    emit_record(s, o, L, pl, op, w, nullptr);
    return (uint32_t)w.pos();
}

// Pass 1, any record: decode + filter; returns the status word and the output length.
template <class Src>
SSE_HD uint32_t plan_record(const Src& s, uint32_t o, uint32_t L, uint32_t op, int64_t start_seq,
                            Plan& pl, uint32_t* nest, uint32_t& out_len) {
    out_len = 0;
    uint32_t st = ST_OK;
    pl.fast = 0;
    if ((op & OP_NO_FAST) || !parse_fast(s, o, L, op, pl)) st = parse_general(s, o, L, pl, nest);
    if (st != ST_OK) return st;
    if (!(op & OP_BRIDGE) && start_seq > 0 && pl.seq <= start_seq) return ST_FILTERED;   // sse_handler.go:143
    Ptr32Dst nd{nullptr};
    Writer<Ptr32Dst, false> w(nd, 0);
    emit_record(s, o, L, pl, op, w, nullptr);
    out_len = (uint32_t)w.pos();
    return pl.done ? (ST_OK | ST_DONE) : ST_OK;
}

// ------------------------------------------------------------ producer side ----
// json.Marshal(TokenMessage{...}) from columnar fields (llm-stream-proxy/main.go:230-239).  A string whose bytes are
// all "plain" (special_mask) is its own JSON spelling: it is measured by its length and copied word-wise; anything
// else goes through the rune loop.  OP_BRIDGE prepends "chat." + raw id + ".tokens" (main.go:240); OP_RESP wraps channel and
// payload in the RESP2 command go-redis writes for rdb.Publish (main.go:244): *3 $7 publish $<n> <channel> $<m> <payload>.
struct EncPlan { uint32_t cid_n, tok_n, len, chan, json; bool cid_plain, tok_plain, resp; };

template <class Src> SSE_HD bool go_string_is_plain(const Src& s, uint32_t b, uint32_t e) {
    // aligned words covering [b, e): one load per word, the bytes outside the string are forced plain in the first
    // and the last word only
    if (b >= e) return true;
    uint32_t a = b & ~3u;
    const uint32_t last = (e - 1) & ~3u;
    const uint32_t head = ~(0xFFFFFFFFu << ((b & 3u) * 8));               // bytes before b
    const uint32_t tail = ((e & 3u) ? (0xFFFFFFFFu << ((e & 3u) * 8)) : 0u);  // bytes from e on
    uint32_t w = s.ld32a(a);
    w = (w & ~head) | (0x20202020u & head);
    uint32_t bad = 0;
    while (a < last) {
        bad |= special_mask(w);
        a += 4;
        w = s.ld32a(a);
    }
    w = (w & ~tail) | (0x20202020u & tail);
    return (bad | special_mask(w)) == 0;
}
// encode.go appendString(escapeHTML=true) over a RAW Go string (not a JSON literal; producer side,
// llm-stream-proxy/main.go:230-239): bytes that do not start a valid UTF-8 sequence become the six characters \\ufffd, one per byte.

// One thread walks the string: runs of plain bytes go out up to four at a time, every other byte takes
// the rune path for that one rune.
template <class Src, class W> SSE_HD void put_go_string_units(const Src& s, uint32_t b, uint32_t e, W& w, bool go122) {
    w.put('"');
    uint32_t pos = b;
    while (pos < e) {
        uint32_t wd = ld32u(s, pos);
        const uint32_t rem = e - pos;
        if (rem < 4) { const uint32_t keep = 0xFFFFFFFFu >> (32 - 8 * rem); wd = (wd & keep) | (0x01010101u & ~keep); }   // bytes past the end stop the run
        const uint32_t m = special_mask(wd);
        const uint32_t np = m ? first_byte_idx(m) : 4u;
        if (np) { w.put_n(wd, np); pos += np; }
        if (m && pos < e) {
            uint32_t r;
            const uint32_t n = utf8_decode(s, pos, e, r);
            if (r == 0xFFFD && n == 1) { w.put('\\'); w.put('u'); w.put('f'); w.put('f'); w.put('f'); w.put('d'); }
            else put_rune_json(w, r, go122);
            pos += n;
        }
    }
    w.put('"');
}
template <class Src> SSE_HD uint32_t go_string_len(const Src& s, uint32_t b, uint32_t e, bool go122) {
    Ptr32Dst nd{nullptr};
    Writer<Ptr32Dst, false> w(nd, 0);
    put_go_string_units(s, b, e, w, go122);
    return w.pos();
}

// ---- whole message -----------------------------------------------------------------------------------------------------
// Lengths of the two quoted strings come from the caller (n+2 for a plain string, else measured); everything else is closed form.
SSE_HD EncPlan encode_plan_from(uint32_t cid_raw, uint32_t cid_n, bool cid_plain, uint32_t tok_n, bool tok_plain, int64_t seq, bool done,
                                int64_t ts, uint32_t op) {
    EncPlan p;
    p.cid_plain = cid_plain; p.tok_plain = tok_plain; p.cid_n = cid_n; p.tok_n = tok_n;
    p.resp = (op & OP_RESP) != 0;
    p.chan = (op & (OP_BRIDGE | OP_RESP)) ? 12 + cid_raw : 0;
    p.json = 62u + p.cid_n + p.tok_n + dec_len(seq) + dec_len(ts) + (done ? 4u : 5u);
    p.len = p.chan + p.json + (p.resp ? 27u + dec_digits(p.chan) + dec_digits(p.json) : 0u);
    return p;
}
template <class Src>
SSE_HD EncPlan encode_plan(const Src& cid, uint32_t cb, uint32_t ce, const Src& tok, uint32_t tb, uint32_t te, int64_t seq, bool done,
                           int64_t ts, uint32_t op) {
    const bool go122 = (op & OP_GO122) != 0;
    const bool cp = go_string_is_plain(cid, cb, ce), tp = go_string_is_plain(tok, tb, te);
    return encode_plan_from(ce - cb, cp ? ce - cb + 2 : go_string_len(cid, cb, ce, go122), cp,
                            tp ? te - tb + 2 : go_string_len(tok, tb, te, go122), tp, seq, done, ts, op);
}
// View of a writer whose first (possibly shared) word is already out: every put is the cheap no-first-word form.
template <class W> struct NfView {
    W& w;
    SSE_HD void put(uint32_t b) { w.put_n_nf(b, 1); }
    SSE_HD void put_n(uint32_t v, uint32_t n) { w.put_n_nf(v, n); }
    SSE_HD void put32(uint32_t v) { w.put32_nf(v); }
};
template <class Src, class W>
SSE_HD void encode_emit(const Src& cid, uint32_t cb, uint32_t ce, const Src& tok, uint32_t tb, uint32_t te, int64_t seq, bool done,
                        int64_t ts, bool go122, const EncPlan& p, W& w) {
    // a record's first word may share its 32-bit word with the previous record: general put; everything after it goes
    // through the no-first-word forms
    NfView<W> v{w};
    if (p.resp) {
        w.put32(w4("*3\r\n"));
        SSE_PUT_LIT_NF(w, "$7\r\npublish\r\n$");
        put_int(v, (int64_t)p.chan);
        SSE_PUT_LIT_NF(w, "\r\nchat.");
        put_copy_t<true>(cid, cb, ce - cb, w);
        SSE_PUT_LIT_NF(w, ".tokens\r\n$");
        put_int(v, (int64_t)p.json);
        SSE_PUT_LIT_NF(w, "\r\n{\"conversation_id\":");
    } else if (p.chan) {
        w.put32(w4("chat"));
        w.put_n_nf('.', 1);
        put_copy_t<true>(cid, cb, ce - cb, w);
        SSE_PUT_LIT_NF(w, ".tokens{\"conversation_id\":");
    } else {
        w.put32(w4("{\"co"));
        SSE_PUT_LIT_NF(w, "nversation_id\":");
    }
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int k = 0; k < 2; k++) {                        // one copy of the string code for both fields (instruction footprint)
        const Src src = k ? tok : cid;
        const uint32_t sb = k ? tb : cb, se = k ? te : ce;
        if (k ? p.tok_plain : p.cid_plain) { w.put_n_nf('"', 1); put_copy_t<true>(src, sb, se - sb, w); w.put_n_nf('"', 1); }
        else put_go_string_units(src, sb, se, v, go122);
        if (k == 0) SSE_PUT_LIT_NF(w, ",\"token\":"); else SSE_PUT_LIT_NF(w, ",\"sequence\":");
    }
    put_int(v, seq);
    if (done) SSE_PUT_LIT_NF(w, ",\"done\":true,\"timestamp\":"); else SSE_PUT_LIT_NF(w, ",\"done\":false,\"timestamp\":");
    put_int(v, ts);
    if (p.resp) w.put_n_nf(w4("}\r\n\0"), 3); else w.put_n_nf('}', 1);
    w.finish();
}

}  // namespace sse
