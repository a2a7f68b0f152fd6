// sse_inspect.cuh -- device-side walk of the inspection DFAs over one record's decoded token text.
//
// Stands in for the commented-out hook of the adapter (src/sse-adapter/sse_handler.go:193-208:
// "result := callInspector(endpoint, token); drop / redact") and for the snippet inspectors
// (docs/security-inspection-patterns.md:205-243; src/spin-functions/nats-subscriber/src/lib.rs:57-102).
// The reference never runs any of it, so the semantics are this project's:
// tables from inspect_compile.hpp; one thread walks its own record's token (a few bytes), table
// lookups hit L1/L2 (the whole forward DFA of Set A+B is 29 KB).
//
// __host__ __device__ like sse_record.cuh: tests/hostcheck runs the very same walk on the CPU.
#pragma once
#include <stdint.h>
#include "sse_record.cuh"

namespace sse {

constexpr uint32_t INSPECT_MAX_ANCHORED = 8;
constexpr uint32_t INSPECT_BEGIN_BUF = 256;          // decoded bytes kept for the leftmost-start search
constexpr int32_t  MATCH_BEGIN_UNKNOWN = (int32_t)0x80000000;
constexpr uint32_t VERDICT_BEGIN_UNKNOWN = 1u << 16;
enum : uint32_t { INSPECT_PER_TOKEN = 0, INSPECT_STREAMING = 1 };
enum : uint32_t { ACT_ALLOW = 0, ACT_REDACT = 1, ACT_DROP = 2, ACT_ALERT = 3 };

// On-chip part of the forward DFA inside the frame kernel: byte-class map (256 B), a bitmap of the states whose
// end-of-text accept mask is not empty (128 B, DFAs of up to 1024 states), and the transition rows of the
// lowest-numbered states.  Subset construction numbers states breadth-first from the start state, so ordinary text
// lives in the first few dozen (measured on the BASELINE token stream: 99.7 % of all steps in states < 32, 99.99 % in
// states < 48); the other rows are read through L1/L2.
constexpr uint32_t DFA_CACHE_ENTRIES = 992;           // u16 transitions kept on chip (with five tiles per SM the shared memory is full to within 8 bytes)
constexpr uint32_t DFA_EOT_WORDS = 32;
constexpr uint32_t DFA_CACHE_BYTES = 256 + 4 * DFA_EOT_WORDS + 2 * DFA_CACHE_ENTRIES;

// Streaming mode: the frame kernel leaves one 32-byte token entry (one DRAM sector) per record for the per-conversation
// walk (sse_stream.cu): word 0 = offset of the token literal's content in the payload arena, word 1 = its raw length
// (24 bits, saturating) | flags << 24, words 2..7 = the first 24 raw bytes of the literal -- every BASELINE token, so the
// walk needs the arena only for longer ones.
enum : uint32_t { ENTRY_OK = 1u, ENTRY_DECODE = 2u, ENTRY_DONE = 4u, ENTRY_HUGE = 8u,
                  ENTRY_CANON = 16u };   // with ENTRY_DECODE: the only escapes are the ones json.Marshal itself writes
constexpr uint32_t ENTRY_LEN_MASK = 0xFFFFFFu;
constexpr uint32_t ENTRY_WORDS = 8, ENTRY_INLINE = 24;
constexpr uint32_t STREAM_BUF_WORDS = 16;            // the walk fetches tokens of up to 56 raw bytes from the arena in one round of word loads

struct DfaDev {
    const uint8_t* cls;        // [256]
    const uint16_t* next;      // [n_states][n_classes], 0xFFFF = dead, bit 15 = target reports a match
    const uint32_t* acc_imm;   // per state
    const uint32_t* acc_del;
    const uint32_t* acc_eot;
    uint32_t n_classes, start0, start1, n_states;
    const uint8_t* dist;       // per state: fewest further bytes before anything can be reported (saturates at 15); may be null
};

struct InspectDev {
    DfaDev fwd;
    DfaDev anchored[INSPECT_MAX_ANCHORED];
    uint32_t n_patterns, mode;
    uint32_t min_len;          // shortest text any pattern can match: a shorter token cannot hold a match of its own
    uint32_t action[32];
    uint32_t fixed_len[32];    // 0 = variable length
    int32_t anchored_of[32];   // index into anchored[] or -1
};

SSE_HD bool word_byte(uint32_t b) { return (b - '0' <= 9u) || ((b | 0x20) - 'a' <= 25u) || b == '_'; }

SSE_HD uint32_t lowest_bit(uint32_t m) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)__ffs((int)m) - 1;
#else
    return (uint32_t)__builtin_ctz(m);
#endif
}

// Calls f(byte) for every decoded UTF-8 byte of the JSON string content [b,e) (decode.go unquoteBytes).
template <class Src, class F> SSE_HD void for_each_decoded_byte(const Src& s, uint32_t b, uint32_t e, F&& f) {
    while (b < e) {
        uint32_t r = next_rune(s, b, e);
        if (r < 0x80) f(r);
        else if (r < 0x800) { f(0xC0 | (r >> 6)); f(0x80 | (r & 0x3F)); }
        else if (r < 0x10000) { f(0xE0 | (r >> 12)); f(0x80 | ((r >> 6) & 0x3F)); f(0x80 | (r & 0x3F)); }
        else { f(0xF0 | (r >> 18)); f(0x80 | ((r >> 12) & 0x3F)); f(0x80 | ((r >> 6) & 0x3F)); f(0x80 | (r & 0x3F)); }
    }
}

struct InspectResult { uint32_t verdict; int32_t begin, end; uint32_t state; };

// On-chip copy of the byte-class map and of the transition rows of the first `n_cached` states of the forward
// DFA (subset construction numbers states breadth-first from the start state, so these are the states ordinary
// text lives in).  On the host it simply points at the full tables.
// next/n_classes/n_states/min_len repeat fields of the (global-memory) descriptor so that the walk finds them in registers.
// "Packed" on-chip rows (per_token mode, DFAs of up to 2048 states) carry, in bits 11..14 of each transition, the target
// state's `dist`: the walk learns from the transition it has just loaded whether the rest of the token is long enough
// for anything to be reported, and stops when it is not.
constexpr uint32_t DFA_PACK_MAX_STATES = 2048, DFA_PACK_SHIFT = 11;
struct DfaCache {
    const uint8_t* cls; const uint16_t* rows; uint32_t n_cached; const uint32_t* eot_bits;
    const uint16_t* next; uint32_t n_classes, n_states, start0, min_len;
    uint32_t cmask, nmask;     // on-chip rows: state bits, dist bits (0 = rows are not packed)
};
SSE_HD bool dfa_rows_packable(const DfaDev& d) { return d.dist != nullptr && d.n_states <= DFA_PACK_MAX_STATES; }
SSE_HD uint16_t dfa_pack_row(const DfaDev& d, uint16_t nx) { return (uint16_t)(nx | ((uint32_t)d.dist[nx & 0x7FFFu] << DFA_PACK_SHIFT)); }
SSE_HD DfaCache make_dfa_cache(const uint16_t* next, uint32_t n_classes, uint32_t n_states, uint32_t start0, uint32_t min_len, const uint8_t* cls,
                               const uint16_t* rows, uint32_t n_cached, const uint32_t* eot_bits, bool packed) {
    return DfaCache{cls, rows, n_cached, eot_bits, next, n_classes, n_states, start0, min_len,
                    packed ? (1u << DFA_PACK_SHIFT) - 1u : 0x7FFFu, packed ? 15u : 0u};
}
SSE_HD DfaCache make_dfa_cache(const InspectDev& t, const uint8_t* cls, const uint16_t* rows, uint32_t n_cached, const uint32_t* eot_bits,
                               bool packed = false) {
    return DfaCache{cls, rows, n_cached, eot_bits, t.fwd.next, t.fwd.n_classes, t.fwd.n_states, t.fwd.start0, t.min_len,
                    packed ? (1u << DFA_PACK_SHIFT) - 1u : 0x7FFFu, packed ? 15u : 0u};
}

// State of one walk.  per_token: the lowest-numbered pattern wins, with its first (= smallest) end; streaming: the
// first match to become decidable wins, ties to the lowest id.
struct Walk { uint32_t st, idx, seen, win, win_end, need; };   // need: bytes that must still follow for anything to be reported (packed rows)

template <bool STREAMING> SSE_HD void walk_event(Walk& w, uint32_t mask, uint32_t e) {
    if (STREAMING) {
        if (w.win == 32 && mask) { w.win = lowest_bit(mask); w.win_end = e; }
    } else {
        const uint32_t fresh = mask & ~w.seen;
        w.seen |= mask;
        if (fresh) { const uint32_t k = lowest_bit(fresh); if (k < w.win) { w.win = k; w.win_end = e; } }
    }
}
template <bool STREAMING> SSE_HD void walk_byte(const DfaDev& d, const DfaCache& c, Walk& w, uint32_t byte) {
    const uint32_t k = w.st * c.n_classes + c.cls[byte];
    uint32_t nx;
    if (STREAMING) {                                               // rows are never packed here, and a carried walk never stops early
        nx = w.st < c.n_cached ? c.rows[k] : c.next[k];
        w.st = nx & 0x7FFF;
    } else if (w.st < c.n_cached) { nx = c.rows[k]; w.need = (nx >> DFA_PACK_SHIFT) & c.nmask; w.st = nx & c.cmask; }
    else { nx = c.next[k]; w.need = 0; w.st = nx & 0x7FFF; }       // the unanchored forward DFA has no dead state
    if (nx & 0x8000) {                                             // the target state reports a match: rare
        walk_event<STREAMING>(w, d.acc_del[w.st], w.idx);
        walk_event<STREAMING>(w, d.acc_imm[w.st], w.idx + 1);
    }
    w.idx++;
}
template <bool STREAMING> SSE_HD void walk_eot(const DfaDev& d, const DfaCache& c, Walk& w) {
    if (c.eot_bits && c.n_states <= 32 * DFA_EOT_WORDS && !((c.eot_bits[w.st >> 5] >> (w.st & 31)) & 1u)) return;
    walk_event<STREAMING>(w, d.acc_eot[w.st], w.idx);
}

// Leftmost start b of a match of the (variable-length) pattern behind dfa `a` that ends exactly at
// `end`, over the decoded bytes buf[0,len) (per_token mode: the text ends at len).  -1: none found.
SSE_HD int32_t leftmost_begin(const DfaDev& a, const uint8_t* buf, uint32_t len, uint32_t end) {
    for (uint32_t b = 0; b < end; b++) {
        uint32_t st = (b > 0 && word_byte(buf[b - 1])) ? a.start1 : a.start0;
        bool hit = false;
        uint32_t i = b;
        for (; i <= end; i++) {                       // one byte past `end` is the look-ahead of a trailing \b
            if (i == len) { if (i == end && a.acc_eot[st]) hit = true; break; }
            uint32_t nx = a.next[st * a.n_classes + a.cls[buf[i]]];
            if (nx == 0xFFFF) break;
            st = nx & 0x7FFF;
            if (i == end) { if (a.acc_del[st]) hit = true; break; }
            if (i + 1 == end && a.acc_imm[st]) { hit = true; break; }
        }
        if (hit) return (int32_t)b;
    }
    return -1;
}

// Walks the forward DFA over the token text.  `state_in` is the carried state (streaming) or ignored
// (per_token: the walk starts from the start state).  `eot`: the text ends with this token.
template <class Src>
SSE_HD_NOINLINE InspectResult inspect_token(const InspectDev& t, const Src& s, uint32_t tok_b, uint32_t tok_e, uint32_t state_in, bool eot) {
    const bool streaming = t.mode == INSPECT_STREAMING;
    const DfaDev& d = t.fwd;
    uint32_t st = streaming ? state_in : d.start0;
    uint32_t idx = 0, seen = 0, win = 32, win_end = 0;
    bool found = false;
    uint8_t buf[INSPECT_BEGIN_BUF];
    auto event = [&](uint32_t mask, uint32_t e) {
        if (streaming) {                              // first decidable match wins; ties: lowest id
            if (!found && mask) { found = true; win = lowest_bit(mask); win_end = e; }
        } else {                                      // lowest id wins; its first (= smallest) end
            uint32_t fresh = mask & ~seen;
            seen |= mask;
            if (fresh) { uint32_t k = lowest_bit(fresh); if (k < win) { win = k; win_end = e; } }
        }
    };
    for_each_decoded_byte(s, tok_b, tok_e, [&](uint32_t c) {
        if (idx < INSPECT_BEGIN_BUF) buf[idx] = (uint8_t)c;
        const uint32_t nx = d.next[st * d.n_classes + d.cls[c]];     // the unanchored forward DFA has no dead state
        st = nx & 0x7FFF;
        if (nx & 0x8000) { event(d.acc_del[st], idx); event(d.acc_imm[st], idx + 1); }
        idx++;
    });
    if (eot) event(d.acc_eot[st], idx);
    InspectResult r;
    r.state = (streaming && eot) ? d.start0 : st;     // a finished conversation frees its slot in the start state
    r.verdict = 0; r.begin = 0; r.end = 0;
    if (win < 32) {
        r.verdict = t.action[win] | (win << 8);
        r.end = (int32_t)win_end;
        if (t.fixed_len[win]) r.begin = (int32_t)win_end - (int32_t)t.fixed_len[win];
        else {
            int32_t b = -1;
            if (!streaming && idx <= INSPECT_BEGIN_BUF && t.anchored_of[win] >= 0)
                b = leftmost_begin(t.anchored[t.anchored_of[win]], buf, idx, win_end);
            if (b >= 0) r.begin = b;
            else { r.begin = MATCH_BEGIN_UNKNOWN; r.verdict |= VERDICT_BEGIN_UNKNOWN; }
        }
    }
    return r;
}

// Walk over a token literal whose only escapes are the ones json.Marshal itself writes: two-character forms and
// u00XX (one byte each), u2028 / u2029 (three bytes); raw UTF-8 passes through.
template <bool STREAMING, class Src>
SSE_HD void walk_canon_escaped(const DfaDev& d, const DfaCache& c, Walk& w, const Src& s, uint32_t tok_b, uint32_t tok_e) {
    for (uint32_t pos = tok_b; pos < tok_e;) {
        uint32_t ch = s.ld8(pos++);
        if (ch == '\\') {
            const uint32_t e = s.ld8(pos++);
            if (e == 'u') {
                const uint32_t r = (uint32_t)hex4(s, pos);
                pos += 4;
                if (r < 0x80) ch = r;
                else { walk_byte<STREAMING>(d, c, w, 0xE0 | (r >> 12)); walk_byte<STREAMING>(d, c, w, 0x80 | ((r >> 6) & 0x3F)); ch = 0x80 | (r & 0x3F); }
            } else ch = e == 'n' ? 10u : e == 'r' ? 13u : e == 't' ? 9u : e == 'b' ? 8u : e == 'f' ? 12u : e;
        }
        walk_byte<STREAMING>(d, c, w, ch);
    }
}

// per_token mode, canonical record (what json.Marshal wrote): the token literal's content [tok_b, tok_e) sits in the
// tile.  Without escapes its decoded text IS those bytes (word loads, funnel-shifted); with escapes (1.5 % of the
// BASELINE tokens) the same walk runs over the decoded bytes.  Class map, end-of-text bitmap and hot rows are on
// chip; the accept masks are touched only when a transition's match bit is set.  Returns false (nothing reported)
// when the general routine must take over: a variable-length pattern won and its leftmost start has to be searched.
template <class Src>
SSE_HD bool inspect_canon_token(const InspectDev& t, const DfaCache& c, const Src& s, uint32_t tok_b, uint32_t tok_e, bool has_esc,
                                InspectResult& r) {
    const DfaDev& d = t.fwd;
    r.verdict = 0; r.begin = 0; r.end = 0; r.state = 0;
    const uint32_t n = tok_e - tok_b;
    if (n < c.min_len) return true;                    // decoded text is never longer than its literal: nothing fits
    Walk w{c.start0, 0, 0, 32, 0};
    if (!has_esc) {
        const uint32_t a = tok_b & ~3u, shb = (tok_b & 3u) * 8;
        uint32_t lo = s.ld32a(a);
        for (uint32_t i = 0; i < n; i += 4) {
            const uint32_t hi = s.ld32a(a + i + 4);
            uint32_t word = funnel_r(lo, hi, shb);
            lo = hi;
            const uint32_t m = n - i < 4 ? n - i : 4;
            for (uint32_t k = 0; k < m; k++, word >>= 8) {
                walk_byte<false>(d, c, w, word & 0xFF);
                if (n - w.idx < w.need) { i = n; break; }          // too few bytes left for any pattern to complete
            }
        }
    } else {
        walk_canon_escaped<false>(d, c, w, s, tok_b, tok_e);
    }
    walk_eot<false>(d, c, w);
    if (w.win < 32) {
        if (!t.fixed_len[w.win]) return false;
        r.verdict = t.action[w.win] | (w.win << 8);
        r.end = (int32_t)w.win_end;
        r.begin = (int32_t)w.win_end - (int32_t)t.fixed_len[w.win];
    }
    return true;
}

// Token entry left by the frame kernel for the streaming walk (layout: ENTRY_* above).
template <class Src>
SSE_HD void make_token_entry(const Src& s, uint32_t tok_b, uint32_t raw_len, uint32_t abs_tok_b, uint32_t flags, uint32_t out[ENTRY_WORDS]) {
    out[0] = abs_tok_b;
    out[1] = (raw_len < ENTRY_LEN_MASK ? raw_len : ENTRY_LEN_MASK) | ((flags | (raw_len >= ENTRY_LEN_MASK ? ENTRY_HUGE : 0u)) << 24);
    // the literal's first 24 bytes (whatever follows a shorter token rides along: the walk stops at raw_len)
    const uint32_t a = tok_b & ~3u, shb = (tok_b & 3u) * 8;
    uint32_t lo = s.ld32a(a);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (uint32_t k = 0; k < ENTRY_INLINE / 4; k++) {
        uint32_t w = 0;
        if ((flags & ENTRY_OK) && 4 * k < raw_len) { const uint32_t hi = s.ld32a(a + 4 * k + 4); w = funnel_r(lo, hi, shb); lo = hi; }
        out[2 + k] = w;
    }
}

// streaming mode: one token of a conversation, given by its entry e[0..7]; `arena` reads the payload arena at
// absolute offsets.  state_in is the conversation's carried state; the returned state is what the next token starts from.
template <class Src>
SSE_HD InspectResult stream_walk(const InspectDev& t, const DfaCache& c, const Src& arena, const uint32_t e[ENTRY_WORDS], uint32_t state_in) {
    const DfaDev& d = t.fwd;
    const uint32_t flags = e[1] >> 24;
    const uint32_t n = e[1] & ENTRY_LEN_MASK;
    const bool eot = (flags & ENTRY_DONE) != 0;
    Walk w{state_in, 0, 0, 32, 0};
    if (!(flags & ENTRY_DECODE) && n <= ENTRY_INLINE) {
        // the common case: the decoded text is the raw bytes and they are all in the entry (registers)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (uint32_t k = 0; k < ENTRY_INLINE / 4; k++) {
            if (4 * k < n) {
                uint32_t word = e[2 + k];
                const uint32_t m = n - 4 * k < 4 ? n - 4 * k : 4;
                for (uint32_t b = 0; b < m; b++, word >>= 8) walk_byte<true>(d, c, w, word & 0xFF);
            }
        }
    } else if (!(flags & ENTRY_HUGE) && n <= 4 * STREAM_BUF_WORDS - 8) {
        // escapes to resolve, or a longer token: raw bytes into a small local buffer (from the entry, or from the arena as
        // independent word loads: one memory round trip instead of one per byte), the general decoder walks that
        uint32_t buf[STREAM_BUF_WORDS];
        uint32_t sh = 0;
        if (n <= ENTRY_INLINE) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (uint32_t k = 0; k < STREAM_BUF_WORDS; k++) buf[k] = k < ENTRY_INLINE / 4 ? e[2 + k] : 0u;
        } else {
            const uint32_t a = e[0] & ~3u, nwords = ((e[0] & 3u) + n + 3) >> 2;
            sh = e[0] & 3u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (uint32_t k = 0; k < STREAM_BUF_WORDS; k++) buf[k] = k < nwords ? arena.ld32a(a + 4 * k) : 0u;
        }
        const PtrSrc local{reinterpret_cast<const uint8_t*>(buf)};
        if (!(flags & ENTRY_DECODE)) for (uint32_t k = 0; k < n; k++) walk_byte<true>(d, c, w, local.ld8(sh + k));
        else if (flags & ENTRY_CANON) walk_canon_escaped<true>(d, c, w, local, sh, sh + n);
        else for_each_decoded_byte(local, sh, sh + n, [&](uint32_t b) { walk_byte<true>(d, c, w, b); });
    } else {
        uint32_t end = e[0] + n;
        if (flags & ENTRY_HUGE) end = scan_string(arena, e[0] - 1, 0xFFFFFFFFu) - 1;     // valid literal: the closing quote exists
        for_each_decoded_byte(arena, e[0], end, [&](uint32_t b) { walk_byte<true>(d, c, w, b); });
    }
    if (eot) walk_eot<true>(d, c, w);
    InspectResult r;
    r.state = eot ? c.start0 : w.st;                   // a finished conversation frees its slot in the start state
    r.verdict = 0; r.begin = 0; r.end = 0;
    if (w.win < 32) {
        r.verdict = t.action[w.win] | (w.win << 8);
        r.end = (int32_t)w.win_end;
        if (t.fixed_len[w.win]) r.begin = (int32_t)w.win_end - (int32_t)t.fixed_len[w.win];
        else { r.begin = MATCH_BEGIN_UNKNOWN; r.verdict |= VERDICT_BEGIN_UNKNOWN; }
    }
    return r;
}

}  // namespace sse
