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

constexpr uint32_t DFA_CACHE_ENTRIES = 2048;          // u16 transitions of the lowest-numbered (nearest-to-start) states kept on chip
constexpr uint32_t DFA_CACHE_BYTES = 256 + 2 * DFA_CACHE_ENTRIES;

struct DfaDev {
    const uint8_t* cls;        // [256]
    const uint16_t* next;      // [n_states][n_classes], 0xFFFF = dead, bit 15 = target reports a match
    const uint32_t* acc_imm;   // per state
    const uint32_t* acc_del;
    const uint32_t* acc_eot;
    uint32_t n_classes, start0, start1, n_states;
};

struct InspectDev {
    DfaDev fwd;
    DfaDev anchored[INSPECT_MAX_ANCHORED];
    uint32_t n_patterns, mode;
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
struct DfaCache { const uint8_t* cls; const uint16_t* rows; uint32_t n_cached; };

SSE_HD uint32_t dfa_step(const DfaDev& d, const DfaCache& c, uint32_t st, uint32_t byte) {
    const uint32_t k = st * d.n_classes + c.cls[byte];
    return st < c.n_cached ? c.rows[k] : d.next[k];
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

// Fast lane for tokens whose decoded text IS the raw bytes (canonical record, no backslash escape in the token):
// word loads from the tile, class map and hot transition rows on chip, accept masks touched only when the
// transition's match bit is set.  Returns false (nothing reported yet) when the slow routine must take over: a
// variable-length pattern won in per_token mode and its leftmost start has to be searched.
template <class Src>
SSE_HD bool inspect_raw(const InspectDev& t, const DfaCache& c, const Src& s, uint32_t tok_b, uint32_t tok_e, uint32_t state_in, bool eot,
                        InspectResult& r) {
    const bool streaming = t.mode == INSPECT_STREAMING;
    const DfaDev& d = t.fwd;
    uint32_t st = streaming ? state_in : d.start0;
    uint32_t seen = 0, win = 32, win_end = 0;
    const uint32_t n = tok_e - tok_b;
    for (uint32_t i = 0; i < n; i += 4) {
        uint32_t w = ld32u(s, tok_b + i);
        const uint32_t m = n - i < 4 ? n - i : 4;
        for (uint32_t k = 0; k < m; k++, w >>= 8) {
            const uint32_t nx = dfa_step(d, c, st, w & 0xFF);
            st = nx & 0x7FFF;
            if (nx & 0x8000) {
                const uint32_t del = d.acc_del[st], imm = d.acc_imm[st], idx = i + k;
                if (streaming) {
                    if (win == 32 && del) { win = lowest_bit(del); win_end = idx; }
                    if (win == 32 && imm) { win = lowest_bit(imm); win_end = idx + 1; }
                } else {
                    uint32_t fresh = del & ~seen; seen |= del;
                    if (fresh && lowest_bit(fresh) < win) { win = lowest_bit(fresh); win_end = idx; }
                    fresh = imm & ~seen; seen |= imm;
                    if (fresh && lowest_bit(fresh) < win) { win = lowest_bit(fresh); win_end = idx + 1; }
                }
            }
        }
    }
    if (eot) {
        const uint32_t e = d.acc_eot[st];
        if (streaming) { if (win == 32 && e) { win = lowest_bit(e); win_end = n; } }
        else { const uint32_t fresh = e & ~seen; if (fresh && lowest_bit(fresh) < win) { win = lowest_bit(fresh); win_end = n; } }
    }
    r.state = (streaming && eot) ? d.start0 : st;
    r.verdict = 0; r.begin = 0; r.end = 0;
    if (win < 32) {
        if (!t.fixed_len[win] && !streaming) return false;            // leftmost start needed: slow routine
        r.verdict = t.action[win] | (win << 8);
        r.end = (int32_t)win_end;
        if (t.fixed_len[win]) r.begin = (int32_t)win_end - (int32_t)t.fixed_len[win];
        else { r.begin = MATCH_BEGIN_UNKNOWN; r.verdict |= VERDICT_BEGIN_UNKNOWN; }
    }
    return true;
}

}  // namespace sse
