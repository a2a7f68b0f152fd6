// sse_interest.cuh -- device-side lookup of a record's subject in the subscriber/interest table.
//
// In the reference the "interest table" is not in the repo: every HTTP connection calls
// nc.Subscribe("chat.<id>.tokens", cb) (src/sse-adapter/sse_handler.go:133-135, :314-316) and the match of
// a published subject (src/redis-nats-bridge/main.go:160-164) against those subscriptions happens inside
// nats.go v1.31.0 / nats-server 2.10 (third-party, not vendored).  Semantics restated here are the NATS
// subject rules: subjects are '.'-separated tokens; a subscription matches literally, '*' matches exactly
// one token, '>' (last token only) matches one or more trailing tokens; empty tokens never match anything.
// Only literal chat.<id>.tokens and the wildcards chat.*.tokens / chat.*.control occur in the reference
// (docs/security-inspection-patterns.md:110,268; kubernetes/base/nats-core/core-cluster.yaml:234,248).
//
// Layout: literal subscriptions of the form chat.<id>.tokens live in an open-addressing hash table keyed
// by the DECODED id bytes (the bridge builds the subject from the decoded id, main.go:160): 64-byte slots
// that carry tag, value and the key itself (ids of up to 48 bytes; a UUID is 36), so that a lookup is ONE
// 64-byte access and one L2 round trip instead of slot -> key arena -> compare; linear probing, <= 50 % load,
// a word-wise multiplicative hash, keys as zero-padded 32-bit words so that hash and compare are word-wise;
// values are spans of subscriber ids in an append-only arena.  The few wildcard subscriptions are a side
// list; the common shapes are pre-classified by the host so that an id without dots never needs the
// token-by-token matcher.  At 100 k subjects the table is 17 MB: far beyond one SM's shared memory (SURVEY.md
// H3), so it is read through L2 (126 MB holds it entirely).
#pragma once
#include <stdint.h>
#include "sse_inspect.cuh"

namespace sse {

constexpr uint32_t INTEREST_INLINE_KEY = 48;   // bytes of id a slot holds itself
struct alignas(64) InterestSlot {
    uint32_t hash;               // 0 = empty (never a valid hash)
    uint32_t key_len;            // bytes; > INTEREST_INLINE_KEY: key[0] is the first word of the key in keys[]
    uint32_t list_off, list_cnt; // span in the subscriber arena; count 0 = every subscriber has left (slot kept until the next rebuild)
    uint32_t key[INTEREST_INLINE_KEY / 4];   // the id, zero-padded little-endian words
};

constexpr uint32_t INTEREST_MAX_WILD = 32;
constexpr uint32_t INTEREST_MAX_ID = 1024;      // longer ids cannot be subscribed to literally

struct InterestDev {
    const InterestSlot* slots;
    const uint32_t* keys;        // zero-padded little-endian words
    uint32_t mask;               // capacity - 1 (capacity is a power of two)
    uint32_t n_wild;
    uint32_t wild_single;        // wildcards that match every valid single-token id   (chat.*.tokens, chat.>, >, ...)
    uint32_t wild_generic;       // wildcards that need the token-by-token matcher for single-token ids too
    const uint8_t* wild_bytes;   // wildcard subscription subjects, back to back
    uint32_t wild_off[INTEREST_MAX_WILD + 1];
};

struct InterestHit { uint32_t list_off, list_cnt, wild_mask; };

SSE_HD uint32_t key_hash_init(uint32_t len) { return 0x9E3779B9u ^ len; }
SSE_HD uint32_t key_hash_step(uint32_t h, uint32_t w) { h = (h ^ w) * 0x85EBCA6Bu; return (h << 13) | (h >> 19); }
SSE_HD uint32_t key_hash_fin(uint32_t h) { h ^= h >> 16; h *= 0xC2B2AE35u; h ^= h >> 13; return h ? h : 1u; }

// Probe for the key given by word(k), k < ceil(len/4) (last word zero-padded).
template <class WordFn>
SSE_HD void lookup_literal(const InterestDev& t, WordFn&& word, uint32_t len, InterestHit& h) {
    if (!t.slots) return;
    const uint32_t nw = (len + 3) >> 2;
    uint32_t hash = key_hash_init(len);
    for (uint32_t k = 0; k < nw; k++) hash = key_hash_step(hash, word(k));
    hash = key_hash_fin(hash);
    for (uint32_t p = hash & t.mask;; p = (p + 1) & t.mask) {
        const InterestSlot& e = t.slots[p];
        const uint32_t eh = e.hash;
        if (eh == 0) return;
        if (eh == hash && e.key_len == len) {
            const uint32_t* key = len <= INTEREST_INLINE_KEY ? e.key : t.keys + e.key[0];
            bool same = true;
            for (uint32_t k = 0; k < nw; k++) same = same && key[k] == word(k);
            if (same) { h.list_off = e.list_off; h.list_cnt = e.list_cnt; return; }
        }
    }
}

// NATS subject match of pattern pat[0,pn) against the subject given by byte accessor sub(i), length sn.
template <class SubFn>
SSE_HD bool nats_match(const uint8_t* pat, uint32_t pn, SubFn&& sub, uint32_t sn) {
    uint32_t pi = 0, si = 0;
    for (;;) {
        uint32_t pe = pi;
        while (pe < pn && pat[pe] != '.') pe++;
        uint32_t se = si;
        while (se < sn && sub(se) != '.') se++;
        const bool p_last = pe == pn, s_last = se == sn;
        if (se == si) return false;                                  // empty subject token
        if (pe - pi == 1 && pat[pi] == '>') return p_last;           // '>' swallows the rest (it must be the last pattern token)
        if (!(pe - pi == 1 && pat[pi] == '*')) {
            if (pe - pi != se - si) return false;
            for (uint32_t k = 0; k < pe - pi; k++) if (pat[pi + k] != sub(si + k)) return false;
        }
        if (p_last || s_last) return p_last && s_last;
        pi = pe + 1; si = se + 1;
    }
}

// General case (out of line in the kernel): ids with escapes, with dots, longer than the fast lane handles, or
// wildcard shapes that need the matcher.  Decodes the id into a local buffer first.
template <class Src>
SSE_HD_NOINLINE InterestHit interest_lookup(const InterestDev& t, const Src& s, uint32_t cid_b, uint32_t cid_e) {
    InterestHit h{0, 0, 0};
    uint32_t idw[INTEREST_MAX_ID / 4];                               // decoded id, zero-padded words
    uint8_t* id = reinterpret_cast<uint8_t*>(idw);
    uint32_t len = 0, prev = '.';
    bool empty_token = false;                                        // "chat..tokens", "chat.a..b.tokens": not a valid NATS subject
    for_each_decoded_byte(s, cid_b, cid_e, [&](uint32_t c) {
        if (len < INTEREST_MAX_ID) id[len] = (uint8_t)c;
        len++;
        if (c == '.' && prev == '.') empty_token = true;
        prev = c;
    });
    if (prev == '.' || empty_token) return h;                        // covers the empty id too
    const bool fits = len <= INTEREST_MAX_ID;
    if (fits) {
        for (uint32_t k = len; k < ((len + 3) & ~3u); k++) id[k] = 0;
        lookup_literal(t, [&](uint32_t k) { return idw[k]; }, len, h);
    }
    if (t.n_wild) {
        const uint32_t sn = 5 + len + 7;
        auto sub = [&](uint32_t i) -> uint32_t {
            if (i < 5) return (uint32_t)(uint8_t)"chat."[i];
            if (i < 5 + len) {
                const uint32_t j = i - 5;
                if (j < INTEREST_MAX_ID) return id[j];
                uint32_t q = 0, r = 0;                                 // rare: re-decode up to byte j
                for_each_decoded_byte(s, cid_b, cid_e, [&](uint32_t c) { if (q == j) r = c; q++; });
                return r;
            }
            return (uint32_t)(uint8_t)".tokens"[i - 5 - len];
        };
        for (uint32_t w = 0; w < t.n_wild; w++)
            if (nats_match(t.wild_bytes + t.wild_off[w], t.wild_off[w + 1] - t.wild_off[w], sub, sn)) h.wild_mask |= 1u << w;
    }
    return h;
}

// Fast lane: the id's decoded text IS its raw bytes (canonical record, no backslash escape in the id), it fits a
// slot, it has no dot (one NATS token) and no wildcard needs the general matcher.  Two steps so that the table's
// latency hides behind the rest of the parse:
//   interest_prepare  (called from parse_fast's hook as soon as the id has been scanned) one pass over the id's
//                     words -- aligned loads from the tile + a funnel shift each -- yields the dot test and the hash,
//                     and the first 16 bytes (tag, length, value) of the first probed slot are requested;
//   interest_finish   (after the parse) tag and length decide; on a match the slot's key words (same 64-byte slot)
//                     are compared with the id's words, read again from the tile.
// The descriptor fields the fast lane needs are held in registers (InterestFast), not re-read from the descriptor.
struct InterestFast { const InterestSlot* slots; uint32_t mask, wild_single, wild_generic; };
SSE_HD InterestFast make_interest_fast(const InterestDev& t) { return InterestFast{t.slots, t.mask, t.wild_single, t.wild_generic}; }
struct InterestPre { uint32_t hash, q0, q1, q2, q3; };     // hash 0 (never a key's hash): the general routine must take over

SSE_HD void slot_head(const InterestSlot* e, uint32_t& q0, uint32_t& q1, uint32_t& q2, uint32_t& q3) {
#if defined(__CUDA_ARCH__)
    const uint4 q = *reinterpret_cast<const uint4*>(e);
    q0 = q.x; q1 = q.y; q2 = q.z; q3 = q.w;
    // the slot's second 32-byte sector (key words 4..11) is pulled next to the SM now: the compare then hits L1
    asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const uint8_t*>(e) + 32));
#else
    q0 = e->hash; q1 = e->key_len; q2 = e->list_off; q3 = e->list_cnt;
#endif
}

template <class Src>
SSE_HD void interest_prepare(const InterestFast& t, const Src& s, uint32_t cid_b, uint32_t cid_e, bool plain, InterestPre& pre) {
    const uint32_t len = cid_e - cid_b;
    pre.hash = 0; pre.q0 = pre.q1 = pre.q2 = pre.q3 = 0;
    if (!plain || len == 0 || len > INTEREST_INLINE_KEY || t.wild_generic) return;
    const uint32_t a = cid_b & ~3u, shb = (cid_b & 3u) * 8, nw = (len + 3) >> 2;
    uint32_t lo = s.ld32a(a), dots = 0, hash = key_hash_init(len);
    for (uint32_t k = 0; k < nw; k++) {
        const uint32_t hi = s.ld32a(a + 4 * k + 4);
        uint32_t w = funnel_r(lo, hi, shb);
        lo = hi;
        if (k + 1 == nw) w &= 0xFFFFFFFFu >> (8 * (4 * nw - len));   // zero padding of the last word
        const uint32_t x = w ^ 0x2E2E2E2Eu;                          // a '.' byte becomes zero (the zero padding does not)
        dots |= (x - 0x01010101u) & ~x & 0x80808080u;
        hash = key_hash_step(hash, w);
    }
    if (dots) return;
    pre.hash = key_hash_fin(hash);
    if (t.slots) slot_head(t.slots + (pre.hash & t.mask), pre.q0, pre.q1, pre.q2, pre.q3);
}

template <class Src>
SSE_HD void interest_finish(const InterestFast& t, const Src& s, uint32_t cid_b, uint32_t cid_e, const InterestPre& pre, InterestHit& h) {
    const uint32_t len = cid_e - cid_b, nw = (len + 3) >> 2;
    h.list_off = 0; h.list_cnt = 0; h.wild_mask = t.wild_single;
    if (!t.slots) return;
    uint32_t q0 = pre.q0, q1 = pre.q1, q2 = pre.q2, q3 = pre.q3;
    for (uint32_t p = pre.hash & t.mask;;) {
        if (q0 == 0) return;                                         // empty slot: nobody subscribed to this id
        if (q0 == pre.hash && q1 == len) {
            // the key words sit in the slot itself (second sector prefetched by slot_head): word-wise against the id in the tile
            const uint32_t a = cid_b & ~3u, shb = (cid_b & 3u) * 8;
            uint32_t lo = s.ld32a(a), diff = 0;
#if defined(__CUDA_ARCH__)
            // all twelve key words requested at once (three 16-byte loads), then compared from registers: a loop of
            // dependent 4-byte loads left one L1 round trip per word exposed
            const uint4* kp = reinterpret_cast<const uint4*>(t.slots[p].key);
            const uint4 ka = kp[0], kb = kp[1], kc = kp[2];
            const uint32_t kw[INTEREST_INLINE_KEY / 4] = {ka.x, ka.y, ka.z, ka.w, kb.x, kb.y, kb.z, kb.w, kc.x, kc.y, kc.z, kc.w};
            const uint32_t rem = len - 4 * (nw - 1);                 // 1..4 bytes in the last word
#pragma unroll
            for (uint32_t k = 0; k < INTEREST_INLINE_KEY / 4; k++) {
                if (k < nw) {                                        // (words past the id are not loaded)
                    const uint32_t hi = s.ld32a(a + 4 * k + 4);
                    uint32_t w = funnel_r(lo, hi, shb);
                    if (k + 1 == nw) w &= 0xFFFFFFFFu >> (8 * (4 - rem));
                    diff |= w ^ kw[k];
                    lo = hi;
                }
            }
#else
            const uint32_t* key = t.slots[p].key;
            uint32_t k = 0;
            for (; k + 1 < nw; k++) {
                const uint32_t hi = s.ld32a(a + 4 * k + 4);
                diff |= funnel_r(lo, hi, shb) ^ key[k];
                lo = hi;
            }
            const uint32_t rem = len - 4 * k;                        // 1..4 bytes in the last word
            diff |= (funnel_r(lo, s.ld32a(a + 4 * k + 4), shb) & (0xFFFFFFFFu >> (8 * (4 - rem)))) ^ key[k];
#endif
            if (diff == 0) { h.list_off = q2; h.list_cnt = q3; return; }
        }
        p = (p + 1) & t.mask;
        slot_head(t.slots + p, q0, q1, q2, q3);
    }
}

}  // namespace sse
