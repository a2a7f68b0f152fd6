// sse_interest.cuh -- device-side lookup of a record's subject in the subscriber/interest table.
//
// In the reference the "interest table" is not in the repo: every HTTP connection calls
// nc.Subscribe("chat.<id>.tokens", cb) (src/sse-adapter/sse_handler.go:133-135, :314-316) and the match of
// a published subject (src/redis-nats-bridge/main.go:160-164) against those subscriptions happens inside
// nats.go v1.31.0 / nats-server 2.10 (third-party, not vendored).  Semantics restated here are the NATS
// subject rules: subjects are '.'-separated tokens; a subscription matches literally, '*' matches exactly
// one token, '>' (last token only) matches one or more trailing tokens; empty tokens never match a wildcard.
// Only literal chat.<id>.tokens and the wildcards chat.*.tokens / chat.*.control occur in the reference
// (docs/security-inspection-patterns.md:110,268; kubernetes/base/nats-core/core-cluster.yaml:234,248).
//
// Layout: literal subscriptions of the form chat.<id>.tokens live in an open-addressing hash table keyed
// by the id bytes (fnv1a64, linear probing, <= 50% load), values are CSR spans of subscriber ids; the few
// wildcard subscriptions are a side list matched token by token.  At 100 k subjects the table is ~4 MB
// (slots) + keys: far beyond one SM's shared memory, so it is read through L2 (126 MB holds it entirely).
#pragma once
#include <stdint.h>
#include "sse_inspect.cuh"

namespace sse {

struct InterestSlot {
    uint64_t hash;               // 0 = empty
    uint32_t key_off, key_len;   // id bytes in keys[]
    uint32_t list_off, list_cnt; // span in the CSR subscriber array
};

constexpr uint32_t INTEREST_MAX_WILD = 32;

struct InterestDev {
    const InterestSlot* slots;
    const uint8_t* keys;
    uint32_t mask;               // capacity - 1 (capacity is a power of two)
    uint32_t n_wild;
    const uint8_t* wild_bytes;   // wildcard subscription subjects, back to back
    uint32_t wild_off[INTEREST_MAX_WILD + 1];
};

struct InterestHit { uint32_t list_off, list_cnt, wild_mask; };

// NATS subject match of pattern pat[0,pn) against the subject "chat." + id + ".tokens" given by byte accessor sub(i), length sn.
template <class SubFn>
SSE_HD bool nats_match(const uint8_t* pat, uint32_t pn, SubFn&& sub, uint32_t sn) {
    uint32_t pi = 0, si = 0;
    for (;;) {
        // one token of each
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

template <class Src>
SSE_HD_NOINLINE InterestHit interest_lookup(const InterestDev& t, const Src& s, uint32_t cid_b, uint32_t cid_e) {
    InterestHit h{0, 0, 0};
    // fnv1a64 over the DECODED id (the bridge builds the subject from the decoded id, main.go:160)
    uint64_t hash = 0xcbf29ce484222325ull;
    uint32_t len = 0;
    uint32_t prev = '.';
    bool empty_token = false;                                        // "chat..tokens", "chat.a..b.tokens": not a valid NATS subject
    for_each_decoded_byte(s, cid_b, cid_e, [&](uint32_t c) {
        hash ^= c; hash *= 0x100000001b3ull; len++;
        if (c == '.' && prev == '.') empty_token = true;
        prev = c;
    });
    if (prev == '.' || empty_token) return h;                        // covers the empty id too
    if (hash == 0) hash = 1;
    if (t.slots) {
        for (uint32_t p = (uint32_t)hash & t.mask;; p = (p + 1) & t.mask) {
            const InterestSlot e = t.slots[p];
            if (e.hash == 0) break;
            if (e.hash == hash && e.key_len == len) {
                uint32_t k = 0;
                bool same = true;
                for_each_decoded_byte(s, cid_b, cid_e, [&](uint32_t c) { same = same && t.keys[e.key_off + k] == c; k++; });
                if (same) { h.list_off = e.list_off; h.list_cnt = e.list_cnt; break; }
            }
        }
    }
    if (t.n_wild) {
        // subject bytes: "chat." + id + ".tokens"; the id is materialised only up to 128 bytes, which
        // covers every wildcard decision except for ids longer than that (then compared by the slow accessor)
        uint8_t idbuf[128];
        uint32_t k = 0;
        for_each_decoded_byte(s, cid_b, cid_e, [&](uint32_t c) { if (k < 128) idbuf[k] = (uint8_t)c; k++; });
        const uint32_t sn = 5 + len + 7;
        auto sub = [&](uint32_t i) -> uint32_t {
            if (i < 5) return (uint32_t)(uint8_t)"chat."[i];
            if (i < 5 + len) {
                uint32_t j = i - 5;
                if (j < 128) return idbuf[j];
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

}  // namespace sse
