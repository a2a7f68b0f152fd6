// interest_table.hpp -- host-side builder of the interest table (see sse_interest.cuh for semantics and layout).
//
// The reference changes a subscription on every HTTP connect and disconnect (nc.Subscribe / sub.Unsubscribe,
// src/sse-adapter/sse_handler.go:133-135, :166-169), so the table is maintained incrementally: the flattened form
// (open-addressing slots + an append-only arena of subscriber-id lists) is rebuilt only when a load limit is hit;
// every other delta touches one slot (reported in `dirty`) and at most appends one list to the arena.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "sse_interest.cuh"

namespace sse_interest {

inline std::vector<std::string> tokens_of(const std::string& s) {
    std::vector<std::string> t;
    size_t i = 0;
    for (;;) {
        size_t e = s.find('.', i);
        if (e == std::string::npos) { t.push_back(s.substr(i)); break; }
        t.push_back(s.substr(i, e - i));
        i = e + 1;
    }
    return t;
}

inline bool has_wildcard(const std::string& s) {
    for (const std::string& t : tokens_of(s)) if (t == "*" || t == ">") return true;
    return false;
}

// How a wildcard subscription behaves on subjects chat.<id>.tokens whose id is ONE token:
// 1 = matches all of them, 0 = matches none of them, 2 = depends on the id (needs the matcher).
inline int classify_wild(const std::string& pat) {
    const std::vector<std::string> p = tokens_of(pat);
    for (const std::string& t : p) if (t.empty()) return 0;
    const char* subj[3] = {"chat", nullptr, "tokens"};             // nullptr = the id token
    for (size_t i = 0; i < p.size(); i++) {
        if (p[i] == ">") return i == p.size() - 1 && i < 3 ? 1 : 0;
        if (i >= 3) return 0;
        if (p[i] == "*") continue;
        if (subj[i] == nullptr) return 2;                          // literal in the id position
        if (p[i] != subj[i]) return 0;
    }
    return p.size() == 3 ? 1 : 0;
}

struct HostTable {
    std::map<std::string, std::vector<uint32_t>> literal;    // subject -> subscriber ids (insertion order)
    std::vector<std::pair<std::string, uint32_t>> wild;      // wildcard subject, subscriber id
    // flattened form (one "generation": valid from one flatten() to the next)
    std::vector<sse::InterestSlot> slots;
    std::vector<uint32_t> keys;                               // zero-padded words of the ids that do not fit a slot
    // Append-only arena of subscriber-id lists.  Reserved once per generation and never grown past that, so a
    // span handed out for this generation stays readable at the same address while later deltas append.
    std::shared_ptr<std::vector<uint32_t>> subs = std::make_shared<std::vector<uint32_t>>();
    std::vector<uint8_t> wild_bytes;
    std::vector<uint32_t> wild_off, wild_ids;
    uint32_t wild_single = 0, wild_generic = 0;
    size_t used_slots = 0, live_ids = 0;
    bool flattened = false;
    std::vector<uint32_t> dirty;                              // slots changed by sync_subject() since the caller last cleared it

    // returns 1 = literal map changed, 2 = wildcard list changed, 0 = nothing changed, -1 = error
    int add(const std::string& subject, uint32_t id, std::string& err) {
        if (has_wildcard(subject)) {
            for (auto& w : wild) if (w.first == subject && w.second == id) return 0;
            if (wild.size() >= sse::INTEREST_MAX_WILD) { err = "more than 32 wildcard subscriptions"; return -1; }
            wild.emplace_back(subject, id);
            return 2;
        }
        auto& v = literal[subject];
        for (uint32_t x : v) if (x == id) return 0;
        v.push_back(id);
        return 1;
    }
    int del(const std::string& subject, uint32_t id) {
        if (has_wildcard(subject)) {
            for (size_t i = 0; i < wild.size(); i++) if (wild[i].first == subject && wild[i].second == id) { wild.erase(wild.begin() + i); return 2; }
            return 0;
        }
        auto it = literal.find(subject);
        if (it == literal.end()) return 0;
        auto& v = it->second;
        int changed = 0;
        for (size_t i = 0; i < v.size(); i++) if (v[i] == id) { v.erase(v.begin() + i); changed = 1; break; }
        if (v.empty()) literal.erase(it);
        return changed;
    }
    static bool id_of(const std::string& s, std::string& id) {        // chat.<id>.tokens with a usable id
        if (!(s.size() >= 12 && s.compare(0, 5, "chat.") == 0 && s.compare(s.size() - 7, 7, ".tokens") == 0)) return false;
        id = s.substr(5, s.size() - 12);
        if (id.empty() || id.size() > sse::INTEREST_MAX_ID) return false;
        if (id.front() == '.' || id.back() == '.' || id.find("..") != std::string::npos) return false;   // empty token: invalid subject
        return true;
    }
    static uint32_t hash_of(const std::string& id, std::vector<uint32_t>& words) {
        const uint32_t len = (uint32_t)id.size(), nw = (len + 3) / 4;
        words.assign(nw, 0);
        memcpy(words.data(), id.data(), len);
        uint32_t h = sse::key_hash_init(len);
        for (uint32_t w : words) h = sse::key_hash_step(h, w);
        return sse::key_hash_fin(h);
    }
    // slot of `id` in the flattened form, or the empty slot its probe sequence ends at (found = false)
    size_t probe(const std::string& id, uint32_t h, const std::vector<uint32_t>& words, bool& found) const {
        const size_t mask = slots.size() - 1;
        const uint32_t len = (uint32_t)id.size();
        for (size_t p = h & mask;; p = (p + 1) & mask) {
            const sse::InterestSlot& e = slots[p];
            if (e.hash == 0) { found = false; return p; }
            if (e.hash == h && e.key_len == len) {
                const uint32_t* key = len <= sse::INTEREST_INLINE_KEY ? e.key : keys.data() + e.key[0];
                if (memcmp(key, words.data(), words.size() * 4) == 0) { found = true; return p; }
            }
        }
    }
    void flatten_wild() {
        wild_bytes.clear(); wild_off.assign(1, 0); wild_ids.clear();
        wild_single = wild_generic = 0;
        for (size_t w = 0; w < wild.size(); w++) {
            wild_bytes.insert(wild_bytes.end(), wild[w].first.begin(), wild[w].first.end());
            wild_off.push_back((uint32_t)wild_bytes.size());
            wild_ids.push_back(wild[w].second);
            const int c = classify_wild(wild[w].first);
            if (c == 1) wild_single |= 1u << w;
            if (c == 2) wild_generic |= 1u << w;
        }
        if (wild_bytes.empty()) wild_bytes.push_back(0);
    }
    // Rebuild the flattened form from the maps: a new generation with room to grow (slots at most a quarter full,
    // arena reserved at twice the live ids plus slack).
    void flatten() {
        keys.clear();
        size_t n_keys = 0, n_ids = 0;
        std::string id;
        for (auto& kv : literal) if (id_of(kv.first, id)) { n_keys++; n_ids += kv.second.size(); }
        size_t cap = 16;
        while (cap < 4 * n_keys + 2) cap <<= 1;
        slots.assign(cap, sse::InterestSlot{});
        subs = std::make_shared<std::vector<uint32_t>>();
        subs->reserve(2 * n_ids + 65536);
        std::vector<uint32_t> words;
        for (auto& kv : literal) {
            if (!id_of(kv.first, id)) continue;                       // cannot match this path's subjects
            const uint32_t h = hash_of(id, words), len = (uint32_t)id.size();
            sse::InterestSlot e{};
            e.hash = h; e.key_len = len; e.list_off = (uint32_t)subs->size(); e.list_cnt = (uint32_t)kv.second.size();
            if (len <= sse::INTEREST_INLINE_KEY) memcpy(e.key, words.data(), words.size() * 4);
            else { e.key[0] = (uint32_t)keys.size(); keys.insert(keys.end(), words.begin(), words.end()); }
            subs->insert(subs->end(), kv.second.begin(), kv.second.end());
            size_t p = (size_t)(h & (uint32_t)(cap - 1));
            while (slots[p].hash != 0) p = (p + 1) & (cap - 1);
            slots[p] = e;
        }
        used_slots = n_keys; live_ids = n_ids;
        flatten_wild();
        if (keys.empty()) keys.push_back(0);
        dirty.clear();
        flattened = true;
    }
    // After add()/del() changed literal[subject]: bring that subject's slot in line.  Returns false when the
    // flattened form cannot take the change in place (slots half full, arena full, id too long for a slot): the
    // caller rebuilds with flatten().
    bool sync_subject(const std::string& subject) {
        std::string id;
        if (!flattened) return false;
        if (!id_of(subject, id)) return true;                           // cannot match this path's subjects: not in the table
        std::vector<uint32_t> words;
        const uint32_t h = hash_of(id, words);
        bool found;
        const size_t p = probe(id, h, words, found);
        auto it = literal.find(subject);
        const std::vector<uint32_t>* list = it == literal.end() ? nullptr : &it->second;
        const uint32_t cnt = list ? (uint32_t)list->size() : 0;
        if (!found) {
            if (cnt == 0) return true;
            if (id.size() > sse::INTEREST_INLINE_KEY || (used_slots + 1) * 2 > slots.size()) return false;
            if (subs->size() + cnt > subs->capacity()) return false;
            sse::InterestSlot e{};
            e.hash = h; e.key_len = (uint32_t)id.size(); e.list_off = (uint32_t)subs->size(); e.list_cnt = cnt;
            memcpy(e.key, words.data(), words.size() * 4);
            subs->insert(subs->end(), list->begin(), list->end());
            slots[p] = e;
            used_slots++; live_ids += cnt;
            dirty.push_back((uint32_t)p);
            return true;
        }
        sse::InterestSlot& e = slots[p];
        const uint32_t* old = subs->data() + e.list_off;
        live_ids += cnt; live_ids -= e.list_cnt;
        if (cnt <= e.list_cnt && (cnt == 0 || memcmp(old, list->data(), cnt * 4) == 0)) {
            e.list_cnt = cnt;                                           // the list lost its tail (or everything): same span, shorter
        } else if (cnt == e.list_cnt + 1 && e.list_off + e.list_cnt == subs->size() && subs->size() < subs->capacity() &&
                   memcmp(old, list->data(), e.list_cnt * 4) == 0) {
            subs->push_back(list->back());                              // the list is the arena's last one: grow it in place
            e.list_cnt = cnt;
        } else {
            if (subs->size() + cnt > subs->capacity()) return false;
            e.list_off = (uint32_t)subs->size(); e.list_cnt = cnt;
            subs->insert(subs->end(), list->begin(), list->end());
        }
        dirty.push_back((uint32_t)p);
        return true;
    }
};

}  // namespace sse_interest
