// interest_table.hpp -- host-side builder of the interest table (see sse_interest.cuh for semantics and layout).
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
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
    // flattened form
    std::vector<sse::InterestSlot> slots;
    std::vector<uint32_t> keys;                               // zero-padded words
    std::vector<uint32_t> subs;                               // CSR subscriber ids
    std::vector<uint8_t> wild_bytes;
    std::vector<uint32_t> wild_off, wild_ids;
    uint32_t wild_single = 0, wild_generic = 0;

    bool add(const std::string& subject, uint32_t id, std::string& err) {
        if (has_wildcard(subject)) {
            for (auto& w : wild) if (w.first == subject && w.second == id) return true;
            if (wild.size() >= sse::INTEREST_MAX_WILD) { err = "more than 32 wildcard subscriptions"; return false; }
            wild.emplace_back(subject, id);
        } else {
            auto& v = literal[subject];
            for (uint32_t x : v) if (x == id) return true;
            v.push_back(id);
        }
        return true;
    }
    void del(const std::string& subject, uint32_t id) {
        if (has_wildcard(subject)) {
            for (size_t i = 0; i < wild.size(); i++) if (wild[i].first == subject && wild[i].second == id) { wild.erase(wild.begin() + i); return; }
        } else {
            auto it = literal.find(subject);
            if (it == literal.end()) return;
            auto& v = it->second;
            for (size_t i = 0; i < v.size(); i++) if (v[i] == id) { v.erase(v.begin() + i); break; }
            if (v.empty()) literal.erase(it);
        }
    }
    static bool id_of(const std::string& s, std::string& id) {        // chat.<id>.tokens with a usable id
        if (!(s.size() >= 12 && s.compare(0, 5, "chat.") == 0 && s.compare(s.size() - 7, 7, ".tokens") == 0)) return false;
        id = s.substr(5, s.size() - 12);
        if (id.empty() || id.size() > sse::INTEREST_MAX_ID) return false;
        if (id.front() == '.' || id.back() == '.' || id.find("..") != std::string::npos) return false;   // empty token: invalid subject
        return true;
    }
    void flatten() {
        keys.clear(); subs.clear(); wild_bytes.clear(); wild_off.assign(1, 0); wild_ids.clear();
        size_t n_keys = 0;
        std::string id;
        for (auto& kv : literal) if (id_of(kv.first, id)) n_keys++;
        size_t cap = 16;
        while (cap < 2 * n_keys + 2) cap <<= 1;
        slots.assign(cap, sse::InterestSlot{});
        for (auto& kv : literal) {
            if (!id_of(kv.first, id)) continue;                       // cannot match this path's subjects
            const uint32_t len = (uint32_t)id.size(), nw = (len + 3) / 4;
            std::vector<uint32_t> words(nw, 0);
            memcpy(words.data(), id.data(), len);
            uint32_t h = sse::key_hash_init(len);
            for (uint32_t w : words) h = sse::key_hash_step(h, w);
            h = sse::key_hash_fin(h);
            sse::InterestSlot e{};
            e.hash = h; e.key_len = len; e.list_off = (uint32_t)subs.size(); e.list_cnt = (uint32_t)kv.second.size();
            if (len <= sse::INTEREST_INLINE_KEY) memcpy(e.key, words.data(), nw * 4);
            else { e.key[0] = (uint32_t)keys.size(); keys.insert(keys.end(), words.begin(), words.end()); }
            subs.insert(subs.end(), kv.second.begin(), kv.second.end());
            size_t p = (size_t)(h & (uint32_t)(cap - 1));
            while (slots[p].hash != 0) p = (p + 1) & (cap - 1);
            slots[p] = e;
        }
        wild_single = wild_generic = 0;
        for (size_t w = 0; w < wild.size(); w++) {
            wild_bytes.insert(wild_bytes.end(), wild[w].first.begin(), wild[w].first.end());
            wild_off.push_back((uint32_t)wild_bytes.size());
            wild_ids.push_back(wild[w].second);
            const int c = classify_wild(wild[w].first);
            if (c == 1) wild_single |= 1u << w;
            if (c == 2) wild_generic |= 1u << w;
        }
        if (keys.empty()) keys.push_back(0);
        if (subs.empty()) subs.push_back(0);
        if (wild_bytes.empty()) wild_bytes.push_back(0);
    }
};

}  // namespace sse_interest
