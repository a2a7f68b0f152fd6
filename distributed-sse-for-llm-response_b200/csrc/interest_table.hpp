// interest_table.hpp -- host-side builder of the interest table (see sse_interest.cuh for semantics and layout).
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "sse_interest.cuh"

namespace sse_interest {

inline bool has_wildcard(const std::string& s) {
    size_t i = 0;
    while (i <= s.size()) {
        size_t e = s.find('.', i);
        if (e == std::string::npos) e = s.size();
        if (e - i == 1 && (s[i] == '*' || s[i] == '>')) return true;
        i = e + 1;
    }
    return false;
}

struct HostTable {
    std::map<std::string, std::vector<uint32_t>> literal;    // subject -> subscriber ids (sorted by insertion)
    std::vector<std::pair<std::string, uint32_t>> wild;      // wildcard subject, subscriber id
    // flattened form
    std::vector<sse::InterestSlot> slots;
    std::vector<uint8_t> keys;
    std::vector<uint32_t> subs;                               // CSR subscriber ids
    std::vector<uint8_t> wild_bytes;
    std::vector<uint32_t> wild_off, wild_ids;

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
    void flatten() {
        keys.clear(); subs.clear(); wild_bytes.clear(); wild_off.assign(1, 0); wild_ids.clear();
        size_t n_keys = 0;
        for (auto& kv : literal) if (kv.first.size() >= 12 && kv.first.compare(0, 5, "chat.") == 0 && kv.first.compare(kv.first.size() - 7, 7, ".tokens") == 0) n_keys++;
        size_t cap = 16;
        while (cap < 2 * n_keys + 2) cap <<= 1;
        slots.assign(cap, sse::InterestSlot{0, 0, 0, 0, 0});
        for (auto& kv : literal) {
            const std::string& s = kv.first;
            if (!(s.size() >= 12 && s.compare(0, 5, "chat.") == 0 && s.compare(s.size() - 7, 7, ".tokens") == 0)) continue;   // cannot match this path's subjects
            const std::string id = s.substr(5, s.size() - 12);
            if (id.empty() || id.front() == '.' || id.back() == '.' || id.find("..") != std::string::npos) continue;   // empty token: invalid subject
            uint64_t h = 0xcbf29ce484222325ull;
            for (unsigned char c : id) { h ^= c; h *= 0x100000001b3ull; }
            if (h == 0) h = 1;
            sse::InterestSlot e{h, (uint32_t)keys.size(), (uint32_t)id.size(), (uint32_t)subs.size(), (uint32_t)kv.second.size()};
            keys.insert(keys.end(), id.begin(), id.end());
            subs.insert(subs.end(), kv.second.begin(), kv.second.end());
            size_t p = (size_t)((uint32_t)h & (uint32_t)(cap - 1));
            while (slots[p].hash != 0) p = (p + 1) & (cap - 1);
            slots[p] = e;
        }
        for (auto& w : wild) {
            wild_bytes.insert(wild_bytes.end(), w.first.begin(), w.first.end());
            wild_off.push_back((uint32_t)wild_bytes.size());
            wild_ids.push_back(w.second);
        }
        if (keys.empty()) keys.push_back(0);
        if (subs.empty()) subs.push_back(0);
        if (wild_bytes.empty()) wild_bytes.push_back(0);
    }
};

}  // namespace sse_interest
