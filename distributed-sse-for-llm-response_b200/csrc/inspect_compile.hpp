// inspect_compile.hpp -- host-side compiler of the content-inspection pattern set into byte-class DFAs.
//
// The reference never executes inspection (the adapter's hook is a comment, src/sse-adapter/
// sse_handler.go:193-208); its pattern sets are illustrative Rust snippets:
//   Set A  docs/security-inspection-patterns.md:205-209 (3 regexes -> Redact), :229-234 (4 literals -> Alert)
//   Set B  src/spin-functions/nats-subscriber/src/lib.rs:59-64, :79-84 (8 literals -> redact / drop)
// The matching semantics are therefore DEFINED by this project:
// byte-oriented, ASCII classes (\d = [0-9], \w = [0-9A-Za-z_], \b on ASCII word bytes, every byte
// >= 0x80 is a non-word byte), ASCII-only case folding for literals.
//
// Pattern language: literals (ASCII case-insensitive) and a regex subset without groups or
// alternation -- a sequence of items, item = atom [quantifier] | \b, atom = literal byte | \x escapes
// (\d \w \s \. ...) | [class] | '.', quantifier = + * ? {n} {n,} {n,m}.  That covers both sets.
//
// Output: one forward multi-pattern DFA (unanchored; the start set is re-injected at every position)
// and, for every variable-length pattern, one anchored DFA of that pattern alone, used to find the
// leftmost start of a match whose end is already known (try each start, accept only at that end).  A word-boundary assertion needs one byte of look-ahead,
// so a DFA state is (NFA node set, previous-byte-is-word, patterns accepted THROUGH an assertion on
// the way in); per state the tables give three accept masks:
//   acc_del  matches that END BEFORE the byte just consumed (found via \b, end = index of that byte)
//   acc_imm  matches that end WITH the byte just consumed   (end = index + 1)
//   acc_eot  matches completed by end-of-text (evaluated when the text ends in this state)
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <set>
#include <tuple>
#include <string>
#include <vector>

namespace sse_inspect {

constexpr uint32_t KIND_LITERAL_CI = 0, KIND_REGEX = 1;
constexpr uint32_t MAX_PATTERNS = 32;
constexpr uint32_t INF = 0xFFFFFFFFu;

struct ByteSet {
    uint64_t w[4] = {0, 0, 0, 0};
    void add(unsigned b) { w[b >> 6] |= 1ull << (b & 63); }
    bool has(unsigned b) const { return (w[b >> 6] >> (b & 63)) & 1; }
    void invert() { for (auto& x : w) x = ~x; }
    bool operator<(const ByteSet& o) const { return std::memcmp(w, o.w, sizeof w) < 0; }
    bool operator==(const ByteSet& o) const { return std::memcmp(w, o.w, sizeof w) == 0; }
};

struct Item {
    bool is_assert = false;     // \b
    ByteSet set;
    uint32_t min = 1, max = 1;  // max == INF: unbounded
};

inline bool is_word(unsigned b) { return (b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || (b >= 'a' && b <= 'z') || b == '_'; }

inline ByteSet class_of_escape(char c, bool& ok) {
    ByteSet s;
    ok = true;
    switch (c) {
    case 'd': for (unsigned b = '0'; b <= '9'; b++) s.add(b); break;
    case 'w': for (unsigned b = 0; b < 128; b++) if (is_word(b)) s.add(b); break;
    case 's': for (char b : {' ', '\t', '\n', '\r', '\f', '\v'}) s.add((unsigned char)b); break;
    case 'n': s.add('\n'); break;
    case 't': s.add('\t'); break;
    case 'r': s.add('\r'); break;
    default:
        if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9')) ok = false;   // unsupported escape
        else s.add((unsigned char)c);                                                                // escaped punctuation
    }
    return s;
}

// Parses the regex subset; returns false (with err) on anything outside it.
inline bool parse_regex(const std::string& re, std::vector<Item>& items, std::string& err) {
    size_t i = 0;
    auto fail = [&](const char* m) { err = std::string(m) + " at offset " + std::to_string(i) + " of '" + re + "'"; return false; };
    while (i < re.size()) {
        Item it;
        char c = re[i];
        if (c == '\\') {
            if (i + 1 >= re.size()) return fail("dangling backslash");
            char e = re[i + 1];
            i += 2;
            if (e == 'b') { it.is_assert = true; items.push_back(it); continue; }
            bool ok;
            it.set = class_of_escape(e, ok);
            if (!ok) return fail("unsupported escape");
        } else if (c == '[') {
            i++;
            bool neg = false;
            if (i < re.size() && re[i] == '^') { neg = true; i++; }
            bool first = true;
            while (i < re.size() && (re[i] != ']' || first)) {
                first = false;
                ByteSet one;
                unsigned lo;
                bool single = true;
                if (re[i] == '\\') {
                    if (i + 1 >= re.size()) return fail("dangling backslash in class");
                    bool ok;
                    one = class_of_escape(re[i + 1], ok);
                    if (!ok) return fail("unsupported escape in class");
                    lo = (unsigned char)re[i + 1];
                    single = !(re[i + 1] == 'd' || re[i + 1] == 'w' || re[i + 1] == 's');
                    i += 2;
                } else { lo = (unsigned char)re[i]; one.add(lo); i++; }
                if (single && i + 1 < re.size() && re[i] == '-' && re[i + 1] != ']') {
                    unsigned hi = (unsigned char)re[i + 1];
                    i += 2;
                    if (hi < lo) return fail("bad range");
                    for (unsigned b = lo; b <= hi; b++) it.set.add(b);
                } else {
                    for (unsigned b = 0; b < 256; b++) if (one.has(b)) it.set.add(b);
                }
            }
            if (i >= re.size()) return fail("unterminated class");
            i++;
            if (neg) it.set.invert();
        } else if (c == '.') {
            for (unsigned b = 0; b < 256; b++) if (b != '\n') it.set.add(b);
            i++;
        } else if (c == '(' || c == ')' || c == '|' || c == '^' || c == '$') {
            return fail("groups, alternation and anchors are outside the supported subset");
        } else if (c == '+' || c == '*' || c == '?' || c == '{') {
            return fail("quantifier without atom");
        } else { it.set.add((unsigned char)c); i++; }
        // quantifier
        if (i < re.size()) {
            char q = re[i];
            if (q == '+') { it.min = 1; it.max = INF; i++; }
            else if (q == '*') { it.min = 0; it.max = INF; i++; }
            else if (q == '?') { it.min = 0; it.max = 1; i++; }
            else if (q == '{') {
                size_t j = i + 1;
                auto num = [&](uint32_t& v) { if (j >= re.size() || re[j] < '0' || re[j] > '9') return false; v = 0; while (j < re.size() && re[j] >= '0' && re[j] <= '9') { v = v * 10 + (re[j] - '0'); if (v > 1000) return false; j++; } return true; };
                uint32_t a, b;
                if (!num(a)) return fail("bad {n}");
                b = a;
                if (j < re.size() && re[j] == ',') { j++; if (j < re.size() && re[j] == '}') b = INF; else if (!num(b)) return fail("bad {n,m}"); }
                if (j >= re.size() || re[j] != '}') return fail("unterminated {}");
                if (b != INF && b < a) return fail("bad {n,m}");
                it.min = a; it.max = b; i = j + 1;
            }
        }
        items.push_back(it);
    }
    return true;
}

inline void literal_items(const std::string& lit, std::vector<Item>& items) {
    for (unsigned char c : lit) {
        Item it;
        it.set.add(c);
        if (c >= 'a' && c <= 'z') it.set.add(c - 32);
        if (c >= 'A' && c <= 'Z') it.set.add(c + 32);
        items.push_back(it);
    }
}

// ------------------------------------------------------------------------ NFA ----
struct Nfa {
    struct Edge { uint32_t set_id; uint32_t to; };
    std::vector<std::vector<Edge>> byte_edges;
    std::vector<std::vector<uint32_t>> eps, asserts;
    std::vector<int> accept_of;          // pattern id or -1
    std::vector<ByteSet> sets;           // distinct byte sets
    uint32_t node() { byte_edges.emplace_back(); eps.emplace_back(); asserts.emplace_back(); accept_of.push_back(-1); return (uint32_t)accept_of.size() - 1; }
    uint32_t set_id(const ByteSet& s) {
        for (uint32_t i = 0; i < sets.size(); i++) if (sets[i] == s) return i;
        sets.push_back(s);
        return (uint32_t)sets.size() - 1;
    }
    // appends the item sequence; returns (start, accept)
    std::pair<uint32_t, uint32_t> add(const std::vector<Item>& items, int pattern) {
        uint32_t start = node(), cur = start;
        for (const Item& it : items) {
            if (it.is_assert) { uint32_t n = node(); asserts[cur].push_back(n); cur = n; continue; }
            uint32_t sid = set_id(it.set);
            for (uint32_t k = 0; k < it.min; k++) { uint32_t n = node(); byte_edges[cur].push_back({sid, n}); cur = n; }
            if (it.max == INF) {
                // zero or more further copies; a fresh node keeps the loop away from what precedes it
                uint32_t loop = node();
                eps[cur].push_back(loop);
                byte_edges[loop].push_back({sid, loop});
                cur = loop;
            } else if (it.max > it.min) {
                std::vector<uint32_t> chain{cur};
                for (uint32_t k = it.min; k < it.max; k++) { uint32_t n = node(); byte_edges[chain.back()].push_back({sid, n}); chain.push_back(n); }
                for (size_t k = 0; k + 1 < chain.size(); k++) eps[chain[k]].push_back(chain.back());
                cur = chain.back();
            }
        }
        accept_of[cur] = pattern;
        return {start, cur};
    }
};

struct PatternInfo {
    uint32_t kind = 0, action = 0;
    uint32_t min_len = 0, max_len = 0;   // max_len == INF: unbounded
    std::vector<Item> items;
};

// One compiled DFA in the flat form the kernels and the host use.
struct Dfa {
    uint32_t n_states = 0, n_classes = 0;
    uint32_t start[2] = {0, 0};                 // start state when the byte before the text is a non-word / word byte
    uint8_t cls[256];                           // byte -> class
    std::vector<uint16_t> next;                 // [state][class]; 0xFFFF = dead; bit 15 = target state reports a match
    std::vector<uint32_t> acc_imm, acc_del, acc_eot;
    // per state: the fewest further bytes after which anything can be reported (0 = at end of text in this state),
    // saturating at 15.  A walk that has fewer bytes left than that can stop.
    std::vector<uint8_t> dist;
};

namespace detail {

typedef std::vector<uint32_t> NodeSet;   // sorted, unique

inline void close_over(const Nfa& n, NodeSet& s, bool with_asserts) {
    std::vector<uint32_t> stack(s.begin(), s.end());
    std::set<uint32_t> seen(s.begin(), s.end());
    while (!stack.empty()) {
        uint32_t v = stack.back();
        stack.pop_back();
        for (uint32_t t : n.eps[v]) if (seen.insert(t).second) stack.push_back(t);
        if (with_asserts) for (uint32_t t : n.asserts[v]) if (seen.insert(t).second) stack.push_back(t);
    }
    s.assign(seen.begin(), seen.end());
}

inline uint32_t accept_mask(const Nfa& n, const NodeSet& s) {
    uint32_t m = 0;
    for (uint32_t v : s) if (n.accept_of[v] >= 0) m |= 1u << n.accept_of[v];
    return m;
}

}  // namespace detail

// Subset construction.  starts: nodes injected at every position (unanchored) or only at the beginning
// (anchored).  max_states bounds the blow-up.
inline bool build_dfa(const Nfa& nfa, const std::vector<uint32_t>& starts, bool unanchored, Dfa& out, uint32_t max_states, std::string& err) {
    using detail::NodeSet;
    // byte classes: bytes that no edge set and the word predicate can tell apart
    {
        std::map<std::vector<bool>, uint8_t> sig2cls;
        for (unsigned b = 0; b < 256; b++) {
            std::vector<bool> sig;
            sig.push_back(is_word(b));
            for (const ByteSet& s : nfa.sets) sig.push_back(s.has(b));
            auto it = sig2cls.find(sig);
            if (it == sig2cls.end()) it = sig2cls.emplace(sig, (uint8_t)sig2cls.size()).first;
            out.cls[b] = it->second;
        }
        out.n_classes = (uint32_t)sig2cls.size();
    }
    std::vector<unsigned> rep(out.n_classes, 0);
    for (int b = 255; b >= 0; b--) rep[out.cls[b]] = (unsigned)b;

    struct Key { NodeSet nodes; bool pw; uint32_t del; bool operator<(const Key& o) const { return std::tie(nodes, pw, del) < std::tie(o.nodes, o.pw, o.del); } };
    std::map<Key, uint32_t> ids;
    std::vector<Key> states;
    auto intern = [&](Key k) {
        auto it = ids.find(k);
        if (it != ids.end()) return it->second;
        uint32_t id = (uint32_t)states.size();
        ids.emplace(k, id);
        states.push_back(std::move(k));
        return id;
    };
    NodeSet s0(starts.begin(), starts.end());
    std::sort(s0.begin(), s0.end());
    detail::close_over(nfa, s0, false);
    out.start[0] = intern({s0, false, 0});
    out.start[1] = intern({s0, true, 0});
    for (uint32_t si = 0; si < states.size(); si++) {
        if (states.size() > max_states) { err = "pattern set needs more than " + std::to_string(max_states) + " DFA states"; return false; }
        const Key cur = states[si];
        out.next.resize((size_t)(si + 1) * out.n_classes, 0xFFFF);
        {   // end of text behaves like a non-word byte
            NodeSet e = cur.nodes;
            if (cur.pw) detail::close_over(nfa, e, true);
            out.acc_eot.push_back(detail::accept_mask(nfa, e) & ~detail::accept_mask(nfa, cur.nodes));
        }
        out.acc_imm.push_back(detail::accept_mask(nfa, cur.nodes));
        out.acc_del.push_back(cur.del);
        for (uint32_t c = 0; c < out.n_classes; c++) {
            const unsigned b = rep[c];
            const bool w = is_word(b);
            NodeSet n1 = cur.nodes;
            if (cur.pw != w) detail::close_over(nfa, n1, true);
            const uint32_t del = detail::accept_mask(nfa, n1) & ~detail::accept_mask(nfa, cur.nodes);
            std::set<uint32_t> nx;
            for (uint32_t v : n1) for (const Nfa::Edge& e : nfa.byte_edges[v]) if (nfa.sets[e.set_id].has(b)) nx.insert(e.to);
            if (unanchored) nx.insert(starts.begin(), starts.end());
            NodeSet n2(nx.begin(), nx.end());
            detail::close_over(nfa, n2, false);
            if (n2.empty() && del == 0) continue;   // dead
            uint32_t id = intern({n2, w, del});
            if (id >= 0xFFFF) { err = "too many DFA states"; return false; }
            out.next[(size_t)si * out.n_classes + c] = (uint16_t)id;
        }
    }
    out.n_states = (uint32_t)states.size();
    out.next.resize((size_t)out.n_states * out.n_classes, 0xFFFF);
    if (out.n_states >= 0x7FFF) { err = "too many DFA states"; return false; }
    // bit 15 of a transition says "the target state reports a match" (acc_imm | acc_del non-zero), so the walk
    // loads the accept masks only when something fired
    for (uint16_t& nx : out.next)
        if (nx != 0xFFFF && (out.acc_imm[nx] | out.acc_del[nx])) nx |= 0x8000;
    std::vector<uint32_t> dist(out.n_states, INF);
    for (uint32_t st = 0; st < out.n_states; st++) if (out.acc_eot[st]) dist[st] = 0;
    for (bool changed = true; changed;) {
        changed = false;
        for (uint32_t st = 0; st < out.n_states; st++)
            for (uint32_t c = 0; c < out.n_classes; c++) {
                const uint16_t nx = out.next[(size_t)st * out.n_classes + c];
                if (nx == 0xFFFF) continue;
                const uint32_t t = nx & 0x7FFF;
                const uint32_t cand = (nx & 0x8000) ? 1u : (dist[t] == INF ? INF : dist[t] + 1);
                if (cand < dist[st]) { dist[st] = cand; changed = true; }
            }
    }
    out.dist.resize(out.n_states);
    for (uint32_t st = 0; st < out.n_states; st++) out.dist[st] = (uint8_t)(dist[st] > 15 ? 15 : dist[st]);
    return true;
}

struct Compiled {
    std::vector<PatternInfo> patterns;
    Dfa forward;
    std::vector<int> anchored_of;    // pattern id -> index into anchored (or -1: fixed length)
    std::vector<Dfa> anchored;
};

struct PatternSpec { uint32_t kind, action; std::string text; };

inline bool compile(const std::vector<PatternSpec>& specs, Compiled& out, std::string& err, uint32_t max_states = 20000) {
    if (specs.empty() || specs.size() > MAX_PATTERNS) { err = "need 1.." + std::to_string(MAX_PATTERNS) + " patterns"; return false; }
    Nfa nfa;
    std::vector<uint32_t> starts;
    out.patterns.clear(); out.anchored.clear(); out.anchored_of.assign(specs.size(), -1);
    for (size_t p = 0; p < specs.size(); p++) {
        PatternInfo pi;
        pi.kind = specs[p].kind; pi.action = specs[p].action;
        if (pi.action < 1 || pi.action > 3) { err = "action must be 1 (redact), 2 (drop) or 3 (alert)"; return false; }
        if (pi.kind == KIND_LITERAL_CI) literal_items(specs[p].text, pi.items);
        else if (pi.kind == KIND_REGEX) { if (!parse_regex(specs[p].text, pi.items, err)) return false; }
        else { err = "unknown pattern kind"; return false; }
        uint64_t mn = 0, mx = 0;
        bool unbounded = false;
        for (const Item& it : pi.items) if (!it.is_assert) { mn += it.min; if (it.max == INF) unbounded = true; else mx += it.max; }
        if (mn == 0) { err = "pattern '" + specs[p].text + "' can match the empty string"; return false; }
        pi.min_len = (uint32_t)mn; pi.max_len = unbounded ? INF : (uint32_t)mx;
        starts.push_back(nfa.add(pi.items, (int)p).first);
        out.patterns.push_back(pi);
    }
    if (!build_dfa(nfa, starts, true, out.forward, max_states, err)) return false;
    for (size_t p = 0; p < specs.size(); p++) {
        const PatternInfo& pi = out.patterns[p];
        if (pi.min_len == pi.max_len) continue;
        Nfa r;
        uint32_t s = r.add(pi.items, 0).first;
        Dfa d;
        if (!build_dfa(r, {s}, false, d, max_states, err)) return false;
        out.anchored_of[p] = (int)out.anchored.size();
        out.anchored.push_back(std::move(d));
    }
    return true;
}

}  // namespace sse_inspect
