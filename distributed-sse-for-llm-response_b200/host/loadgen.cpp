// loadgen.cpp -- synthetic token streams for the SSE hot path ("replay mode" of the reference's
// demo/load-generator: the same TokenMessage content, written straight into a packed arena
// instead of being PUBLISHed to Redis one message at a time).
//
// Mirrors demo/load-generator/main.go:
//   runProducer (:204-240)  TokenMessage{cid, sampleText[i%56], i+1, i==n-1, UnixNano} -> json.Marshal
//   sampleText  (:92-96)    56 words, four of them containing '\n'
//   id format   (:149)      "loadtest-<UnixNano>-<i>"
// and adds the BASELINE.json shapes the reference has no generator for (UUID ids as made by
// uuid.New().String() at src/sse-adapter/sse_handler.go:275, LLM-like token lengths).
//
// Records are laid out in ARRIVAL order: token index j major, conversation i minor (every
// conversation emits its j-th token before any emits its (j+1)-th), which is what a batch cut
// from thousands of concurrent streams looks like.  Everything is a pure function of
// (seed, i, j), so ranks and threads can generate disjoint slices independently.
//
// Host-side C++; no CUDA.  The JSON encoder below is the producer side of the wire format
// (json.Marshal of TokenMessage, escapeHTML on), written independently of the test oracle.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

const char* const kWords[56] = {
    "The", "quick", "brown", "fox", "jumps", "over", "the", "lazy", "dog.\nThis", "is", "a", "sample", "response", "from",
    "a", "large", "language", "model", "that", "simulates", "streaming", "tokens.\nEach", "word", "is", "sent", "as", "a",
    "separate", "token", "with", "timestamps", "for", "latency", "measurement.\nNATS", "provides", "excellent", "fan-out",
    "capabilities", "for", "distributing", "these", "tokens", "to", "edge", "locations.\nThe", "SSE", "adapter", "converts",
    "NATS", "messages", "into", "Server-Sent", "Events", "for", "browser", "clients."};

inline uint64_t splitmix64(uint64_t& x) {
    uint64_t z = (x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline uint64_t mix(uint64_t seed, uint64_t a, uint64_t b) {
    uint64_t x = seed ^ (a * 0xD6E8FEB86659FD93ull) ^ (b * 0xA0761D6478BD642Full);
    return splitmix64(x);
}

struct Out {
    uint8_t* p;       // nullptr = count only
    uint64_t n = 0;
    void put(uint8_t c) { if (p) p[n] = c; n++; }
    void put(const char* s) { while (*s) put((uint8_t)*s++); }
    void put(const uint8_t* s, size_t len) { if (p) memcpy(p + n, s, len); n += len; }
};

// encode.go appendString, escapeHTML = true, Go 1.21 spelling
void put_json_string(Out& o, const uint8_t* s, size_t len) {
    static const char hex[] = "0123456789abcdef";
    o.put('"');
    for (size_t i = 0; i < len; i++) {
        uint8_t b = s[i];
        if (b >= 0x80) {   // generator only emits valid UTF-8; U+2028/9 are never generated
            o.put(b);
            continue;
        }
        if (b >= 0x20 && b != '"' && b != '\\' && b != '<' && b != '>' && b != '&') { o.put(b); continue; }
        o.put('\\');
        switch (b) {
        case '"': case '\\': o.put(b); break;
        case '\n': o.put('n'); break;
        case '\r': o.put('r'); break;
        case '\t': o.put('t'); break;
        default: o.put('u'); o.put('0'); o.put('0'); o.put((uint8_t)hex[b >> 4]); o.put((uint8_t)hex[b & 15]);
        }
    }
    o.put('"');
}

void put_int(Out& o, int64_t v) {
    char tmp[24];
    int k = 0;
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    do { tmp[k++] = (char)('0' + u % 10); u /= 10; } while (u);
    if (v < 0) o.put('-');
    while (k) o.put((uint8_t)tmp[--k]);
}

void put_message(Out& o, const uint8_t* cid, size_t cid_len, const uint8_t* tok, size_t tok_len, int64_t seq, bool done,
                 int64_t ts) {
    o.put("{\"conversation_id\":");
    put_json_string(o, cid, cid_len);
    o.put(",\"token\":");
    put_json_string(o, tok, tok_len);
    o.put(",\"sequence\":");
    put_int(o, seq);
    o.put(done ? ",\"done\":true" : ",\"done\":false");
    o.put(",\"timestamp\":");
    put_int(o, ts);
    o.put('}');
}

const int64_t kClock = 1700000000000000000ll;   // fixed stand-in for time.Now().UnixNano()

size_t loadtest_id(uint8_t* dst, uint32_t i) {   // fmt.Sprintf("loadtest-%d-%d", UnixNano, i)
    Out o{dst};
    o.put("loadtest-");
    put_int(o, kClock);
    o.put('-');
    put_int(o, i);
    return o.n;
}

size_t uuid_id(uint8_t* dst, uint64_t seed, uint32_t i) {   // uuid.New().String(): v4, 36 bytes
    static const char hex[] = "0123456789abcdef";
    uint64_t a = mix(seed, i, 1), b = mix(seed, i, 2);
    uint8_t raw[16];
    memcpy(raw, &a, 8);
    memcpy(raw + 8, &b, 8);
    raw[6] = (raw[6] & 0x0F) | 0x40;
    raw[8] = (raw[8] & 0x3F) | 0x80;
    size_t n = 0;
    for (int k = 0; k < 16; k++) {
        if (k == 4 || k == 6 || k == 8 || k == 10) dst[n++] = '-';
        dst[n++] = (uint8_t)hex[raw[k] >> 4];
        dst[n++] = (uint8_t)hex[raw[k] & 15];
    }
    return n;
}

// "Mistral-7B-like" token text: byte lengths 1..16 with the weights
// below (mean 4.99), leading space with p=0.55, printable ASCII; 1.5% carry one byte that JSON
// must escape, 1% carry one 2-4 byte UTF-8 scalar.
size_t llm_token(uint8_t* dst, uint64_t seed, uint32_t i, uint32_t j) {
    static const uint16_t cum[16] = {60, 160, 320, 500, 650, 760, 840, 900, 940, 960, 970, 980, 985, 990, 995, 1000};
    static const char plain[] = "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789.,;:!?-_'()/ ";
    static const char esc[] = {'"', '\\', '\n', '<', '>', '&'};
    static const char* const utf[] = {"\xC3\xA9", "\xE2\x82\xAC", "\xE4\xB8\xAD", "\xF0\x9F\x98\x80", "\xC3\xBC", "\xE2\x80\x94"};
    uint64_t st = mix(seed, i, j);
    uint32_t r = (uint32_t)(splitmix64(st) % 1000);
    size_t len = 1;
    while (r >= cum[len - 1]) len++;
    size_t n = 0;
    uint64_t bits = splitmix64(st);
    if ((bits & 0xFFFF) < (uint64_t)(0.55 * 65536)) dst[n++] = ' ';
    while (n < len) {
        uint64_t x = splitmix64(st);
        for (int k = 0; k < 8 && n < len; k++, x >>= 8) dst[n++] = (uint8_t)plain[(x & 0xFF) % (sizeof(plain) - 1)];
    }
    uint32_t kind = (uint32_t)(splitmix64(st) % 1000);
    if (kind < 15) {
        dst[(bits >> 16) % len] = (uint8_t)esc[(bits >> 32) % sizeof(esc)];
    } else if (kind < 25) {
        const char* u = utf[(bits >> 32) % 6];
        size_t ul = strlen(u), at = (bits >> 16) % len;
        memmove(dst + at + ul, dst + at + 1, len - at - 1);
        memcpy(dst + at, u, ul);
        n = len - 1 + ul;
    }
    return n;
}

struct Shape {
    int config;
    uint32_t n_conv, n_tok, conv_first;   // conversations [conv_first, conv_first + n_conv) ...
    uint64_t seed_id, seed_tok;
    const uint32_t* conv_map;             // ... or, when set, conv_map[0..n_conv): a hash shard's members
};

uint64_t one_record(const Shape& sh, uint32_t i_local, uint32_t j, uint8_t* dst) {
    uint8_t cid[64], tok[64];
    const uint32_t i = sh.conv_map ? sh.conv_map[i_local] : sh.conv_first + i_local;
    size_t cl, tl;
    const uint8_t* tp = tok;
    if (sh.config == 0) {
        cl = loadtest_id(cid, i);
        tp = reinterpret_cast<const uint8_t*>(kWords[j % 56]);
        tl = strlen(kWords[j % 56]);
    } else {
        cl = uuid_id(cid, sh.seed_id, i);
        tl = llm_token(tok, sh.seed_tok, i, j);
    }
    Out o{dst};
    put_message(o, cid, cl, tp, tl, (int64_t)j + 1, j == sh.n_tok - 1, kClock + (int64_t)j * 50000000ll + i);
    return o.n;
}

}  // namespace

extern "C" {

// Fills `off` (n_conv*n_tok+1 entries) and, when arena != nullptr, the packed payload bytes.
// Call once with arena == nullptr to learn the byte total, then again with the arena.
// Returns the total payload bytes, or 0 if `cap` is too small / the total exceeds 4 GiB-1.
uint64_t sseload_generate(int config, uint32_t n_conv, uint32_t n_tok, uint32_t conv_first, const uint32_t* conv_map,
                          uint64_t seed_id, uint64_t seed_tok, uint8_t* arena, uint64_t cap, uint32_t* off, int threads) {
    const Shape sh{config, n_conv, n_tok, conv_first, seed_id, seed_tok, conv_map};
    const uint64_t n = (uint64_t)n_conv * n_tok;
    if (threads < 1) threads = 1;
    std::vector<uint64_t> part((size_t)threads + 1, 0);
    auto span = [&](int t, uint64_t& lo, uint64_t& hi) { lo = n * t / threads; hi = n * (t + 1) / threads; };
    {   // pass 1: lengths
        std::vector<std::thread> th;
        for (int t = 0; t < threads; t++)
            th.emplace_back([&, t] {
                uint64_t lo, hi, sum = 0;
                span(t, lo, hi);
                for (uint64_t r = lo; r < hi; r++) {
                    uint64_t len = one_record(sh, (uint32_t)(r % n_conv), (uint32_t)(r / n_conv), nullptr);
                    off[r + 1] = (uint32_t)len;      // length for now; prefix-summed below
                    sum += len;
                }
                part[(size_t)t + 1] = sum;
            });
        for (auto& x : th) x.join();
    }
    for (int t = 0; t < threads; t++) part[(size_t)t + 1] += part[(size_t)t];
    const uint64_t total = part[(size_t)threads];
    if (total > 0xFFFFFFF0ull) return 0;
    {   // prefix sum per thread span
        std::vector<std::thread> th;
        for (int t = 0; t < threads; t++)
            th.emplace_back([&, t] {
                uint64_t lo, hi;
                span(t, lo, hi);
                uint64_t acc = part[(size_t)t];
                for (uint64_t r = lo; r < hi; r++) { acc += off[r + 1]; off[r + 1] = (uint32_t)acc; }
            });
        for (auto& x : th) x.join();
    }
    off[0] = 0;
    if (!arena) return total;
    if (cap < total + 16) return 0;
    {   // pass 2: bytes
        std::vector<std::thread> th;
        for (int t = 0; t < threads; t++)
            th.emplace_back([&, t] {
                uint64_t lo, hi;
                span(t, lo, hi);
                for (uint64_t r = lo; r < hi; r++) one_record(sh, (uint32_t)(r % n_conv), (uint32_t)(r / n_conv), arena + off[r]);
            });
        for (auto& x : th) x.join();
    }
    memset(arena + total, 0, 16);
    return total;
}

// The same messages as sseload_generate, as the producer sees them BEFORE marshalling (llm-stream-proxy/main.go:230-237):
// raw conversation ids and token texts packed back to back with offsets, plus sequence / done / timestamp columns.
// cid / tok may be nullptr to learn the byte totals (returned through *cid_total / *tok_total).  Single pass per column,
// serial prefix: this is bench / test input preparation, not a measured path.
int sseload_generate_columns(int config, uint32_t n_conv, uint32_t n_tok, uint32_t conv_first, const uint32_t* conv_map,
                             uint64_t seed_id, uint64_t seed_tok, uint8_t* cid, uint64_t cid_cap, uint32_t* cid_off,
                             uint8_t* tok, uint64_t tok_cap, uint32_t* tok_off, int64_t* seq, uint8_t* done, int64_t* ts,
                             uint64_t* cid_total, uint64_t* tok_total) {
    const uint64_t n = (uint64_t)n_conv * n_tok;
    uint64_t cw = 0, tw = 0;
    for (uint64_t r = 0; r < n; r++) {
        const uint32_t i_local = (uint32_t)(r % n_conv), j = (uint32_t)(r / n_conv);
        const uint32_t i = conv_map ? conv_map[i_local] : conv_first + i_local;
        uint8_t cb[64], tb[64];
        size_t cl, tl;
        const uint8_t* tp = tb;
        if (config == 0) {
            cl = loadtest_id(cb, i);
            tp = reinterpret_cast<const uint8_t*>(kWords[j % 56]);
            tl = strlen(kWords[j % 56]);
        } else {
            cl = uuid_id(cb, seed_id, i);
            tl = llm_token(tb, seed_tok, i, j);
        }
        if (cw + cl > 0xFFFFFFF0ull || tw + tl > 0xFFFFFFF0ull) return -1;
        if (cid) { if (cw + cl > cid_cap) return -2; memcpy(cid + cw, cb, cl); }
        if (tok) { if (tw + tl > tok_cap) return -2; memcpy(tok + tw, tp, tl); }
        if (cid_off) cid_off[r] = (uint32_t)cw;
        if (tok_off) tok_off[r] = (uint32_t)tw;
        if (seq) seq[r] = (int64_t)j + 1;
        if (done) done[r] = j == n_tok - 1;
        if (ts) ts[r] = kClock + (int64_t)j * 50000000ll + i;
        cw += cl; tw += tl;
    }
    if (cid_off) cid_off[n] = (uint32_t)cw;
    if (tok_off) tok_off[n] = (uint32_t)tw;
    if (cid_total) *cid_total = cw;
    if (tok_total) *tok_total = tw;
    return 0;
}

// Conversation id of conversation i under a config (for sharding / interest tables). Returns length.
uint32_t sseload_conversation_id(int config, uint64_t seed_id, uint32_t i, uint8_t* dst) {
    return (uint32_t)(config == 0 ? loadtest_id(dst, i) : uuid_id(dst, seed_id, i));
}

// Members of one GPU's shard: conversations i in [0, n_total) with fnv1a64(id_i) % world == rank
// (the host-side routing rule).  Writes up to cap indices, returns the member count.
uint32_t sseload_shard_members(int config, uint64_t seed_id, uint32_t n_total, uint32_t rank, uint32_t world,
                               uint32_t* out, uint32_t cap) {
    uint32_t m = 0;
    uint8_t id[64];
    for (uint32_t i = 0; i < n_total; i++) {
        size_t len = config == 0 ? loadtest_id(id, i) : uuid_id(id, seed_id, i);
        uint64_t h = 0xcbf29ce484222325ull;
        for (size_t k = 0; k < len; k++) { h ^= id[k]; h *= 0x100000001b3ull; }
        if (world == 0 || h % world == rank) { if (m < cap) out[m] = i; m++; }
    }
    return m;
}

}  // extern "C"
