// hostcheck.cpp -- TEST DOUBLE.  Compiles the product's per-record device logic
// (csrc/sse_record.cuh, all __host__ __device__) for the CPU so that the -m "not gpu"
// suite can diff it against the oracle on this GPU-less build box before GPU minutes are
// spent.  It is built into tests/hostcheck/ only, is never loaded by the product library,
// the bench or smoke(), and is NOT a CPU fallback: the shipped path is CUDA only.
#include <cstdint>
#include <cstring>
#include <vector>
#include "sse_record.cuh"

using namespace sse;

extern "C" int hc_batch(const uint8_t* in, const uint32_t* in_off, uint32_t n, const int64_t* start_seq,
                        uint32_t op, uint8_t* out, uint64_t out_cap, uint64_t* out_off, uint32_t* subj_len,
                        uint32_t* status, uint32_t* n_fast) {
    std::vector<uint32_t> nest(NEST_WORDS);
    PtrSrc src{in};
    PtrDst dst{out};
    uint64_t w = 0;
    uint32_t fast = 0;
    for (uint32_t i = 0; i < n; i++) {
        Plan pl;
        uint32_t len = 0, st = ST_OK;
        const uint32_t o = in_off[i], L = in_off[i + 1] - in_off[i];
        const int64_t start = start_seq ? start_seq[i] : 0;
        // same dispatch as k_frame_tiles: inline canonical path, else the out-of-line general path
        const bool fastp = plan_record_fast(src, o, L, op, start, pl, st, len);
        if (!fastp) st = plan_record(src, o, L, op, start, pl, nest.data(), len);
        out_off[i] = w;
        status[i] = st;
        if (subj_len) subj_len[i] = 0;
        if ((st & ST_CODE_MASK) == ST_OK) {
            if (w + len > out_cap) return -1;
            fast += fastp ? 1 : 0;
            Writer<PtrDst, true> wr(dst, w);
            if (fastp) emit_record_fast(src, o, L, pl, op, wr, subj_len ? &subj_len[i] : nullptr);
            else emit_record(src, o, L, pl, op, wr, subj_len ? &subj_len[i] : nullptr);
            if (wr.pos() != w + len) return -2;   // measured and emitted lengths must agree
            w += len;
        }
    }
    out_off[n] = w;
    if (n_fast) *n_fast = fast;
    return 0;
}

// ---- inspection: the product's pattern compiler + DFA walk on the CPU (test double, see top of file) ----
#include "inspect_compile.hpp"
#include "sse_inspect.cuh"

struct hc_pattern { uint32_t kind, action; const char* text; };

extern "C" int hc_inspect(const hc_pattern* pats, uint32_t n_pats, uint32_t mode, const uint8_t* in, const uint32_t* in_off,
                          uint32_t n, const uint32_t* conv_slot, uint32_t conv_slots, uint32_t* verdict, int32_t* match,
                          uint32_t* status, uint32_t* n_states) {
    std::vector<sse_inspect::PatternSpec> specs;
    for (uint32_t i = 0; i < n_pats; i++) specs.push_back({pats[i].kind, pats[i].action, pats[i].text});
    sse_inspect::Compiled c;
    std::string err;
    if (!sse_inspect::compile(specs, c, err)) return -1;
    InspectDev t{};
    auto fill = [](DfaDev& dd, const sse_inspect::Dfa& d) {
        dd.cls = d.cls; dd.next = d.next.data(); dd.acc_imm = d.acc_imm.data(); dd.acc_del = d.acc_del.data();
        dd.acc_eot = d.acc_eot.data(); dd.n_classes = d.n_classes; dd.start0 = d.start[0]; dd.start1 = d.start[1]; dd.n_states = d.n_states;
        dd.dist = d.dist.data();
    };
    fill(t.fwd, c.forward);
    if (c.anchored.size() > INSPECT_MAX_ANCHORED) return -2;
    for (size_t k = 0; k < c.anchored.size(); k++) fill(t.anchored[k], c.anchored[k]);
    t.n_patterns = n_pats; t.mode = mode;
    for (uint32_t i = 0; i < 32; i++) { t.action[i] = 0; t.fixed_len[i] = 0; t.anchored_of[i] = -1; }
    for (uint32_t i = 0; i < n_pats; i++) {
        t.action[i] = c.patterns[i].action;
        t.fixed_len[i] = c.patterns[i].min_len == c.patterns[i].max_len ? c.patterns[i].min_len : 0;
        t.anchored_of[i] = c.anchored_of[i];
    }
    t.min_len = 0xFFFFFFFFu;
    for (uint32_t i = 0; i < n_pats; i++) t.min_len = c.patterns[i].min_len < t.min_len ? c.patterns[i].min_len : t.min_len;
    if (n_states) *n_states = c.forward.n_states;
    std::vector<uint32_t> state(conv_slots ? conv_slots : 1, c.forward.start[0]);
    std::vector<uint32_t> nest(NEST_WORDS);
    PtrSrc src{in};
    // the kernel keeps the first few states' rows on chip; here the "cache" aliases the full table but is
    // deliberately small so that both the cached and the uncached branch of the walk run
    std::vector<uint32_t> eot_bits(DFA_EOT_WORDS, 0);
    if (c.forward.n_states <= 32 * DFA_EOT_WORDS)
        for (uint32_t st = 0; st < c.forward.n_states; st++) if (c.forward.acc_eot[st]) eot_bits[st >> 5] |= 1u << (st & 31);
    // per_token mode: packed rows as in the frame kernel (the walk stops when the rest of the token is too short for any match)
    const uint32_t n_cached = c.forward.n_states < 20 ? c.forward.n_states : 20;
    const bool packed = mode != INSPECT_STREAMING && dfa_rows_packable(t.fwd);
    std::vector<uint16_t> rows(c.forward.next.begin(), c.forward.next.begin() + (size_t)n_cached * c.forward.n_classes);
    if (packed) for (uint16_t& nx : rows) nx = dfa_pack_row(t.fwd, nx);
    DfaCache cache = make_dfa_cache(t, c.forward.cls, rows.data(), n_cached, eot_bits.data(), packed);
    const bool streaming = mode == INSPECT_STREAMING;
    for (uint32_t i = 0; i < n; i++) {
        Plan pl;
        uint32_t len = 0, st = ST_OK;
        const uint32_t o = in_off[i], L = in_off[i + 1] - in_off[i];
        const bool fastp = plan_record_fast(src, o, L, 0, 0, pl, st, len);
        if (!fastp) st = plan_record(src, o, L, 0, 0, pl, nest.data(), len);
        status[i] = st;
        verdict[i] = 0; match[2 * i] = 0; match[2 * i + 1] = 0;
        InspectResult r{0, 0, 0, 0};
        if (streaming) {
            // as on the GPU: the frame pass leaves a token entry, the per-conversation walk consumes the entries in
            // arrival order (which is record order here)
            const uint32_t flags = ((st & ST_CODE_MASK) == ST_OK ? ENTRY_OK : 0u) | ((fastp && !pl.tok_esc) ? 0u : ENTRY_DECODE) |
                                   ((st & ST_DONE) ? ENTRY_DONE : 0u) | (fastp ? ENTRY_CANON : 0u);
            uint32_t e[ENTRY_WORDS];
            make_token_entry(src, o + pl.tok_b, pl.tok_e - pl.tok_b, o + pl.tok_b, flags, e);
            if (!(flags & ENTRY_OK)) continue;
            const uint32_t slot = conv_slot[i];
            r = stream_walk(t, cache, src, e, state[slot]);
            state[slot] = r.state;
        } else {
            if ((st & ST_CODE_MASK) != ST_OK) continue;
            bool done_fast = false;
            if (fastp) done_fast = inspect_canon_token(t, cache, src, o + pl.tok_b, o + pl.tok_e, pl.tok_esc != 0, r);
            if (!done_fast) r = inspect_token(t, src, o + pl.tok_b, o + pl.tok_e, 0, true);
        }
        verdict[i] = r.verdict; match[2 * i] = r.begin; match[2 * i + 1] = r.end;
    }
    return 0;
}

// ---- interest table: host builder + HD lookup on the CPU (test double) ----
#include "interest_table.hpp"

struct hc_delta { uint32_t op, sub_id; const char* subject; uint32_t subject_len; };

extern "C" int hc_interest(const hc_delta* d, uint32_t nd, const uint8_t* in, const uint32_t* in_off, uint32_t n, uint32_t* hit,
                           uint32_t* status, uint32_t* subs_out, uint32_t subs_cap, uint32_t* n_subs, uint32_t* wild_ids_out, uint32_t* n_wild) {
    sse_interest::HostTable t;
    std::string err;
    for (uint32_t i = 0; i < nd; i++) {
        std::string s(d[i].subject, d[i].subject_len);
        // every other delta goes through the incremental path once a flattened form exists (as ssegpu_interest_apply does)
        const int ch = d[i].op == 0 ? t.add(s, d[i].sub_id, err) : t.del(s, d[i].sub_id);
        if (ch < 0) return -1;
        if (i == nd / 2) t.flatten();
        else if (t.flattened && (ch == 2 || (ch == 1 && !t.sync_subject(s)))) t.flatten();
    }
    if (!t.flattened) t.flatten();
    if (t.subs->size() > subs_cap) return -2;
    memcpy(subs_out, t.subs->data(), t.subs->size() * 4);
    *n_subs = (uint32_t)t.subs->size();
    memcpy(wild_ids_out, t.wild_ids.data(), t.wild_ids.size() * 4);
    *n_wild = (uint32_t)t.wild_ids.size();
    InterestDev dev{};
    dev.slots = t.slots.data(); dev.keys = t.keys.data(); dev.mask = (uint32_t)t.slots.size() - 1;
    dev.n_wild = (uint32_t)t.wild_ids.size(); dev.wild_bytes = t.wild_bytes.data();
    dev.wild_single = t.wild_single; dev.wild_generic = t.wild_generic;
    for (size_t i = 0; i < t.wild_off.size(); i++) dev.wild_off[i] = t.wild_off[i];
    std::vector<uint32_t> nest(NEST_WORDS);
    PtrSrc src{in};
    for (uint32_t i = 0; i < n; i++) {
        Plan pl;
        uint32_t len = 0;
        const uint32_t o = in_off[i], L = in_off[i + 1] - in_off[i];
        uint32_t st = ST_OK;
        const bool fastp = plan_record_fast(src, o, L, OP_BRIDGE, 0, pl, st, len);      // same dispatch as the kernel
        if (!fastp) st = plan_record(src, o, L, OP_BRIDGE, 0, pl, nest.data(), len);
        hit[3 * i] = hit[3 * i + 1] = hit[3 * i + 2] = 0;
        if ((st & ST_CODE_MASK) == ST_OK) {
            InterestHit h{0, 0, 0};
            const InterestFast fast = make_interest_fast(dev);
            InterestPre pre;
            interest_prepare(fast, src, o + pl.cid_b, o + pl.cid_e, fastp && !pl.cid_esc, pre);     // the kernel calls this from parse_fast's hook
            if (pre.hash) interest_finish(fast, src, o + pl.cid_b, o + pl.cid_e, pre, h);
            else h = interest_lookup(dev, src, o + pl.cid_b, o + pl.cid_e);
            hit[3 * i] = h.list_off; hit[3 * i + 1] = h.list_cnt; hit[3 * i + 2] = h.wild_mask;
            if (h.list_cnt == 0 && h.wild_mask == 0) st = ST_NO_INTEREST;
        }
        status[i] = st;
    }
    return 0;
}

// producer-side encoder: the kernel's per-message steps (plain test, lengths, emission) run sequentially; records are
// packed back to back.
extern "C" int hc_encode(const uint8_t* cid, const uint32_t* cid_off, const uint8_t* tok, const uint32_t* tok_off, const int64_t* seq,
                         const uint8_t* done, const int64_t* ts, uint32_t n, uint32_t op, uint8_t* out, uint64_t out_cap,
                         uint64_t* out_off, uint32_t* chan_len) {
    PtrSrc c{cid}, t{tok};
    PtrDst dst{out};
    const bool go122 = (op & OP_GO122) != 0;
    uint64_t w = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t cb = cid_off[i], ce = cid_off[i + 1], tb = tok_off[i], te = tok_off[i + 1];
        const bool cp = go_string_is_plain(c, cb, ce), tp = go_string_is_plain(t, tb, te);
        const EncPlan p = encode_plan_from(ce - cb, cp ? ce - cb + 2 : go_string_len(c, cb, ce, go122), cp,
                                           tp ? te - tb + 2 : go_string_len(t, tb, te, go122), tp, seq[i], done[i] != 0, ts[i], op);
        if (w + p.len + 4 > out_cap) return -1;
        Writer<PtrDst, true> e(dst, w);
        encode_emit(c, cb, ce, t, tb, te, seq[i], done[i] != 0, ts[i], go122, p, e);
        if (e.pos() != w + p.len) return -2;
        chan_len[i] = p.chan;
        out_off[i] = w;
        w += p.len;
    }
    out_off[n] = w;
    return 0;
}

// ---- wire splitter: the product's per-connection walk on the CPU (test double) ----
#include "sse_wire.cuh"
// the word-at-a-time byte search of the wire walkers and the chunk decoder, on its own
extern "C" uint32_t hc_find_byte2(const uint8_t* buf, uint32_t p, uint32_t e, uint32_t c0, uint32_t c1) {
    PtrSrc src{buf};
    return find_byte2(src, p, e, c0, c1);
}
struct hc_wire_msg { uint32_t payload_off, payload_len, subj_off, subj_len; uint64_t sid; uint32_t conn, pad; };
struct hc_wire_conn { uint32_t n_msgs, payload_bytes, consumed, status, pings; };
extern "C" uint32_t hc_wire_split(uint32_t proto, const uint8_t* buf, const uint32_t* buf_off, uint32_t n_conn, hc_wire_msg* msgs, uint32_t cap,
                                  hc_wire_conn* conns) {
    PtrSrc src{buf};
    uint32_t n = 0;
    for (uint32_t c = 0; c < n_conn; c++) {
        const WireConn r = wire_walk(proto, src, buf_off[c], buf_off[c + 1], [&](uint32_t, const WireMsg& m) {
            if (n < cap) msgs[n] = hc_wire_msg{m.payload_off, m.payload_len, m.subj_off, m.subj_len, m.sid, c, 0};
            n++;
        });
        conns[c] = hc_wire_conn{r.n_msgs, r.payload_bytes, r.consumed, r.status, r.pings};
    }
    return n;
}

// ---- vLLM stream-chunk decoder: the product's per-line logic on the CPU (test double) ----
#include "sse_chunk.cuh"
extern "C" uint32_t hc_chunk_decode(const uint8_t* lines, const uint32_t* line_off, uint32_t n_lines, uint8_t* tok, uint32_t* tok_off, uint8_t* done,
                                    uint32_t* msg_line, uint32_t* line_status) {
    std::vector<uint32_t> nest(NEST_WORDS);
    PtrSrc src{lines};
    uint32_t n = 0, w = 0;
    for (uint32_t i = 0; i < n_lines; i++) {
        ChunkEvents ev;
        const uint32_t st = chunk_line(src, line_off[i], line_off[i + 1], ev, nest.data());
        line_status[i] = st;
        if (st & CHUNK_TOO_MANY) continue;
        if ((st & CHUNK_KIND_MASK) == CHUNK_DATA)
            for (uint32_t k = 0; k < ev.len; k++) {
                if (!(ev.ch[k].has_content && ev.ch[k].ce > ev.ch[k].cb)) continue;
                tok_off[n] = w; done[n] = 0; msg_line[n] = i;
                if (chunk_content_bytes(src, ev.ch[k]) == 0) return 0xFFFFFFFFu;
                for_each_decoded_byte(src, ev.ch[k].cb, ev.ch[k].ce, [&](uint32_t b) { tok[w++] = (uint8_t)b; });
                n++;
            }
        if (st & CHUNK_ENDS) { tok_off[n] = w; done[n] = 1; msg_line[n] = i; memcpy(tok + w, "[DONE]", 6); w += 6; n++; }
    }
    tok_off[n] = w;
    return n;
}
