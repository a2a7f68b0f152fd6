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
