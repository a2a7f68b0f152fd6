// sse_kernels.cu -- the fused B200 kernel of the SSE hot path.
//
// k_frame_tiles: decode -> (filter) -> encode -> frame -> exclusive scan of frame lengths ->
// compaction, in ONE pass over HBM.  It stands in for the per-token body of the adapter's two
// hot loops (src/sse-adapter/sse_handler.go:135-157 + :186-233, :316-333 + :400-449) and of
// the bridge loop (src/redis-nats-bridge/main.go:136-181), batched over every conversation
// that has a token pending.
//
// Shape (sm_100a):
//   * persistent CTAs (grid = SMs x resident CTAs), tiles of TILE records handed out by an
//     atomic ticket, claimed two trips ahead by thread 0 (the next tile's byte range is loaded
//     during the current tile's decode: two barriers per tile, none at the top);
//   * a tile's payload bytes arrive in shared memory as ONE TMA bulk copy
//     (cp.async.bulk ... mbarrier::complete_tx), 16-byte aligned, no register staging; the
//     tile's record offsets (and, in ordered mode's emit pass, its plan words) ride along as
//     further bulk copies on the same mbarrier;
//   * one thread per record runs sse_record.cuh (SWAR canonical check / general decoder);
//   * frame lengths: warp-shuffle scan -> CTA scan; the tile's slice of the arena comes from an
//     atomic cursor (arrival order, the default) or, in ordered mode, from a counting pass +
//     a scan over the tile totals (see "ordered placement" below);
//   * frames are assembled in a shared-memory staging tile BEFORE the tile's global offset is
//     needed (the slice comes back from the atomic cursor while the frames are being written),
//     then leave as ONE TMA bulk store into the tile's 16-byte aligned slice (ordered mode's
//     emit pass knows the offset up front, stages at that offset modulo 16 and stores the
//     aligned interior the same way, plus at most 15 head and 15 tail bytes by threads): HBM
//     sees only full-line traffic.
//   No tensor cores: nothing here is a contraction.
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include "sse_kernels.h"
#include "sse_record.cuh"
#include "sse_inspect.cuh"
#include "sse_interest.cuh"

namespace sse {

// ------------------------------------------------------------------ PTX helpers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA, 1-D).  dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global bulk copy (TMA, 1-D).  16-byte aligned on both sides, bytes a multiple of 16.
__device__ __forceinline__ void tma_store_1d(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
// global -> L2 prefetch of a byte range (16-byte aligned address, size a multiple of 16): no shared memory is tied up
__device__ __forceinline__ void l2_prefetch(const void* gsrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ------------------------------------------------------------ ordered placement ----
// SSEGPU_OP_ORDERED lays the frames out back to back in record order.  A tile's offset is then the sum of ALL
// earlier tiles' byte counts, i.e. it depends on the decode of every earlier record.  A single-pass decoupled
// look-back was tried first (round 1, 2.2-2.3 ms per 10 M records): with ~740 persistent CTAs in flight every tile
// waits for the slowest of hundreds of predecessors, the CTAs fall into lock-step waves and the 5 CTAs per SM stop
// hiding one another's latencies (round 2 tried again with a two-level look-back: profiles/r02_ablations.md).  Ordered
// mode is therefore two passes over the payload with a kernel boundary as the global synchronisation: pass A decodes,
// records each tile's byte count and leaves a plan word per canonical record, a scan turns the counts into offsets,
// pass B emits at the known offsets from the plan words (it decodes only what pass A could not plan).  The default
// (arrival-order) mode stays a single pass.
constexpr int MODE_ARRIVAL = 0, MODE_ORDERED_COUNT = 1, MODE_ORDERED_EMIT = 2;

// Tile totals -> exclusive prefixes, in place, in two small launches of one CTA per 1024 tiles: block sums first (parked in
// out_off[0 .. blocks), which nothing has written yet and the emit pass overwrites afterwards), then every CTA adds up
// the sums of the blocks in front of it and scans its own 1024 totals.  (One CTA scanning all 78 125 tiles of the
// BASELINE batch took 82 us between the two passes.)
constexpr uint32_t SCAN_BLOCK = 1024;
__device__ __forceinline__ uint64_t block_reduce_u64(uint64_t v, uint64_t* warp_sums) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    __syncthreads();                                   // warp_sums may still be read from a previous use
    if (lane == 0) warp_sums[warp] = v;
    __syncthreads();
    uint64_t t = 0;
#pragma unroll
    for (uint32_t w = 0; w < SCAN_BLOCK / 32; w++) t += warp_sums[w];
    return t;
}
__global__ void __launch_bounds__(SCAN_BLOCK) k_tile_block_sums(const uint64_t* tile_state, uint32_t num_tiles, uint64_t* block_sums) {
    __shared__ uint64_t warp_sums[SCAN_BLOCK / 32];
    const uint32_t i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    const uint64_t t = block_reduce_u64(i < num_tiles ? tile_state[i] : 0, warp_sums);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = t;
}
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_tile_totals(uint64_t* tile_state, uint32_t num_tiles, const uint64_t* block_sums,
                                                                 uint64_t* out_total, uint32_t* ticket) {
    __shared__ uint64_t warp_sums[SCAN_BLOCK / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t before = 0;
    for (uint32_t b = tid; b < blockIdx.x; b += SCAN_BLOCK) before += block_sums[b];
    before = block_reduce_u64(before, warp_sums);
    const uint32_t i = blockIdx.x * SCAN_BLOCK + tid;
    const uint64_t mine = i < num_tiles ? tile_state[i] : 0;
    uint64_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint64_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (uint32_t)d) incl += v;
    }
    __syncthreads();
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint64_t wpre = 0, total = 0;
#pragma unroll
    for (uint32_t w = 0; w < SCAN_BLOCK / 32; w++) { const uint64_t v = warp_sums[w]; if (w < warp) wpre += v; total += v; }
    if (i < num_tiles) tile_state[i] = before + wpre + incl - mine;
    if (tid == 0 && blockIdx.x == gridDim.x - 1) *out_total = before + total;
    if (tid == 0 && blockIdx.x == 0) *ticket = 0;               // pass B hands the tiles out again from 0
}

// ---------------------------------------------------------------------- kernel ----
#ifndef SSE_PREFETCH_WAVES
#define SSE_PREFETCH_WAVES 1
#endif
constexpr uint32_t PREFETCH_WAVES = SSE_PREFETCH_WAVES;
// A full tile's TILE + 1 record offsets (in_off[r0 .. r0 + TILE]) come into shared memory as a second bulk copy on the
// payload's mbarrier: the threads then read their offsets on chip instead of waiting for a global load at the top of
// every tile.  OFF_BYTES = (TILE + 1) words rounded up to the 16 bytes a bulk copy moves.
constexpr uint32_t OFF_BYTES = ((TILE + 1) * 4 + 15) & ~15u;
// FrameArgs::plan, one word per record, left by the counting pass of ordered mode for the emit pass.  A canonical
// record: where its sequence literal starts (13 bits), where its id ends (11), the literal's digits (4), done, filtered
// -- everything its frame depends on besides the payload itself, so the emit pass neither decodes it again nor loads its
// status and length.  The reserved values:
constexpr uint32_t PLAN_REDO = 0xFFFFFFFFu;             // decode this record again in the emit pass (status / out_len hold the counting pass's result)
constexpr uint32_t PLAN_HOLE = 0xFFFFFFFEu;             // emit pass -> its tail: this record's frame is still to be written
constexpr uint32_t PLAN_PENDING = 0xFFFFFFFDu;          // counting pass -> its tail: not decoded yet
constexpr uint32_t PLAN_BYTES = TILE * 4;               // a full tile's plan words reach the emit pass by bulk copy, like the offsets
constexpr uint32_t DEFER_MAX_PER_WARP = 9;              // more such records than this in a warp: decoded inline, where the tile is on chip
struct TileSmem {
    uint64_t mbar;
    uint64_t base;           // global byte offset of the tile's output
    uint64_t obase;          // ordered emit pass: the same, known as soon as the tile is claimed (k_scan_tile_totals)
    // Per-tile bookkeeping, double-buffered by the trip's parity: thread 0 parks the NEXT tile's entry during the current
    // trip's decode (before barrier (B)), so every thread finds it at the top of the next trip without another barrier.
    // npend: records of the tile left to the tail phase.
    struct Meta { uint32_t tile, tb, te, npend; } meta[2];
    uint32_t in_total;
    uint32_t wsum[2][TILE / 32];   // double-buffered too: the counting pass has one barrier per tile
    // the CTA's share of the launch counters (ok, syntax, type, filtered, done, fast; bytes in / out): kept here, not in
    // eight registers of every thread across the whole persistent loop
    uint32_t ctr[6];
    unsigned long long c_in, c_out;
    // thread 0's loop state.  Tiles are claimed two trips ahead so that nothing on the way to a tile's TMA load waits for
    // memory after the barrier: next_t is the next trip's tile, ntb / nte its byte range (loaded during this trip's
    // decode), next2 the tile after that; my_chain: last tile (+1) of this CTA that left records to the tail phase.
    uint32_t next_t, next2, my_chain, pad_[3];
};

// Out-of-line general path (generic pointers: the tile may sit in shared or in global memory).
// Rare by construction -- canonical records never get here -- so it is kept out of the hot
// loop's instruction footprint, and the 10000-level nesting stack lives in ITS frame only.
__device__ __noinline__ uint32_t plan_slow(const uint8_t* src_base, uint32_t o, uint32_t L, uint32_t op, int64_t start_seq,
                                           Plan* pl, uint32_t* out_len) {
    uint32_t nest[NEST_WORDS];
    PtrSrc src{src_base};
    uint32_t len = 0;
    uint32_t st = plan_record(src, o, L, op, start_seq, *pl, nest, len);
    *out_len = len;
    return st;
}
__device__ __noinline__ uint32_t emit_slow(const uint8_t* src_base, uint32_t o, uint32_t L, const Plan* pl, uint32_t op,
                                           uint8_t* dst_base, uint64_t dst_pos) {
    PtrSrc src{src_base};
    PtrDst d{dst_base};
    Writer<PtrDst, true> w(d, dst_pos);
    uint32_t subj = 0;
    emit_record(src, o, L, *pl, op, w, &subj);
    return subj;
}

// Does a tile's payload come by TMA, and do its record offsets ride along?  (Every thread evaluates this for the
// current tile; thread 0 for the tile it loads.)
struct TileShape { uint32_t abase, bytes; bool smem_in, with_off; };
__device__ __forceinline__ TileShape tile_shape(const FrameArgs& a, uint32_t t, uint32_t tb, uint32_t te) {
    TileShape sh;
    sh.abase = tb & ~15u;
    sh.bytes = (te - sh.abase + 15u) & ~15u;
    sh.smem_in = sh.bytes <= a.in_cap;
    // the offsets ride along when all OFF_BYTES of them exist (every tile but the batch's last one or two)
    sh.with_off = sh.smem_in && sh.bytes != 0 && a.off_tma && (size_t)t * TILE * 4 + OFF_BYTES <= ((size_t)a.n + 1) * 4;
    return sh;
}
// thread 0, during the decode of the current tile: bookkeeping entry of tile t, whose byte range (tb, te) =
// in_off[first record], in_off[one past the last record] the caller has loaded
__device__ __forceinline__ void park_tile(const FrameArgs& a, TileSmem* ts, uint32_t slot, uint32_t t, uint32_t tb, uint32_t te) {
    if (t < a.num_tiles) {
        // offsets come from the host: clamp them so that nothing below can leave the arena, and say so
        const uint32_t in_total = ts->in_total;
        if (te > in_total || tb > te) { atomicExch(a.bad_offsets, 1u); te = min(te, in_total); tb = min(tb, te); }
    }
    ts->meta[slot].tile = t; ts->meta[slot].tb = tb; ts->meta[slot].te = te; ts->meta[slot].npend = 0;
}
// thread 0, once the input buffer is free: TMA load of the parked tile (payload and, when they all exist, its offsets)
__device__ __forceinline__ void launch_tile(const FrameArgs& a, TileSmem* ts, uint8_t* s_in, uint32_t slot, bool want_base, bool want_plan = false) {
    const uint32_t t = ts->meta[slot].tile;
    if (t >= a.num_tiles) return;
    uint64_t ob = 0;
    if (want_base) ob = a.tile_state[t];                    // requested first: in flight while the TMA load is set up
    const uint32_t tb = ts->meta[slot].tb, te = ts->meta[slot].te;
    const TileShape sh = tile_shape(a, t, tb, te);
    if (sh.smem_in && sh.bytes != 0) {
        const bool with_plan = want_plan && sh.with_off && a.plan_tma;
        mbar_expect_tx(&ts->mbar, sh.bytes + (sh.with_off ? OFF_BYTES : 0u) + (with_plan ? PLAN_BYTES : 0u));
        tma_load_1d(s_in, a.in + sh.abase, sh.bytes, &ts->mbar);
        if (sh.with_off) tma_load_1d(reinterpret_cast<uint8_t*>(ts + 1), a.in_off + (size_t)t * TILE, OFF_BYTES, &ts->mbar);
        if (with_plan) tma_load_1d(reinterpret_cast<uint8_t*>(ts + 1) + OFF_BYTES, a.plan + (size_t)t * TILE, PLAN_BYTES, &ts->mbar);
    }
    ts->c_in += te - tb;
    if (want_base) ts->obase = ob;
}

// Inspection of one record's token text, general case (out of line, generic pointers): records that are not
// canonical, or the leftmost-start search of a variable-length match.  per_token mode only: streaming mode is a
// per-conversation walk in sse_stream.cu.
__device__ __noinline__ InspectResult inspect_slow(const InspectDev* t, const uint8_t* src_base, uint32_t tok_b, uint32_t tok_e) {
    PtrSrc src{src_base};
    return inspect_token(*t, src, tok_b, tok_e, 0, true);
}
__device__ __noinline__ InterestHit interest_slow(const FrameArgs& a, const uint8_t* src_base, uint32_t o, uint32_t cid_b, uint32_t cid_e) {
    PtrSrc src{src_base};
    return interest_lookup(*a.interest, src, o + cid_b, o + cid_e);
}
__device__ __noinline__ uint32_t measure_slow(const uint8_t* src_base, uint32_t o, uint32_t L, const Plan* pl, uint32_t op) {
    PtrSrc src{src_base};
    return measure_record(src, o, L, *pl, op);
}
// Token entry of record i for the streaming walk (sb + o = the record's first byte, abs_b its offset in the arena)
__device__ __forceinline__ void store_token_entry(const FrameArgs& a, uint32_t i, const uint8_t* sb, uint32_t o, uint32_t abs_b, const Plan& pl,
                                                  uint32_t st, bool raw_text, bool canonical) {
    PtrSrc src{sb};
    const uint32_t flags = ((st & ST_CODE_MASK) == ST_OK ? ENTRY_OK : 0u) | (raw_text ? 0u : ENTRY_DECODE) | ((st & ST_DONE) ? ENTRY_DONE : 0u) |
                           (canonical ? ENTRY_CANON : 0u);
    uint32_t e[ENTRY_WORDS];
    make_token_entry(src, o + pl.tok_b, pl.tok_e - pl.tok_b, abs_b + pl.tok_b, flags, e);
    a.tok_entry[2 * (size_t)i] = make_uint4(e[0], e[1], e[2], e[3]);
    a.tok_entry[2 * (size_t)i + 1] = make_uint4(e[4], e[5], e[6], e[7]);
}
// The optional stages for one record that went through the GENERAL decoder (out of line: inline in a few places,
// densely in the tail phase).  In: the record's plan, status and length.  Out: verdict/match/hit/entry stored, status
// and length updated (inline drop / redact, nobody interested).
__device__ __noinline__ void ext_stage_slow(const FrameArgs& a, uint32_t i, const uint8_t* sb, uint32_t o, uint32_t L, uint32_t abs_b, Plan* plp,
                                            uint32_t* stp, uint32_t* lenp, bool write_results, bool write_entry) {
    Plan pl = *plp;
    uint32_t st = *stp, len = *lenp;
    const bool ok = (st & ST_CODE_MASK) == ST_OK;
    if (a.insp != nullptr || (a.xflags & XF_PRE_VERDICT)) {
        InspectResult io{0, 0, 0, 0};
        if (ok) {
            if (a.xflags & XF_PRE_VERDICT) io.verdict = a.verdict[i];
            else io = inspect_slow(a.insp, sb, o + pl.tok_b, o + pl.tok_e);
            if (a.op & OP_INSPECT_INLINE) {                                   // sse_handler.go:200-207
                const uint32_t act = io.verdict & 0xFF;
                if (act == ACT_DROP) { st = ST_INSPECT_DROP; len = 0; }
                else if (act == ACT_REDACT) { pl.tok_redacted = 1; pl.fast = 0; len = measure_slow(sb, o, L, &pl, a.op); st |= ST_REDACTED; }
            }
        }
        if (write_results && !(a.xflags & XF_PRE_VERDICT)) {
            a.verdict[i] = io.verdict;
            reinterpret_cast<int2*>(a.match)[i] = make_int2(io.begin, io.end);
        }
    }
    if (a.tok_entry != nullptr && write_entry) store_token_entry(a, i, sb, o, abs_b, *plp, *stp, false, false);
    if (a.interest != nullptr) {
        InterestHit h{0, 0, 0};
        if ((st & ST_CODE_MASK) == ST_OK) {
            h = interest_slow(a, sb, o, pl.cid_b, pl.cid_e);
            if (h.list_cnt == 0 && h.wild_mask == 0) { st = ST_NO_INTEREST; len = 0; }
        }
        if (write_results) { const size_t k = 3 * (size_t)i; a.hit[k] = h.list_off; a.hit[k + 1] = h.list_cnt; a.hit[k + 2] = h.wild_mask; }
    }
    *plp = pl; *stp = st; *lenp = len;
}

// EXT is a compile-time switch so that the plain framing kernel carries none of the registers of the optional
// stages (per-token content inspection, interest lookup, token entries for the streaming walk).
constexpr int X_INSPECT = 1, X_INTEREST = 2;    // compile-time set of optional stages (each variant pays only for its own registers)
template <int EXT, int MODE>
__global__ void __launch_bounds__(TILE, (MODE == MODE_ORDERED_COUNT && EXT == 0) ? 8 : 640 / TILE) k_frame_tiles(const __grid_constant__ FrameArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* s_in = smem;                                   // in_cap + 32 (word reads run up to 24 bytes past a record)
    uint8_t* s_out = smem + a.in_cap + 32;                  // out_cap_s + 48
    TileSmem* ts = reinterpret_cast<TileSmem*>(s_out + a.out_cap_s + 48);

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Launch counters: the variants with optional stages sit at the register cap and keep their counts in shared memory
    // (ts->ctr, one ballot per kind and tile); the plain kernel has the registers and counts per thread.
    constexpr bool SMEM_CTR = EXT != 0;
    uint32_t c_ok = 0, c_syn = 0, c_typ = 0, c_fil = 0, c_done = 0, c_fast = 0;
    uint32_t parity = 0;

    // EXT, fused per-token inspection: on-chip copy of the forward DFA's byte-class map, end-of-text bitmap and hottest rows
    const bool pre_verdict = (EXT & X_INSPECT) && (a.xflags & XF_PRE_VERDICT) != 0;          // launch-uniform
    const bool inspect_fused = (EXT & X_INSPECT) && a.insp != nullptr && !pre_verdict;
    DfaCache dcache{nullptr, nullptr, 0, nullptr, nullptr, 0, 0, 0, 0};
    if (inspect_fused) {
        uint8_t* s_cls = reinterpret_cast<uint8_t*>(ts + 1) + OFF_BYTES;
        uint32_t* s_eot = reinterpret_cast<uint32_t*>(s_cls + 256);
        uint16_t* s_rows = reinterpret_cast<uint16_t*>(s_eot + DFA_EOT_WORDS);
        const DfaDev& d = a.insp->fwd;
        // the descriptor's scalars come from the kernel parameters (a.dfa), not from the descriptor in global memory
        const uint32_t n_states = a.dfa.n_states, n_classes = a.dfa.n_classes;
        const uint32_t nc = min(n_states, DFA_CACHE_ENTRIES / n_classes);
        const bool packed = a.dfa.dist != nullptr && n_states <= DFA_PACK_MAX_STATES;
        for (uint32_t i = tid; i < 256; i += TILE) s_cls[i] = d.cls[i];
        for (uint32_t i = tid; i < nc * n_classes; i += TILE) {
            const uint16_t nx = a.dfa.next[i];
            s_rows[i] = packed ? (uint16_t)(nx | ((uint32_t)a.dfa.dist[nx & 0x7FFFu] << DFA_PACK_SHIFT)) : nx;
        }
        if (n_states <= 32 * DFA_EOT_WORDS)
            for (uint32_t w = tid; w < DFA_EOT_WORDS; w += TILE) {
                uint32_t bits = 0;
                for (uint32_t k = 0; k < 32 && 32 * w + k < n_states; k++) bits |= (d.acc_eot[32 * w + k] != 0 ? 1u : 0u) << k;
                s_eot[w] = bits;
            }
        dcache = make_dfa_cache(a.dfa.next, n_classes, n_states, a.dfa.start0, a.dfa.min_len, s_cls, s_rows, nc, s_eot, packed);
    }
    InterestFast ifast{nullptr, 0, 0, 0};
    if (EXT & X_INTEREST) ifast = InterestFast{a.itab.slots, a.itab.mask, a.itab.wild_single, a.itab.wild_generic};
    if (tid == 0) {
        mbar_init(&ts->mbar, 1);
        ts->in_total = a.in_off[a.n];
        for (int k = 0; k < 6; k++) ts->ctr[k] = 0;
        ts->my_chain = 0;
        ts->c_in = 0; ts->c_out = 0;
        const uint32_t t0 = atomicAdd(a.ticket, 1u);
        uint32_t b0 = 0, e0 = 0;
        if (t0 < a.num_tiles) { b0 = a.in_off[t0 * TILE]; e0 = a.in_off[min((t0 + 1) * TILE, a.n)]; }
        park_tile(a, ts, 0, t0, b0, e0);
        launch_tile(a, ts, s_in, 0, MODE == MODE_ORDERED_EMIT, MODE == MODE_ORDERED_EMIT && !EXT && a.plan != nullptr);
        ts->next_t = atomicAdd(a.ticket, 1u);
    }
    __syncthreads();                                                          // the first tile's entry is visible

    // Barriers per tile: (B) behind the decode, (D) behind the staging (the counting pass of ordered mode: (B) only).
    // Nothing is needed at the top: the tile's entry was parked before the previous trip's (B).  What keeps a fast warp of
    // trip k+1 off a slow warp of trip k:
    //   meta[slot], wsum[slot]   written before (B) of the trip that uses them next; the slot's previous user (two trips
    //                            back) was read before that trip's (B), which every thread has passed since;
    //   s_in, the offsets, plan  overwritten by the TMA load thread 0 issues behind (D) (behind (B) in the counting pass),
    //                            i.e. after every thread has finished reading them; readers of the new contents wait on
    //                            the mbarrier, whose phase cannot complete before thread 0's arrive.expect_tx;
    //   s_out                    written behind (B); the previous tile's bulk store has finished reading it (thread 0's
    //                            bulk_wait_read0 in front of (B));
    //   base / obase             written by thread 0 behind (B) / (D), read behind the following barrier, rewritten only
    //                            after the next (B) / (D), by which time every reader has arrived there.
    for (uint32_t trip = 0;; trip++) {
        const uint32_t slot = trip & 1u;
        const uint32_t tile = ts->meta[slot].tile;
        if (tile >= a.num_tiles) break;
        // thread 0: the tile after next is claimed and the next tile's byte range requested now; the results are parked in
        // shared memory after the decode (nothing waits for them here)
        uint32_t claim = 0, nb = 0, ne = 0, nt = 0;
        if (tid == 0) {
            nt = ts->next_t;
            claim = atomicAdd(a.ticket, 1u);
            if (nt < a.num_tiles) { nb = a.in_off[nt * TILE]; ne = a.in_off[min((nt + 1) * TILE, a.n)]; }
        }
        const uint32_t r0 = tile * TILE;
        const uint32_t nrec = min((uint32_t)TILE, a.n - r0);
        const uint32_t tb = ts->meta[slot].tb, te = ts->meta[slot].te;
        const TileShape shape = tile_shape(a, tile, tb, te);
        const uint32_t abase = shape.abase, in_bytes = shape.bytes;
        const bool smem_in = shape.smem_in;

        // One wave ahead: the payload bytes of tile (tile + gridDim.x) are pulled into L2 now, so that the TMA load of
        // whichever CTA gets that tile pays an L2 hit instead of a DRAM round trip.  The bytes in flight towards DRAM are
        // then no longer limited by the shared-memory buffers (which spend most of their cycle being parsed/copied).
        uint32_t pf_b = 0, pf_e = 0;
        const uint32_t pf_tile = tile + gridDim.x * PREFETCH_WAVES;
        const bool pf = PREFETCH_WAVES != 0 && tid == TILE / 2 && pf_tile < a.num_tiles;
        if (pf) {
            pf_b = a.in_off[pf_tile * TILE];
            pf_e = a.in_off[min((pf_tile + 1) * TILE, a.n)];
        }

        uint32_t my_b = 0, my_e = 0;
        int64_t my_start = 0;
        const bool active = tid < nrec;
        const bool off_smem = shape.with_off;                                 // tile-uniform
        if (active) {
            if (!off_smem) { my_b = a.in_off[r0 + tid]; my_e = a.in_off[r0 + tid + 1]; }
            if (a.start_seq) my_start = a.start_seq[r0 + tid];
        }
        if (smem_in && in_bytes != 0) { mbar_wait(&ts->mbar, parity); parity ^= 1; }
        if (active) {
            if (off_smem) {
                const uint32_t* s_off = reinterpret_cast<const uint32_t*>(ts + 1);
                my_b = s_off[tid]; my_e = s_off[tid + 1];
            }
            if (my_b < tb || my_e > te || my_b > my_e) {                     // malformed offsets: clamp into the tile
                atomicExch(a.bad_offsets, 1u);
                my_e = min(max(my_e, tb), te); my_b = min(max(my_b, tb), my_e);
            }
        }

        // ---- pass 1: decode + filter, output length --------------------------------
        // Arrival mode: a record that is not canonical is left to the tail phase of this CTA, which runs the general
        // decoder (and the general forms of the optional stages) over all such records densely, one per lane.  Inline,
        // ONE such record makes its whole warp -- and at the next barrier its CTA -- wait ten times as long as a
        // canonical tile takes (measured: 1 % of them cost 7x).
        constexpr bool DEFER = MODE == MODE_ARRIVAL;
        Plan pl;
        uint32_t st = ST_OK, len = 0;
        const uint32_t o = my_b - abase, L = my_e - my_b;
        const uint8_t* sb = smem_in ? s_in : a.in + abase;                    // where this tile's bytes are read from
        bool fastp = false;
        bool planned = false;
        // Ordered mode with plan words (no optional stages): both passes leave the records that are not canonical to
        // their tail phases too.  The count pass's tail adds their lengths to the tile totals before the scan; the emit
        // pass places such a record as a hole of the known length and its tail writes the frame into it.
        const bool defer_ordered = (MODE == MODE_ORDERED_COUNT || MODE == MODE_ORDERED_EMIT) && !EXT && a.plan != nullptr;   // launch-uniform
        bool hole = false;
        if (MODE == MODE_ORDERED_EMIT && defer_ordered) {
            uint32_t pw = 0;
            if (active) {                                                     // left by the count pass; on chip with the payload when the tile is full
                if (off_smem && a.plan_tma) pw = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(ts + 1) + OFF_BYTES)[tid];
                else pw = a.plan[r0 + tid];
            }
            const bool redo = active && pw >= PLAN_PENDING;
            const uint32_t nredo = __popc(__ballot_sync(0xFFFFFFFFu, redo));
            if (redo && nredo < DEFER_MAX_PER_WARP) {
                hole = true; planned = true;
                st = a.status[r0 + tid]; len = a.out_len[r0 + tid];           // the counting pass's tail decoded it
                a.plan[r0 + tid] = PLAN_HOLE;
            } else if (active && !redo && smem_in) {
                pl.fast = 1; pl.cid_b = 20; pl.seq_b = pw & 0x1FFFu; pl.cid_e = (pw >> 13) & 0x7FFu; pl.seq_n = (pw >> 24) & 0xFu;
                pl.done = (pw >> 28) & 1u; pl.cid_esc = 0; pl.tok_esc = 0; pl.tok_redacted = 0;
                fast_status_len(a.op, L, pl.cid_e - pl.cid_b, pl.seq_n, pl.done != 0, ((pw >> 29) & 1u) != 0, st, len);
                fastp = true; planned = true;
            }                                                                 // else: decoded again below
            if (lane == 0 && nredo && nredo < DEFER_MAX_PER_WARP) atomicAdd(&ts->meta[slot].npend, nredo);
        }
        InterestPre ipre{0, 0, 0, 0, 0};
        if (active && smem_in && !planned) {
            PtrSrc src{s_in};
            if (EXT & X_INTEREST)                  // the id is hashed and its first table probe requested inside the parse
                fastp = plan_record_fast(src, o, L, a.op, my_start, pl, st, len,
                                         [&](const PtrSrc& ss, uint32_t b, uint32_t e, bool plain) { interest_prepare(ifast, ss, b, e, plain, ipre); });
            else fastp = plan_record_fast(src, o, L, a.op, my_start, pl, st, len);
        }
        // A few such records in a warp: deferred (the tail phase runs them densely).  Many of
        // them: decoded here, where the tile is in shared memory (22 ns per warp however many of its lanes take part).
        const bool slow = active && !fastp && !hole;
        const uint32_t nslow = __popc(__ballot_sync(0xFFFFFFFFu, slow));
        const bool defer = (DEFER || (MODE == MODE_ORDERED_COUNT && defer_ordered)) && nslow < DEFER_MAX_PER_WARP;      // warp-uniform
        if (slow) {
            if (defer) { st = ST_PENDING; len = 0; }
            else {                   // separate Plan object: taking pl's address would push the hot path's plan into local memory
                Plan sp;
                st = plan_slow(sb, o, L, a.op, my_start, &sp, &len);
                if (EXT) ext_stage_slow(a, r0 + tid, sb, o, L, my_b, &sp, &st, &len, MODE != MODE_ORDERED_COUNT,
                                        MODE != MODE_ORDERED_COUNT || (a.xflags & XF_ENTRY_IN_COUNT));
                pl = sp;
            }
        }
        if (defer && lane == 0 && nslow) atomicAdd(&ts->meta[slot].npend, nslow);

        if (pf) {
            const uint32_t in_total = ts->in_total;
            pf_e = min(pf_e, in_total); pf_b = min(pf_b, pf_e) & ~15u;
            const uint32_t bytes = min((pf_e - pf_b + 15u) & ~15u, 64u * 1024u);
            if (bytes) l2_prefetch(a.in + pf_b, bytes);
        }

        // ---- optional stages, canonical records (the others went through ext_stage_slow or wait for the tail) ----
        if (EXT) {
            const bool canon = active && fastp;
            // content inspection of the decoded token text, per_token mode (sse_handler.go:193-208)
            if ((EXT & X_INSPECT) && (a.insp != nullptr || pre_verdict)) {    // launch-uniform
                InspectResult io{0, 0, 0, 0};
                if (canon && (st & ST_CODE_MASK) == ST_OK) {
                    if (pre_verdict) io.verdict = a.verdict[r0 + tid];
                    else {
                        PtrSrc src{s_in};
                        if (!inspect_canon_token(*a.insp, dcache, src, o + pl.tok_b, o + pl.tok_e, pl.tok_esc != 0, io))
                            io = inspect_slow(a.insp, s_in, o + pl.tok_b, o + pl.tok_e);
                    }
                    if (a.op & OP_INSPECT_INLINE) {                           // sse_handler.go:200-207
                        const uint32_t act = io.verdict & 0xFF;
                        if (act == ACT_DROP) { st = ST_INSPECT_DROP; len = 0; }
                        else if (act == ACT_REDACT) {                         // the canonical path skipped values the encoder now needs
                            Plan sp;
                            uint32_t unused;
                            plan_slow(s_in, o, L, a.op | OP_NO_FAST, my_start, &sp, &unused);
                            sp.tok_redacted = 1; sp.fast = 0; fastp = false;
                            sp.cid_esc = pl.cid_esc; sp.tok_esc = pl.tok_esc;  // the general decoder does not note escapes; same literal ranges
                            len = measure_slow(s_in, o, L, &sp, a.op);
                            pl = sp;
                            st = (st & ~ST_HEADER_ONLY) | ST_REDACTED;                 // a re-marshalled record goes out whole
                        }
                    }
                }
                if (!pre_verdict && active && !slow && MODE != MODE_ORDERED_COUNT) {
                    a.verdict[r0 + tid] = io.verdict;
                    reinterpret_cast<int2*>(a.match)[r0 + tid] = make_int2(io.begin, io.end);
                }
            }
            // token entry for the streaming walk
            if ((EXT & X_INSPECT) && a.tok_entry != nullptr && canon && (MODE != MODE_ORDERED_COUNT || (a.xflags & XF_ENTRY_IN_COUNT)))
                store_token_entry(a, r0 + tid, s_in, o, my_b, pl, st, !pl.tok_esc, true);
            // subject -> subscriber/interest table
            if (EXT & X_INTEREST) {
                InterestHit h{0, 0, 0};
                if (canon && (st & ST_CODE_MASK) == ST_OK) {
                    if (ipre.hash) { PtrSrc src{s_in}; interest_finish(ifast, src, o + pl.cid_b, o + pl.cid_e, ipre, h); }
                    else h = interest_slow(a, s_in, o, pl.cid_b, pl.cid_e);
                    if (h.list_cnt == 0 && h.wild_mask == 0) { st = ST_NO_INTEREST; len = 0; }   // nobody subscribed: nothing to frame
                }
                if (active && !slow && MODE != MODE_ORDERED_COUNT) {
                    const size_t k = 3 * (size_t)(r0 + tid);
                    a.hit[k] = h.list_off; a.hit[k + 1] = h.list_cnt; a.hit[k + 2] = h.wild_mask;
                }
            }
        }

        // ---- CTA exclusive scan of lengths --------------------------------------------
        uint32_t incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) ts->wsum[slot][warp] = incl;
        if (tid == 0) {
            park_tile(a, ts, slot ^ 1u, nt, nb, ne);                          // the next trip's tile: visible behind (B)
            ts->next2 = claim;
            bulk_wait_read0();                                                // previous tile's bulk store has finished reading s_out
        }
        __syncthreads();                                                      // (B)
        uint32_t wpre = 0, total = 0;
#pragma unroll
        for (int w = 0; w < TILE / 32; w++) { uint32_t v = ts->wsum[slot][w]; if (w < (int)warp) wpre += v; total += v; }
        if (MODE == MODE_ORDERED_COUNT) {        // pass A of ordered mode: only the tile's byte count is wanted
            if (!EXT && a.plan && active) {      // what the emit pass needs of a canonical record, so that it does not decode it again
                uint32_t pw = PLAN_REDO;
                if (fastp && pl.seq_b < (1u << 13) && pl.cid_e < (1u << 11))
                    pw = pl.seq_b | (pl.cid_e << 13) | (pl.seq_n << 24) | (pl.done << 28) | ((st & ST_CODE_MASK) == ST_FILTERED ? 1u << 29 : 0u);
                else if (st == ST_PENDING) pw = PLAN_PENDING;
                a.plan[r0 + tid] = pw;
                if (pw == PLAN_REDO) { a.status[r0 + tid] = st; a.out_len[r0 + tid] = len; }   // decoded here: what a hole in the emit pass needs
            }
            if (tid == 0) {
                uint64_t w = total;
                if (ts->meta[slot].npend) { w |= (uint64_t)ts->my_chain << 32; ts->my_chain = tile + 1; }   // tiles with records left to the tail, chained
                a.tile_state[tile] = w;
                launch_tile(a, ts, s_in, slot ^ 1u, false); ts->next_t = ts->next2;   // every thread is past (B): the tile has been consumed
            }
            continue;
        }
        constexpr bool ordered = MODE == MODE_ORDERED_EMIT;
        // Ordered emit pass: the tile's arena offset is already known (k_scan_tile_totals), so the staging tile is laid
        // out at the same position modulo 16 as the arena bytes it becomes; its 16-byte aligned interior then leaves as
        // ONE TMA bulk store like an arrival-mode tile, and only the <= 15 bytes at either end are stored by threads.
        const uint64_t obase = ordered ? ts->obase : 0;                       // read before (D); setup_tile rewrites it after (D)
        const uint32_t shift = ordered ? (uint32_t)obase & 15u : 0u;
        const bool smem_out = total + (ordered ? 16u : 0u) <= a.out_cap_smem();   // tile-uniform
        // arrival-order placement: every tile owns a 16-byte aligned slice of the arena, so its staging
        // tile leaves as ONE TMA bulk store; the <=15 pad bytes after its last frame are zeroed.
        const uint32_t alloc = ordered ? total : ((total + 15u) & ~15u);
        if ((DEFER || MODE == MODE_ORDERED_EMIT) && tid == 0 && ts->meta[slot].npend) {   // chain this tile to the CTA's list of tiles with records left over
            a.tile_state[tile] = (uint64_t)ts->meta[slot].npend | ((uint64_t)ts->my_chain << 32);
            ts->my_chain = tile + 1;
        }
        if (tid == 0) {
            ts->c_out += total;
            if (ordered) ts->base = obase;                                    // exclusive prefix from k_scan_tile_totals
            else ts->base = atomicAdd(a.cursor, (unsigned long long)alloc);   // nothing to wait for
        }
        const uint32_t local = wpre + incl - len;
        const uint32_t code = st & ST_CODE_MASK;
        uint32_t subj = 0;

        if (smem_out) {
            // ---- pass 2 into the staging tile; the global offset is not needed yet ------
            if (active && code == ST_OK && !hole) {
                if (fastp) {
                    PtrSrc src{s_in};
                    Ptr32Dst d{s_out};
                    Writer<Ptr32Dst, true> w(d, local + shift);
                    emit_record_fast(src, o, L, pl, a.op, w, &subj);
                } else {
                    Plan sp = pl;
                    subj = emit_slow(sb, o, L, &sp, a.op, s_out, (uint64_t)(local + shift));
                }
            }
            if (!ordered && tid < alloc - total) s_out[total + tid] = 0;
            fence_proxy_async_smem();                                         // staging writes -> visible to the TMA store
            __syncthreads();                                                  // (D) s_in free, staging tile complete, base visible
            if (tid == 0) { launch_tile(a, ts, s_in, slot ^ 1u, ordered, ordered && !EXT && a.plan != nullptr); ts->next_t = ts->next2; }   // next tile's TMA load flies during the copy-out
        } else {
            // ---- oversize tile: emit straight into the arena ---------------------------
            __syncthreads();                                                  // (D')
        }
        const uint64_t base = ts->base;
        const bool fits = base + alloc <= a.out_cap;
        if (!smem_out) {
            if (active && code == ST_OK && fits && !hole) {
                Plan sp = pl;
                subj = emit_slow(sb, o, L, &sp, a.op, a.out, base + local);
            }
            if (fits && tid < alloc - total) a.out[base + total + tid] = 0;       // pad of an aligned tile
            __syncthreads();                                                  // (E') s_in free
            if (tid == 0) { launch_tile(a, ts, s_in, slot ^ 1u, ordered, ordered && !EXT && a.plan != nullptr); ts->next_t = ts->next2; }
        }
        if (active) {
            a.out_off[r0 + tid] = base + local;
            a.out_len[r0 + tid] = len;
            a.status[r0 + tid] = st;
            if (a.subj_len) a.subj_len[r0 + tid] = subj;
        }
        if (!SMEM_CTR) {
            if (active) {
                if (code == ST_OK) { c_ok++; c_fast += fastp ? 1u : 0u; if (st & ST_DONE) c_done++; }
                else if (code == ST_ERR_SYNTAX) c_syn++;
                else if (code == ST_ERR_TYPE) c_typ++;
                else if (code == ST_FILTERED) c_fil++;
            }
        } else {   // counters: one ballot per kind, lane 0 adds the warp's counts to the CTA's words
            const bool okr = active && code == ST_OK;
            const uint32_t m_ok = __ballot_sync(0xFFFFFFFFu, okr), m_fast = __ballot_sync(0xFFFFFFFFu, okr && fastp);
            const uint32_t m_done = __ballot_sync(0xFFFFFFFFu, okr && (st & ST_DONE) != 0);
            const uint32_t m_other = __ballot_sync(0xFFFFFFFFu, active && !okr && code != ST_PENDING);
            if (lane == 0) {
                if (m_ok) atomicAdd(&ts->ctr[0], (uint32_t)__popc(m_ok));
                if (m_fast) atomicAdd(&ts->ctr[5], (uint32_t)__popc(m_fast));
                if (m_done) atomicAdd(&ts->ctr[4], (uint32_t)__popc(m_done));
            }
            if (m_other && active) {                         // rare: dropped / filtered records
                if (code == ST_ERR_SYNTAX) atomicAdd(&ts->ctr[1], 1u);
                else if (code == ST_ERR_TYPE) atomicAdd(&ts->ctr[2], 1u);
                else if (code == ST_FILTERED) atomicAdd(&ts->ctr[3], 1u);
            }
        }
        if (tid == 0 && !fits) atomicExch(a.overflow, 1u);
        if (smem_out && fits && total != 0) {
            if (ordered) {
                const uint64_t g1 = base + total, ga = (base + 15) & ~15ull, gb = g1 & ~15ull;
                if (ga < gb) {
                    const uint32_t head = (uint32_t)(ga - base), tail = (uint32_t)(g1 - gb);
                    if (tid == 0) { tma_store_1d(a.out + ga, s_out + shift + head, (uint32_t)(gb - ga)); bulk_commit(); }
                    if (tid >= 32 && tid - 32 < head) a.out[base + (tid - 32)] = s_out[shift + (tid - 32)];
                    if (tid >= 64 && tid - 64 < tail) a.out[gb + (tid - 64)] = s_out[shift + (total - tail) + (tid - 64)];
                } else {
                    for (uint32_t i = tid; i < total; i += TILE) a.out[base + i] = s_out[shift + i];
                }
            } else if (tid == 0) { tma_store_1d(a.out + base, s_out, alloc); bulk_commit(); }
        }
    }

    if (tid == 0) {
        bulk_wait0();
        if (MODE == MODE_ORDERED_EMIT) asm volatile("fence.proxy.async;" ::: "memory");   // the tail writes into tiles stored by TMA
    }
    unsigned long long* row = a.counters + (size_t)(blockIdx.x % CTR_ROWS) * CTR_COLS;

    // ---- tail: the records this CTA left over --------------------------------------------------
    // Arrival mode.  The tiles that left records behind are chained through tile_state; their record numbers are
    // gathered in the (now idle) staging tile and the general decoder -- with the general forms of the optional
    // stages -- runs over them one record per lane, 128 at a time, straight from and to global memory.  Each warp
    // takes a 16-byte aligned slice of the arena from the same cursor, so such a frame simply lives somewhere else in
    // the arena (out_off/out_len address it like any other).
    if (MODE == MODE_ORDERED_COUNT && !EXT && a.plan != nullptr) {
        // Count pass: lengths and statuses of the records left over; each length is added to its tile's total (the
        // chain link in the upper half of the tile word is cleared as the chain is walked, the scan wants plain totals).
        uint32_t* s_list = reinterpret_cast<uint32_t*>(s_in);
        __syncthreads();
        if (tid == 0) { ts->meta[0].tile = ts->my_chain; ts->meta[0].npend = 0; }
        __syncthreads();
        uint32_t t1 = ts->meta[0].tile;
        const uint32_t in_total = ts->in_total;
        while (t1 != 0) {
            uint32_t cnt = 0;
            while (t1 != 0 && cnt < (uint32_t)TILE) {
                const uint32_t t = t1 - 1, i = t * TILE + tid;
                if (i < a.n && a.plan[i] == PLAN_PENDING) s_list[atomicAdd(&ts->meta[0].npend, 1u)] = i;
                t1 = (uint32_t)(a.tile_state[t] >> 32);
                __syncthreads();
                if (tid == 0) a.tile_state[t] &= 0xFFFFFFFFull;
                cnt = ts->meta[0].npend;
            }
            __syncthreads();                                                  // links cleared before any length is added
            for (uint32_t g = 0; g < cnt; g += TILE) {
                if (g + tid < cnt) {
                    const uint32_t i = s_list[g + tid];
                    const uint32_t b = min(a.in_off[i], in_total), e = min(max(a.in_off[i + 1], b), in_total);
                    Plan pl;
                    uint32_t len = 0;
                    const uint32_t st = plan_slow(a.in, b, e - b, a.op, a.start_seq ? a.start_seq[i] : 0, &pl, &len);
                    a.status[i] = st;
                    a.out_len[i] = len;
                    a.plan[i] = PLAN_REDO;
                    if (len) atomicAdd(reinterpret_cast<unsigned long long*>(&a.tile_state[i / TILE]), (unsigned long long)len);
                }
            }
            __syncthreads();
            if (tid == 0) ts->meta[0].npend = 0;
            __syncthreads();
        }
    }
    if (MODE == MODE_ORDERED_EMIT && !EXT && a.plan != nullptr) {
        // Emit pass: the frames of the records placed as holes, written where the main loop put them.
        uint32_t* s_list = reinterpret_cast<uint32_t*>(s_out);
        __syncthreads();                                                      // the last tile's bulk store is complete (bulk_wait0 above)
        if (tid == 0) { ts->meta[0].tile = ts->my_chain; ts->meta[0].npend = 0; }
        __syncthreads();
        uint32_t t1 = ts->meta[0].tile;
        const uint32_t in_total = ts->in_total;
        while (t1 != 0) {
            uint32_t cnt = 0;
            while (t1 != 0 && cnt < (uint32_t)TILE) {
                const uint32_t t = t1 - 1, i = t * TILE + tid;
                if (i < a.n && a.plan[i] == PLAN_HOLE) s_list[atomicAdd(&ts->meta[0].npend, 1u)] = i;
                t1 = (uint32_t)(a.tile_state[t] >> 32);
                __syncthreads();
                cnt = ts->meta[0].npend;
            }
            for (uint32_t g = 0; g < cnt; g += TILE) {
                if (g + tid < cnt) {
                    const uint32_t i = s_list[g + tid];
                    const uint32_t b = min(a.in_off[i], in_total), e = min(max(a.in_off[i + 1], b), in_total);
                    Plan pl;
                    uint32_t len = 0;
                    const uint32_t st = plan_slow(a.in, b, e - b, a.op, a.start_seq ? a.start_seq[i] : 0, &pl, &len);
                    const uint64_t at = a.out_off[i];
                    if ((st & ST_CODE_MASK) == ST_OK && len != 0 && at + len <= a.out_cap) {
                        const uint32_t subj = emit_slow(a.in, b, e - b, &pl, a.op, a.out, at);
                        if (a.subj_len) a.subj_len[i] = subj;
                    }
                }
            }
            __syncthreads();
            if (tid == 0) ts->meta[0].npend = 0;
            __syncthreads();
        }
    }
    if (MODE == MODE_ARRIVAL) {
        uint32_t* s_list = reinterpret_cast<uint32_t*>(s_out);
        __syncthreads();                                                      // every thread has left the loop
        if (tid == 0) { ts->meta[0].tile = ts->my_chain; ts->meta[0].npend = 0; }
        __syncthreads();
        uint32_t t1 = ts->meta[0].tile;                                               // CTA-uniform
        const uint32_t in_total = ts->in_total;
        while (t1 != 0) {
            uint32_t cnt = 0;
            while (t1 != 0 && cnt < (uint32_t)TILE) {                         // gather until there is a full group
                const uint32_t t = t1 - 1, i = t * TILE + tid;
                if (i < a.n && (a.status[i] & ST_CODE_MASK) == ST_PENDING) s_list[atomicAdd(&ts->meta[0].npend, 1u)] = i;
                t1 = (uint32_t)(a.tile_state[t] >> 32);
                __syncthreads();
                cnt = ts->meta[0].npend;
            }
            for (uint32_t g = 0; g < cnt; g += TILE) {
                const bool act = g + tid < cnt;
                uint32_t i = 0, b = 0, e = 0, st = ST_OK, len = 0;
                Plan pl;
                if (act) {
                    i = s_list[g + tid];
                    b = min(a.in_off[i], in_total); e = min(max(a.in_off[i + 1], b), in_total);
                    st = plan_slow(a.in, b, e - b, a.op, a.start_seq ? a.start_seq[i] : 0, &pl, &len);
                    if (EXT) ext_stage_slow(a, i, a.in, b, e - b, b, &pl, &st, &len, true, true);
                }
                uint32_t incl = len;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                    if (lane >= (uint32_t)d) incl += v;
                }
                const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31), alloc = (total + 15u) & ~15u;
                unsigned long long base = 0;
                if (lane == 0 && alloc) base = atomicAdd(a.cursor, (unsigned long long)alloc);
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                const bool fits = base + alloc <= a.out_cap;
                const uint32_t code = st & ST_CODE_MASK;
                if (act) {
                    uint32_t subj = 0;
                    if (code == ST_OK && fits) subj = emit_slow(a.in, b, e - b, &pl, a.op, a.out, base + incl - len);
                    a.out_off[i] = base + incl - len;
                    a.out_len[i] = len;
                    a.status[i] = st;
                    if (a.subj_len) a.subj_len[i] = subj;
                    if (code == ST_OK) { atomicAdd(&ts->ctr[0], 1u); if (st & ST_DONE) atomicAdd(&ts->ctr[4], 1u); }
                    else if (code == ST_ERR_SYNTAX) atomicAdd(&ts->ctr[1], 1u);
                    else if (code == ST_ERR_TYPE) atomicAdd(&ts->ctr[2], 1u);
                    else if (code == ST_FILTERED) atomicAdd(&ts->ctr[3], 1u);
                }
                if (fits && lane < alloc - total) a.out[base + total + lane] = 0;
                if (lane == 0 && total) atomicAdd(&row[7], (unsigned long long)total);
                if (lane == 0 && !fits) atomicExch(a.overflow, 1u);
            }
            __syncthreads();
            if (tid == 0) ts->meta[0].npend = 0;
            __syncthreads();
        }
    }

    // ---- counters: one row per CTA slot, summed by the host --------------------------
    if (!SMEM_CTR) {
        uint32_t vals[6] = {c_ok, c_syn, c_typ, c_fil, c_done, c_fast};
#pragma unroll
        for (int k = 0; k < 6; k++) {
            uint32_t v = vals[k];
#pragma unroll
            for (int o2 = 16; o2 > 0; o2 >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o2);
            if (lane == 0 && v) atomicAdd(&ts->ctr[k], v);
        }
    }
    __syncthreads();
    if (tid < 6) { const uint32_t v = ts->ctr[tid]; if (v) atomicAdd(&row[tid], (unsigned long long)v); }
    if (tid == 0 && MODE != MODE_ORDERED_COUNT) {
        if (ts->c_in) atomicAdd(&row[6], ts->c_in);
        if (ts->c_out) atomicAdd(&row[7], ts->c_out);
        if (MODE == MODE_ARRIVAL) {     // the CTA that finishes last publishes the arena's total size as out_off[n]
            __threadfence();
            if (atomicAdd(a.done_ctas, 1u) == gridDim.x - 1) {
                __threadfence();
                a.out_off[a.n] = (uint64_t)atomicAdd(a.cursor, 0ull);
            }
        }
    }
}

// ---------------------------------------------------------------------- launch ----
// Shared memory per CTA: a tile of TILE records is ~17 KB in / ~21 KB out at the BASELINE workload; the caps leave
// ~15% slack.  Tiles that exceed them take the direct global-memory path inside the same kernel.
//   cfg 0  160 B in / 184 B out per record                      5 CTAs per SM (the SM's shared memory is then full)
//   cfg 1  "wide": NATS PUB framing adds ~60 bytes per record (subject + protocol line): 224 B out
//   cfg 2  fused per-token inspection of adapter frames: the on-chip DFA part (2.4 KB) has to fit beside five tiles
//          per SM, so the caps shrink to 150 B in / 181 B out per record (~10 % slack at the BASELINE shape)
//   cfg 3  fused inspection of bridge records (subject + payload ~182 B: the narrow cap would overflow): full caps, 4 CTAs per SM
//   cfg 4  fused inspection + NATS PUB framing
constexpr int N_CFG = 5;
static void cfg_caps(int cfg, uint32_t* ic, uint32_t* oc, uint32_t* smem) {
    const bool cache = cfg >= 2;
    *ic = TILE * (cfg == 2 ? 150 : 160);
    *oc = TILE * (cfg == 2 ? 181 : (cfg == 0 ? 184 : ((cfg == 1 || cfg == 4) ? 224 : 192)));
    // behind the offsets: the plan words of ordered mode's emit pass (variants without inspection), or the DFA part
    *smem = *ic + 32 + *oc + 48 + (uint32_t)sizeof(TileSmem) + OFF_BYTES + (cache ? DFA_CACHE_BYTES : PLAN_BYTES);
}
static_assert(sizeof(TileSmem) % 16 == 0, "the record offsets behind TileSmem are the target of a bulk copy");
// The counting pass of ordered mode never stages a frame: launched with out_cap_s = 0 it needs the input tile only, and
// more CTAs fit an SM.
static uint32_t count_smem(int cfg) {
    uint32_t ic, oc, sm;
    cfg_caps(cfg, &ic, &oc, &sm);
    return sm - oc;
}

static std::mutex g_cfg_mu;             // the cache below is filled once per device, from whichever thread launches first
static int g_num_sms[64];
static int g_ctas_per_sm[64][4][N_CFG]; // [device][ext][cfg]; 0 = device not configured yet
static int g_count_ctas_per_sm[64][4][N_CFG];   // the same for the counting pass of ordered mode

template <int EXT, int MODE>
static cudaError_t configure_one(int* per_sm /* [N_CFG] */) {
    uint32_t ic, oc, sm, sm_max = 0;
    for (int c = 0; c < N_CFG; c++) { cfg_caps(c, &ic, &oc, &sm); sm_max = sm > sm_max ? sm : sm_max; }
    cudaError_t e = cudaFuncSetAttribute(k_frame_tiles<EXT, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_max);
    if (e != cudaSuccess) return e;
    for (int c = 0; c < N_CFG; c++) {
        cfg_caps(c, &ic, &oc, &sm);
        int per = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_frame_tiles<EXT, MODE>, TILE, sm);
        if (e != cudaSuccess) return e;
        per = per < 1 ? 1 : per;
        per_sm[c] = per_sm[c] == 0 ? per : min(per_sm[c], per);      // arrival and emit pass: one table
    }
    return cudaSuccess;
}
template <int EXT>
static cudaError_t configure_count(int* per_sm /* [N_CFG] */) {
    uint32_t sm_max = 0;
    for (int c = 0; c < N_CFG; c++) sm_max = max(sm_max, count_smem(c));
    cudaError_t e = cudaFuncSetAttribute(k_frame_tiles<EXT, MODE_ORDERED_COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_max);
    if (e != cudaSuccess) return e;
    for (int c = 0; c < N_CFG; c++) {
        int per = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_frame_tiles<EXT, MODE_ORDERED_COUNT>, TILE, count_smem(c));
        if (e != cudaSuccess) return e;
        per_sm[c] = per < 1 ? 1 : per;
    }
    return cudaSuccess;
}

cudaError_t frame_configure(int device) {
    std::lock_guard<std::mutex> lk(g_cfg_mu);
    if (g_ctas_per_sm[device][0][0] != 0) return cudaSuccess;
    int sms = 0;
    cudaError_t e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    int per[4][N_CFG] = {}, cnt[4][N_CFG] = {};
    if ((e = configure_one<0, MODE_ARRIVAL>(per[0])) != cudaSuccess) return e;
    if ((e = configure_one<0, MODE_ORDERED_EMIT>(per[0])) != cudaSuccess) return e;
    if ((e = configure_count<0>(cnt[0])) != cudaSuccess) return e;
    if ((e = configure_one<1, MODE_ARRIVAL>(per[1])) != cudaSuccess) return e;
    if ((e = configure_one<1, MODE_ORDERED_EMIT>(per[1])) != cudaSuccess) return e;
    if ((e = configure_count<1>(cnt[1])) != cudaSuccess) return e;
    if ((e = configure_one<2, MODE_ARRIVAL>(per[2])) != cudaSuccess) return e;
    if ((e = configure_one<2, MODE_ORDERED_EMIT>(per[2])) != cudaSuccess) return e;
    if ((e = configure_count<2>(cnt[2])) != cudaSuccess) return e;
    if ((e = configure_one<3, MODE_ARRIVAL>(per[3])) != cudaSuccess) return e;
    if ((e = configure_one<3, MODE_ORDERED_EMIT>(per[3])) != cudaSuccess) return e;
    if ((e = configure_count<3>(cnt[3])) != cudaSuccess) return e;
    g_num_sms[device] = sms;
    for (int x = 3; x >= 0; x--) for (int c = N_CFG - 1; c >= 0; c--) g_count_ctas_per_sm[device][x][c] = cnt[x][c];
    for (int x = 3; x >= 0; x--) for (int c = N_CFG - 1; c >= 0; c--) g_ctas_per_sm[device][x][c] = per[x][c];
    return cudaSuccess;
}

struct CountLaunch { int grid; uint32_t smem; };
static cudaError_t choose_variant(int device, FrameArgs& a, int* ext, uint32_t* smem_bytes, int* grid, CountLaunch* cl = nullptr) {
    cudaError_t e = frame_configure(device);
    if (e != cudaSuccess) return e;
    *ext = ((a.insp != nullptr || a.tok_entry != nullptr || (a.xflags & XF_PRE_VERDICT)) ? X_INSPECT : 0) | (a.interest != nullptr ? X_INTEREST : 0);
    const bool cache = a.insp != nullptr && !(a.xflags & XF_PRE_VERDICT);
    const bool wide = (a.op & OP_NATS_PUB) != 0, bridge = (a.op & OP_BRIDGE) != 0;
    const int cfg = cache ? (wide ? 4 : (bridge ? 3 : 2)) : (wide ? 1 : 0);
    cfg_caps(cfg, &a.in_cap, &a.out_cap_s, smem_bytes);
    *grid = g_num_sms[device] * g_ctas_per_sm[device][*ext][cfg];
    a.num_tiles = (a.n + TILE - 1) / TILE;
    if ((int)a.num_tiles < *grid) *grid = (int)a.num_tiles;
    if (cl) {
        cl->grid = min((int)a.num_tiles, g_num_sms[device] * g_count_ctas_per_sm[device][*ext][cfg]);
        cl->smem = count_smem(cfg);
    }
    return cudaSuccess;
}

template <int EXT>
static void launch_variant(const FrameArgs& a, int grid, uint32_t smem_bytes, const CountLaunch& cl, cudaStream_t stream) {
    if (a.op & OP_ORDERED) {
        FrameArgs ac = a;
        ac.out_cap_s = 0;
        k_frame_tiles<EXT, MODE_ORDERED_COUNT><<<cl.grid, TILE, cl.smem, stream>>>(ac);
        const uint32_t nb = (a.num_tiles + SCAN_BLOCK - 1) / SCAN_BLOCK;
        uint64_t* block_sums = a.out_off;                 // free until the emit pass; nb <= n, so out_off[n] is not among them
        k_tile_block_sums<<<nb, SCAN_BLOCK, 0, stream>>>(a.tile_state, a.num_tiles, block_sums);
        k_scan_tile_totals<<<nb, SCAN_BLOCK, 0, stream>>>(a.tile_state, a.num_tiles, block_sums, a.out_off + a.n, a.ticket);
        k_frame_tiles<EXT, MODE_ORDERED_EMIT><<<grid, TILE, smem_bytes, stream>>>(a);
    } else {
        k_frame_tiles<EXT, MODE_ARRIVAL><<<grid, TILE, smem_bytes, stream>>>(a);
    }
}

static cudaError_t clear_scratch(const FrameArgs& a, cudaStream_t stream) {
    // ticket + overflow flag + cursor + counters + per-tile words, zeroed per launch
    uint8_t* lo = static_cast<uint8_t*>(a.zero_base ? a.zero_base : a.scratch_base);
    uint8_t* hi = lo + (a.zero_base ? a.zero_bytes : scratch_bytes(a.n));
    if (!a.keep_bad_conv) return cudaMemsetAsync(lo, 0, (size_t)(hi - lo), stream);
    uint8_t* flag = reinterpret_cast<uint8_t*>(a.bad_conv);
    cudaError_t e = cudaMemsetAsync(lo, 0, (size_t)(flag - lo), stream);
    return e != cudaSuccess ? e : cudaMemsetAsync(flag + 4, 0, (size_t)(hi - flag - 4), stream);
}

cudaError_t frame_launch(int device, FrameArgs a, cudaStream_t stream) {
    uint32_t smem_bytes;
    int grid;
    int ext;
    CountLaunch cl;
    cudaError_t e = choose_variant(device, a, &ext, &smem_bytes, &grid, &cl);
    if (e != cudaSuccess) return e;
    if ((e = clear_scratch(a, stream)) != cudaSuccess) return e;
    if (a.n == 0) return cudaMemsetAsync(a.out_off, 0, sizeof(uint64_t), stream);
    a.off_tma = (reinterpret_cast<uintptr_t>(a.in_off) & 15) == 0 ? 1u : 0u;
    a.plan_tma = (a.off_tma && a.plan != nullptr && (reinterpret_cast<uintptr_t>(a.plan) & 15) == 0) ? 1u : 0u;
    switch (ext) {
    case 0: launch_variant<0>(a, grid, smem_bytes, cl, stream); break;
    case 1: launch_variant<1>(a, grid, smem_bytes, cl, stream); break;
    case 2: launch_variant<2>(a, grid, smem_bytes, cl, stream); break;
    default: launch_variant<3>(a, grid, smem_bytes, cl, stream); break;
    }
    return cudaGetLastError();
}

cudaError_t frame_plan_launch(int device, FrameArgs a, cudaStream_t stream) {
    uint32_t smem_bytes;
    int grid;
    int ext;
    a.xflags |= XF_ENTRY_IN_COUNT;
    CountLaunch cl;
    cudaError_t e = choose_variant(device, a, &ext, &smem_bytes, &grid, &cl);
    if (e != cudaSuccess) return e;
    if ((e = clear_scratch(a, stream)) != cudaSuccess) return e;
    if (a.n == 0) return cudaSuccess;
    a.off_tma = (reinterpret_cast<uintptr_t>(a.in_off) & 15) == 0 ? 1u : 0u;
    a.out_cap_s = 0;
    k_frame_tiles<X_INSPECT, MODE_ORDERED_COUNT><<<cl.grid, TILE, cl.smem, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace sse
