// sse_kernels.cu -- the fused B200 kernel of the SSE hot path.
//
// k_frame_tiles: decode -> (filter) -> encode -> frame -> exclusive scan of frame lengths ->
// compaction, in ONE pass over HBM.  It stands in for the per-token body of the adapter's two
// hot loops (src/sse-adapter/sse_handler.go:135-157 + :186-233, :316-333 + :400-449) and of
// the bridge loop (src/redis-nats-bridge/main.go:136-181), batched over every conversation
// that has a token pending.
//
// Shape (sm_100a):
//   * persistent CTAs (grid = SMs x resident CTAs), tiles of TILE records handed out in order
//     by an atomic ticket, so the decoupled look-back below can never wait on a tile that has
//     not started;
//   * a tile's payload bytes arrive in shared memory as ONE TMA bulk copy
//     (cp.async.bulk ... mbarrier::complete_tx), 16-byte aligned, no register staging;
//   * one thread per record runs sse_record.cuh (SWAR canonical check / general decoder);
//   * frame lengths: warp-shuffle scan -> CTA scan -> single-pass decoupled look-back
//     (64-bit {flag,prefix} words) for the tile's global byte offset;
//   * frames are assembled in a shared-memory staging tile laid out congruent (mod 16) to the
//     output arena, then leave as ONE TMA bulk store (cp.async.bulk.global.shared::cta) plus
//     at most 15 head and 15 tail bytes, so HBM sees only full-line traffic.
//   No tensor cores: nothing here is a contraction.
#include <cuda_runtime.h>
#include <stdint.h>
#include "sse_kernels.h"
#include "sse_record.cuh"

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
// shared -> global bulk copy (TMA, 1-D)
__device__ __forceinline__ void tma_store_1d(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
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

// ---------------------------------------------------------------- tile status ----
constexpr uint64_t FLAG_AGG = 1ull << 62;      // value = this tile's own byte count
constexpr uint64_t FLAG_PFX = 2ull << 62;      // value = inclusive prefix up to and including this tile
constexpr uint64_t VAL_MASK = (1ull << 62) - 1;

// Single-pass decoupled look-back, run by warp 0.  Returns the exclusive prefix of `tile`.
__device__ __forceinline__ uint64_t lookback(uint64_t* tile_state, uint32_t tile, uint64_t total, uint32_t lane) {
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(&tile_state[0], FLAG_PFX | total);
        return 0;
    }
    if (lane == 0) st_relaxed_u64(&tile_state[tile], FLAG_AGG | total);
    uint64_t excl = 0;
    int64_t j = (int64_t)tile - 1;             // window covers tiles j, j-1, ..., j-31
    for (;;) {
        int64_t idx = j - (int64_t)lane;
        uint64_t v;
        do {
            v = idx >= 0 ? ld_relaxed_u64(&tile_state[idx]) : FLAG_PFX;   // before tile 0: prefix 0
        } while (__any_sync(0xFFFFFFFFu, (v >> 62) == 0));
        uint32_t pfx_mask = __ballot_sync(0xFFFFFFFFu, (v >> 62) == 2);
        uint32_t first_pfx = pfx_mask ? (uint32_t)(__ffs((int)pfx_mask) - 1) : 32u;   // nearest tile holding a prefix
        uint64_t contrib = lane <= first_pfx ? (v & VAL_MASK) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xFFFFFFFFu, contrib, o);
        excl += contrib;
        if (pfx_mask) break;
        j -= 32;
    }
    if (lane == 0) st_relaxed_u64(&tile_state[tile], FLAG_PFX | (excl + total));
    return excl;
}

// ---------------------------------------------------------------------- kernel ----
struct TileSmem {
    uint64_t mbar;
    uint64_t base;           // global byte offset of the tile's output
    uint32_t tile, tb, te, pad;
    uint32_t wsum[TILE / 32];
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

__global__ void __launch_bounds__(TILE) k_frame_tiles(const FrameArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* s_in = smem;                                   // in_cap + 16
    uint8_t* s_out = smem + a.in_cap + 16;                  // out_cap_s + 16
    TileSmem* ts = reinterpret_cast<TileSmem*>(s_out + a.out_cap_s + 16);

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t c_ok = 0, c_syn = 0, c_typ = 0, c_fil = 0, c_done = 0, c_fast = 0;
    uint64_t c_in = 0, c_out = 0;
    uint32_t parity = 0;

    if (tid == 0) mbar_init(&ts->mbar, 1);
    __syncthreads();

    for (;;) {
        // ---- S0: claim a tile, start its TMA load --------------------------------
        if (tid == 0) {
            uint32_t t = atomicAdd(a.ticket, 1u);
            ts->tile = t;
            if (t < a.num_tiles) {
                uint32_t r0 = t * TILE;
                uint32_t nrec = min((uint32_t)TILE, a.n - r0);
                uint32_t tb = a.in_off[r0], te = a.in_off[r0 + nrec];
                ts->tb = tb; ts->te = te;
                uint32_t abase = tb & ~15u;
                uint32_t bytes = (te - abase + 15u) & ~15u;
                if (bytes != 0 && bytes <= a.in_cap) {
                    mbar_expect_tx(&ts->mbar, bytes);
                    tma_load_1d(s_in, a.in + abase, bytes, &ts->mbar);
                }
                c_in += te - tb;
            }
        }
        __syncthreads();                                                      // (A)
        const uint32_t tile = ts->tile;
        if (tile >= a.num_tiles) break;
        const uint32_t r0 = tile * TILE;
        const uint32_t nrec = min((uint32_t)TILE, a.n - r0);
        const uint32_t tb = ts->tb, te = ts->te;
        const uint32_t abase = tb & ~15u;
        const uint32_t in_bytes = (te - abase + 15u) & ~15u;
        const bool smem_in = in_bytes <= a.in_cap;

        uint32_t my_b = 0, my_e = 0;
        int64_t my_start = 0;
        const bool active = tid < nrec;
        if (active) {
            my_b = a.in_off[r0 + tid];
            my_e = a.in_off[r0 + tid + 1];
            if (a.start_seq) my_start = a.start_seq[r0 + tid];
        }
        if (smem_in && in_bytes != 0) { mbar_wait(&ts->mbar, parity); parity ^= 1; }

        // ---- pass 1: decode + filter, output length --------------------------------
        Plan pl;
        uint32_t st = ST_OK, len = 0;
        const uint32_t o = my_b - abase, L = my_e - my_b;
        bool fastp = false;
        if (active) {
            if (smem_in) { PtrSrc src{s_in}; fastp = plan_record_fast(src, o, L, a.op, my_start, pl, st, len); }
            if (!fastp) st = plan_slow(smem_in ? s_in : a.in + abase, o, L, a.op, my_start, &pl, &len);
        }

        // ---- CTA exclusive scan of lengths --------------------------------------------
        uint32_t incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) ts->wsum[warp] = incl;
        __syncthreads();                                                      // (B)
        uint32_t wpre = 0, total = 0;
#pragma unroll
        for (int w = 0; w < TILE / 32; w++) { uint32_t v = ts->wsum[w]; if (w < (int)warp) wpre += v; total += v; }
        if (warp == 0) {
            uint64_t excl = lookback(a.tile_state, tile, total, lane);
            if (lane == 0) {
                ts->base = excl;
                c_out += total;
                if (tile == a.num_tiles - 1) a.out_off[a.n] = excl + total;
                bulk_wait_read0();            // previous tile's TMA store has finished reading s_out
            }
        }
        __syncthreads();                                                      // (C)
        const uint64_t base = ts->base;
        const uint32_t local = wpre + incl - len;
        const uint32_t shift = (uint32_t)base & 15u;
        const bool fits = base + total <= a.out_cap;
        const bool smem_out = shift + total <= a.out_cap_smem();

        uint32_t subj = 0;
        if (active) {
            const uint32_t code = st & ST_CODE_MASK;
            if (code == ST_OK) {
                c_ok++; c_fast += fastp ? 1u : 0u; if (st & ST_DONE) c_done++;
                if (fits) {
                    if (fastp && smem_out) {
                        PtrSrc src{s_in};
                        Ptr32Dst d{s_out};
                        Writer<Ptr32Dst, true> w(d, shift + local);
                        emit_record_fast(src, o, L, pl, a.op, w, &subj);
                    } else {
                        subj = emit_slow(smem_in ? s_in : a.in + abase, o, L, &pl, a.op, smem_out ? s_out : a.out,
                                         smem_out ? (uint64_t)(shift + local) : base + local);
                    }
                }
            } else if (code == ST_ERR_SYNTAX) c_syn++;
            else if (code == ST_ERR_TYPE) c_typ++;
            else if (code == ST_FILTERED) c_fil++;
            a.out_off[r0 + tid] = base + local;
            a.status[r0 + tid] = st;
            if (a.subj_len) a.subj_len[r0 + tid] = subj;
        }
        if (!fits && tid == 0) atomicExch(a.overflow, 1u);

        // ---- staging tile -> output arena ----------------------------------------------
        if (smem_out && fits && total != 0) {
            fence_proxy_async_smem();
            __syncthreads();                                                  // (D)
            const uint64_t g0 = base, g1 = base + total;
            const uint64_t ga = (g0 + 15) & ~15ull, gb = g1 & ~15ull;         // 16-byte aligned interior
            if (ga < gb) {
                if (tid == 0) {
                    tma_store_1d(a.out + ga, s_out + (uint32_t)(ga - (g0 - shift)), (uint32_t)(gb - ga));
                    bulk_commit();
                }
                const uint32_t head = (uint32_t)(ga - g0), tail = (uint32_t)(g1 - gb);
                if (tid < head) a.out[g0 + tid] = s_out[shift + tid];
                if (tid >= 32 && tid - 32 < tail) a.out[gb + (tid - 32)] = s_out[(uint32_t)(gb - (g0 - shift)) + (tid - 32)];
            } else {
                for (uint32_t i = tid; i < total; i += TILE) a.out[g0 + i] = s_out[shift + i];
            }
        } else {
            __syncthreads();                                                  // (D')
        }
    }

    if (tid == 0) bulk_wait0();

    // ---- counters: one row per CTA slot, summed by the host --------------------------
    unsigned long long* row = a.counters + (size_t)(blockIdx.x % CTR_ROWS) * CTR_COLS;
    uint32_t vals[6] = {c_ok, c_syn, c_typ, c_fil, c_done, c_fast};
#pragma unroll
    for (int k = 0; k < 6; k++) {
        uint32_t v = vals[k];
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o2);
        if (lane == 0 && v) atomicAdd(&row[k], (unsigned long long)v);
    }
    if (tid == 0) {
        if (c_in) atomicAdd(&row[6], (unsigned long long)c_in);
        if (c_out) atomicAdd(&row[7], (unsigned long long)c_out);
    }
}

// ---------------------------------------------------------------------- launch ----
static int g_num_sms[64];
static int g_ctas_per_sm[64];

cudaError_t frame_configure(int device, uint32_t* in_cap, uint32_t* out_cap, uint32_t* smem_bytes, int* grid) {
    // Shared memory per CTA: a tile of TILE records is ~17 KB in / ~21 KB out at the
    // BASELINE workload; caps leave ~15% slack.  Tiles that exceed them take the direct
    // global-memory path inside the same kernel.
    const uint32_t ic = 20 * 1024, oc = 24 * 1024;
    const uint32_t sm = ic + 16 + oc + 16 + (uint32_t)sizeof(TileSmem);
    cudaError_t e = cudaFuncSetAttribute(k_frame_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e != cudaSuccess) return e;
    if (g_num_sms[device] == 0) {
        int sms = 0, per = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_frame_tiles, TILE, sm);
        if (e != cudaSuccess) return e;
        if (per < 1) per = 1;
        g_num_sms[device] = sms;
        g_ctas_per_sm[device] = per;
    }
    *in_cap = ic; *out_cap = oc; *smem_bytes = sm;
    *grid = g_num_sms[device] * g_ctas_per_sm[device];
    return cudaSuccess;
}

cudaError_t frame_launch(int device, FrameArgs a, cudaStream_t stream) {
    uint32_t smem_bytes;
    int grid;
    cudaError_t e = frame_configure(device, &a.in_cap, &a.out_cap_s, &smem_bytes, &grid);
    if (e != cudaSuccess) return e;
    a.num_tiles = (a.n + TILE - 1) / TILE;
    // scratch: ticket + overflow flag + counters + tile status words, zeroed per launch
    e = cudaMemsetAsync(a.scratch_base, 0, scratch_bytes(a.n), stream);
    if (e != cudaSuccess) return e;
    if (a.n == 0) return cudaMemsetAsync(a.out_off, 0, sizeof(uint64_t), stream);
    if ((int)a.num_tiles < grid) grid = (int)a.num_tiles;
    k_frame_tiles<<<grid, TILE, smem_bytes, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace sse
