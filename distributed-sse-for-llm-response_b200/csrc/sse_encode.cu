// sse_encode.cu -- producer-side batch encoder: TokenMessage columns -> canonical JSON payloads.
//
// Stands in for publishToken() of the LLM stream proxy (src/llm-stream-proxy/main.go:230-245: build the
// TokenMessage, json.Marshal it, derive the channel "chat.<id>.tokens") and for the load generator's producer
// (demo/load-generator/main.go:215-223), batched over every conversation that produced a delta in this interval.
// Output records are exactly what the adapter/bridge kernels recognise as canonical, so the downstream
// fast path always hits.
//
// Persistent CTAs, one tile of 128 messages per trip, tiles strided by the grid (so the next tile is known and can be
// prefetched): the tile's id bytes and token bytes are contiguous ranges of the two input arenas and arrive in shared
// memory as two TMA bulk copies, issued one tile AHEAD into the other half of a double buffer; the per-message scalars
// of the next tile are loaded into registers while the current tile is encoded.  One thread per message: measure,
// CTA scan, one atomic for the tile's 16-byte aligned slice of the arena (tiles land in completion order, like
// k_frame_tiles), records assembled in a staging tile, one TMA bulk store per tile.  (A warp-cooperative, one-byte-per-lane
// escaper for the non-plain strings was built and measured: 1.40 ms against 1.18 ms for the in-lane walk on the BASELINE
// workload, where 2.5 % of the tokens need it -- see profiles/r01_encode_notes.md.)
#include <cuda_runtime.h>
#include <stdint.h>
#include "sse_kernels.h"
#include "sse_record.cuh"

namespace sse {

#ifndef SSE_ENC_TILE
#define SSE_ENC_TILE 128
#endif
constexpr int ENC_TILE = SSE_ENC_TILE;                  // messages per tile == threads per CTA

// staging tile: 128 records x 144 bytes for bare payloads (a config-1 record is ~134), x 224 with a channel / RESP
// wrapper.  Input buffers: 48 id bytes and 24 token bytes per message.  Tiles beyond any of the caps read their strings
// from global memory / write straight to the arena instead (same code, other base pointers).
constexpr uint32_t ENC_STAGE_PLAIN = ENC_TILE * 144, ENC_STAGE_WRAPPED = ENC_TILE * 224;
constexpr uint32_t ENC_CID_CAP = ENC_TILE * 48, ENC_TOK_CAP = ENC_TILE * 24, ENC_SLACK = 48;   // slack: word reads run a little past a string

struct EncSmem {
    uint64_t mbar[2];
    unsigned long long base;
    uint32_t wsum[ENC_TILE / 32];
    uint32_t c0a[2], t0a[2], in_smem[2];
};

__device__ __forceinline__ uint32_t enc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void enc_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(enc_smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void enc_tma_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(enc_smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(enc_smem_u32(bar)) : "memory");
}

// rare (a tile of 128 messages longer than the staging tile): out of line and serial, straight into the arena
__device__ __noinline__ void encode_emit_global(uint32_t op, const uint8_t* cbase, const uint8_t* tbase, uint32_t cb, uint32_t ce, uint32_t tb,
                                                uint32_t te, int64_t seq, bool done, int64_t ts, const EncPlan* p, uint8_t* dst, uint64_t pos) {
    PtrSrc cid{cbase}, tok{tbase};
    PtrDst d{dst};
    Writer<PtrDst, true> w(d, pos);
    encode_emit(cid, cb, ce, tok, tb, te, seq, done, ts, (op & OP_GO122) != 0, *p, w);
}

// byte ranges of a tile's ids and tokens, clamped into the arenas (offsets are caller data)
struct EncRange { uint32_t c0, c1, t0, t1; };
__device__ __forceinline__ EncRange enc_tile_range(const EncodeArgs& a, uint32_t tile, uint32_t c_total, uint32_t t_total) {
    const uint32_t r0 = tile * ENC_TILE, r1 = min(r0 + (uint32_t)ENC_TILE, a.n);
    EncRange r;
    r.c0 = min(a.cid_off[r0], c_total); r.c1 = min(max(a.cid_off[r1], r.c0), c_total);
    r.t0 = min(a.tok_off[r0], t_total); r.t1 = min(max(a.tok_off[r1], r.t0), t_total);
    return r;
}
// tid 0: start the two bulk copies of a tile into buffer `buf` (or mark the tile as read-from-global)
__device__ __forceinline__ void enc_issue_loads(const EncodeArgs& a, EncSmem* es, uint8_t* s_cid, uint8_t* s_tok, uint32_t buf, const EncRange& r) {
    const uint32_t c0a = r.c0 & ~15u, t0a = r.t0 & ~15u;
    const uint32_t cbytes = (r.c1 - c0a + 15u) & ~15u, tbytes = (r.t1 - t0a + 15u) & ~15u;
    const bool fit = cbytes <= ENC_CID_CAP && tbytes <= ENC_TOK_CAP;
    es->c0a[buf] = c0a; es->t0a[buf] = t0a; es->in_smem[buf] = fit ? 1u : 0u;
    const uint32_t tx = fit ? cbytes + tbytes : 0u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(enc_smem_u32(&es->mbar[buf])), "r"(tx) : "memory");
    if (fit && cbytes) enc_tma_load(s_cid + buf * (ENC_CID_CAP + ENC_SLACK), a.cid + c0a, cbytes, &es->mbar[buf]);
    if (fit && tbytes) enc_tma_load(s_tok + buf * (ENC_TOK_CAP + ENC_SLACK), a.tok + t0a, tbytes, &es->mbar[buf]);
}

__global__ void __launch_bounds__(ENC_TILE) k_encode_messages(const EncodeArgs a, const uint32_t stage_bytes, const uint32_t num_tiles) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* s_out = smem;                                               // stage_bytes + 16
    uint8_t* s_cid = smem + stage_bytes + 16;                            // 2 x (ENC_CID_CAP + ENC_SLACK)
    uint8_t* s_tok = s_cid + 2 * (ENC_CID_CAP + ENC_SLACK);              // 2 x (ENC_TOK_CAP + ENC_SLACK)
    EncSmem* es = reinterpret_cast<EncSmem*>(s_tok + 2 * (ENC_TOK_CAP + ENC_SLACK));
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool go122 = (a.op & OP_GO122) != 0;
    const uint32_t G = gridDim.x;
    const uint32_t c_total = a.cid_off[a.n], t_total = a.tok_off[a.n];

    uint32_t tile = blockIdx.x;
    // this tile's scalars (registers), refilled one tile ahead below
    uint32_t cb = 0, ce = 0, tb = 0, te = 0;
    int64_t seq = 0, ts = 0;
    uint32_t done = 0;
    auto load_scalars = [&](uint32_t t, uint32_t& xcb, uint32_t& xce, uint32_t& xtb, uint32_t& xte, int64_t& xseq, int64_t& xts, uint32_t& xdone) {
        const uint32_t i = t * ENC_TILE + tid;
        if (t < num_tiles && i < a.n) {
            xcb = a.cid_off[i]; xce = a.cid_off[i + 1]; xtb = a.tok_off[i]; xte = a.tok_off[i + 1];
            xseq = a.seq[i]; xts = a.ts[i]; xdone = a.done[i];
        }
    };
    load_scalars(tile, cb, ce, tb, te, seq, ts, done);
    EncRange nr{0, 0, 0, 0};                                             // tid 0: byte ranges of the tile after this one
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(enc_smem_u32(&es->mbar[0])), "r"(1) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(enc_smem_u32(&es->mbar[1])), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (tile < num_tiles) enc_issue_loads(a, es, s_cid, s_tok, 0, enc_tile_range(a, tile, c_total, t_total));
        if (tile + G < num_tiles) nr = enc_tile_range(a, tile + G, c_total, t_total);
    }
    __syncthreads();

    uint32_t par0 = 0, par1 = 0;
    for (uint32_t k = 0; tile < num_tiles; tile += G, k++) {
        const uint32_t buf = k & 1u;
        const uint32_t next = tile + G;
        // one tile ahead: bulk copies into the other buffer (every thread finished reading it before the last barrier of
        // the previous trip), scalars into registers
        if (tid == 0 && next < num_tiles) {
            enc_issue_loads(a, es, s_cid, s_tok, buf ^ 1u, nr);
            if (next + G < num_tiles) nr = enc_tile_range(a, next + G, c_total, t_total);
        }
        uint32_t ncb = 0, nce = 0, ntb = 0, nte = 0, ndone = 0;
        int64_t nseq = 0, nts = 0;
        load_scalars(next, ncb, nce, ntb, nte, nseq, nts, ndone);

        enc_mbar_wait(&es->mbar[buf], buf ? par1 : par0);
        if (buf) par1 ^= 1u; else par0 ^= 1u;
        const bool in_smem = es->in_smem[buf] != 0;
        const uint32_t c0a = es->c0a[buf], t0a = es->t0a[buf];
        // generic base pointers: shared memory (shifted so that arena offsets index it directly) or the arenas themselves
        const uint8_t* cbase = in_smem ? s_cid + buf * (ENC_CID_CAP + ENC_SLACK) - c0a : a.cid;
        const uint8_t* tbase = in_smem ? s_tok + buf * (ENC_TOK_CAP + ENC_SLACK) - t0a : a.tok;
        PtrSrc cid{cbase}, tok{tbase};

        const uint32_t i = tile * ENC_TILE + tid;
        const bool active = i < a.n;
        EncPlan p{};
        if (active) {
            // keep every string inside the tile's byte range (what is in shared memory), whatever the offsets say
            const uint32_t clo = c0a, chi = min(c_total, c0a + (in_smem ? ENC_CID_CAP : 0xFFFFFFFFu - c0a));
            const uint32_t tlo = t0a, thi = min(t_total, t0a + (in_smem ? ENC_TOK_CAP : 0xFFFFFFFFu - t0a));
            cb = min(max(cb, clo), chi); ce = min(max(ce, cb), chi);
            tb = min(max(tb, tlo), thi); te = min(max(te, tb), thi);
        }
        // plan: plain strings are their own spelling; the others are measured by their lane
        bool cplain = true, tplain = true;
        if (active) { cplain = go_string_is_plain(cid, cb, ce); tplain = go_string_is_plain(tok, tb, te); }
        uint32_t cid_n = ce - cb + 2, tok_n = te - tb + 2;
        if (!cplain) cid_n = go_string_len(cid, cb, ce, go122);
        if (!tplain) tok_n = go_string_len(tok, tb, te, go122);
        if (active) p = encode_plan_from(ce - cb, cid_n, cplain, tok_n, tplain, seq, done != 0, ts, a.op);
        const uint32_t len = p.len;
        uint32_t incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) es->wsum[warp] = incl;
        if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the previous tile's store has read s_out
        __syncthreads();                                                              // (B)
        uint32_t wpre = 0, total = 0;
#pragma unroll
        for (int w = 0; w < ENC_TILE / 32; w++) { uint32_t v = es->wsum[w]; if (w < (int)warp) wpre += v; total += v; }
        const uint32_t alloc = (total + 15u) & ~15u;
        if (tid == 0) es->base = atomicAdd(a.cursor, (unsigned long long)alloc);        // the answer is not needed until after the emission
        const uint32_t local = wpre + incl - len;
        const bool staged = alloc <= stage_bytes;                                     // tile-uniform
        if (staged) {
            if (active) {
                Ptr32Dst d{s_out};
                Writer<Ptr32Dst, true> w(d, local);
                encode_emit(cid, cb, ce, tok, tb, te, seq, done != 0, ts, go122, p, w);
            }
            if (tid < alloc - total) s_out[total + tid] = 0;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");              // staging writes -> visible to the bulk store
        }
        __syncthreads();                                                              // (D) staging complete, base visible, input buffer consumed
        const uint64_t base = es->base;
        const bool fits = base + alloc <= a.out_cap;
        if (staged) {
            if (tid == 0 && fits && alloc) {                                          // one TMA bulk store for the whole tile
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.out + base), "r"(enc_smem_u32(s_out)), "r"(alloc) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        } else if (fits) {                                                            // oversize tile: straight into the arena
            if (active) { EncPlan pp = p; encode_emit_global(a.op, cbase, tbase, cb, ce, tb, te, seq, done != 0, ts, &pp, a.out, base + local); }
            if (tid < alloc - total) a.out[base + total + tid] = 0;
            __syncthreads();                                                          // input buffer consumed (global-path readers are done)
        }
        if (active) {
            a.out_off[i] = base + local;
            a.out_len[i] = len;
            if (a.chan_len) a.chan_len[i] = p.chan;
        }
        if (tid == 0 && !fits) atomicExch(a.overflow, 1u);
        cb = ncb; ce = nce; tb = ntb; te = nte; seq = nseq; ts = nts; done = ndone;
    }
    if (tid == 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        __threadfence();
        if (atomicAdd(a.done_ctas, 1u) == gridDim.x - 1) {
            __threadfence();
            a.out_off[a.n] = (uint64_t)atomicAdd(a.cursor, 0ull);
        }
    }
}

static int g_enc_grid[64][2];

cudaError_t encode_launch(EncodeArgs a, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(a.scratch_base, 0, SCRATCH_HDR, stream);
    if (e != cudaSuccess) return e;
    if (a.n == 0) return cudaMemsetAsync(a.out_off, 0, sizeof(uint64_t), stream);
    const bool wrapped = (a.op & (OP_BRIDGE | OP_RESP)) != 0;
    const uint32_t stage = wrapped ? ENC_STAGE_WRAPPED : ENC_STAGE_PLAIN;
    const uint32_t smem = stage + 16 + 2 * (ENC_CID_CAP + ENC_SLACK) + 2 * (ENC_TOK_CAP + ENC_SLACK) + (uint32_t)sizeof(EncSmem);
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (g_enc_grid[dev][wrapped] == 0) {
        int sms = 0, per = 0;
        if ((e = cudaFuncSetAttribute(k_encode_messages, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ENC_STAGE_WRAPPED + 16 + 2 * (ENC_CID_CAP + ENC_SLACK) + 2 * (ENC_TOK_CAP + ENC_SLACK) + sizeof(EncSmem)))) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_encode_messages, ENC_TILE, smem)) != cudaSuccess) return e;
        g_enc_grid[dev][wrapped] = sms * (per < 1 ? 1 : per);
    }
    const uint32_t num_tiles = (a.n + ENC_TILE - 1) / ENC_TILE;
    const uint32_t grid = min((uint32_t)g_enc_grid[dev][wrapped], num_tiles);
    k_encode_messages<<<grid, ENC_TILE, smem, stream>>>(a, stage, num_tiles);
    return cudaGetLastError();
}

}  // namespace sse
