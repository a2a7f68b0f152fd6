// sse_encode.cu -- producer-side batch encoder: TokenMessage columns -> canonical JSON payloads.
//
// Stands in for publishToken() of the LLM stream proxy (src/llm-stream-proxy/main.go:230-245: build the
// TokenMessage, json.Marshal it, derive the channel "chat.<id>.tokens") and for the load generator's producer
// (demo/load-generator/main.go:215-223), batched over every conversation that produced a delta in this interval.
// Output records are exactly what the adapter/bridge kernels recognise as canonical, so the downstream
// fast path always hits.  One thread per message: measure, CTA scan, one atomic for the tile's 16-byte aligned
// slice of the arena (tiles land in completion order, like k_frame_tiles), then emit straight into the arena.
#include <cuda_runtime.h>
#include <stdint.h>
#include "sse_kernels.h"
#include "sse_record.cuh"

namespace sse {

constexpr uint32_t ENC_STAGE = 24 * 1024;       // staging tile: 128 records x up to 192 bytes (a config-1 record is ~133)

__device__ __forceinline__ uint32_t enc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(TILE) k_encode_messages(const EncodeArgs a) {
    __shared__ __align__(128) uint8_t s_out[ENC_STAGE + 16];
    __shared__ uint32_t wsum[TILE / 32];
    __shared__ unsigned long long s_base;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t i = blockIdx.x * TILE + tid;
    const bool active = i < a.n;
    const bool go122 = (a.op & OP_GO122) != 0;
    PtrSrc cid{a.cid}, tok{a.tok};
    uint32_t cb = 0, ce = 0, tb = 0, te = 0;
    int64_t seq = 0, ts = 0;
    bool done = false;
    EncPlan p{};
    if (active) {
        cb = a.cid_off[i]; ce = max(cb, a.cid_off[i + 1]);
        tb = a.tok_off[i]; te = max(tb, a.tok_off[i + 1]);
        seq = a.seq[i]; ts = a.ts[i]; done = a.done[i] != 0;
        p = encode_plan(cid, cb, ce, tok, tb, te, seq, done, ts, a.op);
    }
    const uint32_t len = p.len;
    uint32_t incl = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += v;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    uint32_t wpre = 0, total = 0;
#pragma unroll
    for (int w = 0; w < TILE / 32; w++) { uint32_t v = wsum[w]; if (w < (int)warp) wpre += v; total += v; }
    const uint32_t alloc = (total + 15u) & ~15u;
    if (tid == 0) s_base = atomicAdd(a.cursor, (unsigned long long)alloc);     // the answer is not needed until after the emission
    const uint32_t local = wpre + incl - len;
    const bool staged = alloc <= ENC_STAGE;                                      // tile-uniform
    if (staged) {
        if (active) {
            Ptr32Dst d{s_out};
            Writer<Ptr32Dst, true> w(d, local);
            encode_emit(cid, cb, ce, tok, tb, te, seq, done, ts, go122, p, w);
        }
        if (tid < alloc - total) s_out[total + tid] = 0;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");             // staging writes -> visible to the bulk store
    }
    __syncthreads();
    const uint64_t base = s_base;
    const bool fits = base + alloc <= a.out_cap;
    if (staged) {
        if (tid == 0 && fits && alloc) {                                         // one TMA bulk store for the whole tile
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.out + base), "r"(enc_smem_u32(s_out)), "r"(alloc) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    } else if (fits) {                                                           // oversize tile: straight into the arena
        if (active) {
            PtrDst d{a.out};
            Writer<PtrDst, true> w(d, base + local);
            encode_emit(cid, cb, ce, tok, tb, te, seq, done, ts, go122, p, w);
        }
        if (tid < alloc - total) a.out[base + total + tid] = 0;
    }
    if (active) {
        a.out_off[i] = base + local;
        a.out_len[i] = len;
        if (a.chan_len) a.chan_len[i] = p.chan;
    }
    if (tid == 0) {
        if (!fits) atomicExch(a.overflow, 1u);
        __threadfence();
        if (atomicAdd(a.done_ctas, 1u) == gridDim.x - 1) {
            __threadfence();
            a.out_off[a.n] = (uint64_t)atomicAdd(a.cursor, 0ull);
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

cudaError_t encode_launch(EncodeArgs a, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(a.scratch_base, 0, SCRATCH_HDR, stream);
    if (e != cudaSuccess) return e;
    if (a.n == 0) return cudaMemsetAsync(a.out_off, 0, sizeof(uint64_t), stream);
    k_encode_messages<<<(a.n + TILE - 1) / TILE, TILE, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace sse
