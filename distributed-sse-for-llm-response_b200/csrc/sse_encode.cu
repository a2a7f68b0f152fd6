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

__global__ void __launch_bounds__(TILE) k_encode_messages(const EncodeArgs a) {
    __shared__ uint32_t wsum[TILE / 32];
    __shared__ unsigned long long s_base;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t i = blockIdx.x * TILE + tid;
    const bool active = i < a.n;
    const bool go122 = (a.op & OP_GO122) != 0, chan = (a.op & OP_BRIDGE) != 0;
    PtrSrc cid{a.cid}, tok{a.tok};
    uint32_t cb = 0, ce = 0, tb = 0, te = 0, len = 0;
    int64_t seq = 0, ts = 0;
    bool done = false;
    if (active) {
        cb = a.cid_off[i]; ce = max(cb, a.cid_off[i + 1]);
        tb = a.tok_off[i]; te = max(tb, a.tok_off[i + 1]);
        seq = a.seq[i]; ts = a.ts[i]; done = a.done[i] != 0;
        Ptr32Dst nd{nullptr};
        Writer<Ptr32Dst, false> w(nd, 0);
        encode_message(cid, cb, ce, tok, tb, te, seq, done, ts, go122, chan, w);
        len = w.pos();
    }
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
    if (tid == 0) s_base = atomicAdd(a.cursor, (unsigned long long)alloc);
    __syncthreads();
    const uint64_t base = s_base;
    const bool fits = base + alloc <= a.out_cap;
    const uint32_t local = wpre + incl - len;
    if (active) {
        uint32_t chan_len = 0;
        if (fits) {
            PtrDst d{a.out};
            Writer<PtrDst, true> w(d, base + local);
            chan_len = encode_message(cid, cb, ce, tok, tb, te, seq, done, ts, go122, chan, w);
        }
        a.out_off[i] = base + local;
        a.out_len[i] = len;
        if (a.chan_len) a.chan_len[i] = chan_len;
    }
    if (fits && tid < alloc - total) a.out[base + total + tid] = 0;
    if (tid == 0) {
        if (!fits) atomicExch(a.overflow, 1u);
        __threadfence();
        if (atomicAdd(a.done_ctas, 1u) == gridDim.x - 1) {
            __threadfence();
            a.out_off[a.n] = (uint64_t)atomicAdd(a.cursor, 0ull);
        }
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
