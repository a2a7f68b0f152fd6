// sse_stream.cu -- streaming content inspection: one DFA walk per CONVERSATION over the tokens a batch holds for it.
//
// The reference's inspection hook sits in the per-connection loop (src/sse-adapter/sse_handler.go:193-208) and the
// pattern document scans a per-conversation context window of consecutive tokens
// (docs/security-inspection-patterns.md:236): a pattern may begin in one token and end in a later one.  A batch cut by
// size or time may hold any number of tokens of one conversation (bursts, catch-up after a reconnect), so the carried
// DFA state makes the tokens of a conversation a sequential chain while different conversations stay independent.
//
// Pipeline (all on the caller's stream, after the frame kernel has left one 32-byte token entry per record):
//   1. stable LSD radix sort of the record numbers by conversation slot (1-4 passes of <= 10 bits: per-tile digit
//      histogram, exclusive scan, stable scatter with per-warp running offsets and match.any ranking) -- after it every
//      conversation's records are contiguous and in arrival order;
//   2. segment heads are compacted into a list;
//   3. one lane per conversation walks its segment: carried state in a register, forward DFA in shared memory (the
//      whole table when it fits), token bytes from the entry itself (<= 24 raw bytes: every BASELINE token) or from
//      the payload arena; the next record's entry is requested before the current one is walked; only hits are written (verdict/match are cleared first); the state is committed once.
// No atomics on the conversation state, no limit on how often a conversation occurs in a batch.
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include "sse_kernels.h"
#include "sse_record.cuh"
#include "sse_inspect.cuh"

namespace sse {

namespace {

constexpr int SORT_WARPS = 8, SORT_THREADS = SORT_WARPS * 32;
constexpr uint32_t SORT_PER_WARP = 512, SORT_TILE = SORT_WARPS * SORT_PER_WARP;      // keys per CTA
constexpr uint32_t MAX_DIGIT_BITS = 10;
constexpr uint32_t SCAN_BLOCK = 4096;                                                // elements per scan CTA (1024 threads x 4)
constexpr uint32_t WALK_THREADS = 256;
constexpr uint32_t WALK_TABLE_MAX = 96 * 1024;                                       // bytes of transition rows kept on chip by the walk

struct SortPlan { uint32_t bits, passes, digit_bits; };
SortPlan sort_plan(uint32_t conv_slots) {
    uint32_t bits = 1;
    while (bits < 32 && (1ull << bits) < conv_slots) bits++;
    SortPlan p;
    p.bits = bits;
    p.passes = (bits + MAX_DIGIT_BITS - 1) / MAX_DIGIT_BITS;
    p.digit_bits = (bits + p.passes - 1) / p.passes;
    return p;
}

// ---- 1a. per-tile digit histogram: hist[digit * num_tiles + tile] ----
// Keys come from the caller's conv_slot array in the first pass (record number = position) and from the previous
// pass's (key, record number) pairs afterwards: `stride` is 1 or 2 words.
__global__ void __launch_bounds__(SORT_THREADS) k_radix_hist(const uint32_t* __restrict__ keys, uint32_t stride, uint32_t n, uint32_t shift, uint32_t digit_bits,
                                                            uint32_t* __restrict__ hist, uint32_t num_tiles, uint32_t conv_slots,
                                                            uint32_t* bad_conv) {
    extern __shared__ uint32_t s_hist[];
    const uint32_t radix = 1u << digit_bits, mask = radix - 1;
    for (uint32_t d = threadIdx.x; d < radix; d += SORT_THREADS) s_hist[d] = 0;
    __syncthreads();
    const uint32_t lo = blockIdx.x * SORT_TILE, hi = min(n, lo + SORT_TILE);
    bool bad = false;
    for (uint32_t i = lo + threadIdx.x; i < hi; i += SORT_THREADS) {
        const uint32_t k = keys[(size_t)i * stride];
        bad |= k >= conv_slots;
        atomicAdd(&s_hist[(k >> shift) & mask], 1u);
    }
    if (bad_conv && bad) atomicExch(bad_conv, 1u);
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < radix; d += SORT_THREADS) hist[(size_t)d * num_tiles + blockIdx.x] = s_hist[d];
}

// ---- 1b. exclusive scan of a u32 array in place (three small kernels) ----
__device__ __forceinline__ uint32_t block_exclusive_scan_1024(uint32_t v, uint32_t* s_warp, uint32_t* total) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= (uint32_t)d) incl += t; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = s_warp[lane], wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, d); if (lane >= (uint32_t)d) wi += t; }
        s_warp[lane] = wi - w;
        if (lane == 31) s_warp[32] = wi;
    }
    __syncthreads();
    *total = s_warp[32];
    return s_warp[warp] + incl - v;
}
__global__ void __launch_bounds__(1024) k_scan_reduce(const uint32_t* __restrict__ data, size_t m, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t s_warp[33];
    const size_t base = (size_t)blockIdx.x * SCAN_BLOCK + threadIdx.x * 4;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) if (base + k < m) v += data[base + k];
    uint32_t total;
    block_exclusive_scan_1024(v, s_warp, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(1024) k_scan_block_sums(uint32_t* block_sums, uint32_t nb) {
    __shared__ uint32_t s_warp[33];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t b0 = 0; b0 < nb; b0 += 1024) {
        const uint32_t i = b0 + threadIdx.x;
        const uint32_t v = i < nb ? block_sums[i] : 0;
        uint32_t total;
        const uint32_t ex = block_exclusive_scan_1024(v, s_warp, &total);
        const uint32_t carry = s_carry;
        if (i < nb) block_sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + total;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(1024) k_scan_apply(uint32_t* __restrict__ data, size_t m, const uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t s_warp[33];
    const size_t base = (size_t)blockIdx.x * SCAN_BLOCK + threadIdx.x * 4;
    uint32_t x[4], v = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) { x[k] = base + k < m ? data[base + k] : 0; v += x[k]; }
    uint32_t total;
    uint32_t run = block_sums[blockIdx.x] + block_exclusive_scan_1024(v, s_warp, &total);
#pragma unroll
    for (int k = 0; k < 4; k++) { if (base + k < m) data[base + k] = run; run += x[k]; }
}

// ---- 1c. stable scatter: every warp owns SORT_PER_WARP consecutive keys and its own running offsets ----
// Output: (key, record number) pairs, one 8-byte store each (the scattered stores are what bounds this kernel).
__global__ void __launch_bounds__(SORT_THREADS) k_radix_scatter(const uint32_t* __restrict__ keys, const uint2* __restrict__ pairs_in, uint32_t n,
                                                               uint32_t shift, uint32_t digit_bits, const uint32_t* __restrict__ hist,
                                                               uint32_t num_tiles, uint2* __restrict__ pairs_out) {
    extern __shared__ uint32_t s_off[];                      // [SORT_WARPS][radix]
    const uint32_t radix = 1u << digit_bits, mask = radix - 1;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t d = threadIdx.x; d < radix * SORT_WARPS; d += SORT_THREADS) s_off[d] = 0;
    __syncthreads();
    uint32_t* mine = s_off + warp * radix;
    const uint32_t w_lo = blockIdx.x * SORT_TILE + warp * SORT_PER_WARP, w_hi = min(n, w_lo + SORT_PER_WARP);
    for (uint32_t i = w_lo + lane; i < w_hi; i += 32) atomicAdd(&mine[((pairs_in ? pairs_in[i].x : keys[i]) >> shift) & mask], 1u);
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < radix; d += SORT_THREADS) {   // counts -> first output position of (warp, digit)
        uint32_t run = hist[(size_t)d * num_tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) { const uint32_t c = s_off[w * radix + d]; s_off[w * radix + d] = run; run += c; }
    }
    __syncthreads();
    for (uint32_t i0 = w_lo; i0 < w_hi; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool valid = i < w_hi;
        uint2 kv = make_uint2(0u, i);
        if (valid) { if (pairs_in) kv = pairs_in[i]; else kv.x = keys[i]; }
        const uint32_t d = valid ? ((kv.x >> shift) & mask) : 0xFFFFFFFFu;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1));
        uint32_t pos = 0;
        if (valid) pos = mine[d] + rank;
        __syncwarp();
        if (valid && rank == 0) mine[d] += __popc(peers);
        __syncwarp();
        if (valid) pairs_out[pos] = kv;
    }
}

// ---- 2. segment heads ----
__global__ void __launch_bounds__(1024) k_segment_heads(const uint2* __restrict__ pairs, uint32_t n, uint32_t* __restrict__ heads, uint32_t* n_heads) {
    __shared__ uint32_t s_cnt[33];
    const uint32_t i = blockIdx.x * 1024 + threadIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool head = i < n && (i == 0 || pairs[i].x != pairs[i - 1].x);
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, head);
    if (lane == 0) s_cnt[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (warp == 0) {                                  // one atomic per CTA: the list keeps runs of 1024 positions in order
        const uint32_t c = s_cnt[lane];
        uint32_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= (uint32_t)d) incl += t; }
        uint32_t base = 0;
        if (lane == 31 && incl) base = atomicAdd(n_heads, incl);
        base = __shfl_sync(0xFFFFFFFFu, base, 31);
        s_cnt[lane] = base + incl - c;
    }
    __syncthreads();
    if (head) heads[s_cnt[warp] + __popc(m & ((1u << lane) - 1))] = i;
}

// ---- 3. the walk: one lane per conversation ----
__global__ void __launch_bounds__(WALK_THREADS) k_stream_walk(const StreamArgs a, const uint2* __restrict__ pairs,
                                                             const uint32_t* __restrict__ heads, const uint32_t* __restrict__ n_heads,
                                                             uint32_t n_cached) {
    extern __shared__ __align__(16) uint8_t s_tab[];
    uint8_t* s_cls = s_tab;
    uint32_t* s_eot = reinterpret_cast<uint32_t*>(s_cls + 256);
    uint16_t* s_rows = reinterpret_cast<uint16_t*>(s_eot + DFA_EOT_WORDS);
    const uint32_t nh = *n_heads;
    if (blockIdx.x * WALK_THREADS >= nh) return;                          // the grid is sized for the worst case (every record its own conversation)
    const InspectDev& t = *a.insp;
    {   // table -> shared memory (descriptor fields first into registers: the descriptor itself lives in global memory)
        const uint32_t* g_cls = reinterpret_cast<const uint32_t*>(t.fwd.cls);
        const uint32_t* g_next = reinterpret_cast<const uint32_t*>(t.fwd.next);
        const uint32_t* g_eot = t.fwd.acc_eot;
        const uint32_t n_states = t.fwd.n_states, row_words = (n_cached * t.fwd.n_classes + 1) / 2;
        for (uint32_t i = threadIdx.x; i < 64; i += WALK_THREADS) reinterpret_cast<uint32_t*>(s_cls)[i] = g_cls[i];
        for (uint32_t i = threadIdx.x; i < row_words; i += WALK_THREADS) reinterpret_cast<uint32_t*>(s_rows)[i] = g_next[i];
        if (n_states <= 32 * DFA_EOT_WORDS)
            for (uint32_t w = threadIdx.x; w < DFA_EOT_WORDS; w += WALK_THREADS) {
                uint32_t bits = 0;
                for (uint32_t k = 0; k < 32 && 32 * w + k < n_states; k++) bits |= (g_eot[32 * w + k] != 0 ? 1u : 0u) << k;
                s_eot[w] = bits;
            }
    }
    __syncthreads();
    const DfaCache cache = make_dfa_cache(t, s_cls, s_rows, n_cached, s_eot);
    const PtrSrc arena{a.in};
    for (uint32_t h = blockIdx.x * WALK_THREADS + threadIdx.x; h < nh; h += gridDim.x * WALK_THREADS) {
        uint32_t q = heads[h];
        uint2 cur = pairs[q];                                             // (conversation slot, record number)
        const uint32_t slot = cur.x;
        if (slot >= a.conv_slots) continue;                               // reported through bad_conv by the sort
        uint32_t state = a.conv_state[slot];
        // software pipeline: while record q is walked, the entries of q+1 and q+2 and the pair of q+3 are already on
        // their way (the loads are speculative: any position below n holds a valid record number)
        const uint2 none = make_uint2(~slot, 0u);
        uint2 p1 = q + 1 < a.n ? pairs[q + 1] : none, p2 = q + 2 < a.n ? pairs[q + 2] : none;
        uint4 ea = a.tok_entry[2 * (size_t)cur.y], eb = a.tok_entry[2 * (size_t)cur.y + 1];
        uint4 fa = a.tok_entry[2 * (size_t)p1.y], fb = a.tok_entry[2 * (size_t)p1.y + 1];
        for (;;) {
            const uint2 p3 = q + 3 < a.n ? pairs[q + 3] : none;
            const uint4 ga = a.tok_entry[2 * (size_t)p2.y], gb = a.tok_entry[2 * (size_t)p2.y + 1];
            if ((ea.y >> 24) & ENTRY_OK) {                                // else: dropped or filtered before the hook, never inspected
                const uint32_t e[ENTRY_WORDS] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
                const InspectResult r = stream_walk(t, cache, arena, e, state);
                state = r.state;
                if (r.verdict) {
                    a.verdict[cur.y] = r.verdict;
                    reinterpret_cast<int2*>(a.match)[cur.y] = make_int2(r.begin, r.end);
                }
            }
            if (p1.x != slot) break;
            q++; cur = p1; p1 = p2; p2 = p3; ea = fa; eb = fb; fa = ga; fb = gb;
        }
        a.conv_state[slot] = state;
    }
}

std::mutex g_mu;
bool g_configured[64];
int g_sms[64];

uint32_t* carve(uint8_t*& p, size_t words) {
    uint32_t* r = reinterpret_cast<uint32_t*>(p);
    p += ((words * 4 + 255) & ~(size_t)255);
    return r;
}
size_t hist_words(uint32_t n) { return ((size_t)(n + SORT_TILE - 1) / SORT_TILE) << MAX_DIGIT_BITS; }

}  // namespace

// exclusive scan of data[0, m) in place; block_sums: m / 4096 + 2 words of scratch
cudaError_t scan_exclusive_u32(uint32_t* data, size_t m, uint32_t* block_sums, cudaStream_t stream) {
    if (m == 0) return cudaSuccess;
    const uint32_t nb = (uint32_t)((m + SCAN_BLOCK - 1) / SCAN_BLOCK);
    k_scan_reduce<<<nb, 1024, 0, stream>>>(data, m, block_sums);
    k_scan_block_sums<<<1, 1024, 0, stream>>>(block_sums, nb);
    k_scan_apply<<<nb, 1024, 0, stream>>>(data, m, block_sums);
    return cudaGetLastError();
}

size_t stream_scratch_bytes(uint32_t n) {
    const size_t a = ((size_t)n * 4 + 255) & ~(size_t)255;
    const size_t h = (hist_words(n) * 4 + 255) & ~(size_t)255;
    const size_t bs = ((hist_words(n) / SCAN_BLOCK + 2) * 4 + 255) & ~(size_t)255;
    return 5 * a + h + bs + 256;
}

cudaError_t stream_configure(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_configured[device]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k_radix_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(4u << MAX_DIGIT_BITS));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_radix_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((4u * SORT_WARPS) << MAX_DIGIT_BITS));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_stream_walk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(256 + 4 * DFA_EOT_WORDS + WALK_TABLE_MAX));
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&g_sms[device], cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    g_configured[device] = true;
    return cudaSuccess;
}

cudaError_t stream_inspect_launch(int device, const StreamArgs& a, cudaStream_t stream) {
    cudaError_t e = stream_configure(device);
    if (e != cudaSuccess) return e;
    if (a.n == 0) return cudaSuccess;
    const uint32_t n = a.n;
    uint8_t* p = static_cast<uint8_t*>(a.scratch);
    uint2* pA = reinterpret_cast<uint2*>(carve(p, 2 * (size_t)n)); uint2* pB = reinterpret_cast<uint2*>(carve(p, 2 * (size_t)n));
    uint32_t* heads = carve(p, n);
    uint32_t* hist = carve(p, hist_words(n));
    uint32_t* block_sums = carve(p, hist_words(n) / SCAN_BLOCK + 2);
    uint32_t* n_heads = carve(p, 1);
    if ((e = cudaMemsetAsync(a.verdict, 0, (size_t)n * 4, stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(a.match, 0, (size_t)n * 8, stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(n_heads, 0, 4, stream)) != cudaSuccess) return e;
    const SortPlan sp = sort_plan(a.conv_slots);
    const uint32_t num_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    const uint32_t radix = 1u << sp.digit_bits;
    const size_t m = (size_t)radix * num_tiles;
    const uint32_t nb = (uint32_t)((m + SCAN_BLOCK - 1) / SCAN_BLOCK);
    const uint2* in = nullptr;
    uint2* out = pA;
    for (uint32_t pass = 0; pass < sp.passes; pass++) {
        const uint32_t shift = pass * sp.digit_bits;
        const uint32_t* k_in = in ? reinterpret_cast<const uint32_t*>(in) : a.conv_slot;
        k_radix_hist<<<num_tiles, SORT_THREADS, radix * 4, stream>>>(k_in, in ? 2u : 1u, n, shift, sp.digit_bits, hist, num_tiles, a.conv_slots,
                                                                      pass == 0 ? a.bad_conv : nullptr);
        k_scan_reduce<<<nb, 1024, 0, stream>>>(hist, m, block_sums);
        k_scan_block_sums<<<1, 1024, 0, stream>>>(block_sums, nb);
        k_scan_apply<<<nb, 1024, 0, stream>>>(hist, m, block_sums);
        k_radix_scatter<<<num_tiles, SORT_THREADS, radix * 4 * SORT_WARPS, stream>>>(a.conv_slot, in, n, shift, sp.digit_bits, hist, num_tiles, out);
        in = out;
        out = (out == pA) ? pB : pA;
    }
    k_segment_heads<<<(n + 1023) / 1024, 1024, 0, stream>>>(in, n, heads, n_heads);
    uint32_t n_cached = a.dfa_states;
    if ((size_t)n_cached * a.dfa_classes * 2 > WALK_TABLE_MAX) n_cached = WALK_TABLE_MAX / (a.dfa_classes * 2);
    const size_t smem = 256 + 4 * DFA_EOT_WORDS + (((size_t)n_cached * a.dfa_classes * 2 + 15) & ~(size_t)15);
    int grid = (int)((n + WALK_THREADS - 1) / WALK_THREADS);
    const int cap = g_sms[device] * 6;
    if (grid > cap) grid = cap;
    k_stream_walk<<<grid, WALK_THREADS, smem, stream>>>(a, in, heads, n_heads, n_cached);
    return cudaGetLastError();
}

}  // namespace sse
