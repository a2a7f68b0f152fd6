// sse_wire.cu -- batched splitter of NATS / RESP socket bytes into the packed payload arena the frame kernel reads
// (receiving side; semantics in sse_wire.cuh).
//   1. k_wire_count   one lane per connection walks its buffer: number of deliveries, payload bytes, bytes consumed
//   2. two exclusive scans (deliveries, payload bytes) over the connections
//   3. k_wire_emit    the same walk again, now writing every delivery's descriptor at its final index
//   4. k_wire_copy    one warp per delivery copies the payload into the packed arena (in_off[] as the frame kernel wants it)
#include <cuda_runtime.h>
#include <stdint.h>
#include "sse_kernels.h"
#include "sse_wire.cuh"

namespace sse {

namespace {

__global__ void __launch_bounds__(128) k_wire_count(const WireArgs a) {
    const uint32_t c = blockIdx.x * 128 + threadIdx.x;
    if (c >= a.n_conn) return;
    const PtrSrc src{a.buf};
    const WireConn r = wire_walk(a.proto, src, a.buf_off[c], a.buf_off[c + 1], [](uint32_t, const WireMsg&) {});
    a.cnt_msgs[c] = r.n_msgs; a.cnt_bytes[c] = r.payload_bytes;
    a.conn_consumed[c] = r.consumed; a.conn_status[c] = r.status; a.conn_pings[c] = r.pings;
}

// serial-per-block exclusive scan of two u32 arrays at once (n_conn is a few hundred thousand at most: one CTA)
__global__ void __launch_bounds__(1024) k_wire_scan(uint32_t* a0, uint32_t* a1, uint32_t n, uint32_t* totals) {
    __shared__ uint32_t s0[1024], s1[1024];
    const uint32_t tid = threadIdx.x, per = (n + 1023) / 1024;
    const uint32_t lo = min(n, tid * per), hi = min(n, lo + per);
    uint32_t t0 = 0, t1 = 0;
    for (uint32_t i = lo; i < hi; i++) { t0 += a0[i]; t1 += a1[i]; }
    s0[tid] = t0; s1[tid] = t1;
    __syncthreads();
    for (uint32_t d = 1; d < 1024; d <<= 1) {
        const uint32_t v0 = tid >= d ? s0[tid - d] : 0, v1 = tid >= d ? s1[tid - d] : 0;
        __syncthreads();
        s0[tid] += v0; s1[tid] += v1;
        __syncthreads();
    }
    uint32_t r0 = s0[tid] - t0, r1 = s1[tid] - t1;
    for (uint32_t i = lo; i < hi; i++) { const uint32_t v0 = a0[i], v1 = a1[i]; a0[i] = r0; a1[i] = r1; r0 += v0; r1 += v1; }
    if (tid == 1023) { totals[0] = s0[1023]; totals[1] = s1[1023]; }
}

__global__ void __launch_bounds__(128) k_wire_emit(const WireArgs a) {
    const uint32_t c = blockIdx.x * 128 + threadIdx.x;
    const uint32_t total_msgs = a.totals[0], total_bytes = a.totals[1];
    if (c == 0) {
        const bool over = total_msgs > a.max_msgs || (uint64_t)total_bytes + 48 > a.arena_cap;
        a.totals[2] = over ? 1u : 0u;
        a.in_off[min(total_msgs, a.max_msgs)] = over ? 0u : total_bytes;
    }
    if (c >= a.n_conn) return;
    if (total_msgs > a.max_msgs || (uint64_t)total_bytes + 48 > a.arena_cap) return;        // reported, nothing written
    const PtrSrc src{a.buf};
    const uint32_t base = a.cnt_msgs[c];
    uint32_t dst = a.cnt_bytes[c];
    wire_walk(a.proto, src, a.buf_off[c], a.buf_off[c + 1], [&](uint32_t k, const WireMsg& m) {
        const uint32_t i = base + k;
        a.in_off[i] = dst;
        a.msg_src[i] = m.payload_off;
        a.msg_conn[i] = c;
        a.msg_sid[i] = m.sid;
        a.msg_subj[2 * (size_t)i] = m.subj_off; a.msg_subj[2 * (size_t)i + 1] = m.subj_len;
        dst += m.payload_len;
    });
}

__global__ void __launch_bounds__(256) k_wire_copy(const WireArgs a) {
    const uint32_t total_msgs = a.totals[0];
    if (a.totals[2]) return;
    const uint32_t lane = threadIdx.x & 31, warps = gridDim.x * 8;
    for (uint32_t i = blockIdx.x * 8 + (threadIdx.x >> 5); i < total_msgs; i += warps) {
        const uint32_t d0 = a.in_off[i], len = a.in_off[i + 1] - d0, s0 = a.msg_src[i];
        for (uint32_t k = lane; k < len; k += 32) a.arena[d0 + k] = a.buf[s0 + k];
    }
    if (blockIdx.x == 0 && threadIdx.x < 48) a.arena[a.totals[1] + threadIdx.x] = 0;         // readable padding behind the last payload
}

}  // namespace

size_t wire_scratch_bytes(uint32_t n_conn, uint32_t max_msgs) { return ((size_t)n_conn * 2 + (size_t)max_msgs + 16) * 4; }

cudaError_t wire_split_launch(WireArgs a, void* scratch, cudaStream_t stream) {
    uint32_t* p = static_cast<uint32_t*>(scratch);
    a.totals = p; p += 4;
    a.cnt_msgs = p; p += a.n_conn;
    a.cnt_bytes = p; p += a.n_conn;
    a.msg_src = p;
    cudaError_t e = cudaMemsetAsync(a.totals, 0, 16, stream);
    if (e != cudaSuccess) return e;
    if (a.n_conn == 0) return cudaMemsetAsync(a.in_off, 0, 4, stream);
    const uint32_t blocks = (a.n_conn + 127) / 128;
    k_wire_count<<<blocks, 128, 0, stream>>>(a);
    k_wire_scan<<<1, 1024, 0, stream>>>(a.cnt_msgs, a.cnt_bytes, a.n_conn, a.totals);
    k_wire_emit<<<blocks, 128, 0, stream>>>(a);
    int sms = 148;
    k_wire_copy<<<sms * 8, 256, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace sse
