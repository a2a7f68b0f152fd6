// sse_wire.cu -- batched splitter of NATS / RESP socket bytes into the packed payload arena the frame kernel reads
// (receiving side; semantics in sse_wire.cuh).
//   1. k_wire_count   one lane per connection walks its buffer: number of deliveries, payload bytes, bytes consumed
//   2. two exclusive scans (deliveries, payload bytes) over the connections (the multi-CTA scan of sse_stream.cu)
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
    if (c == a.n_conn - 1) { a.totals[4] = r.n_msgs; a.totals[5] = r.payload_bytes; }
    a.conn_consumed[c] = r.consumed; a.conn_status[c] = r.status; a.conn_pings[c] = r.pings;
}

// totals of the two scans = exclusive prefix of the last connection + its own counts (parked by k_wire_count)
__global__ void k_wire_totals(const WireArgs a) {
    a.totals[0] = a.cnt_msgs[a.n_conn - 1] + a.totals[4];
    a.totals[1] = a.cnt_bytes[a.n_conn - 1] + a.totals[5];
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
        // destination words: bytes up to the first 4-byte boundary of the arena, whole words (source read as two aligned
        // words and funnel-shifted: source and destination are not aligned to each other), the bytes behind the last word
        const uint32_t head = min(len, (4u - (d0 & 3u)) & 3u);
        if (lane < head) a.arena[d0 + lane] = a.buf[s0 + lane];
        const uint32_t nw = (len - head) >> 2, sb = s0 + head, sh = (sb & 3u) * 8;
        const uint32_t* src_w = reinterpret_cast<const uint32_t*>(a.buf + (sb & ~3u));
        uint32_t* dst_w = reinterpret_cast<uint32_t*>(a.arena + d0 + head);
        for (uint32_t k = lane; k < nw; k += 32) dst_w[k] = __funnelshift_r(src_w[k], src_w[k + 1], sh);
        const uint32_t done = head + 4 * nw;
        if (lane < len - done) a.arena[d0 + done + lane] = a.buf[s0 + done + lane];
    }
    if (blockIdx.x == 0 && threadIdx.x < 48) a.arena[a.totals[1] + threadIdx.x] = 0;         // readable padding behind the last payload
}

}  // namespace

size_t wire_scratch_bytes(uint32_t n_conn, uint32_t max_msgs) { return ((size_t)n_conn * 2 + (size_t)max_msgs + (size_t)n_conn / 4096 + 32) * 4; }

cudaError_t wire_split_launch(WireArgs a, void* scratch, cudaStream_t stream) {
    uint32_t* p = static_cast<uint32_t*>(scratch);
    a.totals = p; p += 8;
    a.cnt_msgs = p; p += a.n_conn;
    a.cnt_bytes = p; p += a.n_conn;
    uint32_t* block_sums = p; p += a.n_conn / 4096 + 2;
    a.msg_src = p;
    cudaError_t e = cudaMemsetAsync(a.totals, 0, 32, stream);
    if (e != cudaSuccess) return e;
    if (a.n_conn == 0) return cudaMemsetAsync(a.in_off, 0, 4, stream);
    const uint32_t blocks = (a.n_conn + 127) / 128;
    k_wire_count<<<blocks, 128, 0, stream>>>(a);
    if ((e = scan_exclusive_u32(a.cnt_msgs, a.n_conn, block_sums, stream)) != cudaSuccess) return e;
    if ((e = scan_exclusive_u32(a.cnt_bytes, a.n_conn, block_sums, stream)) != cudaSuccess) return e;
    k_wire_totals<<<1, 1, 0, stream>>>(a);
    k_wire_emit<<<blocks, 128, 0, stream>>>(a);
    int sms = 148;
    k_wire_copy<<<sms * 8, 256, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace sse
