// sse_chunk.cu -- batched decode of vLLM stream lines into the encoder's token columns (row N3, decode half;
// semantics in sse_chunk.cuh, reference src/llm-stream-proxy/main.go:188-226).
//   1. k_chunk_count   one lane per line: status, number of messages it publishes, their decoded token bytes
//   2. exclusive scans of both counts
//   3. k_chunk_emit    the same decode again, writing token bytes / offsets / done flags at their final places
#include <cuda_runtime.h>
#include <stdint.h>
#include "sse_kernels.h"
#include "sse_chunk.cuh"

namespace sse {

namespace {

__device__ __noinline__ uint32_t chunk_line_dev(const uint8_t* base, uint32_t b, uint32_t e, ChunkEvents* ev) {
    uint32_t nest[NEST_WORDS];
    PtrSrc src{base};
    return chunk_line(src, b, e, *ev, nest);
}

__global__ void __launch_bounds__(128) k_chunk_count(const ChunkArgs a) {
    const uint32_t i = blockIdx.x * 128 + threadIdx.x;
    if (i >= a.n_lines) return;
    ChunkEvents ev;
    const uint32_t st = chunk_line_dev(a.lines, a.line_off[i], a.line_off[i + 1], &ev);
    const PtrSrc src{a.lines};
    uint32_t msgs = 0, bytes = 0, cb = 0, ce = 0;
    if (!(st & CHUNK_TOO_MANY)) {
        if ((st & CHUNK_KIND_MASK) == CHUNK_DATA)
            for (uint32_t k = 0; k < ev.len; k++)
                if (ev.ch[k].has_content && ev.ch[k].ce > ev.ch[k].cb) { msgs++; bytes += chunk_content_bytes(src, ev.ch[k]); cb = ev.ch[k].cb; ce = ev.ch[k].ce; }
        // the usual line carries one delta: its content span is left for the emit pass, which then does not decode the line again
        if (msgs != 1) cb = ce = 0;
        if (st & CHUNK_ENDS) { msgs++; bytes += 6; }
    }
    a.span[2 * (size_t)i] = cb; a.span[2 * (size_t)i + 1] = ce;
    a.line_status[i] = st; a.cnt_msgs[i] = msgs; a.cnt_bytes[i] = bytes;
    if (i == a.n_lines - 1) { a.totals[3] = msgs; a.totals[4] = bytes; }
}

__global__ void k_chunk_totals(const ChunkArgs a) {
    const uint32_t n = a.cnt_msgs[a.n_lines - 1] + a.totals[3], bytes = a.cnt_bytes[a.n_lines - 1] + a.totals[4];
    a.totals[0] = n; a.totals[1] = bytes;
    const bool over = n > a.max_msgs || (uint64_t)bytes + 48 > a.tok_cap;
    a.totals[2] = over ? 1u : 0u;
    if (!over) { a.tok_off[n] = bytes; for (uint32_t k = 0; k < 48; k++) a.tok[bytes + k] = 0; }
}

__global__ void __launch_bounds__(128) k_chunk_emit(const ChunkArgs a) {
    const uint32_t i = blockIdx.x * 128 + threadIdx.x;
    if (i >= a.n_lines || a.totals[2]) return;
    const uint32_t st = a.line_status[i];
    if ((st & CHUNK_TOO_MANY) || !((st & CHUNK_ENDS) || (st & CHUNK_KIND_MASK) == CHUNK_DATA)) return;
    uint32_t m = a.cnt_msgs[i], w = a.cnt_bytes[i];
    const PtrSrc src{a.lines};
    const uint32_t cb = a.span[2 * (size_t)i], ce = a.span[2 * (size_t)i + 1];
    if ((st & CHUNK_KIND_MASK) == CHUNK_DATA && ce > cb) {              // one delta with content: span known from the count pass
        a.tok_off[m] = w; a.done[m] = 0; a.msg_line[m] = i;
        for_each_decoded_byte(src, cb, ce, [&](uint32_t b) { a.tok[w++] = (uint8_t)b; });
        m++;
    } else if ((st & CHUNK_KIND_MASK) == CHUNK_DATA) {
        ChunkEvents ev;
        chunk_line_dev(a.lines, a.line_off[i], a.line_off[i + 1], &ev);
        for (uint32_t k = 0; k < ev.len; k++) {
            if (!(ev.ch[k].has_content && ev.ch[k].ce > ev.ch[k].cb)) continue;
            a.tok_off[m] = w; a.done[m] = 0; a.msg_line[m] = i;
            for_each_decoded_byte(src, ev.ch[k].cb, ev.ch[k].ce, [&](uint32_t b) { a.tok[w++] = (uint8_t)b; });
            m++;
        }
    }
    if (st & CHUNK_ENDS) {
        a.tok_off[m] = w; a.done[m] = 1; a.msg_line[m] = i;
        const char* dn = "[DONE]";
        for (uint32_t k = 0; k < 6; k++) a.tok[w++] = (uint8_t)dn[k];
    }
}

}  // namespace

size_t chunk_scratch_bytes(uint32_t n_lines) { return ((size_t)n_lines * 4 + (size_t)n_lines / 4096 + 32) * 4; }

cudaError_t chunk_decode_launch(ChunkArgs a, void* scratch, cudaStream_t stream) {
    uint32_t* p = static_cast<uint32_t*>(scratch);
    a.totals = p; p += 8;
    a.cnt_msgs = p; p += a.n_lines;
    a.cnt_bytes = p; p += a.n_lines;
    a.span = p; p += 2 * (size_t)a.n_lines;
    a.block_sums = p;
    cudaError_t e = cudaMemsetAsync(a.totals, 0, 32, stream);
    if (e != cudaSuccess) return e;
    if (a.n_lines == 0) return cudaMemsetAsync(a.tok_off, 0, 4, stream);
    const uint32_t blocks = (a.n_lines + 127) / 128;
    k_chunk_count<<<blocks, 128, 0, stream>>>(a);
    if ((e = scan_exclusive_u32(a.cnt_msgs, a.n_lines, a.block_sums, stream)) != cudaSuccess) return e;
    if ((e = scan_exclusive_u32(a.cnt_bytes, a.n_lines, a.block_sums, stream)) != cudaSuccess) return e;
    k_chunk_totals<<<1, 1, 0, stream>>>(a);
    k_chunk_emit<<<blocks, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace sse
