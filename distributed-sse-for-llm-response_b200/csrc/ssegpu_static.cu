// ssegpu_static.cu -- ssegpu_frame_launch_static(): the frame kernel for callers that must not
// allocate GPU memory.
//
// Who that is: a service that shares its GPU with an inference server (vLLM and friends reserve
// almost all of HBM for weights and KV cache at start-up, so a later cudaMalloc can fail or
// fragment the little that is left), and latency-critical host loops that cannot afford the
// device-wide synchronisation cudaMalloc/cudaFree imply.  Such a caller brings its own input and
// output buffers; the only other memory the kernel needs is its per-launch scratch (tile ticket,
// arena cursor, counter rows, one word per tile), and here that scratch is a fixed
// array that is part of the library image instead of a cudaMalloc'ed block.
//
// The array below is a `static __device__` object: every GPU on which this library is loaded
// carries its own copy (SSEGPU_STATIC_SCRATCH_BYTES, about 70 KB) from module load until the
// process exits.  In host code the NAME of such an array is only a handle -- it is not a device
// address, and handing it to a kernel makes the kernel fault (or, where the GPU can reach host
// memory, scribble on the wrong memory).  The device address is obtained once per device, with
// that device current, through cudaGetSymbolAddress(), and that address is what the kernel gets.
//
// This file calls no cudaMalloc, cudaFree, cudaHostAlloc or any library routine that allocates.
#include <cuda_runtime.h>
#include <cstdio>
#include <mutex>

#include "../../include/ssegpu.h"
#include "sse_kernels.h"

namespace {

constexpr size_t kStaticTiles = (SSEGPU_STATIC_MAX_RECORDS + sse::TILE - 1) / sse::TILE;
constexpr size_t kStaticBytes = sse::SCRATCH_HDR + (size_t)sse::CTR_ROWS * sse::CTR_COLS * 8 + kStaticTiles * 8;
static_assert(kStaticBytes <= SSEGPU_STATIC_SCRATCH_BYTES, "static scratch array is too small for SSEGPU_STATIC_MAX_RECORDS");

// The scratch itself.  One copy per GPU that loads the library, kept until exit.
static __device__ __align__(256) unsigned char g_static_scratch[SSEGPU_STATIC_SCRATCH_BYTES];

__global__ void k_publish_bad_offsets(const uint32_t* bad_offsets, uint64_t* out_total) {
    if (*bad_offsets) *out_total |= SSEGPU_STATIC_BAD_OFFSETS;
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_addr[kMaxDevices];      // device address of g_static_scratch on device d (nullptr = not set up yet)
thread_local char g_err[256] = "";

// Set-up, once per device, with that device current: resolve the array's DEVICE address.
int static_setup(int device, void** out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_scratch_addr[device] == nullptr) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            snprintf(g_err, sizeof g_err, "cudaGetSymbolAddress(g_static_scratch) on device %d: %s", device, cudaGetErrorString(e));
            return SSEGPU_E_CUDA;
        }
        g_scratch_addr[device] = ptr;
    }
    *out = g_scratch_addr[device];
    return SSEGPU_OK;
}

}  // namespace

extern "C" {

const char* ssegpu_static_last_error(void) { return g_err; }

int ssegpu_frame_launch_static(int device, const void* d_in, const uint32_t* d_in_off, const int64_t* d_start_seq,
                               uint32_t n, uint32_t op, void* d_out, uint64_t out_cap, uint64_t* d_out_off,
                               uint32_t* d_out_len, uint32_t* d_subj_len, uint32_t* d_status, void* stream) {
    g_err[0] = 0;
    if (device < 0 || device >= kMaxDevices || !d_in_off || !d_out_off || (n && (!d_in || !d_out || !d_status || !d_out_len))) {
        snprintf(g_err, sizeof g_err, "invalid argument");
        return SSEGPU_E_INVAL;
    }
    if ((op & SSEGPU_OP_RESP) || ((op & SSEGPU_OP_NATS_PUB) && !(op & SSEGPU_OP_BRIDGE))) {
        snprintf(g_err, sizeof g_err, "SSEGPU_OP_NATS_PUB needs SSEGPU_OP_BRIDGE; SSEGPU_OP_RESP belongs to ssegpu_encode_launch");
        return SSEGPU_E_INVAL;
    }
    if (op & (SSEGPU_OP_INSPECT | SSEGPU_OP_INSPECT_INLINE | SSEGPU_OP_MATCH)) {
        // pattern tables and the interest table belong to a context; this entry has none, and running the plain kernel
        // instead would hand back unredacted, undropped frames with status OK
        snprintf(g_err, sizeof g_err, "SSEGPU_OP_INSPECT / SSEGPU_OP_INSPECT_INLINE / SSEGPU_OP_MATCH need a context's tables: "
                 "use ssegpu_frame_launch or the batch ring");
        return SSEGPU_E_INVAL;
    }
    if (n > SSEGPU_STATIC_MAX_RECORDS) {
        snprintf(g_err, sizeof g_err, "n_records = %u exceeds SSEGPU_STATIC_MAX_RECORDS = %u: the built-in scratch array holds "
                 "one scratch word per tile for at most that many records; split the batch or use ssegpu_frame_launch",
                 n, (unsigned)SSEGPU_STATIC_MAX_RECORDS);
        return SSEGPU_E_TOOBIG;
    }
    if ((reinterpret_cast<uintptr_t>(d_in) | reinterpret_cast<uintptr_t>(d_out)) & 15) {
        snprintf(g_err, sizeof g_err, "d_in and d_out must be 16-byte aligned (tiles move as 16-byte granular bulk copies)");
        return SSEGPU_E_INVAL;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e)); return SSEGPU_E_CUDA; }
    void* scratch = nullptr;
    int rc = static_setup(device, &scratch);         // first call on this device resolves the address
    if (rc != SSEGPU_OK) return rc;

    sse::FrameArgs a{};
    a.in = static_cast<const uint8_t*>(d_in); a.in_off = d_in_off; a.start_seq = d_start_seq;
    a.n = n; a.op = op;
    a.out = static_cast<uint8_t*>(d_out); a.out_cap = out_cap; a.out_off = d_out_off; a.out_len = d_out_len;
    a.subj_len = d_subj_len; a.status = d_status;
    sse::scratch_carve(scratch, a);                  // the DEVICE address, never the symbol's name
    // frame_launch clears scratch_bytes(n) of the array on `stream` (ticket, cursor, counters, per-tile words
    // must start at zero; the array still holds the previous call's values) and launches the kernel.
    e = sse::frame_launch(device, a, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "launch: %s", cudaGetErrorString(e)); return SSEGPU_E_CUDA; }
    // There is no result call on this entry, so the one data-level error the kernel can find -- offsets in in_off that
    // are not monotonic or run past in_off[n]; such records are clamped, never read out of bounds -- travels in bit 63
    // of out_off[n].
    k_publish_bad_offsets<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(a.bad_offsets, d_out_off + n);
    e = cudaGetLastError();
    if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "launch: %s", cudaGetErrorString(e)); return SSEGPU_E_CUDA; }
    return SSEGPU_OK;
}

}  // extern "C"
