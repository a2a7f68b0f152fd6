// ssegpu_api.cu -- C ABI of libssegpu.so (include/ssegpu.h): contexts, the pinned batch ring,
// the device-resident launch entry, counters.  Host side only; all arithmetic of the hot path
// is in sse_kernels.cu.  There is no CPU implementation behind this ABI: without a usable CUDA
// device ssegpu_init() fails.
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/ssegpu.h"
#include "sse_kernels.h"

namespace {

thread_local char g_cuda_err[256] = "";

int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_cuda_err, sizeof g_cuda_err, "%s: %s", what, cudaGetErrorString(e));
    return SSEGPU_E_CUDA;
}
#define CU(call)                                              \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
    } while (0)

enum SlotState { FREE = 0, ACQUIRED, SUBMITTED, DONE };

}  // namespace

struct ssegpu_batch {
    ssegpu_ctx* ctx = nullptr;
    int state = FREE;
    uint32_t n = 0, op = 0;
    uint64_t spec_out = 0;       // output bytes whose D2H copy was enqueued at submit
    uint64_t total_out = 0;
    // pinned host
    uint8_t* h_in = nullptr; uint32_t* h_in_off = nullptr; int64_t* h_start = nullptr; uint32_t* h_conv = nullptr;
    uint8_t* h_out = nullptr; uint64_t* h_out_off = nullptr; uint32_t* h_out_len = nullptr; uint32_t* h_status = nullptr; uint32_t* h_subj = nullptr;
    uint8_t* h_hdr = nullptr;    // scratch header + counter rows read back
    // device
    uint8_t* d_in = nullptr; uint32_t* d_in_off = nullptr; int64_t* d_start = nullptr;
    uint8_t* d_out = nullptr; uint64_t* d_out_off = nullptr; uint32_t* d_out_len = nullptr; uint32_t* d_status = nullptr; uint32_t* d_subj = nullptr;
    void* d_scratch = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_done = nullptr;
};

struct ssegpu_ctx {
    ssegpu_config cfg{};
    uint64_t out_cap = 0;
    cudaStream_t stream = nullptr;
    void* d_launch_scratch = nullptr;     // for ssegpu_frame_launch
    uint8_t* h_launch_hdr = nullptr;
    std::vector<ssegpu_batch*> slots;
    std::mutex mu;
    ssegpu_counters totals{};
    uint64_t launches = 0;
};

namespace {

void add_rows(const uint8_t* hdr, ssegpu_counters* c, uint32_t n_records) {
    const unsigned long long* rows = reinterpret_cast<const unsigned long long*>(hdr + sse::SCRATCH_HDR);
    unsigned long long s[sse::CTR_COLS] = {0};
    for (int r = 0; r < sse::CTR_ROWS; r++)
        for (int k = 0; k < sse::CTR_COLS; k++) s[k] += rows[r * sse::CTR_COLS + k];
    c->records += n_records;
    c->frames += s[0]; c->err_syntax += s[1]; c->err_type += s[2]; c->filtered += s[3];
    c->done += s[4]; c->fast_path += s[5]; c->bytes_in += s[6]; c->bytes_out += s[7];
    c->batches += 1;
}

void free_batch(ssegpu_batch* b) {
    if (!b) return;
    cudaFreeHost(b->h_in); cudaFreeHost(b->h_in_off); cudaFreeHost(b->h_start); cudaFreeHost(b->h_conv);
    cudaFreeHost(b->h_out); cudaFreeHost(b->h_out_off); cudaFreeHost(b->h_out_len); cudaFreeHost(b->h_status); cudaFreeHost(b->h_subj);
    cudaFreeHost(b->h_hdr);
    cudaFree(b->d_in); cudaFree(b->d_in_off); cudaFree(b->d_start);
    cudaFree(b->d_out); cudaFree(b->d_out_off); cudaFree(b->d_out_len); cudaFree(b->d_status); cudaFree(b->d_subj);
    cudaFree(b->d_scratch);
    if (b->ev_done) cudaEventDestroy(b->ev_done);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
}

int alloc_batch(ssegpu_ctx* ctx, ssegpu_batch** out) {
    ssegpu_batch* b = new (std::nothrow) ssegpu_batch();
    if (!b) return SSEGPU_E_NOMEM;
    *out = b;
    b->ctx = ctx;
    const size_t n = ctx->cfg.max_records;
    const size_t in_bytes = ctx->cfg.max_in_bytes + 64;
    CU(cudaHostAlloc(&b->h_in, in_bytes, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_in_off, (n + 1) * 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_start, n * 8 + 8, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_conv, n * 4 + 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_out, ctx->out_cap + 64, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_out_off, (n + 1) * 8, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_out_len, n * 4 + 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_status, n * 4 + 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_subj, n * 4 + 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_hdr, sse::scratch_readback_bytes(), cudaHostAllocDefault));
    memset(b->h_in, 0, in_bytes);
    memset(b->h_start, 0, n * 8 + 8);
    memset(b->h_conv, 0, n * 4 + 4);
    b->h_in_off[0] = 0;
    CU(cudaMalloc(&b->d_in, in_bytes));
    CU(cudaMalloc(&b->d_in_off, (n + 1) * 4));
    CU(cudaMalloc(&b->d_start, n * 8 + 8));
    CU(cudaMalloc(&b->d_out, ctx->out_cap + 64));
    CU(cudaMalloc(&b->d_out_off, (n + 1) * 8));
    CU(cudaMalloc(&b->d_out_len, n * 4 + 4));
    CU(cudaMalloc(&b->d_status, n * 4 + 4));
    CU(cudaMalloc(&b->d_subj, n * 4 + 4));
    CU(cudaMalloc(&b->d_scratch, sse::scratch_bytes((uint32_t)n)));
    CU(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&b->ev_done, cudaEventDisableTiming));
    return SSEGPU_OK;
}

}  // namespace

extern "C" {

uint32_t ssegpu_abi_version(void) { return SSEGPU_ABI_VERSION; }

const char* ssegpu_last_cuda_error(void) { return g_cuda_err; }

const char* ssegpu_strerror(int err) {
    switch (err) {
    case SSEGPU_OK: return "ok";
    case SSEGPU_E_INVAL: return "invalid argument";
    case SSEGPU_E_CUDA: return "CUDA error (see ssegpu_last_cuda_error); this library has no CPU fallback";
    case SSEGPU_E_NOMEM: return "out of memory";
    case SSEGPU_E_AGAIN: return "no free batch slot";
    case SSEGPU_E_TOOBIG: return "batch exceeds configured capacity";
    case SSEGPU_E_OVERFLOW: return "output arena too small for this batch";
    case SSEGPU_E_STATE: return "batch used out of order";
    case SSEGPU_E_TIMEOUT: return "timed out";
    default: return "unknown error";
    }
}

int ssegpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

uint32_t ssegpu_shard(const uint8_t* id, size_t len, uint32_t n_gpus) {
    uint64_t h = 0xcbf29ce484222325ull;          // hash/fnv New64a
    for (size_t i = 0; i < len; i++) { h ^= id[i]; h *= 0x100000001b3ull; }
    return n_gpus ? (uint32_t)(h % n_gpus) : 0;
}

// fmt.Sscanf(lastEventID, "%d", &startSequence); startSequence++   (sse_handler.go:92-96)
int64_t ssegpu_start_sequence(const char* id) {
    if (!id || !*id) return 0;
    const unsigned char* s = reinterpret_cast<const unsigned char*>(id);
    while (*s == ' ' || *s == '\t' || *s == '\v' || *s == '\f' || (*s == '\r' && s[1] != '\n')) s++;
    bool neg = false, ok = true;
    if (*s == '+' || *s == '-') { neg = (*s == '-'); s++; }
    uint64_t v = 0;
    int nd = 0;
    for (; (*s >= '0' && *s <= '9') || *s == '_'; s++, nd++) {
        if (*s == '_') { ok = false; continue; }            // strconv.ParseInt(base 10) rejects '_'
        uint64_t d = (uint64_t)(*s - '0');
        if (v > (UINT64_MAX - d) / 10) { ok = false; continue; }
        v = v * 10 + d;
    }
    if (nd == 0) ok = false;
    if (ok && !neg && v > (uint64_t)INT64_MAX) ok = false;
    if (ok && neg && v > (1ull << 63)) ok = false;
    uint64_t start = ok ? (neg ? 0 - v : v) : 0;
    return (int64_t)(start + 1);
}

int ssegpu_init(const ssegpu_config* cfg, ssegpu_ctx** out) {
    if (!cfg || !out || cfg->max_records == 0) return SSEGPU_E_INVAL;
    if (cfg->max_in_bytes > 0xFFFFFFF0ull) return SSEGPU_E_INVAL;      // in_off is uint32
    *out = nullptr;
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (cfg->device < 0 || cfg->device >= ndev || cfg->device >= 64) {
        snprintf(g_cuda_err, sizeof g_cuda_err, "device %d not present (%d CUDA devices)", cfg->device, ndev);
        return SSEGPU_E_CUDA;
    }
    CU(cudaSetDevice(cfg->device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major < 10) {
        snprintf(g_cuda_err, sizeof g_cuda_err, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device,
                 prop.major, prop.minor);
        return SSEGPU_E_CUDA;
    }
    ssegpu_ctx* ctx = new (std::nothrow) ssegpu_ctx();
    if (!ctx) return SSEGPU_E_NOMEM;
    ctx->cfg = *cfg;
    ctx->out_cap = cfg->max_out_bytes ? cfg->max_out_bytes : 2 * cfg->max_in_bytes + 64ull * cfg->max_records;
    int rc = SSEGPU_OK;
    do {
        uint32_t ic, oc, sm; int grid;
        cudaError_t e = sse::frame_configure(cfg->device, &ic, &oc, &sm, &grid);
        if (e != cudaSuccess) { rc = cuda_fail(e, "frame_configure"); break; }
        if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = cuda_fail(e, "cudaStreamCreate"); break; }
        if ((e = cudaMalloc(&ctx->d_launch_scratch, sse::scratch_bytes(cfg->max_records))) != cudaSuccess) { rc = cuda_fail(e, "cudaMalloc scratch"); break; }
        if ((e = cudaHostAlloc(&ctx->h_launch_hdr, sse::scratch_readback_bytes(), cudaHostAllocDefault)) != cudaSuccess) { rc = cuda_fail(e, "cudaHostAlloc"); break; }
        for (uint32_t i = 0; i < cfg->ring_slots; i++) {
            ssegpu_batch* b = nullptr;
            rc = alloc_batch(ctx, &b);
            if (b) ctx->slots.push_back(b);
            if (rc != SSEGPU_OK) break;
        }
    } while (0);
    if (rc != SSEGPU_OK) { ssegpu_destroy(ctx); return rc; }
    *out = ctx;
    return SSEGPU_OK;
}

void ssegpu_destroy(ssegpu_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->cfg.device);
    cudaDeviceSynchronize();
    for (ssegpu_batch* b : ctx->slots) free_batch(b);
    cudaFree(ctx->d_launch_scratch);
    cudaFreeHost(ctx->h_launch_hdr);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int ssegpu_batch_acquire(ssegpu_ctx* ctx, ssegpu_batch** out) {
    if (!ctx || !out) return SSEGPU_E_INVAL;
    std::lock_guard<std::mutex> lk(ctx->mu);
    for (ssegpu_batch* b : ctx->slots)
        if (b->state == FREE) { b->state = ACQUIRED; *out = b; return SSEGPU_OK; }
    *out = nullptr;
    return SSEGPU_E_AGAIN;
}

uint8_t* ssegpu_in_bytes(ssegpu_batch* b) { return b ? b->h_in : nullptr; }
uint32_t* ssegpu_in_off(ssegpu_batch* b) { return b ? b->h_in_off : nullptr; }
int64_t* ssegpu_start_seq(ssegpu_batch* b) { return b ? b->h_start : nullptr; }
uint32_t* ssegpu_conv_slot(ssegpu_batch* b) { return b ? b->h_conv : nullptr; }
const uint8_t* ssegpu_out_bytes(const ssegpu_batch* b) { return b ? b->h_out : nullptr; }
const uint64_t* ssegpu_out_off(const ssegpu_batch* b) { return b ? b->h_out_off : nullptr; }
const uint32_t* ssegpu_out_len(const ssegpu_batch* b) { return b ? b->h_out_len : nullptr; }
const uint32_t* ssegpu_status(const ssegpu_batch* b) { return b ? b->h_status : nullptr; }
const uint32_t* ssegpu_subj_len(const ssegpu_batch* b) { return b ? b->h_subj : nullptr; }

int ssegpu_batch_submit(ssegpu_ctx* ctx, ssegpu_batch* b, uint32_t n, uint32_t op, int use_filter) {
    if (!ctx || !b || b->ctx != ctx) return SSEGPU_E_INVAL;
    if (b->state != ACQUIRED && b->state != DONE) return SSEGPU_E_STATE;   // DONE: the owner re-submits the slot
    if (n > ctx->cfg.max_records) return SSEGPU_E_TOOBIG;
    const uint64_t in_total = n ? b->h_in_off[n] : 0;
    if (in_total > ctx->cfg.max_in_bytes) return SSEGPU_E_TOOBIG;
    CU(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = b->stream;
    b->n = n; b->op = op;
    // 32 readable bytes after the last payload byte (TMA tiles are 16-byte granular, word reads overrun a little)
    memset(b->h_in + in_total, 0, 48);
    CU(cudaMemcpyAsync(b->d_in, b->h_in, ((in_total + 15) & ~15ull) + 32, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(b->d_in_off, b->h_in_off, ((size_t)n + 1) * 4, cudaMemcpyHostToDevice, st));
    if (use_filter && n) CU(cudaMemcpyAsync(b->d_start, b->h_start, (size_t)n * 8, cudaMemcpyHostToDevice, st));

    sse::FrameArgs a{};
    a.in = b->d_in; a.in_off = b->d_in_off; a.start_seq = use_filter ? b->d_start : nullptr;
    a.n = n; a.op = op;
    a.out = b->d_out; a.out_cap = ctx->out_cap; a.out_off = b->d_out_off; a.out_len = b->d_out_len;
    a.subj_len = (op & SSEGPU_OP_BRIDGE) ? b->d_subj : nullptr;
    a.status = b->d_status;
    sse::scratch_carve(b->d_scratch, a);
    CU(sse::frame_launch(ctx->cfg.device, a, st));

    CU(cudaMemcpyAsync(b->h_out_off, b->d_out_off, ((size_t)n + 1) * 8, cudaMemcpyDeviceToHost, st));
    if (n) CU(cudaMemcpyAsync(b->h_out_len, b->d_out_len, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (n) CU(cudaMemcpyAsync(b->h_status, b->d_status, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if ((op & SSEGPU_OP_BRIDGE) && n) CU(cudaMemcpyAsync(b->h_subj, b->d_subj, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(b->h_hdr, b->d_scratch, sse::scratch_readback_bytes(), cudaMemcpyDeviceToHost, st));
    // Speculative copy of the frames: a canonical record grows by 26 + digits(seq) bytes
    // (28 + len("chat..tokens")=40ish in bridge mode); anything beyond is fetched in wait().
    uint64_t spec = in_total + (uint64_t)n * ((op & SSEGPU_OP_BRIDGE) ? 80 : 32) + 16ull * ((n + sse::TILE - 1) / sse::TILE);
    if (spec > ctx->out_cap) spec = ctx->out_cap;
    b->spec_out = spec;
    if (spec) CU(cudaMemcpyAsync(b->h_out, b->d_out, spec, cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(b->ev_done, st));
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        b->state = SUBMITTED;
        ctx->launches += n ? 1 : 0;
    }
    return SSEGPU_OK;
}

int ssegpu_batch_wait(ssegpu_ctx* ctx, ssegpu_batch* b, uint32_t timeout_us) {
    if (!ctx || !b || b->ctx != ctx) return SSEGPU_E_INVAL;
    if (b->state == DONE) return SSEGPU_OK;
    if (b->state != SUBMITTED) return SSEGPU_E_STATE;
    CU(cudaSetDevice(ctx->cfg.device));
    if (timeout_us == 0) {
        CU(cudaEventSynchronize(b->ev_done));
    } else {
        auto t0 = std::chrono::steady_clock::now();
        for (;;) {
            cudaError_t e = cudaEventQuery(b->ev_done);
            if (e == cudaSuccess) break;
            if (e != cudaErrorNotReady) return cuda_fail(e, "cudaEventQuery");
            if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(timeout_us)) return SSEGPU_E_TIMEOUT;
            std::this_thread::yield();
        }
    }
    const uint32_t overflow = reinterpret_cast<const uint32_t*>(b->h_hdr)[1];
    b->total_out = b->h_out_off[b->n];
    int rc = SSEGPU_OK;
    if (overflow) rc = SSEGPU_E_OVERFLOW;
    else if (b->total_out > b->spec_out) {
        CU(cudaMemcpyAsync(b->h_out + b->spec_out, b->d_out + b->spec_out, b->total_out - b->spec_out, cudaMemcpyDeviceToHost, b->stream));
        CU(cudaStreamSynchronize(b->stream));
    }
    std::lock_guard<std::mutex> lk(ctx->mu);
    add_rows(b->h_hdr, &ctx->totals, b->n);
    b->state = DONE;
    return rc;
}

int ssegpu_batch_release(ssegpu_ctx* ctx, ssegpu_batch* b) {
    if (!ctx || !b || b->ctx != ctx) return SSEGPU_E_INVAL;
    if (b->state == SUBMITTED) { int rc = ssegpu_batch_wait(ctx, b, 0); if (rc != SSEGPU_OK && rc != SSEGPU_E_OVERFLOW) return rc; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    b->state = FREE;
    return SSEGPU_OK;
}

int ssegpu_counters_get(ssegpu_ctx* ctx, ssegpu_counters* out) {
    if (!ctx || !out) return SSEGPU_E_INVAL;
    std::lock_guard<std::mutex> lk(ctx->mu);
    *out = ctx->totals;
    return SSEGPU_OK;
}

int ssegpu_frame_launch(ssegpu_ctx* ctx, const void* d_in, const uint32_t* d_in_off, const int64_t* d_start_seq,
                        uint32_t n, uint32_t op, void* d_out, uint64_t out_cap, uint64_t* d_out_off,
                        uint32_t* d_out_len, uint32_t* d_subj_len, uint32_t* d_status, void* stream) {
    if (!ctx || !d_in_off || !d_out_off || (n && (!d_in || !d_out || !d_status || !d_out_len))) return SSEGPU_E_INVAL;
    if (n > ctx->cfg.max_records) return SSEGPU_E_TOOBIG;
    if (reinterpret_cast<uintptr_t>(d_in) & 15) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    sse::FrameArgs a{};
    a.in = static_cast<const uint8_t*>(d_in); a.in_off = d_in_off; a.start_seq = d_start_seq;
    a.n = n; a.op = op;
    a.out = static_cast<uint8_t*>(d_out); a.out_cap = out_cap; a.out_off = d_out_off; a.out_len = d_out_len;
    a.subj_len = d_subj_len; a.status = d_status;
    sse::scratch_carve(ctx->d_launch_scratch, a);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    CU(sse::frame_launch(ctx->cfg.device, a, st));
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->launches += n ? 1 : 0;
    return SSEGPU_OK;
}

int ssegpu_frame_launch_result(ssegpu_ctx* ctx, ssegpu_counters* out) {
    if (!ctx) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaMemcpy(ctx->h_launch_hdr, ctx->d_launch_scratch, sse::scratch_readback_bytes(), cudaMemcpyDeviceToHost));
    if (out) add_rows(ctx->h_launch_hdr, out, 0);
    return reinterpret_cast<const uint32_t*>(ctx->h_launch_hdr)[1] ? SSEGPU_E_OVERFLOW : SSEGPU_OK;
}

uint64_t ssegpu_launch_count(ssegpu_ctx* ctx) {
    if (!ctx) return 0;
    std::lock_guard<std::mutex> lk(ctx->mu);
    return ctx->launches;
}

}  // extern "C"
