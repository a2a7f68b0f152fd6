// ssegpu_api.cu -- C ABI of libssegpu.so (include/ssegpu.h): contexts, the pinned batch ring,
// the device-resident launch entry, counters.  Host side only; all arithmetic of the hot path
// is in sse_kernels.cu.  There is no CPU implementation behind this ABI: without a usable CUDA
// device ssegpu_init() fails.
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/ssegpu.h"
#include "inspect_compile.hpp"
#include "interest_table.hpp"
#include "sse_inspect.cuh"
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

// Device tables live in generations that are reference-counted by whoever may still read them: the context (the current
// one) and every submitted batch until it is released.  Swapping a generation in needs no device-wide synchronisation:
// the old one is freed, stream-ordered, when its last user lets go (cudaFreeAsync on the context's table stream).
struct TableGen {                        // interest table
    cudaStream_t stream = nullptr;
    sse::InterestSlot* d_slots = nullptr; uint32_t* d_keys = nullptr; uint8_t* d_wild = nullptr; sse::InterestDev* d_desc = nullptr;
    uint32_t* d_subs = nullptr;                      // device copy of the id arena (ssegpu_fanout_launch expands spans from it)
    uint32_t* d_wild_ids = nullptr;                  // 32 words
    sse::InterestUniform uni{};                      // host copy of the descriptor's scalars, passed to the frame kernel by value
    std::shared_ptr<std::vector<uint32_t>> subs;     // host arena the spans of this generation index (append-only, fixed address)
    std::vector<uint32_t> wild_ids;
    ~TableGen() {
        if (d_subs) cudaFreeAsync(d_subs, stream);
        if (d_wild_ids) cudaFreeAsync(d_wild_ids, stream);
        if (d_slots) cudaFreeAsync(d_slots, stream);
        if (d_keys) cudaFreeAsync(d_keys, stream);
        if (d_wild) cudaFreeAsync(d_wild, stream);
        if (d_desc) cudaFreeAsync(d_desc, stream);
    }
};
struct InspGen {                         // inspection pattern tables
    cudaStream_t stream = nullptr;
    void* d_blob = nullptr; sse::InspectDev* d_desc = nullptr;
    uint32_t mode = 0, start_state = 0, n_states = 0, n_classes = 0;
    sse::DfaUniform dfa{};               // host copy of the forward DFA's descriptor scalars, passed to the frame kernel by value
    ~InspGen() {
        if (d_blob) cudaFreeAsync(d_blob, stream);
        if (d_desc) cudaFreeAsync(d_desc, stream);
    }
};
struct SlotOp { uint32_t index, pad[15]; sse::InterestSlot image; };     // one changed slot of an incremental table update

struct ssegpu_batch {
    ssegpu_ctx* ctx = nullptr;
    std::shared_ptr<TableGen> tab;       // generations this batch was launched on (held until release)
    std::shared_ptr<InspGen> insp;
    uint64_t table_version = 0;
    int state = FREE;
    uint32_t n = 0, op = 0;
    uint64_t spec_out = 0;       // output bytes whose D2H copy was enqueued at submit
    uint64_t total_out = 0;
    // pinned host
    uint8_t* h_in = nullptr; uint32_t* h_in_off = nullptr; int64_t* h_start = nullptr; uint32_t* h_conv = nullptr;
    uint8_t* h_out = nullptr; uint64_t* h_out_off = nullptr; uint32_t* h_out_len = nullptr; uint32_t* h_status = nullptr; uint32_t* h_subj = nullptr;
    uint32_t* h_verdict = nullptr; int32_t* h_match = nullptr; uint32_t* h_hit = nullptr; uint32_t* d_hit = nullptr;
    uint8_t* h_hdr = nullptr;    // scratch header + counter rows read back
    // Everything the kernel reports per batch lives in ONE device block and ONE pinned block of the same layout,
    // [per-tile words | header + counters | out_off | out_len | status | subj_len | verdict | match | sub_span], laid out per
    // submit for that batch's record count: one clear, one D2H copy.  The h_*/d_* result pointers above point into them.
    uint8_t* d_block = nullptr; uint8_t* h_block = nullptr;
    size_t ts_cap = 0;           // bytes reserved for the per-tile words in front of the header
    size_t meta_bytes = 0;       // header .. end of the last array of the current layout
    size_t plan_off = 0;         // ordered mode's plan words inside d_block
    // device
    uint8_t* d_in = nullptr; uint32_t* d_in_off = nullptr; int64_t* d_start = nullptr;
    uint8_t* d_out = nullptr; uint64_t* d_out_off = nullptr; uint32_t* d_out_len = nullptr; uint32_t* d_status = nullptr; uint32_t* d_subj = nullptr;
    uint32_t* d_conv = nullptr; uint32_t* d_verdict = nullptr; int32_t* d_match = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_done = nullptr;
};

struct ssegpu_ctx {
    ssegpu_config cfg{};
    uint64_t out_cap = 0;
    cudaStream_t stream = nullptr;
    void* d_launch_scratch = nullptr;     // for ssegpu_frame_launch
    uint8_t* h_launch_hdr = nullptr;
    void* d_encode_scratch = nullptr;     // for ssegpu_encode_launch (header only)
    std::vector<ssegpu_batch*> slots;
    std::mutex mu;
    ssegpu_counters totals{};
    uint64_t launches = 0;
    // content inspection
    std::shared_ptr<InspGen> insp;        // current pattern tables (nullptr: none loaded)
    uint32_t* d_conv_state = nullptr;
    // streaming mode: token entries + sort scratch of the per-conversation walk, shared by every batch of the context
    // (streaming batches run one after the other on the GPU: ev_stream orders them in submit order)
    uint4* d_tok_entry = nullptr; void* d_stream_scratch = nullptr;
    cudaEvent_t ev_stream = nullptr;
    // interest table
    sse_interest::HostTable interest;     // authoritative subscriptions + host mirror of the current generation
    std::shared_ptr<TableGen> tab;        // current device generation (nullptr: ssegpu_interest_apply not called yet)
    uint64_t table_version = 0;           // successful ssegpu_interest_apply calls so far
    // table updates are GPU work on their own stream, ordered against the batches by events: an update waits for the
    // batches in flight that read the tables, later batches wait for the update (ev_table)
    void* d_wire_scratch = nullptr; size_t wire_scratch_cap = 0;     // ssegpu_split_launch
    void* d_chunk_scratch = nullptr; size_t chunk_scratch_cap = 0;   // ssegpu_chunk_decode_launch
    void* d_fanout_scratch = nullptr;                                // ssegpu_fanout_launch
    cudaStream_t table_stream = nullptr;
    cudaEvent_t ev_table = nullptr;
    static constexpr int kOpChunks = 64, kOpsPerChunk = 256, kLaunchEvents = 8;
    SlotOp* h_ops = nullptr; SlotOp* d_ops = nullptr;            // pinned staging ring + device copy, kOpChunks x kOpsPerChunk
    cudaEvent_t ev_chunk[kOpChunks] = {};
    int next_chunk = 0;
    cudaEvent_t ev_launch[kLaunchEvents] = {};                   // device-resident launches that read the tables
    int next_launch_ev = 0;
    uint32_t* bound_hit = nullptr;
    const uint32_t* bound_conv = nullptr; uint32_t* bound_verdict = nullptr; int32_t* bound_match = nullptr;   // ssegpu_inspect_bind
};

namespace {

// kernels per frame launch: ordered = count pass, tile scan (two launches), emit pass; arrival = the tile kernel
uint32_t launches_of(uint32_t op) { return (op & SSEGPU_OP_ORDERED) ? 4 : 1; }

// Frame kernels plus, in streaming inspection mode, the per-conversation walk.  Streaming batches of one context run in
// submit order (each waits for the previous one's walk): the carried DFA state of a conversation that has tokens in
// two consecutive batches must be committed by the first before the second reads it.
// The caller holds ctx->mu whenever the op reads tables (inspection, interest), from the snapshot of the generations to
// the last enqueue: lock order is then GPU order, for batches among themselves and against table updates.
int launch_path(ssegpu_ctx* ctx, sse::FrameArgs a, const uint32_t* d_conv, cudaStream_t st, const InspGen* insp, uint32_t* n_launches) {
    const bool inspect = (a.op & (SSEGPU_OP_INSPECT | SSEGPU_OP_INSPECT_INLINE)) != 0;
    *n_launches = a.n ? launches_of(a.op) : 0;
    if (!inspect || insp->mode != sse::INSPECT_STREAMING) {
        CU(sse::frame_launch(ctx->cfg.device, a, st));
        return SSEGPU_OK;
    }
    CU(cudaStreamWaitEvent(st, ctx->ev_stream, 0));
    sse::StreamArgs w{};
    w.insp = a.insp; w.in = a.in; w.tok_entry = ctx->d_tok_entry; w.conv_slot = d_conv; w.n = a.n; w.conv_slots = ctx->cfg.conv_slots;
    w.conv_state = ctx->d_conv_state; w.verdict = a.verdict; w.match = a.match; w.scratch = ctx->d_stream_scratch;
    w.dfa_states = insp->n_states; w.dfa_classes = insp->n_classes;
    sse::FrameArgs f = a;
    f.insp = nullptr; f.tok_entry = ctx->d_tok_entry;
    uint32_t bits = 1;
    while (bits < 32 && (1ull << bits) < ctx->cfg.conv_slots) bits++;
    const uint32_t walk_launches = a.n ? 5 * ((bits + 10) / 11) + 2 : 0;
    if (a.op & SSEGPU_OP_INSPECT_INLINE) {
        // verdicts change the frames: planning pass (token entries only), walk, then the frame pass applies the verdicts
        sse::FrameArgs plan = f;
        plan.interest = nullptr;
        CU(sse::frame_plan_launch(ctx->cfg.device, plan, st));
        w.bad_conv = plan.bad_conv;
        // (the frame pass below clears its scratch but leaves that flag word alone: keep_bad_conv)
        CU(sse::stream_inspect_launch(ctx->cfg.device, w, st));
        f.tok_entry = nullptr; f.xflags = sse::XF_PRE_VERDICT;
        f.keep_bad_conv = 1;
        CU(sse::frame_launch(ctx->cfg.device, f, st));
        *n_launches += (a.n ? 1 : 0) + walk_launches;
    } else {
        CU(sse::frame_launch(ctx->cfg.device, f, st));
        w.bad_conv = f.bad_conv;
        CU(sse::stream_inspect_launch(ctx->cfg.device, w, st));
        *n_launches += walk_launches;
    }
    CU(cudaEventRecord(ctx->ev_stream, st));
    return SSEGPU_OK;
}

// SSEGPU_OP_NATS_PUB is a bridge-mode framing; SSEGPU_OP_RESP belongs to the encoder
bool bad_frame_op(uint32_t op) { return (op & SSEGPU_OP_RESP) || ((op & SSEGPU_OP_NATS_PUB) && !(op & SSEGPU_OP_BRIDGE)); }

// With ctx->mu held: the generations `op` needs (SSEGPU_E_STATE if none has been loaded), and `st` made to wait for
// the last table update.
int acquire_tables(ssegpu_ctx* ctx, uint32_t op, cudaStream_t st, std::shared_ptr<TableGen>& tab, std::shared_ptr<InspGen>& insp) {
    if (op & (SSEGPU_OP_INSPECT | SSEGPU_OP_INSPECT_INLINE)) {
        if (!ctx->insp) return SSEGPU_E_STATE;                        // ssegpu_patterns_load first
        if (ctx->insp->mode == sse::INSPECT_STREAMING && (!ctx->d_conv_state || !ctx->d_tok_entry)) return SSEGPU_E_STATE;   // conv_slots was 0 at init
        insp = ctx->insp;
    }
    if (op & SSEGPU_OP_MATCH) {
        if (!ctx->tab) return SSEGPU_E_STATE;                         // ssegpu_interest_apply first
        tab = ctx->tab;
    }
    CU(cudaStreamWaitEvent(st, ctx->ev_table, 0));
    return SSEGPU_OK;
}

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
    cudaFreeHost(b->h_out); cudaFreeHost(b->h_block);
    cudaFree(b->d_conv);
    cudaFree(b->d_in); cudaFree(b->d_in_off); cudaFree(b->d_start);
    cudaFree(b->d_out); cudaFree(b->d_block);
    if (b->ev_done) cudaEventDestroy(b->ev_done);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
}

constexpr uint32_t kPackInputsMax = 16384;      // batches up to this many records send their inputs in one H2D copy

// Result arrays of a batch of n records inside the slot's blocks (16-byte aligned pieces; arrays an op does not produce take no room)
void layout_results(ssegpu_batch* b, uint32_t n, uint32_t op) {
    uint8_t* dh = b->d_block + b->ts_cap;        // device: header + counters sit right behind the per-tile words
    uint8_t* hh = b->h_block;
    size_t off = sse::scratch_readback_bytes();
    auto take = [&](size_t bytes) { const size_t o = off; off = (off + bytes + 15) & ~(size_t)15; return o; };
    const size_t o_off = take(((size_t)n + 1) * 8), o_len = take((size_t)n * 4 + 4), o_st = take((size_t)n * 4 + 4);
    const size_t o_subj = (op & SSEGPU_OP_BRIDGE) ? take((size_t)n * 4 + 4) : o_st;
    const bool insp = (op & (SSEGPU_OP_INSPECT | SSEGPU_OP_INSPECT_INLINE)) != 0;
    const size_t o_ver = insp ? take((size_t)n * 4 + 4) : o_st, o_mat = insp ? take((size_t)n * 8 + 8) : o_st;
    const size_t o_hit = (op & SSEGPU_OP_MATCH) ? take((size_t)n * 12 + 12) : o_st;
    b->meta_bytes = off;
    b->h_hdr = hh;
    b->d_out_off = reinterpret_cast<uint64_t*>(dh + o_off); b->h_out_off = reinterpret_cast<uint64_t*>(hh + o_off);
    b->d_out_len = reinterpret_cast<uint32_t*>(dh + o_len); b->h_out_len = reinterpret_cast<uint32_t*>(hh + o_len);
    b->d_status = reinterpret_cast<uint32_t*>(dh + o_st); b->h_status = reinterpret_cast<uint32_t*>(hh + o_st);
    b->d_subj = reinterpret_cast<uint32_t*>(dh + o_subj); b->h_subj = reinterpret_cast<uint32_t*>(hh + o_subj);
    b->d_verdict = reinterpret_cast<uint32_t*>(dh + o_ver); b->h_verdict = reinterpret_cast<uint32_t*>(hh + o_ver);
    b->d_match = reinterpret_cast<int32_t*>(dh + o_mat); b->h_match = reinterpret_cast<int32_t*>(hh + o_mat);
    b->d_hit = reinterpret_cast<uint32_t*>(dh + o_hit); b->h_hit = reinterpret_cast<uint32_t*>(hh + o_hit);
}

int alloc_batch(ssegpu_ctx* ctx, ssegpu_batch** out) {
    ssegpu_batch* b = new (std::nothrow) ssegpu_batch();
    if (!b) return SSEGPU_E_NOMEM;
    *out = b;
    b->ctx = ctx;
    const size_t n = ctx->cfg.max_records;
    const size_t in_bytes = ctx->cfg.max_in_bytes + 64 + kPackInputsMax * 12 + 128;   // room for offsets + start sequences of a small batch
    CU(cudaHostAlloc(&b->h_in, in_bytes, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_in_off, (n + 1) * 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_start, n * 8 + 8, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_conv, n * 4 + 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&b->h_out, ctx->out_cap + 64, cudaHostAllocDefault));
    CU(cudaMalloc(&b->d_conv, n * 4 + 4));
    b->ts_cap = (((n + sse::TILE - 1) / sse::TILE) * 8 + 15) & ~(size_t)15;
    const size_t meta_cap = sse::scratch_readback_bytes() + (n + 1) * 8 + n * (4 + 4 + 4 + 4 + 8 + 12) + 8 * 16;
    CU(cudaMalloc(&b->d_block, b->ts_cap + meta_cap + n * 4 + 32));      // + the ordered-mode plan words (device only)
    b->plan_off = (b->ts_cap + meta_cap + 15) & ~(size_t)15;     // 16-byte aligned: the emit pass fetches a tile's plan words by bulk copy
    CU(cudaHostAlloc(&b->h_block, meta_cap, cudaHostAllocDefault));
    memset(b->h_block, 0, meta_cap);
    layout_results(b, (uint32_t)n, ~0u);
    memset(b->h_in, 0, in_bytes);
    memset(b->h_start, 0, n * 8 + 8);
    memset(b->h_conv, 0, n * 4 + 4);
    b->h_in_off[0] = 0;
    CU(cudaMalloc(&b->d_in, in_bytes));
    CU(cudaMalloc(&b->d_in_off, (n + 1) * 4));
    CU(cudaMalloc(&b->d_start, n * 8 + 8));
    CU(cudaMalloc(&b->d_out, ctx->out_cap + 64));
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
    case SSEGPU_E_CONVSLOT: return "streaming inspection: a record's conversation slot is not below conv_slots";
    case SSEGPU_E_PATTERN: return "unsupported or malformed inspection pattern (see ssegpu_last_cuda_error for the reason)";
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
    if (prop.major != 10 || prop.minor != 0) {           // sm_100a code is not forward compatible (it does not load on sm_103 / sm_120)
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
        cudaError_t e = sse::frame_configure(cfg->device);
        if (e != cudaSuccess) { rc = cuda_fail(e, "frame_configure"); break; }
        if ((e = sse::stream_configure(cfg->device)) != cudaSuccess) { rc = cuda_fail(e, "stream_configure"); break; }
        if ((e = cudaEventCreateWithFlags(&ctx->ev_stream, cudaEventDisableTiming)) != cudaSuccess) { rc = cuda_fail(e, "cudaEventCreate"); break; }
        if ((e = cudaStreamCreateWithFlags(&ctx->table_stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = cuda_fail(e, "cudaStreamCreate"); break; }
        if ((e = cudaEventCreateWithFlags(&ctx->ev_table, cudaEventDisableTiming)) != cudaSuccess) { rc = cuda_fail(e, "cudaEventCreate"); break; }
        for (int k = 0; k < ssegpu_ctx::kOpChunks && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&ctx->ev_chunk[k], cudaEventDisableTiming);
        for (int k = 0; k < ssegpu_ctx::kLaunchEvents && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&ctx->ev_launch[k], cudaEventDisableTiming);
        if (e != cudaSuccess) { rc = cuda_fail(e, "cudaEventCreate"); break; }
        const size_t ops_bytes = sizeof(SlotOp) * ssegpu_ctx::kOpChunks * ssegpu_ctx::kOpsPerChunk;
        if ((e = cudaHostAlloc(&ctx->h_ops, ops_bytes, cudaHostAllocDefault)) != cudaSuccess) { rc = cuda_fail(e, "cudaHostAlloc"); break; }
        if ((e = cudaMalloc(&ctx->d_ops, ops_bytes)) != cudaSuccess) { rc = cuda_fail(e, "cudaMalloc"); break; }
        if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = cuda_fail(e, "cudaStreamCreate"); break; }
        if ((e = cudaMalloc(&ctx->d_launch_scratch, sse::scratch_bytes_with_plan(cfg->max_records))) != cudaSuccess) { rc = cuda_fail(e, "cudaMalloc scratch"); break; }
        if ((e = cudaMalloc(&ctx->d_encode_scratch, sse::SCRATCH_HDR)) != cudaSuccess) { rc = cuda_fail(e, "cudaMalloc scratch"); break; }
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
    for (ssegpu_batch* b : ctx->slots) if (b->stream) cudaStreamSynchronize(b->stream);   // own streams only: the GPU may be shared
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->table_stream) cudaStreamSynchronize(ctx->table_stream);
    for (ssegpu_batch* b : ctx->slots) { b->tab.reset(); b->insp.reset(); }
    ctx->tab.reset(); ctx->insp.reset();                                  // stream-ordered frees of the table generations
    if (ctx->table_stream) cudaStreamSynchronize(ctx->table_stream);
    for (ssegpu_batch* b : ctx->slots) free_batch(b);
    ctx->slots.clear();
    cudaFree(ctx->d_launch_scratch);
    cudaFree(ctx->d_encode_scratch);
    cudaFree(ctx->d_conv_state); cudaFree(ctx->d_tok_entry); cudaFree(ctx->d_stream_scratch);
    cudaFree(ctx->d_ops); cudaFreeHost(ctx->h_ops); cudaFree(ctx->d_wire_scratch); cudaFree(ctx->d_chunk_scratch); cudaFree(ctx->d_fanout_scratch);
    if (ctx->ev_stream) cudaEventDestroy(ctx->ev_stream);
    if (ctx->ev_table) cudaEventDestroy(ctx->ev_table);
    for (cudaEvent_t e : ctx->ev_chunk) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->ev_launch) if (e) cudaEventDestroy(e);
    if (ctx->table_stream) cudaStreamDestroy(ctx->table_stream);
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
const uint32_t* ssegpu_verdict(const ssegpu_batch* b) { return b ? b->h_verdict : nullptr; }
const int32_t* ssegpu_match(const ssegpu_batch* b) { return b ? b->h_match : nullptr; }
const uint32_t* ssegpu_sub_span(const ssegpu_batch* b) { return b ? b->h_hit : nullptr; }

int ssegpu_batch_submit(ssegpu_ctx* ctx, ssegpu_batch* b, uint32_t n, uint32_t op, int use_filter) {
    if (!ctx || !b || b->ctx != ctx) return SSEGPU_E_INVAL;
    if (b->state != ACQUIRED && b->state != DONE) return SSEGPU_E_STATE;   // DONE: the owner re-submits the slot
    if (bad_frame_op(op)) return SSEGPU_E_INVAL;
    if (n > ctx->cfg.max_records) return SSEGPU_E_TOOBIG;
    const uint64_t in_total = n ? b->h_in_off[n] : 0;
    if (in_total > ctx->cfg.max_in_bytes) return SSEGPU_E_TOOBIG;
    CU(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = b->stream;
    b->n = n; b->op = op;
    layout_results(b, n, op);
    // 32 readable bytes after the last payload byte (TMA tiles are 16-byte granular, word reads overrun a little)
    memset(b->h_in + in_total, 0, 48);
    const size_t pay = ((in_total + 15) & ~15ull) + 32;
    const uint32_t* d_in_off = b->d_in_off;
    const int64_t* d_start = b->d_start;
    if (n <= kPackInputsMax) {
        // small batch (the latency case): offsets and start sequences ride behind the payload, one H2D copy instead of three
        const size_t o_io = (pay + 15) & ~(size_t)15, o_ss = o_io + ((((size_t)n + 1) * 4 + 15) & ~(size_t)15);
        memcpy(b->h_in + o_io, b->h_in_off, ((size_t)n + 1) * 4);
        if (use_filter && n) memcpy(b->h_in + o_ss, b->h_start, (size_t)n * 8);
        CU(cudaMemcpyAsync(b->d_in, b->h_in, o_ss + (use_filter ? (size_t)n * 8 : 0), cudaMemcpyHostToDevice, st));
        d_in_off = reinterpret_cast<const uint32_t*>(b->d_in + o_io);
        d_start = reinterpret_cast<const int64_t*>(b->d_in + o_ss);
    } else {
        CU(cudaMemcpyAsync(b->d_in, b->h_in, pay, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(b->d_in_off, b->h_in_off, ((size_t)n + 1) * 4, cudaMemcpyHostToDevice, st));
        if (use_filter && n) CU(cudaMemcpyAsync(b->d_start, b->h_start, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    }

    sse::FrameArgs a{};
    a.in = b->d_in; a.in_off = d_in_off; a.start_seq = use_filter ? d_start : nullptr;
    a.n = n; a.op = op;
    a.out = b->d_out; a.out_cap = ctx->out_cap; a.out_off = b->d_out_off; a.out_len = b->d_out_len;
    a.subj_len = (op & SSEGPU_OP_BRIDGE) ? b->d_subj : nullptr;
    a.status = b->d_status;
    sse::scratch_carve_split(b->d_block + b->ts_cap, n, a);
    a.plan = reinterpret_cast<uint32_t*>(b->d_block + b->plan_off);
    // ops that read tables hold the context lock from the snapshot of the generations to the last enqueue (see launch_path)
    const bool uses_tables = (op & (SSEGPU_OP_INSPECT | SSEGPU_OP_INSPECT_INLINE | SSEGPU_OP_MATCH)) != 0;
    std::unique_lock<std::mutex> lk(ctx->mu, std::defer_lock);
    b->tab.reset(); b->insp.reset();
    if (uses_tables) {
        lk.lock();
        const int rc = acquire_tables(ctx, op, st, b->tab, b->insp);
        if (rc != SSEGPU_OK) return rc;
        b->table_version = ctx->table_version;
        if (b->insp) {
            if (b->insp->mode == sse::INSPECT_STREAMING && n) CU(cudaMemcpyAsync(b->d_conv, b->h_conv, (size_t)n * 4, cudaMemcpyHostToDevice, st));
            a.insp = b->insp->d_desc; a.dfa = b->insp->dfa; a.verdict = b->d_verdict; a.match = b->d_match;
        }
        if (b->tab) { a.interest = b->tab->d_desc; a.itab = b->tab->uni; a.hit = b->d_hit; }
    }
    uint32_t n_launches = 0;
    { const int rc = launch_path(ctx, a, b->d_conv, st, b->insp.get(), &n_launches); if (rc != SSEGPU_OK) return rc; }
    // header, counters and every per-record result array of this batch: one copy
    CU(cudaMemcpyAsync(b->h_block, b->d_block + b->ts_cap, b->meta_bytes, cudaMemcpyDeviceToHost, st));
    // Speculative copy of the frames: a canonical record grows by 26 + digits(seq) bytes
    // (28 + len("chat..tokens")=40ish in bridge mode); anything beyond is fetched in wait().
    uint64_t spec = ((op & SSEGPU_OP_HEADER_ONLY) ? 0 : in_total) + (uint64_t)n * ((op & SSEGPU_OP_BRIDGE) ? ((op & SSEGPU_OP_NATS_PUB) ? 96 : 80) : 32) +
                    16ull * ((n + sse::TILE - 1) / sse::TILE);
    if (spec > ctx->out_cap) spec = ctx->out_cap;
    b->spec_out = spec;
    if (spec) CU(cudaMemcpyAsync(b->h_out, b->d_out, spec, cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(b->ev_done, st));
    if (!uses_tables) lk.lock();
    b->state = SUBMITTED;
    ctx->launches += n_launches;
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
    if (reinterpret_cast<const uint32_t*>(b->h_hdr)[6]) rc = SSEGPU_E_INVAL;           // in_off was not monotonic
    else if (reinterpret_cast<const uint32_t*>(b->h_hdr)[5]) rc = SSEGPU_E_CONVSLOT;
    else if (overflow) rc = SSEGPU_E_OVERFLOW;
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
    if (b->state == SUBMITTED) { int rc = ssegpu_batch_wait(ctx, b, 0); if (b->state == SUBMITTED) return rc; }   // still in flight: a CUDA error; data-level errors free the slot
    std::lock_guard<std::mutex> lk(ctx->mu);
    b->state = FREE;
    b->tab.reset(); b->insp.reset();
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
    if (bad_frame_op(op)) return SSEGPU_E_INVAL;
    if (n > ctx->cfg.max_records) return SSEGPU_E_TOOBIG;
    if ((reinterpret_cast<uintptr_t>(d_in) | reinterpret_cast<uintptr_t>(d_out)) & 15) return SSEGPU_E_INVAL;   // bulk copies are 16-byte granular
    CU(cudaSetDevice(ctx->cfg.device));
    sse::FrameArgs a{};
    a.in = static_cast<const uint8_t*>(d_in); a.in_off = d_in_off; a.start_seq = d_start_seq;
    a.n = n; a.op = op;
    a.out = static_cast<uint8_t*>(d_out); a.out_cap = out_cap; a.out_off = d_out_off; a.out_len = d_out_len;
    a.subj_len = d_subj_len; a.status = d_status;
    sse::scratch_carve(ctx->d_launch_scratch, a);
    a.plan = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(ctx->d_launch_scratch) + ((sse::scratch_bytes(ctx->cfg.max_records) + 15) & ~(size_t)15));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    const bool uses_tables = (op & (SSEGPU_OP_INSPECT | SSEGPU_OP_INSPECT_INLINE | SSEGPU_OP_MATCH)) != 0;
    std::unique_lock<std::mutex> lk(ctx->mu, std::defer_lock);
    std::shared_ptr<TableGen> tab;
    std::shared_ptr<InspGen> insp;
    if (uses_tables) {
        lk.lock();
        const int rc = acquire_tables(ctx, op, st, tab, insp);
        if (rc != SSEGPU_OK) return rc;
        if (insp) {
            if (!ctx->bound_verdict || !ctx->bound_match || (insp->mode == sse::INSPECT_STREAMING && !ctx->bound_conv)) return SSEGPU_E_STATE;
            a.insp = insp->d_desc; a.dfa = insp->dfa; a.verdict = ctx->bound_verdict; a.match = ctx->bound_match;
        }
        if (tab) {
            if (!ctx->bound_hit) return SSEGPU_E_STATE;
            a.interest = tab->d_desc; a.itab = tab->uni; a.hit = ctx->bound_hit;
        }
    }
    uint32_t n_launches = 0;
    { const int rc = launch_path(ctx, a, ctx->bound_conv, st, insp.get(), &n_launches); if (rc != SSEGPU_OK) return rc; }
    if (uses_tables) {
        // a table update issued after this call waits for this launch; the generations themselves stay alive at least until
        // the next update has been ordered behind that event (they are only ever freed on the table stream)
        cudaEvent_t ev = ctx->ev_launch[ctx->next_launch_ev++ % ssegpu_ctx::kLaunchEvents];
        CU(cudaEventRecord(ev, st));
    } else lk.lock();
    ctx->launches += n_launches;
    return SSEGPU_OK;
}

int ssegpu_frame_launch_result(ssegpu_ctx* ctx, ssegpu_counters* out) {
    if (!ctx) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaMemcpy(ctx->h_launch_hdr, ctx->d_launch_scratch, sse::scratch_readback_bytes(), cudaMemcpyDeviceToHost));
    if (out) add_rows(ctx->h_launch_hdr, out, 0);
    if (reinterpret_cast<const uint32_t*>(ctx->h_launch_hdr)[6]) return SSEGPU_E_INVAL;
    if (reinterpret_cast<const uint32_t*>(ctx->h_launch_hdr)[5]) return SSEGPU_E_CONVSLOT;
    return reinterpret_cast<const uint32_t*>(ctx->h_launch_hdr)[1] ? SSEGPU_E_OVERFLOW : SSEGPU_OK;
}

// ---- producer-side encode --------------------------------------------------------------------
int ssegpu_encode_launch(ssegpu_ctx* ctx, const void* d_cid, const uint32_t* d_cid_off, const void* d_tok, const uint32_t* d_tok_off,
                         const int64_t* d_seq, const uint8_t* d_done, const int64_t* d_ts, uint32_t n, uint32_t op,
                         void* d_out, uint64_t out_cap, uint64_t* d_out_off, uint32_t* d_out_len, uint32_t* d_chan_len, void* stream) {
    if (!ctx || !d_out_off || (op & ~(uint32_t)(SSEGPU_OP_BRIDGE | SSEGPU_OP_GO122 | SSEGPU_OP_RESP))) return SSEGPU_E_INVAL;   // (no SSEGPU_OP_HEADER_ONLY here: the encoder's output IS the payload)
    if (n && (!d_cid || !d_cid_off || !d_tok || !d_tok_off || !d_seq || !d_done || !d_ts || !d_out || !d_out_len)) return SSEGPU_E_INVAL;
    if (n > ctx->cfg.max_records) return SSEGPU_E_TOOBIG;
    if ((reinterpret_cast<uintptr_t>(d_cid) | reinterpret_cast<uintptr_t>(d_tok) | reinterpret_cast<uintptr_t>(d_out)) & 15) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    sse::EncodeArgs a{};
    a.cid = static_cast<const uint8_t*>(d_cid); a.cid_off = d_cid_off; a.tok = static_cast<const uint8_t*>(d_tok); a.tok_off = d_tok_off;
    a.seq = d_seq; a.done = d_done; a.ts = d_ts; a.n = n; a.op = op;
    a.out = static_cast<uint8_t*>(d_out); a.out_cap = out_cap; a.out_off = d_out_off; a.out_len = d_out_len; a.chan_len = d_chan_len;
    sse::encode_scratch_carve(ctx->d_encode_scratch, a);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    CU(sse::encode_launch(a, st));
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->launches += n ? 1 : 0;
    return SSEGPU_OK;
}

int ssegpu_encode_launch_result(ssegpu_ctx* ctx, uint64_t* out_bytes) {
    if (!ctx) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    uint32_t hdr[sse::SCRATCH_HDR / 4];
    CU(cudaMemcpy(hdr, ctx->d_encode_scratch, sizeof hdr, cudaMemcpyDeviceToHost));
    if (out_bytes) memcpy(out_bytes, &hdr[2], 8);
    return hdr[1] ? SSEGPU_E_OVERFLOW : SSEGPU_OK;
}

// ---- table updates ---------------------------------------------------------------------------
// With ctx->mu held: the table stream waits for every batch and device-resident launch in flight that reads tables.
static int table_stream_wait_users(ssegpu_ctx* ctx) {
    for (ssegpu_batch* b : ctx->slots)
        if (b->state == SUBMITTED && (b->tab || b->insp)) CU(cudaStreamWaitEvent(ctx->table_stream, b->ev_done, 0));
    for (cudaEvent_t e : ctx->ev_launch) CU(cudaStreamWaitEvent(ctx->table_stream, e, 0));
    CU(cudaStreamWaitEvent(ctx->table_stream, ctx->ev_stream, 0));
    return SSEGPU_OK;
}

__global__ void k_interest_apply(const SlotOp* __restrict__ ops, uint32_t n, sse::InterestSlot* slots, const uint32_t* __restrict__ new_ids,
                                 uint32_t n_ids, uint32_t* subs_tail) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t k = i; k < n_ids; k += gridDim.x * blockDim.x) subs_tail[k] = new_ids[k];      // the arena's new tail (nothing refers to it yet)
    if (i >= n) return;
    const uint4* src = reinterpret_cast<const uint4*>(&ops[i].image);
    uint4* dst = reinterpret_cast<uint4*>(slots + ops[i].index);
    dst[1] = src[1]; dst[2] = src[2]; dst[3] = src[3];
    dst[0] = src[0];                                              // tag and value last (no reader runs beside this kernel anyway)
}

// ---- content inspection --------------------------------------------------------------------
int ssegpu_patterns_load(ssegpu_ctx* ctx, const ssegpu_pattern* pats, uint32_t n, uint32_t mode) {
    if (!ctx || !pats || mode > 1) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    std::vector<sse_inspect::PatternSpec> specs;
    for (uint32_t i = 0; i < n; i++) specs.push_back({pats[i].kind, pats[i].action, pats[i].text ? pats[i].text : ""});
    sse_inspect::Compiled c;
    std::string err;
    if (!sse_inspect::compile(specs, c, err)) { snprintf(g_cuda_err, sizeof g_cuda_err, "%s", err.c_str()); return SSEGPU_E_PATTERN; }
    if (c.anchored.size() > sse::INSPECT_MAX_ANCHORED) { snprintf(g_cuda_err, sizeof g_cuda_err, "more than %u variable-length patterns", sse::INSPECT_MAX_ANCHORED); return SSEGPU_E_PATTERN; }
    // flatten every DFA into one blob: cls[256] | next u16[] | acc_imm | acc_del | acc_eot, 16-byte aligned pieces
    std::vector<uint8_t> blob;
    auto put = [&](const void* p, size_t bytes) { size_t off = (blob.size() + 15) & ~(size_t)15; blob.resize(off + bytes); memcpy(blob.data() + off, p, bytes); return off; };
    struct Off { size_t cls, next, imm, del, eot, dist; };
    std::vector<const sse_inspect::Dfa*> dfas{&c.forward};
    for (auto& d : c.anchored) dfas.push_back(&d);
    std::vector<Off> offs;
    for (const sse_inspect::Dfa* d : dfas)
        offs.push_back({put(d->cls, 256), put(d->next.data(), d->next.size() * 2), put(d->acc_imm.data(), d->acc_imm.size() * 4),
                        put(d->acc_del.data(), d->acc_del.size() * 4), put(d->acc_eot.data(), d->acc_eot.size() * 4),
                        put(d->dist.data(), d->dist.size())});
    blob.resize((blob.size() + 31) & ~(size_t)15);
    // A new generation is built beside the live one (uploads on the table stream, which nothing else is using yet); on
    // any failure it is dropped and the old set stays in force.  No device-wide synchronisation anywhere.
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto gen = std::make_shared<InspGen>();
    gen->stream = ctx->table_stream;
    gen->mode = mode; gen->start_state = c.forward.start[0]; gen->n_states = c.forward.n_states; gen->n_classes = c.forward.n_classes;
    CU(cudaMallocAsync(&gen->d_blob, blob.size(), ctx->table_stream));
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_desc), sizeof(sse::InspectDev), ctx->table_stream));
    CU(cudaMemcpyAsync(gen->d_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice, ctx->table_stream));
    sse::InspectDev h{};
    auto fill = [&](sse::DfaDev& dd, const sse_inspect::Dfa& d, const Off& o) {
        const uint8_t* base = static_cast<const uint8_t*>(gen->d_blob);
        dd.cls = base + o.cls; dd.next = reinterpret_cast<const uint16_t*>(base + o.next);
        dd.acc_imm = reinterpret_cast<const uint32_t*>(base + o.imm); dd.acc_del = reinterpret_cast<const uint32_t*>(base + o.del);
        dd.acc_eot = reinterpret_cast<const uint32_t*>(base + o.eot);
        dd.dist = base + o.dist;
        dd.n_classes = d.n_classes; dd.start0 = d.start[0]; dd.start1 = d.start[1]; dd.n_states = d.n_states;
    };
    fill(h.fwd, c.forward, offs[0]);
    for (size_t k = 0; k < c.anchored.size(); k++) fill(h.anchored[k], c.anchored[k], offs[k + 1]);
    h.n_patterns = n; h.mode = mode; h.min_len = 0xFFFFFFFFu;
    for (uint32_t i = 0; i < n; i++) h.min_len = c.patterns[i].min_len < h.min_len ? c.patterns[i].min_len : h.min_len;
    for (uint32_t i = 0; i < 32; i++) { h.action[i] = 0; h.fixed_len[i] = 0; h.anchored_of[i] = -1; }
    for (uint32_t i = 0; i < n; i++) {
        h.action[i] = c.patterns[i].action;
        h.fixed_len[i] = c.patterns[i].min_len == c.patterns[i].max_len ? c.patterns[i].min_len : 0;
        h.anchored_of[i] = c.anchored_of[i];
    }
    gen->dfa = sse::DfaUniform{h.fwd.next, h.fwd.dist, h.fwd.n_classes, h.fwd.n_states, h.fwd.start0, h.min_len};
    CU(cudaMemcpyAsync(gen->d_desc, &h, sizeof h, cudaMemcpyHostToDevice, ctx->table_stream));
    CU(cudaStreamSynchronize(ctx->table_stream));                 // the host buffers above are about to go out of scope (own stream only)
    if (ctx->cfg.conv_slots) {
        // a new pattern set restarts every conversation (the carried states belong to the old DFA): once the batches in
        // flight that still use the old one are through
        if (!ctx->d_conv_state) CU(cudaMalloc(&ctx->d_conv_state, (size_t)ctx->cfg.conv_slots * 4));
        if (mode == sse::INSPECT_STREAMING && !ctx->d_tok_entry) {     // token entries + sort scratch of the per-conversation walk
            CU(cudaMalloc(&ctx->d_tok_entry, (size_t)ctx->cfg.max_records * 32 + 32));
            CU(cudaMalloc(&ctx->d_stream_scratch, sse::stream_scratch_bytes(ctx->cfg.max_records)));
        }
        { const int rc = table_stream_wait_users(ctx); if (rc != SSEGPU_OK) return rc; }
        std::vector<uint32_t> init(ctx->cfg.conv_slots, gen->start_state);
        CU(cudaMemcpyAsync(ctx->d_conv_state, init.data(), init.size() * 4, cudaMemcpyHostToDevice, ctx->table_stream));
        CU(cudaStreamSynchronize(ctx->table_stream));
    }
    CU(cudaEventRecord(ctx->ev_table, ctx->table_stream));
    ctx->insp = gen;                                                  // the old generation goes when its last batch is released
    return SSEGPU_OK;
}

int ssegpu_conv_reset(ssegpu_ctx* ctx) {
    if (!ctx) return SSEGPU_E_INVAL;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->d_conv_state || !ctx->insp) return SSEGPU_OK;
    CU(cudaSetDevice(ctx->cfg.device));
    { const int rc = table_stream_wait_users(ctx); if (rc != SSEGPU_OK) return rc; }
    std::vector<uint32_t> init(ctx->cfg.conv_slots, ctx->insp->start_state);
    CU(cudaMemcpyAsync(ctx->d_conv_state, init.data(), init.size() * 4, cudaMemcpyHostToDevice, ctx->table_stream));
    CU(cudaStreamSynchronize(ctx->table_stream));
    CU(cudaEventRecord(ctx->ev_table, ctx->table_stream));
    return SSEGPU_OK;
}

int ssegpu_inspect_bind(ssegpu_ctx* ctx, const uint32_t* d_conv_slot, uint32_t* d_verdict, int32_t* d_match) {
    if (!ctx) return SSEGPU_E_INVAL;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->bound_conv = d_conv_slot; ctx->bound_verdict = d_verdict; ctx->bound_match = d_match;
    return SSEGPU_OK;
}

// ---- subscriber / interest table ---------------------------------------------------------------
// A new device generation from the host table's flattened form (with ctx->mu held).
static int interest_rebuild(ssegpu_ctx* ctx) {
    sse_interest::HostTable& t = ctx->interest;
    t.flatten();
    auto gen = std::make_shared<TableGen>();
    gen->stream = ctx->table_stream;
    gen->subs = t.subs; gen->wild_ids = t.wild_ids;
    cudaStream_t ts = ctx->table_stream;
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_slots), t.slots.size() * sizeof(sse::InterestSlot), ts));
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_keys), t.keys.size() * 4, ts));
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_wild), t.wild_bytes.size(), ts));
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_desc), sizeof(sse::InterestDev), ts));
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_subs), (t.subs->capacity() + 4) * 4, ts));
    CU(cudaMallocAsync(reinterpret_cast<void**>(&gen->d_wild_ids), sse::INTEREST_MAX_WILD * 4, ts));
    if (!t.subs->empty()) CU(cudaMemcpyAsync(gen->d_subs, t.subs->data(), t.subs->size() * 4, cudaMemcpyHostToDevice, ts));
    {
        uint32_t w[sse::INTEREST_MAX_WILD] = {0};
        for (size_t k = 0; k < t.wild_ids.size(); k++) w[k] = t.wild_ids[k];
        CU(cudaMemcpyAsync(gen->d_wild_ids, w, sizeof w, cudaMemcpyHostToDevice, ts));
        CU(cudaStreamSynchronize(ts));
    }
    CU(cudaMemcpyAsync(gen->d_slots, t.slots.data(), t.slots.size() * sizeof(sse::InterestSlot), cudaMemcpyHostToDevice, ts));
    CU(cudaMemcpyAsync(gen->d_keys, t.keys.data(), t.keys.size() * 4, cudaMemcpyHostToDevice, ts));
    CU(cudaMemcpyAsync(gen->d_wild, t.wild_bytes.data(), t.wild_bytes.size(), cudaMemcpyHostToDevice, ts));
    sse::InterestDev h{};
    h.slots = gen->d_slots; h.keys = gen->d_keys;
    h.wild_single = t.wild_single; h.wild_generic = t.wild_generic;
    h.mask = (uint32_t)t.slots.size() - 1; h.n_wild = (uint32_t)t.wild_ids.size();
    h.wild_bytes = gen->d_wild;
    for (size_t i = 0; i < t.wild_off.size(); i++) h.wild_off[i] = t.wild_off[i];
    gen->uni = sse::InterestUniform{h.slots, h.mask, h.wild_single, h.wild_generic};
    CU(cudaMemcpyAsync(gen->d_desc, &h, sizeof h, cudaMemcpyHostToDevice, ts));
    CU(cudaStreamSynchronize(ts));                                // own stream only: the uploads came from pageable host memory
    CU(cudaEventRecord(ctx->ev_table, ts));
    ctx->tab = gen;                                               // batches in flight keep the old generation alive until they are released
    return SSEGPU_OK;
}

int ssegpu_interest_apply(ssegpu_ctx* ctx, const ssegpu_sub_delta* d, uint32_t n) {
    if (!ctx || (n && !d)) return SSEGPU_E_INVAL;
    for (uint32_t i = 0; i < n; i++) if (!d[i].subject || d[i].op > 1) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    std::lock_guard<std::mutex> lk(ctx->mu);
    sse_interest::HostTable& t = ctx->interest;
    std::string err;
    bool rebuild = !ctx->tab || !t.flattened;
    const size_t arena_before = t.subs->size();
    for (uint32_t i = 0; i < n; i++) {
        std::string subj(d[i].subject, d[i].subject_len);
        const int ch = d[i].op == 0 ? t.add(subj, d[i].sub_id, err) : t.del(subj, d[i].sub_id);
        if (ch < 0) {                                             // the deltas before this one stay applied
            snprintf(g_cuda_err, sizeof g_cuda_err, "%s", err.c_str());
            if (interest_rebuild(ctx) == SSEGPU_OK) ctx->table_version++;
            return SSEGPU_E_TOOBIG;
        }
        if (ch == 2) rebuild = true;                              // the wildcard list lives in the descriptor: new generation
        else if (ch == 1 && !rebuild && !t.sync_subject(subj)) rebuild = true;
    }
    // slot images and the arena's new tail travel together in one pinned staging chunk
    const size_t new_ids = rebuild ? 0 : t.subs->size() - arena_before;
    if (!rebuild && t.dirty.size() * sizeof(SlotOp) + new_ids * 4 > sizeof(SlotOp) * (size_t)ssegpu_ctx::kOpsPerChunk) rebuild = true;
    if (rebuild) {
        const int rc = interest_rebuild(ctx);
        if (rc == SSEGPU_OK) ctx->table_version++;
        return rc;
    }
    if (!t.dirty.empty()) {
        // in place: the changed slots go to the device through a pinned staging chunk and a small kernel on the table
        // stream, which first waits for the batches in flight; batches submitted from now on wait for it (ev_table)
        const int chunk = ctx->next_chunk++ % ssegpu_ctx::kOpChunks;
        CU(cudaEventSynchronize(ctx->ev_chunk[chunk]));           // the chunk's previous use (kOpChunks updates ago) has been consumed
        SlotOp* h = ctx->h_ops + (size_t)chunk * ssegpu_ctx::kOpsPerChunk;
        SlotOp* dv = ctx->d_ops + (size_t)chunk * ssegpu_ctx::kOpsPerChunk;
        const uint32_t m = (uint32_t)t.dirty.size();
        for (uint32_t k = 0; k < m; k++) { h[k].index = t.dirty[k]; h[k].image = t.slots[t.dirty[k]]; }
        t.dirty.clear();
        if (new_ids) memcpy(h + m, t.subs->data() + arena_before, new_ids * 4);
        { const int rc = table_stream_wait_users(ctx); if (rc != SSEGPU_OK) return rc; }
        CU(cudaMemcpyAsync(dv, h, sizeof(SlotOp) * m + new_ids * 4, cudaMemcpyHostToDevice, ctx->table_stream));
        k_interest_apply<<<(m + 127) / 128, 128, 0, ctx->table_stream>>>(dv, m, ctx->tab->d_slots, reinterpret_cast<const uint32_t*>(dv + m), (uint32_t)new_ids,
                                                                        ctx->tab->d_subs + arena_before);
        CU(cudaGetLastError());
        CU(cudaEventRecord(ctx->ev_chunk[chunk], ctx->table_stream));
        CU(cudaEventRecord(ctx->ev_table, ctx->table_stream));
    }
    ctx->table_version++;
    return SSEGPU_OK;
}

const uint32_t* ssegpu_interest_subs(ssegpu_ctx* ctx, uint32_t* count) {
    if (!ctx) return nullptr;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (count) *count = (uint32_t)ctx->interest.subs->size();
    return ctx->interest.subs->data();
}

const uint32_t* ssegpu_interest_wild_ids(ssegpu_ctx* ctx, uint32_t* count) {
    if (!ctx) return nullptr;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (count) *count = (uint32_t)ctx->interest.wild_ids.size();
    return ctx->interest.wild_ids.data();
}

uint64_t ssegpu_interest_version(ssegpu_ctx* ctx) {
    if (!ctx) return 0;
    std::lock_guard<std::mutex> lk(ctx->mu);
    return ctx->table_version;
}

const uint32_t* ssegpu_batch_subs(const ssegpu_batch* b, uint32_t* count) {
    if (!b || !b->tab) { if (count) *count = 0; return nullptr; }
    if (count) *count = (uint32_t)b->tab->subs->size();
    return b->tab->subs->data();
}

const uint32_t* ssegpu_batch_wild_ids(const ssegpu_batch* b, uint32_t* count) {
    if (!b || !b->tab) { if (count) *count = 0; return nullptr; }
    if (count) *count = (uint32_t)b->tab->wild_ids.size();
    return b->tab->wild_ids.data();
}

uint64_t ssegpu_batch_table_version(const ssegpu_batch* b) { return b ? b->table_version : 0; }

int ssegpu_interest_bind(ssegpu_ctx* ctx, uint32_t* d_sub_span) {
    if (!ctx) return SSEGPU_E_INVAL;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->bound_hit = d_sub_span;
    return SSEGPU_OK;
}

// ---- fan-out expansion ----------------------------------------------------------------------------
// (record, subscriber) pairs from the spans OP_MATCH left: counts -> exclusive scan -> one thread per PAIR (binary search
// for its record in the offsets), so a record with 1 024 subscribers costs no more per pair than one with a single one.
__global__ void k_fanout_count(const uint32_t* __restrict__ span, uint32_t n, uint32_t* __restrict__ cnt, uint32_t* last) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = span[3 * (size_t)i + 1] + __popc(span[3 * (size_t)i + 2]);
    cnt[i] = c;
    if (i == n - 1) *last = c;
}
__global__ void k_fanout_totals(const uint32_t* rec_off, uint32_t n, uint64_t max_pairs, uint32_t* totals, uint32_t* rec_off_n) {
    const uint32_t total = rec_off[n - 1] + totals[3];
    totals[0] = total; totals[2] = total > max_pairs ? 1u : 0u;
    *rec_off_n = total;
}
__global__ void k_fanout_emit(const uint32_t* __restrict__ span, const uint32_t* __restrict__ rec_off, uint32_t n, const uint32_t* __restrict__ subs,
                              const uint32_t* __restrict__ wild_ids, const uint32_t* totals, uint32_t* __restrict__ pair_sub, uint32_t* __restrict__ pair_rec) {
    if (totals[2]) return;
    const uint32_t total = totals[0];
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = n;                                  // last record whose offset is <= t
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (rec_off[mid] <= t) lo = mid; else hi = mid; }
        const uint32_t k = t - rec_off[lo], off = span[3 * (size_t)lo], cnt = span[3 * (size_t)lo + 1];
        uint32_t id;
        if (k < cnt) id = subs[off + k];
        else {                                                    // the (k - cnt)-th set bit of the wildcard mask
            uint32_t mask = span[3 * (size_t)lo + 2];
            for (uint32_t j = k - cnt; j; j--) mask &= mask - 1;
            id = wild_ids[__ffs((int)mask) - 1];
        }
        pair_sub[t] = id; pair_rec[t] = lo;
    }
}

int ssegpu_fanout_launch(ssegpu_ctx* ctx, const uint32_t* d_sub_span, uint32_t n, uint32_t* d_pair_sub, uint32_t* d_pair_rec, uint64_t max_pairs,
                         uint32_t* d_rec_off, void* stream) {
    if (!ctx || !d_rec_off || (n && (!d_sub_span || !d_pair_sub || !d_pair_rec))) return SSEGPU_E_INVAL;
    if (n > ctx->cfg.max_records) return SSEGPU_E_TOOBIG;
    CU(cudaSetDevice(ctx->cfg.device));
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->tab) return SSEGPU_E_STATE;
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    if (!ctx->d_fanout_scratch) CU(cudaMalloc(&ctx->d_fanout_scratch, ((size_t)ctx->cfg.max_records / 4096 + 16) * 4 + 64));
    uint32_t* totals = static_cast<uint32_t*>(ctx->d_fanout_scratch);
    uint32_t* block_sums = totals + 8;
    CU(cudaStreamWaitEvent(st, ctx->ev_table, 0));
    CU(cudaMemsetAsync(totals, 0, 32, st));
    if (n == 0) { CU(cudaMemsetAsync(d_rec_off, 0, 4, st)); return SSEGPU_OK; }
    k_fanout_count<<<(n + 255) / 256, 256, 0, st>>>(d_sub_span, n, d_rec_off, totals + 3);
    CU(sse::scan_exclusive_u32(d_rec_off, n, block_sums, st));
    k_fanout_totals<<<1, 1, 0, st>>>(d_rec_off, n, max_pairs, totals, d_rec_off + n);
    k_fanout_emit<<<148 * 8, 256, 0, st>>>(d_sub_span, d_rec_off, n, ctx->tab->d_subs, ctx->tab->d_wild_ids, totals, d_pair_sub, d_pair_rec);
    CU(cudaGetLastError());
    cudaEvent_t ev = ctx->ev_launch[ctx->next_launch_ev++ % ssegpu_ctx::kLaunchEvents];     // a later table update waits for this reader
    CU(cudaEventRecord(ev, st));
    ctx->launches += 6;
    return SSEGPU_OK;
}

int ssegpu_fanout_result(ssegpu_ctx* ctx, uint64_t* n_pairs) {
    if (!ctx || !ctx->d_fanout_scratch) return SSEGPU_E_STATE;
    CU(cudaSetDevice(ctx->cfg.device));
    uint32_t t[4];
    CU(cudaMemcpy(t, ctx->d_fanout_scratch, sizeof t, cudaMemcpyDeviceToHost));
    if (n_pairs) *n_pairs = t[0];
    return t[2] ? SSEGPU_E_OVERFLOW : SSEGPU_OK;
}

// ---- receiving side of the wire protocols ------------------------------------------------------
int ssegpu_split_launch(ssegpu_ctx* ctx, uint32_t proto, const void* d_buf, const uint32_t* d_buf_off, uint32_t n_conn, void* d_arena,
                        uint64_t arena_cap, uint32_t* d_in_off, uint32_t max_msgs, uint32_t* d_msg_conn, uint64_t* d_msg_sid,
                        uint32_t* d_msg_subj, uint32_t* d_conn_consumed, uint32_t* d_conn_status, uint32_t* d_conn_pings, void* stream) {
    if (!ctx || proto > SSEGPU_WIRE_RESP || !d_buf_off || !d_in_off || !d_arena || (reinterpret_cast<uintptr_t>(d_arena) & 15)) return SSEGPU_E_INVAL;
    if (n_conn && (!d_buf || !d_msg_conn || !d_msg_sid || !d_msg_subj || !d_conn_consumed || !d_conn_status || !d_conn_pings)) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    std::lock_guard<std::mutex> lk(ctx->mu);
    const size_t need = sse::wire_scratch_bytes(n_conn, max_msgs);
    if (need > ctx->wire_scratch_cap) {                      // grows on demand (first call, or a larger box of connections)
        cudaStream_t st0 = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
        CU(cudaStreamSynchronize(st0));
        cudaFree(ctx->d_wire_scratch); cudaFree(ctx->d_chunk_scratch); cudaFree(ctx->d_fanout_scratch);
        ctx->d_wire_scratch = nullptr; ctx->wire_scratch_cap = 0;
        CU(cudaMalloc(&ctx->d_wire_scratch, need * 2));
        ctx->wire_scratch_cap = need * 2;
    }
    sse::WireArgs a{};
    a.proto = proto; a.buf = static_cast<const uint8_t*>(d_buf); a.buf_off = d_buf_off; a.n_conn = n_conn; a.max_msgs = max_msgs;
    a.arena = static_cast<uint8_t*>(d_arena); a.arena_cap = arena_cap; a.in_off = d_in_off;
    a.msg_conn = d_msg_conn; a.msg_sid = d_msg_sid; a.msg_subj = d_msg_subj;
    a.conn_consumed = d_conn_consumed; a.conn_status = d_conn_status; a.conn_pings = d_conn_pings;
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    CU(sse::wire_split_launch(a, ctx->d_wire_scratch, st));
    ctx->launches += n_conn ? 4 : 0;
    return SSEGPU_OK;
}

int ssegpu_split_result(ssegpu_ctx* ctx, uint32_t* n_msgs, uint64_t* arena_bytes) {
    if (!ctx || !ctx->d_wire_scratch) return SSEGPU_E_STATE;
    CU(cudaSetDevice(ctx->cfg.device));
    uint32_t t[4];
    CU(cudaMemcpy(t, ctx->d_wire_scratch, sizeof t, cudaMemcpyDeviceToHost));
    if (n_msgs) *n_msgs = t[0];
    if (arena_bytes) *arena_bytes = t[1];
    return t[2] ? SSEGPU_E_OVERFLOW : SSEGPU_OK;
}

// ---- decode half of row N3 ----------------------------------------------------------------------
int ssegpu_chunk_decode_launch(ssegpu_ctx* ctx, const void* d_lines, const uint32_t* d_line_off, uint32_t n_lines, void* d_tok, uint64_t tok_cap,
                               uint32_t* d_tok_off, uint8_t* d_done, uint32_t* d_msg_line, uint32_t max_msgs, uint32_t* d_line_status, void* stream) {
    if (!ctx || !d_line_off || !d_tok || !d_tok_off || (reinterpret_cast<uintptr_t>(d_tok) & 15)) return SSEGPU_E_INVAL;
    if (n_lines && (!d_lines || !d_done || !d_msg_line || !d_line_status)) return SSEGPU_E_INVAL;
    CU(cudaSetDevice(ctx->cfg.device));
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    const size_t need = sse::chunk_scratch_bytes(n_lines);
    if (need > ctx->chunk_scratch_cap) {
        CU(cudaStreamSynchronize(st));
        cudaFree(ctx->d_chunk_scratch);
        ctx->d_chunk_scratch = nullptr; ctx->chunk_scratch_cap = 0;
        CU(cudaMalloc(&ctx->d_chunk_scratch, need * 2));
        ctx->chunk_scratch_cap = need * 2;
    }
    sse::ChunkArgs a{};
    a.lines = static_cast<const uint8_t*>(d_lines); a.line_off = d_line_off; a.n_lines = n_lines; a.max_msgs = max_msgs;
    a.tok = static_cast<uint8_t*>(d_tok); a.tok_cap = tok_cap; a.tok_off = d_tok_off; a.done = d_done; a.msg_line = d_msg_line;
    a.line_status = d_line_status;
    CU(sse::chunk_decode_launch(a, ctx->d_chunk_scratch, st));
    ctx->launches += n_lines ? 9 : 0;
    return SSEGPU_OK;
}

int ssegpu_chunk_decode_result(ssegpu_ctx* ctx, uint32_t* n_msgs, uint64_t* tok_bytes) {
    if (!ctx || !ctx->d_chunk_scratch) return SSEGPU_E_STATE;
    CU(cudaSetDevice(ctx->cfg.device));
    uint32_t t[4];
    CU(cudaMemcpy(t, ctx->d_chunk_scratch, sizeof t, cudaMemcpyDeviceToHost));
    if (n_msgs) *n_msgs = t[0];
    if (tok_bytes) *tok_bytes = t[1];
    return t[2] ? SSEGPU_E_OVERFLOW : SSEGPU_OK;
}

uint64_t ssegpu_launch_count(ssegpu_ctx* ctx) {
    if (!ctx) return 0;
    std::lock_guard<std::mutex> lk(ctx->mu);
    return ctx->launches;
}

}  // extern "C"
