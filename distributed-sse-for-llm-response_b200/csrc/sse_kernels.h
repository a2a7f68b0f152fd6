// sse_kernels.h -- host-visible launch interface of the fused frame kernel.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace sse {

struct InspectDev;               // sse_inspect.cuh
struct InterestDev;              // sse_interest.cuh

#ifndef SSE_TILE
#define SSE_TILE 128
#endif
constexpr int TILE = SSE_TILE;   // records per tile == threads per CTA
constexpr int CTR_ROWS = 64;     // counter rows (CTA i adds into row i % CTR_ROWS; the host sums rows)
constexpr int CTR_COLS = 8;      // ok, syntax, type, filtered, done, fast, bytes_in, bytes_out
constexpr size_t SCRATCH_HDR = 64;   // [0] ticket u32, [4] overflow u32, [8] cursor u64, [16] finished CTAs u32, [20] conversation-slot-out-of-range flag u32, [24] bad-offsets flag u32

enum : uint32_t { XF_PRE_VERDICT = 1u,       // verdict[] already holds this batch's verdicts: apply them (inline mode), do not inspect
                  XF_ENTRY_IN_COUNT = 2u };  // write tok_entry in ordered mode's counting pass too (it runs alone, as a planning pass)

// Scalars of the interest table's descriptor, passed by value: kernel-parameter (constant) space instead of registers
// held across the whole persistent loop.  Constant for the lifetime of a table generation.
struct InterestSlot;
struct InterestUniform { const InterestSlot* slots; uint32_t mask, wild_single, wild_generic; };

// The same for the forward DFA of the content inspection (insp->fwd, insp->min_len).
struct DfaUniform { const uint16_t* next; const uint8_t* dist; uint32_t n_classes, n_states, start0, min_len; };

struct FrameArgs {
    const uint8_t* in;           // packed payload arena, 16-byte aligned, 32 readable bytes after the last record
    const uint32_t* in_off;      // n+1
    const int64_t* start_seq;    // n or nullptr
    uint32_t n, op;
    uint8_t* out;                // packed output arena
    uint64_t out_cap;
    uint64_t* out_off;           // n+1 (entry n = total output bytes)
    uint32_t* out_len;           // n
    uint32_t* subj_len;          // n or nullptr
    uint32_t* status;            // n
    // content inspection.  per_token mode is fused into the frame kernel (insp != nullptr).  streaming mode is a
    // per-conversation walk in its own kernels (sse_stream.cu): the frame kernel then only leaves a 32-byte token
    // entry per record (tok_entry != nullptr) and, for inline verdicts, a second frame pass reads the walk's
    // verdicts back (XF_PRE_VERDICT) instead of inspecting.
    const InspectDev* insp;
    DfaUniform dfa;              // with insp: what insp->fwd and insp->min_len say (filled by the caller from its host copy)
    uint4* tok_entry;            // 2n (one 32-byte entry per record) or nullptr
    uint32_t xflags;
    uint32_t* verdict;           // n
    int32_t* match;              // 2n: begin, end (byte offsets into the decoded token text)
    // subscriber/interest table (nullptr = off): per record (CSR offset, count, wildcard mask)
    const InterestDev* interest;
    InterestUniform itab;        // with interest: the same fields of *interest (filled by the caller from its host copy)
    uint32_t* hit;               // 3n
    // scratch: one device allocation of scratch_bytes(n), zeroed by frame_launch
    void* scratch_base;
    uint32_t* ticket;
    uint32_t* overflow;
    unsigned long long* cursor;  // arrival-order placement: next free output byte
    uint32_t* done_ctas;
    uint32_t* bad_conv;          // streaming inspection: a record's conversation slot is >= conv_slots (set by sse_stream.cu)
    uint32_t* bad_offsets;       // set when in_off was not monotonic / ran past in_off[n] (records are clamped, never read out of bounds)
    unsigned long long* counters;
    uint64_t* tile_state;         // ordered mode: per-tile byte totals / prefix; arrival mode: count of records left to the tail phase | previous such tile of the CTA << 32
    uint32_t* plan = nullptr;    // n words or nullptr: ordered mode's count pass leaves each canonical record's plan here for the emit pass
    void* zero_base = nullptr;   // what frame_launch clears before the launch (default: scratch_base, scratch_bytes(n))
    size_t zero_bytes = 0;
    uint32_t keep_bad_conv = 0;  // the clear leaves the bad_conv word alone (set by the streaming walk in front of this frame pass)
    uint32_t plan_tma = 0;       // filled by frame_launch: the same for the plan words (ordered mode's emit pass)
    uint32_t off_tma = 0;        // filled by frame_launch: in_off is 16-byte aligned, a tile's offsets may ride along with its payload (TMA)
    // filled by frame_launch
    uint32_t num_tiles, in_cap, out_cap_s;
    __host__ __device__ uint32_t out_cap_smem() const { return out_cap_s; }
};

inline size_t scratch_bytes(uint32_t n) {
    size_t tiles = ((size_t)n + TILE - 1) / TILE;
    return SCRATCH_HDR + (size_t)CTR_ROWS * CTR_COLS * 8 + tiles * 8;
}
inline size_t scratch_bytes_with_plan(uint32_t n) { return ((scratch_bytes(n) + 15) & ~(size_t)15) + (size_t)n * 4; }   // + FrameArgs::plan (never cleared)
inline size_t scratch_readback_bytes() { return SCRATCH_HDR + (size_t)CTR_ROWS * CTR_COLS * 8; }
inline void scratch_carve(void* base, FrameArgs& a) {
    uint8_t* p = static_cast<uint8_t*>(base);
    a.scratch_base = base;
    a.ticket = reinterpret_cast<uint32_t*>(p);
    a.overflow = reinterpret_cast<uint32_t*>(p + 4);
    a.cursor = reinterpret_cast<unsigned long long*>(p + 8);
    a.done_ctas = reinterpret_cast<uint32_t*>(p + 16);
    a.bad_conv = reinterpret_cast<uint32_t*>(p + 20);
    a.bad_offsets = reinterpret_cast<uint32_t*>(p + 24);
    a.counters = reinterpret_cast<unsigned long long*>(p + SCRATCH_HDR);
    a.tile_state = reinterpret_cast<uint64_t*>(p + SCRATCH_HDR + (size_t)CTR_ROWS * CTR_COLS * 8);
}

// producer-side encoder (sse_encode.cu)
struct EncodeArgs {
    const uint8_t* cid; const uint32_t* cid_off;   // conversation ids, packed, n+1 offsets
    const uint8_t* tok; const uint32_t* tok_off;   // token texts, packed, n+1 offsets
    const int64_t* seq; const uint8_t* done; const int64_t* ts;
    uint32_t n, op;                                // OP_BRIDGE: prepend the channel "chat.<id>.tokens"; OP_GO122
    uint8_t* out; uint64_t out_cap; uint64_t* out_off; uint32_t* out_len; uint32_t* chan_len;
    void* scratch_base; uint32_t* overflow; unsigned long long* cursor; uint32_t* done_ctas;
};
inline void encode_scratch_carve(void* base, EncodeArgs& a) {
    uint8_t* p = static_cast<uint8_t*>(base);
    a.scratch_base = base;
    a.overflow = reinterpret_cast<uint32_t*>(p + 4);
    a.cursor = reinterpret_cast<unsigned long long*>(p + 8);
    a.done_ctas = reinterpret_cast<uint32_t*>(p + 16);
}
cudaError_t encode_launch(EncodeArgs a, cudaStream_t stream);

// Batch slots keep header + counters directly in front of the per-record result arrays (one D2H copy for all of
// them) and the per-tile words directly in front of the header (one clear for both).
inline void scratch_carve_split(void* hdr_base, uint32_t n, FrameArgs& a) {
    scratch_carve(hdr_base, a);
    const size_t tiles = ((size_t)n + TILE - 1) / TILE;
    a.tile_state = reinterpret_cast<uint64_t*>(static_cast<uint8_t*>(hdr_base) - tiles * 8);
    a.zero_base = a.tile_state;
    a.zero_bytes = tiles * 8 + scratch_readback_bytes();
}

// ---- receiving side of the wire protocols (sse_wire.cu): socket bytes -> packed payload arena ----
struct WireArgs {
    uint32_t proto;              // WIRE_NATS / WIRE_RESP
    const uint8_t* buf;          // every connection's unread socket bytes, back to back
    const uint32_t* buf_off;     // n_conn + 1
    uint32_t n_conn, max_msgs;
    uint8_t* arena; uint64_t arena_cap;   // out: packed payloads (+48 readable bytes), 16-byte aligned: the frame kernel's input
    uint32_t* in_off;            // out: max_msgs + 1
    uint32_t* msg_conn;          // out per delivery: connection index
    uint64_t* msg_sid;           // out per delivery: NATS subscription id (0 for RESP)
    uint32_t* msg_subj;          // out per delivery: (offset into buf, length) of the subject / channel
    uint32_t* conn_consumed; uint32_t* conn_status; uint32_t* conn_pings;   // out per connection
    // scratch (carved by wire_split_launch)
    uint32_t* totals;            // [0] deliveries, [1] payload bytes, [2] overflow flag, [4..5] the last connection's own counts
    uint32_t* cnt_msgs; uint32_t* cnt_bytes; uint32_t* msg_src;
};
size_t wire_scratch_bytes(uint32_t n_conn, uint32_t max_msgs);
cudaError_t wire_split_launch(WireArgs a, void* scratch, cudaStream_t stream);

cudaError_t scan_exclusive_u32(uint32_t* data, size_t m, uint32_t* block_sums, cudaStream_t stream);   // sse_stream.cu

// ---- decode half of row N3 (sse_chunk.cu): vLLM stream lines -> token columns for the encoder ----
struct ChunkArgs {
    const uint8_t* lines;        // SSE lines of the vLLM response bodies, back to back (each with 8 readable bytes behind it is not needed: reads stay inside the line)
    const uint32_t* line_off;    // n_lines + 1
    uint32_t n_lines, max_msgs;
    uint8_t* tok; uint64_t tok_cap;   // out: decoded token texts, packed (+32 readable bytes): the encoder's d_tok
    uint32_t* tok_off;           // out: max_msgs + 1
    uint8_t* done;               // out per message: 1 for the "[DONE]" sentinel
    uint32_t* msg_line;          // out per message: the line it came from
    uint32_t* line_status;       // out per line: CHUNK_* word
    // scratch (carved by chunk_decode_launch)
    uint32_t* totals;            // [0] messages, [1] token bytes, [2] overflow flag, [3..4] the last line's own counts
    uint32_t* cnt_msgs; uint32_t* cnt_bytes; uint32_t* block_sums;
    uint32_t* span;              // 2 per line: content span of a line with exactly one delta (count pass -> emit pass), else 0, 0
};
size_t chunk_scratch_bytes(uint32_t n_lines);
cudaError_t chunk_decode_launch(ChunkArgs a, void* scratch, cudaStream_t stream);

cudaError_t frame_configure(int device);
cudaError_t frame_launch(int device, FrameArgs a, cudaStream_t stream);
// the counting pass of ordered mode on its own: decodes, leaves token entries (XF_ENTRY_IN_COUNT) and emits nothing --
// the planning pass in front of the streaming walk when its verdicts must be applied inline
cudaError_t frame_plan_launch(int device, FrameArgs a, cudaStream_t stream);

// ---- streaming inspection: per-conversation walk over the token entries (sse_stream.cu) ----
struct StreamArgs {
    const InspectDev* insp;      // device descriptor
    const uint8_t* in;           // payload arena (tokens longer than 8 bytes or with escapes are read from it)
    const uint4* tok_entry;      // 2n, written by the frame kernel
    const uint32_t* conv_slot;   // n
    uint32_t n, conv_slots;
    uint32_t* conv_state;        // [conv_slots] carried DFA state
    uint32_t* verdict;           // n   (cleared here; only hits are written)
    int32_t* match;              // 2n
    uint32_t* bad_conv;          // flag word
    void* scratch;               // stream_scratch_bytes(n) bytes
    uint32_t dfa_states, dfa_classes;   // of the forward DFA (sizes the on-chip table)
};
size_t stream_scratch_bytes(uint32_t n);
cudaError_t stream_configure(int device);
cudaError_t stream_inspect_launch(int device, const StreamArgs& a, cudaStream_t stream);

}  // namespace sse
