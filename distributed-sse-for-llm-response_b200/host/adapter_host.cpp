// adapter_host.cpp -- host-side mirror of the SSE Adapter's per-connection flow on top of the C ABI
// (include/ssegpu.h).  It is what the patched Go handler does, written in C++ because no Go toolchain exists in
// this build image; INTEGRATION.md shows the same logic in Go.
//
// Reference flow being mirrored (src/sse-adapter/sse_handler.go):
//   HandleStream   :80-234   GET /stream/{id}: Last-Event-ID -> startSequence (:90-96), Subscribe (:133-157),
//                            ": connected to <id>\n\n" (:178), then per token Marshal + 3x Fprintf + Flush
//                            (:211-215), return after the frame whose done is true (:220-226)
//   HandleChat     :239-450  POST /chat: same loop (:400-449) after "event: connected" (:354-356), no filter
//   NATS callback  :135-157  decode, filter, non-blocking hand-off to the handler goroutine
//
// What changes with the library: the callback APPENDS the raw payload to the current batch (no decode), a batcher
// cuts batches and runs them through ssegpu_batch_submit/wait, and every connection then gets ALL of its frames of
// the batch as ONE contiguous write (the reference pays one Write+Flush syscall per token,
// sse_handler.go:212-215; chunk boundaries are not part of SSE semantics, the byte stream is unchanged).
// A connection stops receiving after its done frame, exactly where the reference handler returns.
//
// Control frames (": connected", keep-alives, error events) are produced here on the host as in the reference;
// they never go through the GPU.
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/ssegpu.h"

namespace {

struct Conn {
    std::string conversation_id;
    int64_t start_seq = 0;       // sse_handler.go:92-96
    bool chat = false;           // POST /chat connection (no sequence filter)
    bool closed = false;         // handler returned (done frame delivered, or Disconnect)
    std::string pending;         // bytes to write with the next Write+Flush
    uint64_t frames = 0, writes = 0;
};

}  // namespace

struct ssehost {
    ssegpu_ctx* ctx = nullptr;
    uint32_t max_records = 0;
    uint64_t max_in_bytes = 0;
    std::unordered_map<uint64_t, Conn> conns;
    ssegpu_batch* cur = nullptr;
    uint32_t n = 0;
    uint64_t used = 0;
    std::vector<uint64_t> owner;   // record -> connection id
    uint64_t delivered = 0;        // sse_messages_delivered_total
    uint64_t dropped_bad = 0, filtered = 0, after_close = 0;
    std::string err;
    bool header_only = false;      // SSEGPU_OP_HEADER_ONLY egress (ssehost_set_header_only)
};

extern "C" {

ssehost* ssehost_new(ssegpu_ctx* ctx, uint32_t max_records, uint64_t max_in_bytes) {
    if (!ctx || !max_records) return nullptr;
    ssehost* h = new ssehost();
    h->ctx = ctx; h->max_records = max_records; h->max_in_bytes = max_in_bytes;
    h->owner.resize(max_records);
    return h;
}

void ssehost_free(ssehost* h) {
    if (!h) return;
    if (h->cur) ssegpu_batch_release(h->ctx, h->cur);
    delete h;
}

// GET /stream/{id} with an optional Last-Event-ID header (chat == 0), or POST /chat (chat == 1).
// Queues the connection's greeting exactly as the reference writes it.
int ssehost_connect(ssehost* h, uint64_t conn_id, const char* conversation_id, const char* last_event_id, int chat) {
    if (!h || !conversation_id) return SSEGPU_E_INVAL;
    Conn& c = h->conns[conn_id];
    c = Conn();
    c.conversation_id = conversation_id;
    c.chat = chat != 0;
    if (!c.chat) {
        c.start_seq = (last_event_id && *last_event_id) ? ssegpu_start_sequence(last_event_id) : 0;
        c.pending = ": connected to " + c.conversation_id + "\n\n";                                           // sse_handler.go:178
    } else {
        c.pending = "event: connected\ndata: {\"conversation_id\":\"" + c.conversation_id + "\"}\n\n";      // :354-355 (id not escaped there either)
    }
    return SSEGPU_OK;
}

int ssehost_disconnect(ssehost* h, uint64_t conn_id) {
    if (!h) return SSEGPU_E_INVAL;
    auto it = h->conns.find(conn_id);
    if (it != h->conns.end()) it->second.closed = true;     // r.Context().Done(): sse_handler.go:188-190
    return SSEGPU_OK;
}

// keep-alive tick of one connection (sse_handler.go:228-231, :444-447)
int ssehost_keepalive(ssehost* h, uint64_t conn_id) {
    auto it = h->conns.find(conn_id);
    if (it == h->conns.end() || it->second.closed) return SSEGPU_E_INVAL;
    it->second.pending += ": keep-alive\n\n";
    return SSEGPU_OK;
}

// The NATS callback of connection conn_id received msg.Data.  Returns SSEGPU_E_AGAIN when the batch is full:
// call ssehost_flush and retry.
int ssehost_on_message(ssehost* h, uint64_t conn_id, const uint8_t* data, uint32_t len) {
    if (!h) return SSEGPU_E_INVAL;
    auto it = h->conns.find(conn_id);
    if (it == h->conns.end()) return SSEGPU_E_INVAL;
    if (!h->cur) {
        int rc = ssegpu_batch_acquire(h->ctx, &h->cur);
        if (rc != SSEGPU_OK) return rc;
        h->n = 0; h->used = 0;
        ssegpu_in_off(h->cur)[0] = 0;
    }
    if (h->n >= h->max_records || h->used + len > h->max_in_bytes) return SSEGPU_E_AGAIN;
    memcpy(ssegpu_in_bytes(h->cur) + h->used, data, len);
    h->used += len;
    ssegpu_start_seq(h->cur)[h->n] = it->second.start_seq;
    h->owner[h->n] = conn_id;
    h->n++;
    ssegpu_in_off(h->cur)[h->n] = (uint32_t)h->used;
    return SSEGPU_OK;
}

// Runs the pending batch through the GPU and appends every connection's frames to its pending write.
int ssehost_flush(ssehost* h) {
    if (!h) return SSEGPU_E_INVAL;
    if (!h->cur || h->n == 0) return SSEGPU_OK;
    int rc = ssegpu_batch_submit(h->ctx, h->cur, h->n, SSEGPU_OP_ADAPTER | (h->header_only ? SSEGPU_OP_HEADER_ONLY : 0u), /*use_filter=*/1);
    if (rc == SSEGPU_OK) rc = ssegpu_batch_wait(h->ctx, h->cur, 0);
    if (rc == SSEGPU_OK) {
        const uint8_t* out = ssegpu_out_bytes(h->cur);
        const uint64_t* off = ssegpu_out_off(h->cur);
        const uint32_t* len = ssegpu_out_len(h->cur);
        const uint32_t* st = ssegpu_status(h->cur);
        for (uint32_t i = 0; i < h->n; i++) {
            Conn& c = h->conns[h->owner[i]];
            if (c.closed) { h->after_close++; continue; }     // handler already returned: the subscription is gone
            const uint32_t code = st[i] & SSEGPU_ST_CODE_MASK;
            if (code == SSEGPU_ST_FILTERED) { h->filtered++; continue; }          // sse_handler.go:143-145
            if (code != SSEGPU_ST_OK) { h->dropped_bad++; continue; }             // :137-140 log + drop
            c.pending.append(reinterpret_cast<const char*>(out + off[i]), len[i]);
            if (st[i] & SSEGPU_ST_HEADER_ONLY) {      // Go-marshalled record: its JSON is the payload, which never left this batch's pinned input
                const uint32_t* in_off = ssegpu_in_off(h->cur);
                c.pending.append(reinterpret_cast<const char*>(ssegpu_in_bytes(h->cur) + in_off[i]), in_off[i + 1] - in_off[i]);
                c.pending.append("\n\n", 2);
            }
            c.frames++;
            h->delivered++;                                                       // messagesDelivered.Inc(), :217
            if (st[i] & SSEGPU_ST_DONE) c.closed = true;                          // :220-226 return after the done frame
        }
    }
    ssegpu_batch_release(h->ctx, h->cur);
    h->cur = nullptr; h->n = 0; h->used = 0;
    return rc;
}

// The handler goroutine's write: everything queued for the connection leaves as ONE Write+Flush.
// Copies up to cap bytes into buf; *len gets the full pending size; *closed tells whether the handler returns
// after this write.  The pending buffer is cleared when it fitted.
int ssehost_take(ssehost* h, uint64_t conn_id, uint8_t* buf, uint64_t cap, uint64_t* len, int* closed) {
    if (!h || !len) return SSEGPU_E_INVAL;
    auto it = h->conns.find(conn_id);
    if (it == h->conns.end()) return SSEGPU_E_INVAL;
    Conn& c = it->second;
    *len = c.pending.size();
    if (closed) *closed = c.closed ? 1 : 0;
    if (c.pending.size() <= cap) {
        if (!c.pending.empty()) { memcpy(buf, c.pending.data(), c.pending.size()); c.writes++; }
        c.pending.clear();
    }
    return SSEGPU_OK;
}

// Egress without the bytes the host already has (SSEGPU_OP_HEADER_ONLY): the GPU returns only each frame's header, the
// coalescing loop above appends the payload from the pinned input slot and the blank line.
int ssehost_set_header_only(ssehost* h, int on) {
    if (!h) return SSEGPU_E_INVAL;
    h->header_only = on != 0;
    return SSEGPU_OK;
}

// counters: [0] delivered frames, [1] dropped (bad JSON / type), [2] filtered, [3] arrived after close
int ssehost_stats(ssehost* h, uint64_t out[4]) {
    if (!h || !out) return SSEGPU_E_INVAL;
    out[0] = h->delivered; out[1] = h->dropped_bad; out[2] = h->filtered; out[3] = h->after_close;
    return SSEGPU_OK;
}

}  // extern "C"
