/*
 * ssegpu.h -- C ABI of the B200 SSE hot-path library (libssegpu.so).
 *
 * The reference (ccie7599/distributed-sse-for-llm-response) has NO plugin, operator or FFI
 * interface: both services are built with CGO_ENABLED=0 (src/sse-adapter/Dockerfile:12,
 * src/redis-nats-bridge/Dockerfile:10).  This ABI is therefore cut at the narrowest seams
 * of the reference's own hot loops, and each entry point cites the Go code it replaces.
 * The cgo binding a maintainer adds is shown in INTEGRATION.md (and go/ssegpu/).
 *
 * Conventions
 *   - plain C types only; every function returns 0 on success or a negative ssegpu_err;
 *   - no callbacks into the host language, no retained host pointers: batch buffers are
 *     library-owned pinned allocations exposed as raw pointers (cgo: unsafe.Slice);
 *   - callable from any OS thread (cgo hops threads): every entry point selects its
 *     device itself and uses the context's own streams;
 *   - one context per GPU; conversations shard across contexts by
 *     ssegpu_shard(conversation_id) = fnv1a64(id) % n_gpus, no cross-GPU exchange;
 *   - there is NO CPU fallback: without a CUDA device ssegpu_init fails with SSEGPU_E_CUDA.
 *
 * Record status word (uint32 per record):
 *   low 4 bits   SSEGPU_ST_* code; a non-OK record has a zero-length output slice, which is
 *                the reference's "log and drop" (sse_handler.go:137-140, :143-145)
 *   SSEGPU_ST_DONE      decoded done==true on an emitted record: the adapter closes the
 *                stream after writing this frame (sse_handler.go:220-226, :436-442)
 */
#ifndef SSEGPU_H
#define SSEGPU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SSEGPU_ABI_VERSION 1

/* ---- error codes ------------------------------------------------------------------ */
typedef enum {
    SSEGPU_OK = 0,
    SSEGPU_E_INVAL = -1,      /* bad argument */
    SSEGPU_E_CUDA = -2,       /* CUDA runtime error / no device (see ssegpu_last_cuda_error) */
    SSEGPU_E_NOMEM = -3,      /* allocation failed */
    SSEGPU_E_AGAIN = -4,      /* ring full: no free batch slot */
    SSEGPU_E_TOOBIG = -5,     /* n_records / bytes exceed the configured capacity */
    SSEGPU_E_OVERFLOW = -6,   /* output arena too small for this batch (nothing truncated silently) */
    SSEGPU_E_STATE = -7,      /* batch used out of order (submit before acquire, ...) */
    SSEGPU_E_TIMEOUT = -8,    /* ssegpu_batch_wait timed out */
    SSEGPU_E_CONVSLOT = -9,   /* streaming inspection: a record's conversation slot is not below conv_slots */
    SSEGPU_E_PATTERN = -10    /* unsupported or malformed inspection pattern */
} ssegpu_err;

/* ---- per-record status ------------------------------------------------------------- */
#define SSEGPU_ST_OK            0u
#define SSEGPU_ST_ERR_SYNTAX    1u  /* json.Unmarshal syntax error: dropped (sse_handler.go:137-140) */
#define SSEGPU_ST_ERR_TYPE      2u  /* json.Unmarshal type error: dropped                        */
#define SSEGPU_ST_FILTERED      3u  /* Last-Event-ID filter (sse_handler.go:142-145)             */
#define SSEGPU_ST_INSPECT_DROP  4u  /* inline inspection verdict "drop" (sse_handler.go:200-204) */
#define SSEGPU_ST_NO_INTEREST   5u  /* SSEGPU_OP_MATCH: no subscription matches the record's subject    */
#define SSEGPU_ST_CODE_MASK     0xFu
#define SSEGPU_ST_DONE          0x100u
#define SSEGPU_ST_REDACTED      0x200u
#define SSEGPU_ST_HEADER_ONLY   0x400u /* SSEGPU_OP_HEADER_ONLY: the output slice holds only what precedes the payload */

/* ---- op flags for submit / launch -------------------------------------------------- */
#define SSEGPU_OP_ADAPTER   0u   /* SSE frame: event/id/data lines (sse_handler.go:211-214, :427-430) */
#define SSEGPU_OP_BRIDGE    1u   /* subject "chat.<cid>.tokens" + re-marshalled JSON (bridge main.go:160-161) */
#define SSEGPU_OP_GO122     2u   /* encode 0x08/0x0C as \b \f (Go >= 1.22) instead of Go 1.21's   */
#define SSEGPU_OP_NO_FAST   4u   /* testing: force every record through the general decoder */
#define SSEGPU_OP_INSPECT   16u  /* scan the decoded token text with the loaded pattern set; verdicts only (the
                                    reference's "async" placement, docs/security-inspection-patterns.md:84-140) */
#define SSEGPU_OP_INSPECT_INLINE 32u /* ... and act on them before framing (sse_handler.go:200-207): verdict drop
                                    suppresses the frame (status SSEGPU_ST_INSPECT_DROP), redact replaces the token
                                    text by "[REDACTED]" (status bit SSEGPU_ST_REDACTED), alert changes nothing */
#define SSEGPU_OP_MATCH     64u  /* look the record's subject chat.<cid>.tokens up in the interest table: per record a span of
                                    subscriber ids (ssegpu_sub_span); records nobody subscribed to are not framed
                                    (status SSEGPU_ST_NO_INTEREST) */
#define SSEGPU_OP_NATS_PUB  128u /* with SSEGPU_OP_BRIDGE: each record is a complete NATS client-protocol message as nats.go's
                                  * publish() writes it, "PUB <subject> <#bytes>\r\n<payload>\r\n" (what nc.Publish puts on the
                                  * wire for main.go:164); with SSEGPU_OP_ORDERED the whole arena is one write() to the NATS
                                  * connection.  subj_len[i] is still the subject's length (it starts 4 bytes into the record).
                                  * The host keeps nats.go's checks that concern the connection (MaxPayload, closed/draining). */
#define SSEGPU_OP_RESP      256u /* ssegpu_encode_launch only: each record is the RESP2 command go-redis writes for
                                  * rdb.Publish(ctx, channel, payload) (llm-stream-proxy/main.go:244):
                                  * "*3\r\n$7\r\npublish\r\n$<n>\r\n<channel>\r\n$<m>\r\n<payload>\r\n"; implies the channel. */
#define SSEGPU_OP_HEADER_ONLY 512u /* egress without the bytes the host already has.  A record that is exactly what json.Marshal writes
                                  * (the proxy's and the load generator's output, counted in fast_path) re-marshals to itself, so its
                                  * frame is header + the payload itself + trailer.  With this flag such a record's output slice is only the
                                  * header -- "event: token\nid: <seq>\ndata: " (bridge: the subject, with SSEGPU_OP_NATS_PUB the whole
                                  * "PUB <subject> <#bytes>\r\n" line) -- and its status carries SSEGPU_ST_HEADER_ONLY: the host writes
                                  * out[off:off+len] + payload (still in the batch's in_bytes) + "\n\n" (bridge: nothing, PUB: "\r\n").
                                  * Every other record (re-marshalled, redacted) comes back whole, without the status bit.  Cuts the
                                  * device-to-host traffic of a Go-marshalled stream from ~180 to ~45 bytes per record. */
#define SSEGPU_OP_ORDERED   8u   /* lay the frames out back to back in record order (the arena then equals the
                                    sequential Go output byte for byte); costs a second decode pass.  Default: frames
                                    land in completion order -- the Go-marshalled records of a tile of 128 consecutive
                                    records as one 16-byte aligned slice, records that needed the general decoder
                                    possibly elsewhere (<=15 zero bytes between slices); every frame is still
                                    out[out_off[i] .. out_off[i]+out_len[i]) */

typedef struct ssegpu_ctx ssegpu_ctx;
typedef struct ssegpu_batch ssegpu_batch;

typedef struct {
    int32_t  device;          /* CUDA ordinal */
    uint32_t max_records;     /* records per batch */
    uint64_t max_in_bytes;    /* payload bytes per batch */
    uint64_t max_out_bytes;   /* output bytes per batch; 0 = 2*max_in_bytes + 64*max_records */
    uint32_t ring_slots;      /* batches in flight (pinned host + device buffers each); 0 = none */
    uint32_t conv_slots;      /* per-conversation inspection state slots; 0 = none */
    uint32_t flags;           /* reserved, 0 */
} ssegpu_config;

/* Batch counters; the host sums them over GPUs into the reference's Prometheus metrics
 * (sse_messages_delivered_total, sse_handler.go:32-35, incremented per frame at :217, :433). */
typedef struct {
    uint64_t records, frames, bytes_in, bytes_out;
    uint64_t err_syntax, err_type, filtered, done;
    uint64_t fast_path;       /* records proven canonical (JSON out == payload) */
    uint64_t batches;
} ssegpu_counters;

/* ---- context ----------------------------------------------------------------------- */
/* Replaces nothing in the reference (which has no accelerator state); called once per GPU
 * from the service's main() next to connectNATS (src/sse-adapter/main.go:177-200). */
int  ssegpu_init(const ssegpu_config *cfg, ssegpu_ctx **out);
void ssegpu_destroy(ssegpu_ctx *ctx);
int  ssegpu_device_count(void);
const char *ssegpu_strerror(int err);
const char *ssegpu_last_cuda_error(void);
uint32_t ssegpu_abi_version(void);

/* fnv1a64(conversation_id) % n_gpus: which context owns a conversation (Go: hash/fnv New64a). */
uint32_t ssegpu_shard(const uint8_t *conversation_id, size_t len, uint32_t n_gpus);

/* fmt.Sscanf(lastEventID,"%d",&s); s++ (sse_handler.go:92-96): the per-connection
 * startSequence the host stores in ssegpu_start_seq() for every record of that connection. */
int64_t ssegpu_start_sequence(const char *last_event_id);

/* ---- pinned batch ring (end-to-end path) ------------------------------------------- */
/* Adapter ingress: the NATS callback (sse_handler.go:135-157, :316-333) appends msg.Data to the
 * current batch instead of calling json.Unmarshal; bridge ingress: msg.Payload
 * (redis-nats-bridge/main.go:142-157). */
int       ssegpu_batch_acquire(ssegpu_ctx *ctx, ssegpu_batch **out);
uint8_t  *ssegpu_in_bytes(ssegpu_batch *b);     /* max_in_bytes + 48 writable bytes          */
uint32_t *ssegpu_in_off(ssegpu_batch *b);       /* n+1 offsets into in_bytes, in_off[0] = 0   */
int64_t  *ssegpu_start_seq(ssegpu_batch *b);    /* n; 0 = filter off (ignored unless use_filter) */
uint32_t *ssegpu_conv_slot(ssegpu_batch *b);    /* n; inspection state slot of the conversation */

/* Starts H2D copy -> kernels -> D2H copy on the batch's own stream and returns at once. */
int ssegpu_batch_submit(ssegpu_ctx *ctx, ssegpu_batch *b, uint32_t n_records, uint32_t op_flags,
                        int use_filter);
/* Blocks until the results are in host memory.  timeout_us = 0 waits forever. */
int ssegpu_batch_wait(ssegpu_ctx *ctx, ssegpu_batch *b, uint32_t timeout_us);

/* Adapter egress: the handler goroutine writes out[out_off[i] : out_off[i]+out_len[i]] with one
 * w.Write + Flush instead of three Fprintf (sse_handler.go:212-215, :428-431); bridge egress:
 * nc.Publish(out[off:off+subj_len], out[off+subj_len:off+out_len]) (redis-nats-bridge/main.go:164).
 * out_off[n] is the total number of output bytes of the batch. */
const uint8_t  *ssegpu_out_bytes(const ssegpu_batch *b);
const uint64_t *ssegpu_out_off(const ssegpu_batch *b);   /* n+1 */
const uint32_t *ssegpu_out_len(const ssegpu_batch *b);   /* n   */
const uint32_t *ssegpu_status(const ssegpu_batch *b);    /* n   */
const uint32_t *ssegpu_subj_len(const ssegpu_batch *b);  /* n, bridge mode only */
int ssegpu_batch_release(ssegpu_ctx *ctx, ssegpu_batch *b);

/* Counters accumulated over all completed batches of this context (host-side gather). */
int ssegpu_counters_get(ssegpu_ctx *ctx, ssegpu_counters *out);

/* ---- content inspection ------------------------------------------------------------ */
/* The reference's inspection hook is a comment (sse_handler.go:193-208: callInspector(endpoint, token) ->
 * {Action, RedactedContent}); its pattern sets are snippets: docs/security-inspection-patterns.md:205-209, :229-234
 * ("Set A": 3 regexes -> redact, 4 literals -> alert) and src/spin-functions/nats-subscriber/src/lib.rs:59-64,
 * :79-84 ("Set B": 8 literals -> redact / drop).  Semantics (defined by this project): the DECODED
 * token text is scanned bytewise; \d [0-9], \w [0-9A-Za-z_], \b on ASCII word bytes; literals fold ASCII case.
 * kind 0 = literal, 1 = regex subset (atoms, classes, + * ? {n,m}, \b; no groups/alternation).  Up to 32 patterns.
 * mode 0 = per_token: every token is scanned on its own; the lowest-numbered matching pattern wins.
 * mode 1 = streaming: the DFA state is carried per conversation slot (ssegpu_conv_slot) across tokens and across
 *          batches, so patterns split over token boundaries are found; the first match to become decidable wins;
 *          a conversation may occur any number of times in a batch (its records are walked in arrival order);
 *          the done token resets the slot.  Streaming batches of one context execute in submit order.
 * verdict word: action (0 allow, 1 redact, 2 drop, 3 alert) | pattern index << 8 | SSEGPU_VERDICT_BEGIN_UNKNOWN.
 * match[2i], match[2i+1]: begin / end byte offsets into the decoded token text (begin may be negative in streaming
 * mode; INT32_MIN with SSEGPU_VERDICT_BEGIN_UNKNOWN for variable-length patterns whose start is not recoverable). */
typedef struct { uint32_t kind; uint32_t action; const char *text; } ssegpu_pattern;
#define SSEGPU_VERDICT_BEGIN_UNKNOWN 0x10000u
/* Safe while batches are in flight: the new tables become a new generation, batches already submitted finish on the old
 * one (freed when the last of them is released), later ones use the new one.  Restarts every conversation slot. */
int ssegpu_patterns_load(ssegpu_ctx *ctx, const ssegpu_pattern *patterns, uint32_t n, uint32_t mode);
int ssegpu_conv_reset(ssegpu_ctx *ctx);                        /* every conversation slot back to the start state */
const uint32_t *ssegpu_verdict(const ssegpu_batch *b);         /* n  */
const int32_t  *ssegpu_match(const ssegpu_batch *b);           /* 2n */
/* device buffers used by ssegpu_frame_launch when op_flags has SSEGPU_OP_INSPECT[_INLINE] */
int ssegpu_inspect_bind(ssegpu_ctx *ctx, const uint32_t *d_conv_slot, uint32_t *d_verdict, int32_t *d_match);

/* ---- subscriber / interest table ---------------------------------------------------- */
/* Replaces, for this path, the subscription matching that the reference delegates to nats.go / nats-server:
 * nc.Subscribe("chat.<id>.tokens", cb) per connection (sse_handler.go:133-135, :314-316), wildcard consumers such
 * as chat.*.tokens (docs/security-inspection-patterns.md:110,268).  NATS subject rules: '.'-separated tokens, '*' = one
 * token, '>' = the rest.  ssegpu_interest_apply adds (op 0) / removes (op 1) (subject, subscriber id) pairs.  The
 * reference does one such change per HTTP connect / disconnect (sse_handler.go:133-135, :166-169), so a call is cheap:
 * a literal subscription changes ONE slot of the device table in place (a small copy + kernel on the context's table
 * stream, ordered by events: after the batches in flight that read the table, before the batches submitted later --
 * no device-wide synchronisation, the call does not wait for the GPU); the table is rebuilt as a new generation only
 * when its slots are half full, its id arena is full, or the wildcard list changes.  Safe to call from any thread
 * while other threads submit: every batch sees exactly the deltas of the calls that returned before its submit
 * (ssegpu_batch_table_version = number of successful calls it saw).
 * With SSEGPU_OP_MATCH every record gets three words in ssegpu_sub_span(): [3i] offset and [3i+1] count of its subscriber
 * ids inside ssegpu_batch_subs(), [3i+2] bit mask over ssegpu_batch_wild_ids() of the wildcard subscriptions that
 * match.  At most 32 wildcard subscriptions. */
typedef struct { uint32_t op; uint32_t sub_id; const char *subject; uint32_t subject_len; } ssegpu_sub_delta;
int ssegpu_interest_apply(ssegpu_ctx *ctx, const ssegpu_sub_delta *deltas, uint32_t n);
uint64_t ssegpu_interest_version(ssegpu_ctx *ctx);                           /* successful applies so far */
/* The subscriber-id arena / wildcard ids of the table generation a batch was matched against: valid from wait to release
 * of that batch, whatever ssegpu_interest_apply does meanwhile. */
const uint32_t *ssegpu_batch_subs(const ssegpu_batch *b, uint32_t *count);
const uint32_t *ssegpu_batch_wild_ids(const ssegpu_batch *b, uint32_t *count);
uint64_t ssegpu_batch_table_version(const ssegpu_batch *b);
/* The same for the CURRENT generation (for ssegpu_frame_launch users): valid until the next ssegpu_interest_apply. */
const uint32_t *ssegpu_interest_subs(ssegpu_ctx *ctx, uint32_t *count);
const uint32_t *ssegpu_interest_wild_ids(ssegpu_ctx *ctx, uint32_t *count);
const uint32_t *ssegpu_sub_span(const ssegpu_batch *b);                      /* 3n */
int ssegpu_interest_bind(ssegpu_ctx *ctx, uint32_t *d_sub_span);             /* buffer used by ssegpu_frame_launch */
/* Fan-out expansion (BASELINE config 3/4: one Subscribe per connection, sse_handler.go:133-135, so a popular subject has
 * many): turns the spans of n records (d_sub_span as SSEGPU_OP_MATCH left it, device memory) into the flat list of
 * deliveries, d_pair_sub[k] = subscriber id, d_pair_rec[k] = record index, sorted by record (literal subscribers in span
 * order, then the matching wildcard subscriptions); d_rec_off[0..n] = first pair of every record.  One thread per pair, so
 * a record with 1 024 subscribers is as cheap per delivery as one with a single one.  Uses the CURRENT table generation:
 * call it before the next ssegpu_interest_apply that follows the match launch.  Asynchronous on stream. */
int ssegpu_fanout_launch(ssegpu_ctx *ctx, const uint32_t *d_sub_span, uint32_t n_records, uint32_t *d_pair_sub, uint32_t *d_pair_rec,
                         uint64_t max_pairs, uint32_t *d_rec_off, void *stream);
int ssegpu_fanout_result(ssegpu_ctx *ctx, uint64_t *n_pairs);               /* after sync; SSEGPU_E_OVERFLOW if max_pairs was too small */

/* ---- device-resident launch (kernel-level path, used by the bench and the parity tests) -- */
/* Same computation as submit, on buffers that already live in this context's GPU memory.
 * d_in and d_out must be 16-byte aligned, d_in with 32 readable bytes after the last payload byte.
 * d_subj_len may be NULL in adapter mode, d_start_seq NULL switches the filter off.
 * stream is a cudaStream_t (NULL = the context's stream).  Asynchronous. Not re-entrant per
 * context (the launch scratch is per context). */
int ssegpu_frame_launch(ssegpu_ctx *ctx, const void *d_in, const uint32_t *d_in_off,
                        const int64_t *d_start_seq, uint32_t n_records, uint32_t op_flags,
                        void *d_out, uint64_t out_cap, uint64_t *d_out_off, uint32_t *d_out_len,
                        uint32_t *d_subj_len, uint32_t *d_status, void *stream);
/* After the launch's stream has been synchronised: adds the launch's counters into *out and
 * returns SSEGPU_E_OVERFLOW if out_cap was too small. */
int ssegpu_frame_launch_result(ssegpu_ctx *ctx, ssegpu_counters *out);

/* ---- producer-side batch encode (row N3) ------------------------------------------------ */
/* Replaces, for a batch of n messages, publishToken() of src/llm-stream-proxy/main.go:230-245 (and the
 * load generator's producer, demo/load-generator/main.go:215-223): message i is
 * json.Marshal(TokenMessage{cid_i, tok_i, seq[i], done[i] != 0, ts[i]}) -- field order, escapeHTML and the
 * \ufffd replacement of invalid UTF-8 exactly as encoding/json -- written at d_out + d_out_off[i],
 * d_out_len[i] bytes.  cid_i / tok_i are the RAW string bytes d_cid[d_cid_off[i] .. d_cid_off[i+1]) and likewise
 * for tok; both arenas and d_out must be 16-byte aligned, the arenas with 32 readable bytes after their last byte.
 * With SSEGPU_OP_BRIDGE each record is preceded by its channel "chat.<cid>.tokens" (the PUBLISH
 * channel, main.go:240) and d_chan_len[i] (may be NULL otherwise) receives the channel's length.  SSEGPU_OP_GO122
 * selects the Go >= 1.22 spelling of \b and \f.  Records of one 128-message tile are contiguous and tiles
 * start on 16-byte boundaries in completion order (as in the default frame launch); d_out_off[n] = bytes used.
 * All pointers are device memory of this context's GPU.  Asynchronous on stream (NULL = the context's). */
int ssegpu_encode_launch(ssegpu_ctx *ctx, const void *d_cid, const uint32_t *d_cid_off,
                         const void *d_tok, const uint32_t *d_tok_off, const int64_t *d_seq,
                         const uint8_t *d_done, const int64_t *d_ts, uint32_t n_messages,
                         uint32_t op_flags, void *d_out, uint64_t out_cap, uint64_t *d_out_off,
                         uint32_t *d_out_len, uint32_t *d_chan_len, void *stream);
/* After the stream has been synchronised: bytes of d_out used; SSEGPU_E_OVERFLOW if out_cap was too small. */
int ssegpu_encode_launch_result(ssegpu_ctx *ctx, uint64_t *out_bytes);

/* ---- receiving side of the wire protocols (row N4): socket bytes -> the packed arena the frame launch reads ----------- */
/* Replaces, batched over every connection of the box, the read side of the two client libraries the reference uses:
 * nats.go's parser, which turns "MSG <subject> <sid> [reply] <#bytes>\r\n<payload>\r\n" into the msg.Data the adapter's
 * callback gets (sse_handler.go:135, :316), and go-redis's reader, which turns the RESP push
 * "*4 $8 pmessage $.. <pattern> $.. <channel> $.. <payload>" (or *3 message) into the msg.Payload of the bridge loop
 * (redis-nats-bridge/main.go:115, :127, :142).  Neither library is vendored: layouts restated from the protocol specs, PARITY
 * UNPINNED.  Connection c's unread bytes are d_buf[d_buf_off[c] .. d_buf_off[c+1]).  Every COMPLETE delivery's payload
 * is copied, in connection order, into d_arena (16-byte aligned; d_in_off[0..n_msgs] are its offsets: exactly the input
 * of ssegpu_frame_launch); per delivery d_msg_conn = connection index, d_msg_sid = NATS subscription id (0 for RESP),
 * d_msg_subj[2i], [2i+1] = offset (into d_buf) and length of the subject / channel.  Per connection d_conn_consumed =
 * bytes used (an incomplete trailing message stays in the socket buffer), d_conn_status = 0, or bit 0: protocol
 * violation at that offset, bit 1: a -ERR line was seen; d_conn_pings = PING lines (the host owes that many PONGs).
 * All pointers are device memory of this context's GPU; d_buf with 8 readable bytes behind the last connection's bytes
 * (payloads are copied word-wise).  Asynchronous on stream (NULL = the context's). */
#define SSEGPU_WIRE_NATS 0u
#define SSEGPU_WIRE_RESP 1u
int ssegpu_split_launch(ssegpu_ctx *ctx, uint32_t proto, const void *d_buf, const uint32_t *d_buf_off, uint32_t n_conn,
                        void *d_arena, uint64_t arena_cap, uint32_t *d_in_off, uint32_t max_msgs, uint32_t *d_msg_conn,
                        uint64_t *d_msg_sid, uint32_t *d_msg_subj, uint32_t *d_conn_consumed, uint32_t *d_conn_status,
                        uint32_t *d_conn_pings, void *stream);
/* After the stream has been synchronised: deliveries found and arena bytes used; SSEGPU_E_OVERFLOW (nothing written)
 * if max_msgs or arena_cap was too small. */
int ssegpu_split_result(ssegpu_ctx *ctx, uint32_t *n_msgs, uint64_t *arena_bytes);

/* ---- decode half of row N3: the proxy's read of the vLLM stream ---------------------------------------------------- */
/* Replaces, batched over the lines of many chats' response bodies, the loop body of streamFromLLM
 * (src/llm-stream-proxy/main.go:188-226): strings.TrimSpace, the "data: " filter, the "[DONE]" line, json.Unmarshal into
 * StreamChunk (main.go:45-53; a line that does not decode is skipped), then per choice: a non-empty delta.content is a
 * token message, a non-nil finish_reason ends the stream with the "[DONE]" sentinel.  Line i is d_lines[d_line_off[i] ..
 * d_line_off[i+1]).  Out: d_line_status[i] = kind (0 skipped, 1 chunk, 2 "[DONE]" line) | SSEGPU_CHUNK_ENDS (the reference
 * stops reading that chat after this line: the host drops the chat's later lines) | SSEGPU_CHUNK_TOO_MANY (more than 16
 * choices in one line: not decoded, told apart from a skip); the messages the lines publish, in line order: decoded token
 * text d_tok[d_tok_off[m] .. d_tok_off[m+1]) (16-byte aligned arena), d_done[m] (1 = sentinel), d_msg_line[m] -- d_tok,
 * d_tok_off and d_done are columns of ssegpu_encode_launch; conversation id, sequence (sequence++ per token, sentinel =
 * sequence + 1) and timestamp are per-chat state the host fills in.  PARITY UNPINNED (no test in the reference). */
#define SSEGPU_CHUNK_ENDS     0x100u
#define SSEGPU_CHUNK_TOO_MANY 0x200u
int ssegpu_chunk_decode_launch(ssegpu_ctx *ctx, const void *d_lines, const uint32_t *d_line_off, uint32_t n_lines, void *d_tok,
                               uint64_t tok_cap, uint32_t *d_tok_off, uint8_t *d_done, uint32_t *d_msg_line, uint32_t max_msgs,
                               uint32_t *d_line_status, void *stream);
/* After the stream has been synchronised: messages and token bytes; SSEGPU_E_OVERFLOW (nothing written) if too small. */
int ssegpu_chunk_decode_result(ssegpu_ctx *ctx, uint32_t *n_msgs, uint64_t *tok_bytes);

/* Number of kernel launches this context has issued so far (bench.py "gpu_launches"). */
uint64_t ssegpu_launch_count(ssegpu_ctx *ctx);

/* ---- launch without any allocation ------------------------------------------------- */
/* No counterpart in the reference (which has no accelerator code at all); added for callers that cannot
 * allocate GPU memory: a service sharing its GPU with an inference server that has already reserved nearly
 * all of HBM, or a latency-critical loop that cannot take the device-wide synchronisation of cudaMalloc.
 * Same computation as ssegpu_frame_launch, but context-free: the kernel's scratch (tile ticket, arena
 * cursor, counter rows, one word per tile) lives in a fixed array inside the library image
 * (SSEGPU_STATIC_SCRATCH_BYTES per GPU that loads the library, held until the process exits) and the call
 * performs no cudaMalloc / cudaFree / pinned allocation.  The caller supplies every input and output
 * buffer (same requirements as ssegpu_frame_launch).
 *   - n_records above SSEGPU_STATIC_MAX_RECORDS is refused with SSEGPU_E_TOOBIG (see ssegpu_static_last_error);
 *   - the first call on a device resolves the array's device address; make that first call outside any
 *     stream capture;
 *   - all calls on one device share the one array: it is cleared on `stream` at the start of every call, and
 *     two calls whose GPU work overlaps (different streams, or different threads) race on it -- serialise them;
 *   - there is no result call: out_off[n_records] holds the arena bytes needed, so the batch overflowed the
 *     output arena exactly when out_off[n_records] > out_cap after the stream has been synchronised; bit 63 of
 *     that word (SSEGPU_STATIC_BAD_OFFSETS) is set when d_in_off was not monotonic or ran past d_in_off[n_records]
 *     (such records are clamped into the arena, the context entries report the same as SSEGPU_E_INVAL);
 *   - SSEGPU_OP_INSPECT, SSEGPU_OP_INSPECT_INLINE and SSEGPU_OP_MATCH are refused with SSEGPU_E_INVAL: their tables
 *     belong to a context and this entry has none;
 *   - the text behind ssegpu_static_last_error (and ssegpu_last_cuda_error) is kept per OS thread: fetch it on the
 *     thread that made the failing call (cgo: runtime.LockOSThread around the pair).
 * d_in and d_out must be 16-byte aligned.  stream is a cudaStream_t (NULL = the legacy default stream).  Asynchronous. */
#define SSEGPU_STATIC_BAD_OFFSETS   (1ull << 63)
#define SSEGPU_STATIC_MAX_RECORDS   (1u << 20)
#define SSEGPU_STATIC_SCRATCH_BYTES (72u * 1024u)
int ssegpu_frame_launch_static(int device, const void *d_in, const uint32_t *d_in_off,
                               const int64_t *d_start_seq, uint32_t n_records, uint32_t op_flags,
                               void *d_out, uint64_t out_cap, uint64_t *d_out_off, uint32_t *d_out_len,
                               uint32_t *d_subj_len, uint32_t *d_status, void *stream);
const char *ssegpu_static_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
