/*
 * sse_oracle.h -- CPU restatement of the reference's per-token hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (include/, the CUDA library,
 * the host layer) may include, link or call this.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs use it, and only as the checker
 * or as the timed CPU baseline.
 *
 * PARITY STATUS
 *   decode / filter / encode / frame / bridge subject:  PARITY UNPINNED.
 *     The arithmetic of this path lives in the Go standard library (encoding/json,
 *     fmt; toolchain pinned to golang:1.21 by src/sse-adapter/Dockerfile:2), which is
 *     not vendored in /root/reference and no Go toolchain exists in this image.  The
 *     reference holds no test or golden vector for the path.  This file restates the
 *     published algorithm of encoding/json 1.21 (scanner.go, decode.go, encode.go,
 *     fold.go) and is anchored on the reference's call sites cited per function.
 *   inspection Set B:  pinned by the reference's three known-answer tests
 *     (src/spin-functions/nats-subscriber/src/lib.rs:108-124).
 *   inspection Set A, interest matching:  PARITY UNPINNED (doc snippets / third-party).
 */
#ifndef SSE_ORACLE_H
#define SSE_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* per-record status word; the numeric values are part of the product ABI too
 * (include/ssegpu.h) and a test asserts that the two headers agree. */
#define SSEO_OK            0u
#define SSEO_ERR_SYNTAX    1u   /* json.Unmarshal syntax error  -> message dropped   */
#define SSEO_ERR_TYPE      2u   /* json.Unmarshal type error    -> message dropped   */
#define SSEO_FILTERED      3u   /* /stream Last-Event-ID filter -> message skipped   */
#define SSEO_INSPECT_DROP  4u   /* inline inspection said drop  -> message skipped   */
#define SSEO_NO_INTEREST   5u   /* no subscription matches the subject               */
#define SSEO_CODE_MASK     0xFu
#define SSEO_DONE          0x100u /* decoded done==true and the record was emitted    */
#define SSEO_REDACTED      0x200u /* inline inspection replaced the token text       */

/* TokenMessage (src/sse-adapter/sse_handler.go:46-52) after json.Unmarshal. */
typedef struct {
    const uint8_t *cid; uint32_t cid_len;   /* ConversationID, decoded UTF-8 */
    const uint8_t *tok; uint32_t tok_len;   /* Token, decoded UTF-8          */
    int64_t seq;                            /* Sequence                      */
    int64_t ts;                             /* Timestamp                     */
    int     done;                           /* Done                          */
} sseo_msg;

/* json.Unmarshal(data, &TokenMessage{})  (sse_handler.go:136-140, :317-321;
 * redis-nats-bridge/main.go:153-157).  scratch must hold 6*n+16 bytes; the decoded
 * strings point into it.  Returns SSEO_OK, SSEO_ERR_SYNTAX or SSEO_ERR_TYPE. */
int sseo_unmarshal(const uint8_t *data, size_t n, sseo_msg *m, uint8_t *scratch);

/* json.Marshal(token) (sse_handler.go:211, :427; bridge main.go:161).
 * go122!=0 selects the Go>=1.22 spelling of 0x08/0x0C ("\b","\f").  Returns length. */
size_t sseo_marshal(const sseo_msg *m, uint8_t *dst, int go122);
size_t sseo_marshal_bound(const sseo_msg *m);

/* Producer side (llm-stream-proxy/main.go:230-245, demo/load-generator/main.go:215-223): n messages from columns,
 * packed back to back; returns bytes written or (size_t)-1 when out_cap is too small. */
size_t sseo_encode_batch(const uint8_t *cid, const uint32_t *cid_off, const uint8_t *tok, const uint32_t *tok_off,
                         const int64_t *seq, const uint8_t *done, const int64_t *ts, uint32_t n, int go122, int with_channel,
                         uint8_t *out, size_t out_cap, uint64_t *off, uint32_t *chan_len);

/* json.Unmarshal every payload into the encoder's columns (tests, encoder bench). 0 = ok. */
int64_t sseo_decode_columns(const uint8_t *in, const uint32_t *in_off, uint32_t n,
                            uint8_t *cid, size_t cid_cap, uint32_t *cid_off, uint8_t *tok, size_t tok_cap, uint32_t *tok_off,
                            int64_t *seq, uint8_t *done, int64_t *ts);

/* Bridge records as NATS client-protocol messages "PUB <subject> <#bytes>\r\n<payload>\r\n" (what nc.Publish writes
 * for redis-nats-bridge/main.go:164; nats.go v1.31.0, not vendored -- stated from the NATS protocol). */
int sseo_bridge_pub_batch(const uint8_t *in, const uint32_t *in_off, uint32_t n, int go122,
                          uint8_t *out, uint64_t out_cap, uint64_t *out_off, uint32_t *subj_len, uint32_t *status);
/* Producer records as the RESP2 PUBLISH command go-redis v9.3.0 writes (llm-stream-proxy/main.go:244; not vendored). */
size_t sseo_encode_resp_batch(const uint8_t *cid, const uint32_t *cid_off, const uint8_t *tok, const uint32_t *tok_off,
                              const int64_t *seq, const uint8_t *done, const int64_t *ts, uint32_t n, int go122,
                              uint8_t *out, size_t out_cap, uint64_t *off, uint32_t *chan_len);

/* The three Fprintf calls (sse_handler.go:212-214, :428-430). Returns frame length. */
size_t sseo_frame(const sseo_msg *m, uint8_t *dst, int go122);

/* fmt.Sscanf(lastEventID,"%d",&s); s++  (sse_handler.go:92-96).  id==NULL or "" -> 0. */
int64_t sseo_start_sequence(const char *last_event_id);

/* Whole-batch adapter path: decode -> filter -> encode -> frame, frames packed
 * back to back.  in_off has n+1 entries; start_seq may be NULL (filter off) and holds
 * the per-record startSequence (0 = off).  out_off gets n+1 entries.  A dropped or
 * filtered record has a zero-length slice.  Returns 0, or -1 if out_cap is too small. */
int sseo_adapter_batch(const uint8_t *in, const uint32_t *in_off, uint32_t n,
                       const int64_t *start_seq, int go122,
                       uint8_t *out, uint64_t out_cap, uint64_t *out_off,
                       uint32_t *status);

/* Whole-batch bridge path (redis-nats-bridge/main.go:153-164): per record the slice is
 * subject "chat.<cid>.tokens" immediately followed by the re-marshalled JSON;
 * subj_len[i] gives the split point. */
int sseo_bridge_batch(const uint8_t *in, const uint32_t *in_off, uint32_t n, int go122,
                      uint8_t *out, uint64_t out_cap, uint64_t *out_off,
                      uint32_t *subj_len, uint32_t *status);

/* Timed CPU baseline: runs the adapter path over the batch on `threads` pthreads
 * (contiguous record ranges, private output arenas), `reps` times; returns the best
 * wall-clock seconds of one repetition; *frames / *out_bytes are per repetition. */
double sseo_bench_adapter(const uint8_t *in, const uint32_t *in_off, uint32_t n,
                          int threads, int reps, uint64_t *frames, uint64_t *out_bytes);
/* the same with buffers allocated and touched once, one timed pass per call (bench.py --impl reference) */
typedef struct sseo_bench sseo_bench;
sseo_bench *sseo_bench_new(const uint8_t *in, const uint32_t *in_off, uint32_t n, int threads);
double sseo_bench_run(sseo_bench *b, uint64_t *frames, uint64_t *out_bytes);
void sseo_bench_free(sseo_bench *b);

/* ---- content inspection (oracle/sse_inspect_oracle.c; semantics in its header comment) ---- */
typedef struct { uint32_t kind;   /* 0 literal, ASCII case-insensitive; 1 regex subset */
                 uint32_t action; /* 1 redact, 2 drop, 3 alert */
                 const char *text; } sseo_pattern;
typedef struct sseo_inspector sseo_inspector;
sseo_inspector *sseo_inspect_new(const sseo_pattern *pats, uint32_t n);   /* NULL: unsupported pattern */
void sseo_inspect_free(sseo_inspector *I);
/* verdict = action | pattern_id<<8 | (begin unknown ? 1<<16 : 0); 0 = allow */
void sseo_inspect_token(const sseo_inspector *I, const uint8_t *text, uint32_t len,
                        uint32_t *verdict, int32_t *begin, int32_t *end);
void sseo_inspect_stream(const sseo_inspector *I, const uint8_t *stream, uint32_t s, uint32_t t, int done,
                         uint32_t *verdict, int32_t *begin, int32_t *end);

/* fnv1a64 of the conversation id: the shard function. */
uint64_t sseo_fnv1a64(const uint8_t *p, size_t n);

/* ---- receiving side of the wire protocols (oracle/sse_wire_oracle.c; row N4, parity unpinned) ---- */
typedef struct { uint32_t payload_off, payload_len, subj_off, subj_len; uint64_t sid; uint32_t conn, pad; } sseo_wire_msg;
typedef struct { uint32_t n_msgs, payload_bytes, consumed, status, pings; } sseo_wire_conn;
uint32_t sseo_wire_split(uint32_t proto, const uint8_t *buf, const uint32_t *buf_off, uint32_t n_conn, sseo_wire_msg **msgs,
                         uint32_t *unused, sseo_wire_conn *conns);
void sseo_free(void *p);

/* ---- row N3, decode half: vLLM stream lines -> token messages (src/llm-stream-proxy/main.go:188-226; parity unpinned) ---- */
uint32_t sseo_chunk_decode(const uint8_t *lines, const uint32_t *line_off, uint32_t n_lines, uint8_t *tok, uint32_t *tok_off,
                           uint8_t *done, uint32_t *msg_line, uint32_t *line_status);

#ifdef __cplusplus
}
#endif
#endif
