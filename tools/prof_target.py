#!/usr/bin/env python3
"""Short single-kernel runs for ncu: a few launches of ONE mode of the frame kernel on a device-resident batch.

  python tools/prof_target.py inspect|inspect_ab|stream|match|fanout|bridge|ordered|plain|wire|chunk [--launches K] [--conv N] [--tok T]

inspect     config-2 shape, Set A, per-token      (k_frame_tiles<true,0>)
inspect_ab  config-2 shape, Set A+B, per-token
stream      config-2 shape, Set A+B, streaming, whole batch in one launch
match       config-3 shape: bridge + interest table at 100 000 subjects (k_frame_tiles<true,0>)
bridge      config-3 shape without the table
ordered     config-1 shape, SSEGPU_OP_ORDERED (count pass, tile scan, emit pass)
plain       config-1 shape, arrival mode
wire        100 000 NATS connections x 8 MSG deliveries of config-1 payloads: socket bytes -> packed arena (ssegpu_split_launch)
chunk       1 000 000 vLLM stream lines (one delta each) -> token columns (ssegpu_chunk_decode_launch)
Prints the mean CUDA-event time per launch; no checks (the parity tests do those).
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_package  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("what")
ap.add_argument("--launches", type=int, default=5)
ap.add_argument("--conv", type=int, default=0)
ap.add_argument("--tok", type=int, default=0)
a = ap.parse_args()

import torch  # noqa: E402
from bench_configs import DevBatch  # noqa: E402
from inspect_sets import SET_A, SET_B  # noqa: E402

pkg = load_package()
stream = torch.cuda.Stream()


def timed(fn, launches):
    evs = []
    for _ in range(3 + launches):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        evs.append((e0, e1))
    stream.synchronize()
    return [x.elapsed_time(y) for x, y in evs[3:]]


if a.what == "wire":
    n_conn, per = a.conv or 100_000, a.tok or 8
    arena, off = pkg.workload.generate(1, n_conn, per)
    raw = arena.tobytes()
    chunks, lens = [], []
    for c in range(n_conn):
        parts = []
        for j in range(per):
            i = j * n_conn + c
            p = raw[int(off[i]):int(off[i + 1])]
            parts.append(b"MSG chat.%s.tokens %d %d\r\n%s\r\n" % (p[20:56], c + 1, len(p), p))
        chunks.append(b"".join(parts))
        lens.append(len(chunks[-1]))
    buf = np.frombuffer(b"".join(chunks) + b"\0" * 64, dtype=np.uint8).copy()
    boff = np.zeros(n_conn + 1, dtype=np.uint32)
    boff[1:] = np.cumsum(lens)
    n_msgs = n_conn * per
    eng = pkg.SseGpu(device=0, max_records=n_msgs, max_in_bytes=1 << 20)
    d_buf, d_boff = torch.from_numpy(buf).cuda(), torch.from_numpy(boff.view(np.int32)).cuda()
    z = lambda m, dt: torch.zeros(m, dtype=dt, device="cuda")                              # noqa: E731
    cap = int(off[n_msgs]) + 4096
    d = [z(cap, torch.uint8), z(n_msgs + 1, torch.int32), z(n_msgs, torch.int32), z(n_msgs, torch.int64), z(2 * n_msgs, torch.int32),
         z(n_conn, torch.int32), z(n_conn, torch.int32), z(n_conn, torch.int32)]
    ms = timed(lambda: eng.split_launch(pkg.WIRE_NATS, d_buf.data_ptr(), d_boff.data_ptr(), n_conn, d[0].data_ptr(), cap, d[1].data_ptr(), n_msgs,
                                        d[2].data_ptr(), d[3].data_ptr(), d[4].data_ptr(), d[5].data_ptr(), d[6].data_ptr(), d[7].data_ptr(),
                                        stream.cuda_stream), a.launches)
    rc, n, used = eng.split_result()
    print("wire: rc=%d deliveries=%d socket bytes=%d payload bytes=%d  ms mean %.4f min %.4f  %.1f M deliveries/s, %.0f GB/s (socket bytes in + payload out)" % (
        rc, n, int(boff[-1]), used, float(np.mean(ms)), min(ms), n / min(ms) / 1e3, (int(boff[-1]) + used) / min(ms) / 1e6))
    eng.close()
    sys.exit(0)
if a.what == "chunk":
    import json
    n_lines = a.conv or 1_000_000
    cid, cid_off, tok, tok_off, seq, done, ts = pkg.workload.generate_columns(1, n_lines // 100, 100)
    traw = tok.tobytes()
    lines = []
    for i in range(n_lines):
        t = traw[int(tok_off[i]):int(tok_off[i + 1])].decode("utf-8")
        lines.append(b"data: " + json.dumps({"id": "cmpl-1", "object": "chat.completion.chunk", "created": 1700000000, "model": "mistral",
                                             "choices": [{"index": 0, "delta": {"content": t}, "logprobs": None, "finish_reason": None}]},
                                            separators=(",", ":")).encode() + b"\n")
    loff = np.zeros(n_lines + 1, dtype=np.uint32)
    loff[1:] = np.cumsum([len(x) for x in lines])
    buf = np.frombuffer(b"".join(lines) + b"\0" * 64, dtype=np.uint8).copy()
    eng = pkg.SseGpu(device=0, max_records=n_lines, max_in_bytes=1 << 20)
    d_buf, d_off = torch.from_numpy(buf).cuda(), torch.from_numpy(loff.view(np.int32)).cuda()
    z = lambda m, dt: torch.zeros(m, dtype=dt, device="cuda")                              # noqa: E731
    d = [z(32 << 20, torch.uint8), z(n_lines + 1, torch.int32), z(n_lines, torch.uint8), z(n_lines, torch.int32), z(n_lines, torch.int32)]
    ms = timed(lambda: eng.chunk_decode_launch(d_buf.data_ptr(), d_off.data_ptr(), n_lines, d[0].data_ptr(), 32 << 20, d[1].data_ptr(), d[2].data_ptr(),
                                               d[3].data_ptr(), n_lines, d[4].data_ptr(), stream.cuda_stream), a.launches)
    rc, n, used = eng.chunk_decode_result()
    print("chunk: rc=%d lines=%d messages=%d line bytes=%d token bytes=%d  ms mean %.4f min %.4f  %.1f M lines/s, %.0f GB/s of line bytes" % (
        rc, n_lines, n, int(loff[-1]), used, float(np.mean(ms)), min(ms), n_lines / min(ms) / 1e3, int(loff[-1]) / min(ms) / 1e6))
    eng.close()
    sys.exit(0)
shape = {"fanout": (100_000, 16), "inspect": (100_000, 32), "inspect_ab": (100_000, 32), "inspect_none": (100_000, 32), "plain32": (100_000, 32), "stream": (100_000, 32), "match": (100_000, 16), "bridge": (100_000, 16),
         "ordered": (10_000, 1000), "plain": (10_000, 1000)}[a.what]
n_conv, n_tok = a.conv or shape[0], a.tok or shape[1]
arena, off = pkg.workload.generate(1, n_conv, n_tok)
n = n_conv * n_tok
eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=1 << 20, conv_slots=n_conv)
db = DevBatch(arena, off, extra_cap=96)
d_conv = 0
op = 0
if a.what in ("inspect", "inspect_ab", "stream", "inspect_none"):
    # inspect_none: one pattern longer than any token, so that every walk is skipped -- what the variant costs before it walks a byte
    eng.load_patterns([(0, 3, b"a pattern that is longer than every token of the workload")] if a.what == "inspect_none" else SET_A if a.what == "inspect" else SET_A + SET_B, pkg.INSPECT_STREAMING if a.what == "stream" else pkg.INSPECT_PER_TOKEN)
    op = pkg.OP_INSPECT
    if a.what == "stream":
        conv = torch.from_numpy(np.tile(np.arange(n_conv, dtype=np.int32), n_tok)).cuda()
        d_conv = conv.data_ptr()
elif a.what in ("match", "fanout"):
    rng = np.random.default_rng(0x5EED0004)
    ids = [pkg.workload.conversation_id(1, i) for i in range(n_conv)]
    fan = np.maximum(1, (1024.0 / np.arange(1, n_conv + 1) ** 1.1)).astype(np.int64)
    rng.shuffle(fan)
    listened = rng.random(n_conv) >= 0.10
    deltas, sid = [], 1
    for i in range(n_conv):
        if listened[i]:
            for _ in range(int(fan[i])):
                deltas.append((0, sid, b"chat." + ids[i] + b".tokens"))
                sid += 1
    deltas += [(0, sid, b"chat.*.tokens"), (0, sid + 1, b"chat.*.control")]
    eng.interest_apply(deltas)
    op = pkg.OP_BRIDGE | pkg.OP_MATCH
elif a.what == "bridge":
    op = pkg.OP_BRIDGE
elif a.what == "ordered":
    op = pkg.OP_ORDERED

if a.what == "fanout":
    # config 3: the spans the match launch left -> flat (subscriber, record) delivery list
    db.launch(eng, op, stream, d_conv)
    stream.synchronize()
    hit = db.d_hit.cpu().numpy().astype(np.uint32).reshape(-1, 3)
    pairs = int(hit[:, 1].sum()) + int(sum(bin(int(m)).count("1") for m in np.unique(hit[:, 2])) * 0) + int(np.vectorize(lambda m: bin(int(m)).count("1"))(hit[:, 2]).sum())
    d_sub = torch.empty(pairs + 16, dtype=torch.int32, device="cuda")
    d_rec = torch.empty(pairs + 16, dtype=torch.int32, device="cuda")
    d_roff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    ms = timed(lambda: eng.fanout_launch(db.d_hit.data_ptr(), n, d_sub.data_ptr(), d_rec.data_ptr(), pairs + 16, d_roff.data_ptr(), stream.cuda_stream), a.launches)
    print("fanout: records=%d deliveries=%d  ms mean %.4f min %.4f  %.2f G deliveries/s, %.0f GB/s (12 B read per record + 8 B written per delivery)" % (
        n, pairs, float(np.mean(ms)), min(ms), pairs / min(ms) / 1e6, (12 * n + 8 * pairs) / min(ms) / 1e6))
    eng.close()
    sys.exit(0)
evs = []
for _ in range(3 + a.launches):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    db.launch(eng, op, stream, d_conv)
    e1.record(stream)
    evs.append((e0, e1))
stream.synchronize()
rc, c = eng.launch_result()
ms = [x.elapsed_time(y) for x, y in evs[3:]]
alg = db.in_bytes + c["bytes_out"] + (32 if op & (pkg.OP_INSPECT | pkg.OP_MATCH) else 16) * n
print("%s: rc=%d n=%d frames/launch=%d  ms mean %.4f min %.4f  alg %.0f GB/s" % (
    a.what, rc, n, c["frames"], float(np.mean(ms)), min(ms), alg / min(ms) / 1e6))
eng.close()
