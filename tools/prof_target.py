#!/usr/bin/env python3
"""Short single-kernel runs for ncu: a few launches of ONE mode of the frame kernel on a device-resident batch.

  python tools/prof_target.py inspect|inspect_ab|stream|match|bridge|ordered|plain [--launches K] [--conv N] [--tok T]

inspect     config-2 shape, Set A, per-token      (k_frame_tiles<true,0>)
inspect_ab  config-2 shape, Set A+B, per-token
stream      config-2 shape, Set A+B, streaming, whole batch in one launch
match       config-3 shape: bridge + interest table at 100 000 subjects (k_frame_tiles<true,0>)
bridge      config-3 shape without the table
ordered     config-1 shape, SSEGPU_OP_ORDERED (count pass, tile scan, emit pass)
plain       config-1 shape, arrival mode
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
shape = {"inspect": (100_000, 32), "inspect_ab": (100_000, 32), "stream": (100_000, 32), "match": (100_000, 16), "bridge": (100_000, 16),
         "ordered": (10_000, 1000), "plain": (10_000, 1000)}[a.what]
n_conv, n_tok = a.conv or shape[0], a.tok or shape[1]
arena, off = pkg.workload.generate(1, n_conv, n_tok)
n = n_conv * n_tok
eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=1 << 20, conv_slots=n_conv)
db = DevBatch(arena, off, extra_cap=96)
d_conv = 0
op = 0
if a.what in ("inspect", "inspect_ab", "stream"):
    eng.load_patterns(SET_A if a.what == "inspect" else SET_A + SET_B, pkg.INSPECT_STREAMING if a.what == "stream" else pkg.INSPECT_PER_TOKEN)
    op = pkg.OP_INSPECT
    if a.what == "stream":
        conv = torch.from_numpy(np.tile(np.arange(n_conv, dtype=np.int32), n_tok)).cuda()
        d_conv = conv.data_ptr()
elif a.what == "match":
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
