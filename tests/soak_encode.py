"""One-off differential soak of the producer-side encoder on a GPU box: seeded hostile/ordinary message mixes, every op mode,
GPU (ssegpu_encode_launch) against the C oracle, byte for byte.  python tests/soak_encode.py [--seeds 10] [--messages 100000]"""
import argparse
import os
import random
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import encode_case as ec  # noqa: E402
import oracle_lib  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from test_gpu_encode import gpu_encode  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--seeds", type=int, default=10)
ap.add_argument("--messages", type=int, default=100000)
a = ap.parse_args()
pkg = load_package()
oracle = oracle_lib.load()
eng = pkg.SseGpu(device=0, max_records=a.messages, max_in_bytes=64 << 20, ring_slots=1)
t0 = time.time()
checked = 0
for seed in range(a.seeds):
    rng = random.Random(7000 + seed)
    msgs = ec.random_messages(a.messages, seed=7000 + seed)
    if seed % 3 == 1:                                    # long strings: more than one 32-byte trip, tiles past the staging caps
        for k in range(0, len(msgs), 97):
            c, t, s, d, ts = msgs[k]
            msgs[k] = (c * rng.choice([1, 9]), t * rng.choice([3, 40]), s, d, ts)
    cols = ec.columns(msgs)
    for op in (0, 2, 1, 256, 256 | 2):
        got, chan, rc = gpu_encode(eng, cols, op)
        assert rc == 0
        if op & 256:
            want, want_chan = oracle.encode_resp(cols, go122=bool(op & 2))
        else:
            want, want_chan = oracle.encode(cols, go122=bool(op & 2), with_channel=bool(op & 1))
        assert got == want, (seed, op)
        assert np.array_equal(chan, want_chan), (seed, op)
        checked += len(msgs)
    print("seed %d ok (%.0f s)" % (seed, time.time() - t0), flush=True)
print("encoder soak ok: %d message checks" % checked)
eng.close()
