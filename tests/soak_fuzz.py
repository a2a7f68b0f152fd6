"""One-off differential soak on a GPU box: many seeds of the test suite's payload fuzzers (canonical, structured,
mutated), every op mode, GPU (through the batch API) against the C oracle, bit-exact.  Test infrastructure, like tests/.
python tests/soak_fuzz.py [--seeds 10] [--records 200000]"""
import argparse
import os
import random
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # this file lives there: it uses the oracle, which only tests may
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_golden as mg  # noqa: E402
import oracle_lib  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import run_batch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--seeds", type=int, default=10)
ap.add_argument("--records", type=int, default=200000)
ap.add_argument("--first-seed", type=int, default=0)
a = ap.parse_args()
pkg = load_package()
oracle = oracle_lib.load()
eng = pkg.SseGpu(device=0, max_records=a.records, max_in_bytes=256 << 20, ring_slots=1)
base = mg.canon("7f3c2b1a-1111-4222-8333-944455556666", " hello", 42, False, 1700000000000000000)
t0 = time.time()
checked = 0
for seed in range(a.first_seed, a.first_seed + a.seeds):
    rng = random.Random(1000 + seed)
    mix = [(0.9, 0.05), (0.4, 0.3), (0.1, 0.3), (0.98, 0.01), (0.995, 0.0), (0.7, 0.0)][seed % 6]      # (canonical, structured) shares; the rest mutated
    payloads = []
    for k in range(a.records):
        r = rng.random()
        if r < mix[0]:
            payloads.append(mg.canon(mg.rand_text(rng, rng.randrange(0, 40), 0.1), mg.rand_text(rng, rng.randrange(0, 24), 0.2),
                                     rng.randrange(-5, 200000), rng.random() < 0.05, rng.randrange(-(1 << 63), 1 << 63)))
        elif r < mix[0] + mix[1]:
            payloads.append(mg.structured(rng))
        else:
            payloads.append(mg.mutate(rng, rng.choice([base, payloads[-1] if payloads else base])))
    arena, off = oracle.pack(payloads)
    start = np.array([rng.choice([0, 0, 3, 100]) for _ in payloads], dtype=np.int64)
    for op in (0, 2, 4, 8, 1, 1 | 128, 1 | 128 | 8):
        go122 = bool(op & 2)
        if op & 128:
            want_out, want_off, want_subj, want_status = oracle.bridge_pub(arena, off, go122=go122)
        elif op & 1:
            want_out, want_off, want_subj, want_status = oracle.bridge(arena, off, go122=go122)
        else:
            want_out, want_off, want_status = oracle.adapter(arena, off, start, go122=go122)
            want_subj = None
        out, out_off, status, subj = run_batch(eng, arena, off, None if op & 1 else start, op)
        assert np.array_equal(status, want_status), (seed, op, "status")
        assert np.array_equal(out_off, want_off), (seed, op, "offsets")
        assert np.array_equal(out, want_out), (seed, op, "bytes")
        if want_subj is not None:
            assert np.array_equal(subj, want_subj), (seed, op, "subject lengths")
        checked += len(payloads)
    print("seed %d mix %s ok (%.0f s)" % (seed, mix, time.time() - t0), flush=True)
print("soak ok: %d record checks" % checked)
eng.close()
