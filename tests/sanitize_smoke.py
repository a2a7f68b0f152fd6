#!/usr/bin/env python3
"""Small all-modes run for compute-sanitizer (memcheck / racecheck): golden fixtures (adversarial JSON, oversize tiles),
a canonical workload, bridge, ordered, inspection (per-token + streaming + inline) and the interest table.
  compute-sanitizer --tool memcheck python tests/sanitize_smoke.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np  # noqa: E402
import interest_case  # noqa: E402
import oracle_lib  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import normalise  # noqa: E402
from inspect_sets import SET_A, SET_B  # noqa: E402

pkg = load_package()
oracle = oracle_lib.load()
golden = oracle_lib.load_golden()
payloads = [bytes.fromhex(c["in"]) for c in golden if len(c["in"]) < 60000][:1500]
a1, o1 = pkg.workload.generate(1, 300, 5)
payloads += [a1[o1[i]:o1[i + 1]].tobytes() for i in range(len(o1) - 1)]
arena, off = oracle.pack(payloads)
n = len(off) - 1
eng = pkg.SseGpu(device=0, max_records=8192, max_in_bytes=8 << 20, ring_slots=1, conv_slots=n)
eng.load_patterns(SET_A + SET_B, pkg.INSPECT_PER_TOKEN)
deltas, live, _, _ = interest_case.build(100)
eng.interest_apply(deltas)
want_out, want_off, want_status = oracle.adapter(arena, off)
for op in (0, pkg.OP_ORDERED, pkg.OP_NO_FAST, pkg.OP_BRIDGE, pkg.OP_INSPECT, pkg.OP_INSPECT_INLINE, pkg.OP_BRIDGE | pkg.OP_MATCH):
    b = eng.acquire()
    b.fill(arena, off, None, np.arange(n, dtype=np.uint32))
    b.submit(n, op)
    b.wait()
    out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), bool(op & pkg.OP_ORDERED))
    if op in (0, pkg.OP_ORDERED, pkg.OP_NO_FAST, pkg.OP_INSPECT):
        assert np.array_equal(out, want_out) and np.array_equal(b.status, want_status)
    b.release()
eng.load_patterns(SET_A, pkg.INSPECT_STREAMING)
b = eng.acquire()
b.fill(arena, off, None, np.arange(n, dtype=np.uint32))
b.submit(n, pkg.OP_INSPECT)
b.wait()
b.release()
eng.close()
print("sanitize_smoke ok: %d records x 8 modes" % n)
