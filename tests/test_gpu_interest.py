"""GPU parity test of the subscriber/interest table through the C ABI, against the NATS subject-matching model;
bridge pipeline (BASELINE config 3 in miniature): decode -> subject match -> reframe, records nobody listens to are
not framed."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import interest_case  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import normalise, slices  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("generic", [True, False])
def test_interest_lookup_and_bridge_reframe(oracle, generic):
    pkg = load_package()
    deltas, live, payloads, subjects = interest_case.build(generic_wildcard=generic)
    arena, off = oracle.pack(payloads)
    n = len(off) - 1
    eng = pkg.SseGpu(device=0, max_records=1 << 13, max_in_bytes=4 << 20, ring_slots=1)
    try:
        half = len(deltas) // 2
        eng.interest_apply(deltas[:half])            # incremental: two applies must equal one
        eng.interest_apply(deltas[half:])
        b = eng.acquire()
        b.fill(arena, off)
        b.submit(n, pkg.OP_BRIDGE | pkg.OP_MATCH)
        b.wait()
        hit, status = b.sub_span.copy(), b.status.copy()
        out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), False)
        subj_len = b.subj_len.copy()
        b.release()
        n_int = interest_case.check(hit, status, eng.interest_subs(), eng.interest_wild_ids(), live, subjects)
        assert n_int > 1000
        # interested records carry exactly the oracle's bridge output; the others are empty
        want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
        got, want = slices(out, out_off), slices(want_out, want_off)
        for i in range(n):
            if int(status[i]) & 0xF == 0:
                assert got[i] == want[i] and int(subj_len[i]) == int(want_subj[i])
                assert got[i][: int(subj_len[i])] == subjects[i]
            else:
                assert got[i] == b""
        # unsubscribing everything leaves no interest at all
        eng.interest_apply([(1, i, s) for s, i in live])
        b = eng.acquire()
        b.fill(arena, off)
        b.submit(n, pkg.OP_BRIDGE | pkg.OP_MATCH)
        b.wait()
        assert int(b.out_off[n]) == 0 and not (b.status & 0xF == 0).any()
        b.release()
    finally:
        eng.close()
