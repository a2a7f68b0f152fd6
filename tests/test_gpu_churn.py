"""GPU test of subscription churn: one thread applies add/del deltas as fast as it can -- the reference changes one
subscription per HTTP connect and disconnect (nc.Subscribe / Unsubscribe, src/sse-adapter/sse_handler.go:133-135, :166-169)
-- while two threads submit SSEGPU_OP_MATCH batches on the same context.  Every batch must see exactly the table of the
version it reports (ssegpu_batch_table_version): the deltas are a function of the call number, so the expected subscriber
set of any subject at any version is computable.  Also covered: the in-place path, rebuilds forced by growth and by a
wildcard subscription arriving mid-run, and that a batch's id arena stays readable after later updates and rebuilds."""
import os
import sys
import threading
import time

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as mg  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pytestmark = pytest.mark.gpu

S = 300                 # subjects the batches ask about
BASE = 1_000_000
WILD_AT = 400           # churn call that adds the wildcard consumer
N_REC = 2000


def subject(j):
    return b"chat.churn-%04d.tokens" % j


def churn_call(c):
    """deltas of churn call c: a new subscriber for subject c % S, the one added 2S calls earlier leaves, now and then a
    brand-new subject appears (growth -> rebuild), once a wildcard consumer arrives (descriptor change -> rebuild)"""
    d = [(0, BASE + c, subject(c % S))]
    if c >= 2 * S:
        d.append((1, BASE + c - 2 * S, subject(c % S)))
    if c % 7 == 0:
        d.append((0, 5, b"chat.extra-%06d.tokens" % c))
    if c == WILD_AT:
        d.append((0, 7, b"chat.*.tokens"))
    return d


def expected_ids(j, m):
    """subscriber ids of subject j after the initial apply and m churn calls"""
    ids = [j + 1]
    lo = max(0, m - 2 * S)
    c = lo + ((j - lo) % S)
    while c < m:
        ids.append(BASE + c)
        c += S
    return sorted(ids)


def test_updates_race_submits_and_every_batch_sees_one_version(oracle):
    pkg = load_package()
    eng = pkg.SseGpu(device=0, max_records=4096, max_in_bytes=1 << 20, ring_slots=4)
    stop = threading.Event()
    errors, checked, versions_seen = [], [0], set()
    try:
        eng.interest_apply([(0, j + 1, subject(j)) for j in range(S)])          # version 1
        assert eng.interest_version() == 1
        calls_done = [0]

        def applier():
            try:
                c = 0
                while not stop.is_set():
                    eng.interest_apply(churn_call(c))
                    c += 1
                    calls_done[0] = c
            except Exception as e:                                                # noqa: BLE001
                errors.append(("applier", repr(e)))

        def submitter(seed):
            try:
                rng = np.random.default_rng(seed)
                while not stop.is_set():
                    subj = rng.integers(0, S + 20, size=N_REC)                    # a few subjects nobody ever subscribed to
                    payloads = [mg.canon("churn-%04d" % j, " t", k, False, 5) for k, j in enumerate(subj)]
                    arena, off = oracle.pack(payloads)
                    b = eng.acquire()
                    b.fill(arena, off)
                    b.submit(N_REC, pkg.OP_BRIDGE | pkg.OP_MATCH)
                    b.wait()
                    m = b.table_version() - 1                                     # churn calls this batch saw
                    hit, status, subs, wild = b.sub_span.copy(), b.status.copy(), b.subs(), b.wild_ids()
                    time.sleep(0.002)                                             # later updates must not disturb what this batch reads
                    subs2 = b.subs()
                    assert np.array_equal(subs, subs2[: len(subs)])
                    b.release()
                    versions_seen.add(m)
                    want_wild = [7] if m > WILD_AT else []
                    for i in range(0, N_REC, 7):
                        j = int(subj[i])
                        o, c, mask = (int(x) for x in hit[i])
                        got = sorted(subs[o:o + c].tolist())
                        got_w = [int(wild[w]) for w in range(len(wild)) if mask >> w & 1]
                        want = expected_ids(j, m) if j < S else []
                        if got != want or got_w != want_wild or (int(status[i]) & 0xF) != (0 if (want or want_wild) else 5):
                            errors.append(("batch", m, j, got[:6], want[:6], got_w, want_wild, int(status[i])))
                            stop.set()
                            return
                    checked[0] += 1
            except Exception as e:                                                # noqa: BLE001
                errors.append(("submitter", repr(e)))
                stop.set()

        threads = [threading.Thread(target=applier)] + [threading.Thread(target=submitter, args=(s,)) for s in (1, 2)]
        for t in threads:
            t.start()
        time.sleep(4.0)
        stop.set()
        for t in threads:
            t.join(timeout=60)
        assert not errors, errors[:3]
        assert checked[0] >= 10 and len(versions_seen) >= 5, (checked, len(versions_seen))
        assert calls_done[0] > WILD_AT + 50, calls_done            # the run got past the wildcard rebuild
        rate = calls_done[0] / 4.0
        print("churn: %d update calls (%.0f/s) beside %d checked batches, %d distinct versions seen" % (calls_done[0], rate, checked[0], len(versions_seen)))
    finally:
        stop.set()
        eng.close()


def test_patterns_reload_while_batches_are_in_flight(oracle):
    """ssegpu_patterns_load swaps generations without a device-wide synchronisation: a batch submitted before the reload
    finishes on the old tables, the next one uses the new ones."""
    from inspect_sets import SET_A, SET_B
    pkg = load_package()
    eng = pkg.SseGpu(device=0, max_records=1 << 14, max_in_bytes=4 << 20, ring_slots=4)
    try:
        payloads = [mg.canon("c%d" % i, "my password is x" if i % 3 == 0 else " hello 123-45-6789", i, False, 5) for i in range(9000)]
        arena, off = oracle.pack(payloads)
        eng.load_patterns(SET_B, pkg.INSPECT_PER_TOKEN)
        slots = []
        for k in range(3):
            b = eng.acquire()
            b.fill(arena, off)
            b.submit(len(payloads), pkg.OP_INSPECT)
            slots.append(b)
            if k == 0:
                eng.load_patterns(SET_A, pkg.INSPECT_PER_TOKEN)          # while batch 0 may still be running
        for k, b in enumerate(slots):
            b.wait()
            v = b.verdict.copy()
            if k == 0:      # Set B: the password records hit (redact = 1, pattern 0), the SSN records do not
                assert (v[0::3] == 1).all() and (v[1::3] == 0).all()
            else:           # Set A: the SSN records hit (redact, pattern 0), the password records do not
                assert (v[1::3] == 1).all() and (v[0::3] == 0).all()
            b.release()
    finally:
        eng.close()
