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
import pymodel  # noqa: E402
from gpu_util import normalise, run_batch, slices  # noqa: E402
from inspect_sets import SET_A, SET_B  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--seeds", type=int, default=10)
ap.add_argument("--records", type=int, default=200000)
ap.add_argument("--first-seed", type=int, default=0)
a = ap.parse_args()
pkg = load_package()
oracle = oracle_lib.load()
eng = pkg.SseGpu(device=0, max_records=a.records, max_in_bytes=256 << 20, ring_slots=1)
eng.load_patterns(SET_A + SET_B, pkg.INSPECT_PER_TOKEN)
insp = oracle.inspector(SET_A + SET_B)
INS, INL, MAT, ORD, HDR = pkg.OP_INSPECT, pkg.OP_INSPECT_INLINE, pkg.OP_MATCH, pkg.OP_ORDERED, pkg.OP_HEADER_ONLY


def run_ext(arena, off, start, op):
    """one batch through the ring with every result array copied out (frames normalised to record order)"""
    b = eng.acquire()
    b.fill(arena, off, start)
    b.submit(len(off) - 1, op, use_filter=start is not None)
    b.wait()
    out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), bool(op & ORD))
    r = dict(out=out, off=out_off, status=b.status.copy(), verdict=b.verdict.copy(), match=b.match.copy(), hit=b.sub_span.copy(), subs=b.subs(), wild=b.wild_ids())
    b.release()
    return r
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
    # ---- the optional stages, in both arena layouts, and the header-only egress (these variants had no soak) ----
    want_out, want_off, want_status = oracle.adapter(arena, off, start)
    want = slices(want_out, want_off)
    sample = range(0, len(payloads), 97)
    decoded = {i: pymodel.unmarshal(payloads[i]) for i in sample}
    base_v = None
    for op in (INS, INS | ORD, INL, INL | ORD):
        r = run_ext(arena, off, start, op)
        if base_v is None:
            base_v = (r["verdict"], r["match"])
            for i in sample:                                  # verdicts against the brute-force matcher on the decoded token
                st_i, m_i = decoded[i]
                if st_i == 0 and (int(want_status[i]) & 0xF) == 0:
                    assert (int(r["verdict"][i]), int(r["match"][i, 0]), int(r["match"][i, 1])) == insp.token(m_i["tok"].encode("utf-8")), (seed, op, i)
        assert np.array_equal(r["verdict"], base_v[0]) and np.array_equal(r["match"], base_v[1]), (seed, op, "verdicts differ between variants")
        got = slices(r["out"], r["off"])
        if not op & INL:
            assert np.array_equal(r["status"], want_status) and got == want, (seed, op, "verdict-only mode changed the frames")
        else:
            act = r["verdict"] & 0xFF
            ok = (want_status & 0xF) == 0
            assert np.array_equal(r["status"][~ok], want_status[~ok])
            assert np.array_equal(r["status"][ok & (act == 2)] & 0xF, np.full(int((ok & (act == 2)).sum()), 4, dtype=np.uint32))
            assert (r["status"][ok & (act == 1)] & 0x200).all()
            for i in np.nonzero(ok & ((act == 0) | (act == 3)))[0][::13]:
                assert got[int(i)] == want[int(i)], (seed, op, int(i))
            for i in np.nonzero(ok & (act == 2))[0]:
                assert got[int(i)] == b"", (seed, op, int(i))
        checked += len(payloads)
    # interest table over a sample of the batch's own (decoded) conversation ids
    cids = sorted({m_i["cid"] for st_i, m_i in decoded.values() if st_i == 0 and m_i["cid"] and all(ch.isalnum() or ch in "_-" for ch in m_i["cid"])})[:1500]   # ("*", ">", "a.b" would be wildcards / several tokens)
    eng.interest_apply([(0, k + 1, b"chat." + c.encode("utf-8") + b".tokens") for k, c in enumerate(cids)])
    b_out, b_off, b_subj, b_status = oracle.bridge(arena, off)
    b_want = slices(b_out, b_off)
    base_h = None
    for op in (1 | MAT, 1 | MAT | ORD):
        r = run_ext(arena, off, None, op)
        if base_h is None:
            base_h = r["hit"]
            idx = {c: k + 1 for k, c in enumerate(cids)}
            for i in sample:
                st_i, m_i = decoded[i]
                if st_i == 0 and (int(b_status[i]) & 0xF) == 0:
                    o, c, _ = (int(x) for x in r["hit"][i])
                    assert r["subs"][o:o + c].tolist() == ([idx[m_i["cid"]]] if m_i["cid"] in idx else []), (seed, op, i)
        assert np.array_equal(r["hit"][:, 1], base_h[:, 1]), (seed, op)
        got = slices(r["out"], r["off"])
        keep = (r["status"] & 0xF) == 0
        for i in range(0, len(payloads), 7):
            assert got[i] == (b_want[i] if keep[i] else b""), (seed, op, i)
        checked += len(payloads)
    eng.interest_apply([(1, k + 1, b"chat." + c.encode("utf-8") + b".tokens") for k, c in enumerate(cids)])
    for op in (0, ORD, 1):
        r = run_ext(arena, off, None if op & 1 else start, op | HDR)
        got = slices(r["out"], r["off"])
        w = b_want if op & 1 else want
        for i in range(0, len(payloads), 5):
            tail = payloads[i] + (b"" if op & 1 else b"\n\n") if int(r["status"][i]) & 0x400 else b""
            assert got[i] + tail == w[i], (seed, op, i)
        checked += len(payloads)
    print("seed %d mix %s ok (%.0f s)" % (seed, mix, time.time() - t0), flush=True)
print("soak ok: %d record checks" % checked)
eng.close()
