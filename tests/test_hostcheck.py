"""CPU tests of the product's per-record logic (sse_record.cuh compiled for the host as a test
double) against the hand vectors, the committed fixtures and the C oracle."""
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hand_vectors as hv  # noqa: E402
import hostcheck_lib as hc  # noqa: E402
import make_golden as mg  # noqa: E402


@pytest.fixture(scope="module")
def lib():
    return hc.load()


def _slices(out, out_off):
    b = out.tobytes()
    return [b[int(out_off[i]):int(out_off[i + 1])] for i in range(len(out_off) - 1)]


def test_hand_vectors(lib, oracle):
    arena, off = oracle.pack([v[1] for v in hv.ADAPTER])
    start = np.array([v[2] for v in hv.ADAPTER], dtype=np.int64)
    for op in (0, hc.OP_NO_FAST):
        out, out_off, _, status, _ = hc.run(lib, arena, off, start, op)
        for (name, _, _, st, frame), g, s in zip(hv.ADAPTER, _slices(out, out_off), status):
            assert int(s) == st, name
            assert g == frame, name


@pytest.mark.parametrize("op", [0, hc.OP_NO_FAST, hc.OP_GO122, hc.OP_BRIDGE, hc.OP_BRIDGE | hc.OP_NO_FAST])
def test_golden_fixtures(lib, oracle, golden, op):
    arena, off = oracle.pack([bytes.fromhex(c["in"]) for c in golden])
    start = np.array([c["start"] for c in golden], dtype=np.int64)
    out, out_off, subj, status, nfast = hc.run(lib, arena, off, start, op)
    got = _slices(out, out_off)
    for c, g, sl, s in zip(golden, got, subj, status):
        if op & hc.OP_BRIDGE:
            assert int(s) == c["bstatus"], c["name"]
            assert g[: int(sl)].hex() == c["subject"] and g[int(sl):].hex() == c["json"], c["name"]
        else:
            assert int(s) == c["status"], c["name"]
            want = c.get("frame122", c["frame"]) if op & hc.OP_GO122 else c["frame"]
            assert g.hex() == want, c["name"]
    if not op & hc.OP_NO_FAST:
        assert nfast > 300          # the canonical path is actually exercised


def test_differential_fuzz_vs_oracle(lib, oracle):
    rng = random.Random(1234)
    base = mg.canon("7f3c2b1a-1111-4222-8333-944455556666", " hello", 42, False, 1700000000000000000)
    payloads = []
    for k in range(20000):
        r = rng.random()
        if r < 0.3:
            payloads.append(mg.canon(mg.rand_text(rng, rng.randrange(0, 40), 0.15), mg.rand_text(rng, rng.randrange(0, 20), 0.3),
                                     rng.randrange(-5, 5000), rng.random() < 0.05, rng.randrange(-(1 << 63), 1 << 63)))
        elif r < 0.6:
            payloads.append(mg.structured(rng))
        else:
            payloads.append(mg.mutate(rng, rng.choice([base, payloads[-1] if payloads else base])))
    arena, off = oracle.pack(payloads)
    start = np.array([rng.choice([0, 0, 3, 100]) for _ in payloads], dtype=np.int64)
    for op, go122 in ((0, False), (hc.OP_GO122, True)):
        want_out, want_off, want_status = oracle.adapter(arena, off, start, go122=go122)
        out, out_off, _, status, _ = hc.run(lib, arena, off, start, op)
        assert np.array_equal(status, want_status)
        assert np.array_equal(out_off, want_off)
        assert np.array_equal(out, want_out)
    want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
    out, out_off, subj, status, _ = hc.run(lib, arena, off, None, hc.OP_BRIDGE)
    assert np.array_equal(status, want_status) and np.array_equal(out_off, want_off)
    assert np.array_equal(subj, want_subj) and np.array_equal(out, want_out)


def test_nats_pub_framing(lib, oracle, golden):
    """SSEGPU_OP_NATS_PUB: bridge records as complete "PUB <subject> <#bytes>\\r\\n<payload>\\r\\n" messages
    (what nc.Publish writes for redis-nats-bridge/main.go:164)."""
    arena, off = oracle.pack([v[0] for v in hv.BRIDGE_PUB])
    for op in (hc.OP_BRIDGE | hc.OP_NATS_PUB, hc.OP_BRIDGE | hc.OP_NATS_PUB | hc.OP_NO_FAST):
        out, out_off, subj, status, _ = hc.run(lib, arena, off, None, op)
        assert _slices(out, out_off) == [v[2] for v in hv.BRIDGE_PUB]
        assert [int(s) for s in status] == [v[1] for v in hv.BRIDGE_PUB]
    o_out, o_off, o_subj, o_status = oracle.bridge_pub(arena, off)
    assert _slices(o_out, o_off) == [v[2] for v in hv.BRIDGE_PUB]
    # the committed fixtures: PUB framing == "PUB " + subject + " " + len + CRLF + json + CRLF of the plain bridge output
    arena, off = oracle.pack([bytes.fromhex(c["in"]) for c in golden])
    p_out, p_off, p_subj, p_status = oracle.bridge(arena, off)
    want = []
    for rec, sl, st in zip(_slices(p_out, p_off), p_subj, p_status):
        sl = int(sl)
        want.append(b"PUB " + rec[:sl] + b" %d\r\n" % (len(rec) - sl) + rec[sl:] + b"\r\n" if int(st) & 0xF == 0 else b"")
    for op in (hc.OP_BRIDGE | hc.OP_NATS_PUB, hc.OP_BRIDGE | hc.OP_NATS_PUB | hc.OP_NO_FAST, hc.OP_BRIDGE | hc.OP_NATS_PUB | hc.OP_GO122):
        go122 = bool(op & hc.OP_GO122)
        if go122:
            p_out, p_off, p_subj, p_status = oracle.bridge(arena, off, go122=True)
            want = []
            for rec, sl, st in zip(_slices(p_out, p_off), p_subj, p_status):
                sl = int(sl)
                want.append(b"PUB " + rec[:sl] + b" %d\r\n" % (len(rec) - sl) + rec[sl:] + b"\r\n" if int(st) & 0xF == 0 else b"")
        out, out_off, subj, status, _ = hc.run(lib, arena, off, None, op)
        assert _slices(out, out_off) == want
        assert (subj == p_subj).all() and (status == p_status).all()
        o_out, o_off, o_subj, o_status = oracle.bridge_pub(arena, off, go122=go122)
        assert _slices(o_out, o_off) == want and (o_subj == p_subj).all()
