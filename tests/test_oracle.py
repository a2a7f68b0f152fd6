"""CPU tests: the C oracle against the hand-typed vectors, the committed fixtures and the Python model."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hand_vectors as hv  # noqa: E402
import pymodel  # noqa: E402


def _frames(out, out_off):
    b = out.tobytes()
    return [b[int(out_off[i]):int(out_off[i + 1])] for i in range(len(out_off) - 1)]


def test_hand_vectors_c_oracle(oracle):
    payloads = [v[1] for v in hv.ADAPTER]
    arena, off = oracle.pack(payloads)
    start = np.array([v[2] for v in hv.ADAPTER], dtype=np.int64)
    out, out_off, status = oracle.adapter(arena, off, start)
    got = _frames(out, out_off)
    for (name, _, _, st, frame), g, s in zip(hv.ADAPTER, got, status):
        assert int(s) == st, name
        assert g == frame, name


def test_hand_vectors_python_model():
    for name, payload, start, st, frame in hv.ADAPTER:
        s, f = pymodel.adapter(payload, start)
        assert (s, f) == (st, frame), name


def test_go122_switch(oracle):
    payload = [v[1] for v in hv.ADAPTER if v[0] == "G5_escapes_go121"]
    arena, off = oracle.pack(payload)
    out, out_off, _ = oracle.adapter(arena, off, None, go122=True)
    assert _frames(out, out_off)[0] == hv.G5_GO122_FRAME
    assert pymodel.adapter(payload[0], 0, go122=True)[1] == hv.G5_GO122_FRAME


def test_hand_bridge(oracle):
    arena, off = oracle.pack([b[0] for b in hv.BRIDGE])
    out, out_off, subj_len, status = oracle.bridge(arena, off)
    got = _frames(out, out_off)
    for (payload, st, subj, js), g, sl, s in zip(hv.BRIDGE, got, subj_len, status):
        assert int(s) == st
        assert g[: int(sl)] == subj and g[int(sl):] == js
        assert pymodel.bridge(payload) == (st, subj, js)


def test_start_sequence(oracle):
    for s, want in hv.START_SEQUENCE:
        assert oracle.start_sequence(s) == want, s
        assert pymodel.start_sequence(s) == want, s


def test_golden_fixtures_c_oracle(oracle, golden):
    payloads = [bytes.fromhex(c["in"]) for c in golden]
    arena, off = oracle.pack(payloads)
    start = np.array([c["start"] for c in golden], dtype=np.int64)
    out, out_off, status = oracle.adapter(arena, off, start)
    got = _frames(out, out_off)
    for c, g, s in zip(golden, got, status):
        assert int(s) == c["status"], c["name"]
        assert g.hex() == c["frame"], c["name"]
    out, out_off, _ = oracle.adapter(arena, off, start, go122=True)
    for c, g in zip(golden, _frames(out, out_off)):
        assert g.hex() == c.get("frame122", c["frame"]), c["name"]
    out, out_off, subj_len, status = oracle.bridge(arena, off)
    for c, g, sl, s in zip(golden, _frames(out, out_off), subj_len, status):
        assert int(s) == c["bstatus"], c["name"]
        assert g[: int(sl)].hex() == c["subject"] and g[int(sl):].hex() == c["json"], c["name"]


def test_frame_length_formula(oracle, golden):
    """len(frame) = 26 + digits(seq) + len(json)."""
    for c in golden[:400]:
        if c["status"] & 0xF == 0:
            fr = bytes.fromhex(c["frame"])
            js = bytes.fromhex(c["json"])
            seq = fr[len(b"event: token\nid: "):fr.index(b"\ndata: ")]
            assert len(fr) == 26 + len(seq) + len(js)
            assert fr.endswith(js + b"\n\n")


def test_fnv1a64(oracle):
    # FNV-1a 64 published test values
    assert oracle.fnv1a64(b"") == 0xCBF29CE484222325
    assert oracle.fnv1a64(b"a") == 0xAF63DC4C8601EC8C
    assert oracle.fnv1a64(b"foobar") == 0x85944171F73967E8
