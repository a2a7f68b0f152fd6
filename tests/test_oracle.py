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


def _cpython_view(payload):
    """What CPython's json module (an implementation nobody on this project wrote) says about a payload, translated to the Go
    rules where the two languages differ only in typing: -> None when the comparison does not apply, "syntax" when it is not
    JSON, else (status, fields) with status 0 / 2 and the five decoded fields."""
    import json
    try:
        text = payload.decode("utf-8")          # Go replaces invalid UTF-8 inside strings, Python cannot read it at all
    except UnicodeDecodeError:
        return None
    if "NaN" in text or "Infinity" in text:     # CPython extension
        return None
    depth = run = 0
    for ch in text:
        run += ch in "[{"
        run -= ch in "]}"
        depth = max(depth, run)
    if depth > 200:
        return None
    try:
        doc = json.loads(text, parse_float=lambda s: ("float", s), parse_int=lambda s: ("int", s),
                         object_pairs_hook=lambda pairs: ("object", pairs))
    except (ValueError, RecursionError):
        return "syntax"
    msg = {"cid": "", "tok": "", "seq": 0, "done": False, "ts": 0}
    if doc is None:
        return 0, msg
    if not (isinstance(doc, tuple) and doc[0] == "object"):
        return 2, None
    names = {"conversation_id": "cid", "token": "tok", "sequence": "seq", "done": "done", "timestamp": "ts"}
    bad = False
    for key, val in doc[1]:
        field = names.get(key)
        if field is None:                        # encoding/json folds case (and two non-ASCII look-alikes) when no exact match exists
            folded = key.replace("ſ", "s").replace("K", "k").lower()
            field = names.get(folded)
        if field is None or val is None:
            continue
        want_str, want_bool = field in ("cid", "tok"), field == "done"
        if want_str:
            if isinstance(val, str):
                msg[field] = val
            else:
                bad = True
        elif want_bool:
            if isinstance(val, bool):
                msg[field] = val
            else:
                bad = True
        else:
            if isinstance(val, tuple) and val[0] == "int" and -(1 << 63) <= int(val[1]) < (1 << 63):
                msg[field] = int(val[1])
            else:
                bad = True
    return (2 if bad else 0), msg


def test_oracle_against_cpython_json(oracle, golden):
    """Third witness for the decoder: CPython's json parser decides what is JSON and what the values are, a thin layer maps its
    result onto TokenMessage with Go's typing rules; syntax verdicts must agree on every comparable payload, status and frame
    on every payload whose strings carry no lone surrogates (Go replaces those, Python keeps them)."""
    import random
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden as mg
    rng = random.Random(2024)
    payloads = [bytes.fromhex(c["in"]) for c in golden]
    base = mg.canon("7f3c2b1a-1111-4222-8333-944455556666", " hello", 42, False, 1700000000000000000)
    for _ in range(20000):
        r = rng.random()
        payloads.append(mg.structured(rng) if r < 0.5 else mg.mutate(rng, rng.choice([base, payloads[-1]])))
    arena, off = oracle.pack(payloads)
    out, out_off, status = oracle.adapter(arena, off)
    raw = out.tobytes()
    compared = syntax_errors = framed = 0
    for i, p in enumerate(payloads):
        view = _cpython_view(p)
        if view is None:
            continue
        compared += 1
        code = int(status[i]) & 0xF
        if view == "syntax":
            assert code == 1, (p, code)
            syntax_errors += 1
            continue
        assert code != 1, (p, "oracle says syntax error, CPython parses it")
        want_code, msg = view
        assert code == want_code, (p, code, want_code)
        if code == 0:
            try:
                (msg["cid"] + msg["tok"]).encode("utf-8")
            except UnicodeEncodeError:
                continue                          # lone surrogate from a \\ud800 escape: Go -> U+FFFD, Python keeps it
            assert raw[int(out_off[i]):int(out_off[i + 1])] == pymodel.frame(msg), p
            framed += 1
    assert compared > 10000 and syntax_errors > 1000 and framed > 2000, (compared, syntax_errors, framed)


def test_go_crosscheck_inputs_are_in_sync():
    """oracle/go_crosscheck/main.go replays tests/golden/hand_vectors.json; it must say what hand_vectors.py says."""
    import json
    import hand_vectors as hv
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hand_vectors.json")) as f:
        recs = {r["name"]: r for r in json.load(f)}
    for name, payload, start, status, frame in hv.ADAPTER:
        r = recs["hand_" + name]
        assert (bytes.fromhex(r["in"]), r["start"], r["status"], bytes.fromhex(r["frame"])) == (payload, start, status, frame)
    for k, (payload, status, subject, js) in enumerate(hv.BRIDGE):
        r = recs["hand_bridge_%d" % k]
        assert (bytes.fromhex(r["in"]), r["bstatus"], bytes.fromhex(r["subject"]), bytes.fromhex(r["json"])) == (payload, status, subject, js)
    assert len(recs) == len(hv.ADAPTER) + len(hv.BRIDGE)
