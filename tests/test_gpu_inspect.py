"""GPU parity tests of content inspection through the C ABI: verdicts and match offsets against the brute-force
oracle (per-token and streaming), inline drop/redact against the Python model, duplicate-conversation detection."""
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hand_vectors as hv  # noqa: E402
import make_golden as mg  # noqa: E402
import pymodel  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import normalise, slices  # noqa: E402
from inspect_sets import ACTION_NAME, SET_A, SET_B  # noqa: E402
from test_inspect_cpu import rand_token  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    return load_package()


def run(eng, arena, off, op, conv=None):
    b = eng.acquire()
    b.fill(arena, off, None, conv)
    b.submit(len(off) - 1, op)
    b.wait()
    out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), False)
    res = dict(out=out, off=out_off, status=b.status.copy(), verdict=b.verdict.copy(), match=b.match.copy())
    b.release()
    return res


def odd_payload(rng, cid, tok, seq, done, ts):
    """the same message, not as json.Marshal would write it (key order, a space, maybe other escapes in the token)"""
    return (b'{ "done":' + (b"true" if done else b"false") + b',"token":' + mg.json_str_variants(rng, tok) + b',"timestamp":%d' % ts
            + b',"sequence":%d,"conversation_id":' % seq + mg.json_str_variants(rng, cid) + b'}')


def test_reference_known_answers_set_b(pkg, oracle):
    eng = pkg.SseGpu(device=0, max_records=4096, max_in_bytes=1 << 20, ring_slots=1)
    try:
        eng.load_patterns(SET_B, pkg.INSPECT_PER_TOKEN)
        payloads = [mg.canon("c", t.decode(), 1, False, 2) for t, _ in hv.SET_B_KNOWN]
        arena, off = oracle.pack(payloads)
        r = run(eng, arena, off, pkg.OP_INSPECT)
        assert [ACTION_NAME[int(v) & 0xFF] for v in r["verdict"]] == [want for _, want in hv.SET_B_KNOWN]
    finally:
        eng.close()


@pytest.mark.parametrize("pats", [SET_A, SET_A + SET_B])
def test_per_token_matches_oracle(pkg, oracle, pats):
    rng = random.Random(21)
    insp = oracle.inspector(pats)
    toks = [rand_token(rng) for _ in range(20000)]
    payloads = [mg.canon("c%d" % i, t.decode("utf-8", "replace"), i, False, 5) if i % 5 else
                b'{"token":' + mg.json_str_variants(rng, t.decode("utf-8", "replace")) + b',"sequence":3}' for i, t in enumerate(toks)]
    payloads += [b'{"token":1}', b"{}x", b"null"]
    arena, off = oracle.pack(payloads)
    eng = pkg.SseGpu(device=0, max_records=1 << 16, max_in_bytes=8 << 20, ring_slots=1)
    try:
        eng.load_patterns(pats, pkg.INSPECT_PER_TOKEN)
        r = run(eng, arena, off, pkg.OP_INSPECT)
        want_out, want_off, want_status = oracle.adapter(arena, off)
        assert np.array_equal(r["status"], want_status) and np.array_equal(r["out"], want_out)       # verdict-only mode leaves frames alone
        for i, p in enumerate(payloads):
            st, m = pymodel.unmarshal(p)
            want = insp.token(m["tok"].encode("utf-8")) if st == 0 else (0, 0, 0)
            assert (int(r["verdict"][i]), int(r["match"][i, 0]), int(r["match"][i, 1])) == want, (i, p)
    finally:
        eng.close()


def test_streaming_matches_oracle_and_carries_state_across_batches(pkg, oracle):
    pats = SET_A + SET_B
    rng = random.Random(33)
    insp = oracle.inspector(pats)
    n_conv, n_tok = 500, 30
    eng = pkg.SseGpu(device=0, max_records=4096, max_in_bytes=1 << 20, ring_slots=1, conv_slots=n_conv)
    try:
        eng.load_patterns(pats, pkg.INSPECT_STREAMING)
        hist = [b""] * n_conv
        hits = 0
        for j in range(n_tok):                      # one batch per arrival round: a conversation occurs at most once in it
            convs = [c for c in range(n_conv) if rng.random() < 0.9]
            toks = [rand_token(rng) for _ in convs]
            dones = [j == n_tok - 1 or rng.random() < 0.03 for _ in convs]
            # a few records per warp are not Go-marshalled: they take the general decoder in the kernel's tail phase
            payloads = [mg.canon("conv-%d" % c, t.decode("utf-8", "replace"), j + 1, d, 9) if rng.random() < 0.9 else
                        odd_payload(rng, "conv-%d" % c, t.decode("utf-8", "replace"), j + 1, d, 9) for c, t, d in zip(convs, toks, dones)]
            arena, off = oracle.pack(payloads)
            r = run(eng, arena, off, pkg.OP_INSPECT, np.array(convs, dtype=np.uint32))
            for i, (c, t, d) in enumerate(zip(convs, toks, dones)):
                t = t.decode("utf-8", "replace").encode("utf-8")
                s = len(hist[c])
                hist[c] += t
                want = insp.stream(hist[c], s, len(hist[c]), d)
                assert (int(r["verdict"][i]), int(r["match"][i, 0]), int(r["match"][i, 1])) == want, (j, c, hist[c])
                hits += want[0] != 0
                if d:
                    hist[c] = b""
        assert hits > 200
        # a conversation twice in one batch: its tokens are walked in arrival order ("pass" + "word" = password)
        arena, off = oracle.pack([mg.canon("c", "my pass", 1, False, 2), mg.canon("c", "word is", 2, True, 2)])
        r = run(eng, arena, off, pkg.OP_INSPECT, np.array([7, 7], dtype=np.uint32))
        assert int(r["verdict"][0]) == 0 and int(r["verdict"][1]) == insp.stream(b"my password is", 7, 14, True)[0] != 0
        # a conversation slot the context does not have is an error, not a wild write
        b = eng.acquire()
        b.fill(arena, off, None, np.array([7, n_conv], dtype=np.uint32))
        b.submit(2, pkg.OP_INSPECT)
        with pytest.raises(pkg.SseGpuError) as e:
            b.wait()
        assert e.value.code == -9
        b.release()
    finally:
        eng.close()


def test_inline_drop_and_redact(pkg, oracle):
    """sse_handler.go:200-207: drop skips the token, redact replaces token.Token before the frame is encoded."""
    rng = random.Random(5)
    insp = oracle.inspector(SET_B)
    toks = [rand_token(rng) for _ in range(5000)]
    payloads = [mg.canon("c%d" % i, t.decode("utf-8", "replace"), i, i % 50 == 0, 5) if i % 17 else
                odd_payload(rng, "c%d" % i, t.decode("utf-8", "replace"), i, i % 50 == 0, 5) for i, t in enumerate(toks)]
    arena, off = oracle.pack(payloads)
    eng = pkg.SseGpu(device=0, max_records=1 << 16, max_in_bytes=4 << 20, ring_slots=1)
    try:
        eng.load_patterns(SET_B, pkg.INSPECT_PER_TOKEN)
        r = run(eng, arena, off, pkg.OP_INSPECT_INLINE)
        frames = slices(r["out"], r["off"])
        n_drop = n_red = 0
        for i, p in enumerate(payloads):
            st, m = pymodel.unmarshal(p)
            v, _, _ = insp.token(m["tok"].encode("utf-8"))
            act = v & 0xFF
            if act == 2:
                assert int(r["status"][i]) == 4 and frames[i] == b""
                n_drop += 1
            elif act == 1:
                m["tok"] = "[REDACTED]"
                assert int(r["status"][i]) == (0x200 | (0x100 if m["done"] else 0)) and frames[i] == pymodel.frame(m)
                n_red += 1
            else:
                assert int(r["status"][i]) == (0x100 if m["done"] else 0) and frames[i] == pymodel.frame(m)
        assert n_drop > 50 and n_red > 50
    finally:
        eng.close()


def test_bad_pattern_is_rejected(pkg):
    eng = pkg.SseGpu(device=0, max_records=16, max_in_bytes=4096)
    try:
        for bad in ([(1, 1, b"(a|b)")], [(1, 1, b"a*")], [(0, 9, b"x")], [(1, 1, b"[a-")]):
            with pytest.raises(pkg.SseGpuError) as e:
                eng.load_patterns(bad)
            assert e.value.code == -10
    finally:
        eng.close()
