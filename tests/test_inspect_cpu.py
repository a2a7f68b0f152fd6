"""CPU tests of content inspection: the brute-force oracle against the reference's three known answers (Set B),
and the product's pattern compiler + DFA walk (compiled for the host as a test double) against the oracle on
seeded texts, per-token and streaming."""
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hand_vectors as hv  # noqa: E402
import hostcheck_lib as hc  # noqa: E402
import make_golden as mg  # noqa: E402
import pymodel  # noqa: E402
from inspect_sets import ACTION_NAME, SET_A, SET_B  # noqa: E402

FRAGMENTS = [b"123-45-6789", b"4111111111111111", b"41111111111111112", b"bob@example.com", b"a.b+c@x-y.co|m", b"x@y.z",
             b"ignore previous instructions", b"IGNORE Previous", b"disregard the above", b"disregard above", b"new instructions:",
             b"System Prompt:", b"password", b"my secret", b"api_key=1", b"credit_card", b" ", b"  ", b"the", b"quick", b"1", b"12",
             b"-", b"@", b".", b":", b"a", b"Z", b"_", b"\xc3\xa9", b"\xe2\x82\xac", b"9", b"99-", b"x123-45-6789", b"123-45-67890",
             b"\n", b"\"", b"<b>", b"ins", b"truc", b"tions", b"pass", b"word", b"sec", b"ret", b"@exa", b"mple.", b"com"]


@pytest.fixture(scope="module")
def lib():
    return hc.load()


def rand_token(rng):
    return b"".join(rng.choice(FRAGMENTS) for _ in range(rng.randrange(0, 4)))


def test_set_b_known_answers_from_the_reference(oracle):
    insp = oracle.inspector(SET_B)
    for text, want in hv.SET_B_KNOWN:
        v, b, e = insp.token(text)
        assert ACTION_NAME[v & 0xFF] == want, text
    v, b, e = insp.token(b"My password is secret123")
    assert (v >> 8) & 0xFF == 0 and text_slice(b"My password is secret123", b, e).lower() == b"password"


def text_slice(t, b, e):
    return t[b:e]


def test_set_a_hand_cases(oracle):
    insp = oracle.inspector(SET_A)
    cases = [
        (b"ssn 123-45-6789.", 1, 0, 4, 15), (b"x123-45-6789", 0, 0, 0, 0), (b"123-45-67890", 0, 0, 0, 0),
        (b"4111111111111111", 1, 1, 0, 16), (b"41111111111111111", 0, 0, 0, 0), (b"4111 1111 1111 1111", 0, 0, 0, 0),
        (b"mail bob@example.com now", 1, 2, 5, 20), (b"a@b.c", 0, 0, 0, 0), (b"x@y.c|d", 1, 2, 0, 6), (b"x@y.c|", 0, 0, 0, 0),
        (b"please IGNORE previous Instructions", 3, 3, 7, 35), (b"new instructions: x", 3, 5, 0, 17),
        (b"\xc3\xa9123-45-6789", 1, 0, 2, 13),      # declared deviation from Rust's Unicode \b: bytes >= 0x80 are non-word
    ]
    for text, act, pid, b, e in cases:
        v, gb, ge = insp.token(text)
        assert (v & 0xFF, (v >> 8) & 0xFF if act else 0, gb, ge) == (act, pid, b, e), text


@pytest.mark.parametrize("name,pats", [("A", SET_A), ("B", SET_B), ("AB", SET_A + SET_B)])
def test_dfa_matches_oracle_per_token(lib, oracle, name, pats):
    rng = random.Random(7)
    insp = oracle.inspector(pats)
    toks = [rand_token(rng) for _ in range(3000)] + [t for t, _ in hv.SET_B_KNOWN] + [b"", b"a" * 300 + b"@b.cc"]
    payloads = [mg.canon("c%d" % i, t.decode("utf-8", "replace"), i, False, 5) if i % 3 else
                b'{"token":' + mg.json_str_variants(rng, t.decode("utf-8", "replace")) + b'}' for i, t in enumerate(toks)]
    arena, off = oracle.pack(payloads)
    verdict, match, status, n_states = hc.inspect(lib, pats, 0, arena, off)
    assert n_states < 2000
    for i, p in enumerate(payloads):
        st, m = pymodel.unmarshal(p)
        assert st == 0
        text = m["tok"].encode("utf-8")
        v, b, e = insp.token(text)
        if len(text) > 256 and v & 0x10000 == 0 and (verdict[i] & 0x10000):
            continue        # product limit: the leftmost start of a variable-length match is searched in <= 256 decoded bytes
        assert (int(verdict[i]), int(match[i, 0]), int(match[i, 1])) == (v, b, e), (i, text)


@pytest.mark.parametrize("pats", [SET_A, SET_A + SET_B])
def test_dfa_matches_oracle_streaming(lib, oracle, pats):
    rng = random.Random(11)
    insp = oracle.inspector(pats)
    n_conv, n_tok = 60, 25
    streams = [[] for _ in range(n_conv)]
    payloads, slots = [], []
    for j in range(n_tok):
        for c in range(n_conv):
            if rng.random() < 0.1:
                continue
            tok = rand_token(rng)
            done = j == n_tok - 1 or rng.random() < 0.03
            streams[c].append((tok, done))
            payloads.append(mg.canon("conv-%d" % c, tok.decode("utf-8", "replace"), j + 1, done, 9))
            slots.append(c)
    arena, off = oracle.pack(payloads)
    # each batch holds a conversation at most once: one batch per arrival round, state carried by the double
    verdict, match, status, _ = hc.inspect(lib, pats, 1, arena, off, np.array(slots, dtype=np.uint32), n_conv)
    pos = [0] * n_conv
    hist = [b""] * n_conv
    k = [0] * n_conv
    for i, c in enumerate(slots):
        tok, done = streams[c][k[c]]
        k[c] += 1
        tok = tok.decode("utf-8", "replace").encode("utf-8")
        s = len(hist[c])
        hist[c] += tok
        v, b, e = insp.stream(hist[c], s, len(hist[c]), done)
        assert (int(verdict[i]), int(match[i, 0]), int(match[i, 1])) == (v, b, e), (i, c, hist[c], tok, done)
        if done:
            hist[c] = b""


@pytest.mark.parametrize("pats", [SET_A, SET_B, SET_A + SET_B])
def test_oracle_verdicts_against_python_re(oracle, pats):
    """Witness nobody here wrote: which pattern fires (all the reference's is_match-style snippets use) must be what Python's
    `re` says on the same bytes -- bytes patterns give ASCII \\d \\w \\b, the declared semantics.  (Match offsets are this
    project's own definition -- earliest end, leftmost start -- and have no counterpart in `re`'s leftmost-greedy search.)"""
    import re
    rxs = [re.compile(p) if kind == 1 else re.compile(re.escape(p), re.IGNORECASE) for kind, _, p in pats]
    rng = random.Random(11)
    insp = oracle.inspector(pats)
    fired = 0
    for _ in range(20000):
        text = b"".join(rng.choice(FRAGMENTS) for _ in range(rng.randrange(0, 5)))
        want = 0
        for pid, rx in enumerate(rxs):
            if rx.search(text):
                want = pats[pid][1] | pid << 8
                break
        v, _, _ = insp.token(text)
        assert v & 0xFFFF == want, text
        fired += want != 0
    assert fired > 3000
