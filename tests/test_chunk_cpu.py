"""CPU tests of the vLLM stream-chunk decoder (row N3, decode half): oracle and product logic (host test double)
against the hand vectors, each other and a Python model on seeded lines."""
import os
import random
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import chunk_case  # noqa: E402
import hostcheck_lib as hc  # noqa: E402


def test_hand_vectors(oracle):
    lib = hc.load()
    buf, off = chunk_case.pack([v[0] for v in chunk_case.HAND])
    for name, fn in (("oracle", oracle.lib.sseo_chunk_decode), ("product", lib.hc_chunk_decode)):
        st, msgs = chunk_case.run_c(fn, buf, off)
        for i, (line, want_st, want_msgs) in enumerate(chunk_case.HAND):
            if name == "oracle" and want_st & 0x200:
                continue                                   # the oracle has no limit on the number of choices
            assert st[i] == want_st, (name, i, line)
            assert [(t, d) for ln, t, d in msgs if ln == i] == want_msgs, (name, i, line)


def test_product_matches_oracle_and_python_model(oracle):
    lib = hc.load()
    rng = random.Random(11)
    lines = [chunk_case.rand_line(rng, k) for k in range(20000)]
    buf, off = chunk_case.pack(lines)
    st_o, msgs_o = chunk_case.run_c(oracle.lib.sseo_chunk_decode, buf, off)
    st_p, msgs_p = chunk_case.run_c(lib.hc_chunk_decode, buf, off)
    assert st_p == st_o
    assert msgs_p == msgs_o
    by_line = {}
    for ln, t, d in msgs_o:
        by_line.setdefault(ln, []).append((t, d))
    n_cmp = 0
    for i, line in enumerate(lines):
        try:
            line.decode("utf-8")
        except UnicodeDecodeError:
            continue                                       # CPython's codec is strict where Go replaces: not comparable
        if b"\\ud" in line.lower():
            continue                                       # lone surrogate escapes: same reason
        want_st, want_msgs = chunk_case.model_line(line)
        assert st_o[i] == want_st, (i, line)
        assert by_line.get(i, []) == want_msgs, (i, line)
        n_cmp += 1
    assert n_cmp > 15000 and len(msgs_o) > 8000 and sum(1 for s in st_o if s & 0x100) > 500
