"""CPU tests of the wire splitter (row N4, receiving side): the oracle against hand vectors typed from the protocol
specifications, and the product's per-connection walk (compiled for the host as a test double) against the oracle on
seeded streams with cuts and violations."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hostcheck_lib as hc  # noqa: E402
import wire_case  # noqa: E402


def test_oracle_hand_vectors(oracle):
    for k, (proto, data, want_msgs, consumed, status, pings) in enumerate(wire_case.HAND):
        buf = np.frombuffer(data + b"\0" * 64, dtype=np.uint8).copy()
        off = np.array([0, len(data)], dtype=np.uint32)
        msgs, conns = wire_case.oracle_split(oracle.lib, proto, buf, off)
        assert [(s, i, p) for _, s, i, p in msgs] == want_msgs, k
        assert conns[0] == (len(want_msgs), consumed, status, pings), (k, conns[0])


@pytest.mark.parametrize("proto", [0, 1])
def test_product_walk_matches_oracle(oracle, proto):
    lib = hc.load()
    buf, off, want = wire_case.build(proto, 3000, 41 + proto)
    msgs, conns = wire_case.oracle_split(oracle.lib, proto, buf, off)
    # uncut, undamaged connections: the oracle must return exactly what the generator put in
    k = 0
    for c, w in enumerate(want):
        got = []
        while k < len(msgs) and msgs[k][0] == c:
            got.append(msgs[k][1:])
            k += 1
        if w is not None:
            assert got == w, c
            assert conns[c][1] == int(off[c + 1] - off[c]) and conns[c][2] & 1 == 0
    u8p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)
    lib.hc_wire_split.restype = C.c_uint32
    lib.hc_wire_split.argtypes = [C.c_uint32, u8p, u32p, C.c_uint32, C.POINTER(wire_case.OMsg), C.c_uint32, C.POINTER(wire_case.OConn)]
    n_conn = len(off) - 1
    out = (wire_case.OMsg * 40000)()
    oc = (wire_case.OConn * n_conn)()
    n = lib.hc_wire_split(proto, buf.ctypes.data_as(u8p), off.ctypes.data_as(u32p), n_conn, out, 40000, oc)
    raw = buf.tobytes()
    got = [(out[i].conn, raw[out[i].subj_off:out[i].subj_off + out[i].subj_len], int(out[i].sid),
            raw[out[i].payload_off:out[i].payload_off + out[i].payload_len]) for i in range(n)]
    assert got == msgs
    assert [(oc[c].n_msgs, oc[c].consumed, oc[c].status, oc[c].pings) for c in range(n_conn)] == conns
    assert n > 5000 and sum(1 for x in conns if x[2] & 1) > 50


def test_find_byte2_every_alignment_and_neighbour_value():
    """The SWAR search under the walkers (sse_record.cuh find_byte2) against a byte loop: every start / end alignment, the
    wanted bytes at every position, and the byte values that trip inexact zero-byte tests (c +- 1, 0x00, 0x01, 0x80, 0xFF)."""
    import random
    lib = hc.load()
    lib.hc_find_byte2.restype = C.c_uint32
    lib.hc_find_byte2.argtypes = [C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    rng = random.Random(7)
    for c0, c1 in ((10, 10), (32, 9), (34, 92), (0x80, 0x01)):
        tricky = [c0, c1, c0 ^ 1, (c0 + 1) & 255, (c0 - 1) & 255, (c1 + 1) & 255, (c1 - 1) & 255, 0, 1, 0x7F, 0x80, 0x81, 0xFF, 65]
        for trial in range(300):
            n = 40
            data = bytes(rng.choice(tricky) if rng.random() < 0.5 else rng.randrange(256) for _ in range(n)) + b"\xff" * 8
            if trial % 3 == 0:                                   # long runs without a match
                data = bytes(b if b not in (c0, c1) else 65 for b in data[:n]) + data[n:]
            buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
            for p in range(0, 9):
                for e in (p, p + 1, p + 2, p + 3, p + 5, 17, 31, n):
                    if e < p:
                        continue
                    want = next((q for q in range(p, e) if data[q] in (c0, c1)), e)
                    assert lib.hc_find_byte2(buf, p, e, c0, c1) == want, (c0, c1, data.hex(), p, e)
