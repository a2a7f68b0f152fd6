"""Replay test of the host-side adapter mirror (load-generator replay mode): synthetic producers
feed /stream and /chat connections through the batched GPU path; every connection's byte stream must equal what the
reference handler would have written -- greeting, then one 3-line frame per delivered token (Last-Event-ID filter,
bad messages dropped), ending with the done frame -- while each connection gets at most one write per batch."""
import os
import random
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as mg  # noqa: E402
import pymodel  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("header_only", [False, True])
def test_replay_streams_match_reference_model(header_only):
    """header_only: SSEGPU_OP_HEADER_ONLY egress -- the GPU returns only the frame headers of Go-marshalled records and the
    coalescing loop appends the payload it still holds; the byte streams must not change."""
    pkg = load_package()
    rng = random.Random(3)
    n_conn, n_tok = 120, 40
    eng = pkg.SseGpu(device=0, max_records=512, max_in_bytes=512 * 160, ring_slots=2)
    host = pkg.host.AdapterHost(eng, 512, 512 * 160)
    host.set_header_only(header_only)
    try:
        conns = []
        for c in range(n_conn):
            cid = "loadtest-1700000000000000000-%d" % c
            chat = c % 3 == 0
            leid = None if chat or c % 3 == 1 else rng.choice(["3", "0", "abc", "17", "-2"])
            host.connect(c, cid.encode(), leid.encode() if leid else None, chat)
            greeting = (b'event: connected\ndata: {"conversation_id":"' + cid.encode() + b'"}\n\n') if chat else \
                (b": connected to " + cid.encode() + b"\n\n")
            conns.append(dict(cid=cid, chat=chat, start=0 if chat else pymodel.start_sequence(leid), want=bytearray(greeting),
                              got=bytearray(), closed=False, model_closed=False, writes=0))
        flushes = 0
        for j in range(n_tok):
            order = list(range(n_conn))
            rng.shuffle(order)
            for c in order:
                st = conns[c]
                done = j == n_tok - 1 or (c % 7 == 0 and j == 20)        # some streams end early; later tokens must not be delivered
                r = rng.random()
                if r < 0.03:
                    payload = b'{"sequence":"oops"}'                       # type error -> dropped
                elif r < 0.06:
                    payload = b'{"token": "x", "sequence": %d, "conversation_id": "%s"}' % (j + 1, st["cid"].encode())   # non-canonical
                else:
                    payload = mg.canon(st["cid"], mg.WORDS[j % 56], j + 1, done, 1700000000000000000 + j)
                host.on_message(c, payload)
                if not st["model_closed"]:
                    status, frame = pymodel.adapter(payload, st["start"])
                    st["want"] += frame
                    if status & 0x100:
                        st["model_closed"] = True
            if j % 3 == 2 or j == n_tok - 1:
                host.flush()
                flushes += 1
                for c in range(n_conn):
                    data, closed = host.take(c)
                    if data:
                        conns[c]["got"] += data
                        conns[c]["writes"] += 1
                    conns[c]["closed"] = closed
        for c, st in enumerate(conns):
            assert bytes(st["got"]) == bytes(st["want"]), c
            assert st["closed"] == st["model_closed"]                 # the handler returns exactly after a delivered done frame
            assert st["writes"] <= flushes                              # coalesced: at most one write per batch per connection
        assert sum(st["closed"] for st in conns) > n_conn * 0.8
        s = host.stats()
        assert s["delivered"] > 3000 and s["dropped"] > 50 and s["filtered"] > 50 and s["after_close"] > 100
    finally:
        host.close()
        eng.close()
