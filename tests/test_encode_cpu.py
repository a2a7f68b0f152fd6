"""CPU tests of the producer-side encoder (row N3): oracle and the product's encode_message() (host test double) against
the hand-written known answers and an independent Python model, and the round trip into the adapter path."""
import numpy as np
import pytest

import encode_case as ec
import hostcheck_lib as hc


@pytest.fixture(scope="module")
def lib():
    return hc.load()


def test_model_against_hand_vectors():
    for msg, want in ec.HAND:
        assert ec.model_encode(msg) == want, msg
    for msg, want in ec.HAND_GO122:
        assert ec.model_encode(msg, go122=True) == want, msg


def test_oracle_hand_vectors(oracle):
    got, _ = oracle.encode(ec.columns([m for m, _ in ec.HAND]))
    assert got == [w for _, w in ec.HAND]
    got, _ = oracle.encode(ec.columns([m for m, _ in ec.HAND_GO122]), go122=True)
    assert got == [w for _, w in ec.HAND_GO122]
    got, chan = oracle.encode(ec.columns([m for m, _, _ in ec.HAND_CHANNEL]), with_channel=True)
    assert got == [c + w for _, c, w in ec.HAND_CHANNEL]
    assert list(chan) == [len(c) for _, c, _ in ec.HAND_CHANNEL]


def test_hostcheck_hand_vectors(lib):
    got, _ = hc.encode(lib, ec.columns([m for m, _ in ec.HAND]))
    assert got == [w for _, w in ec.HAND]
    got, _ = hc.encode(lib, ec.columns([m for m, _ in ec.HAND_GO122]), hc.OP_GO122)
    assert got == [w for _, w in ec.HAND_GO122]
    got, chan = hc.encode(lib, ec.columns([m for m, _, _ in ec.HAND_CHANNEL]), hc.OP_BRIDGE)
    assert got == [c + w for _, c, w in ec.HAND_CHANNEL]
    assert list(chan) == [len(c) for _, c, _ in ec.HAND_CHANNEL]


@pytest.mark.parametrize("go122", [False, True])
def test_random_messages_three_ways(lib, oracle, go122):
    msgs = ec.random_messages(4000, seed=11 + go122)
    cols = ec.columns(msgs)
    want = [ec.model_encode(m, go122) for m in msgs]
    got_o, _ = oracle.encode(cols, go122=go122)
    got_h, _ = hc.encode(lib, cols, hc.OP_GO122 if go122 else 0)
    assert got_o == want
    assert got_h == want
    got_o, chan_o = oracle.encode(cols, go122=go122, with_channel=True)
    got_h, chan_h = hc.encode(lib, cols, hc.OP_BRIDGE | (hc.OP_GO122 if go122 else 0))
    assert got_o == [ec.model_channel(m) + w for m, w in zip(msgs, want)]
    assert got_h == got_o
    assert (chan_o == chan_h).all() and list(chan_o) == [12 + len(m[0]) for m in msgs]


def test_empty_batch(lib, oracle):
    cols = ec.columns([])
    assert oracle.encode(cols)[0] == []
    assert hc.encode(lib, cols)[0] == []


def test_encoded_payloads_take_the_adapter_fast_path(lib, oracle):
    """whatever the producer emits is canonical: the adapter recognises it and frames it verbatim -- except that a
    \\ufffd escape (an invalid byte at the producer) decodes to U+FFFD and is re-marshalled as its three UTF-8 bytes."""
    def clean(raw):
        return "".join(ch for ch, _ in ec.go_decode(raw)).encode("utf-8")

    msgs = ec.random_messages(3000, seed=3)
    payloads, _ = hc.encode(lib, ec.columns(msgs))
    arena, off = oracle.pack(payloads)
    out, out_off, _, status, nfast = hc.run(lib, arena, off, None, 0)
    b = out.tobytes()
    for i, m in enumerate(msgs):
        assert int(status[i]) & 0xF == 0
        assert bool(int(status[i]) & 0x100) == bool(m[3])
        assert b[int(out_off[i]):int(out_off[i + 1])] == b"event: token\nid: %d\ndata: " % m[2] + ec.model_encode((clean(m[0]), clean(m[1])) + m[2:]) + b"\n\n"
    o_out, o_off, o_status = oracle.adapter(arena, off)
    assert o_out.tobytes() == b and (o_status == status).all()
    # ordinary traffic (valid UTF-8, sequences of up to 8 digits) goes through the canonical fast lane, every record
    plain = [m for m in msgs if m[0].startswith(b"conv-")]
    payloads, _ = hc.encode(lib, ec.columns(plain))
    arena, off = oracle.pack(payloads)
    out, out_off, _, status, nfast = hc.run(lib, arena, off, None, 0)
    assert nfast == len(plain) > 1000
    assert [out.tobytes()[int(out_off[i]):int(out_off[i + 1])] for i in range(len(plain))] == \
        [b"event: token\nid: %d\ndata: " % m[2] + p + b"\n\n" for m, p in zip(plain, payloads)]


def test_resp_publish_framing(lib, oracle):
    """SSEGPU_OP_RESP: the command go-redis writes for rdb.Publish (llm-stream-proxy/main.go:244)."""
    for m, want in ec.HAND_RESP:
        assert ec.model_resp(m) == want, m
    cols = ec.columns([m for m, _ in ec.HAND_RESP])
    assert oracle.encode_resp(cols)[0] == [w for _, w in ec.HAND_RESP]
    assert hc.encode(lib, cols, hc.OP_RESP)[0] == [w for _, w in ec.HAND_RESP]
    msgs = ec.random_messages(3000, seed=21)
    cols = ec.columns(msgs)
    for go122 in (False, True):
        want = [ec.model_resp(m, go122) for m in msgs]
        got_o, chan_o = oracle.encode_resp(cols, go122=go122)
        got_h, chan_h = hc.encode(lib, cols, hc.OP_RESP | (hc.OP_GO122 if go122 else 0))
        assert got_o == want and got_h == want
        assert list(chan_o) == list(chan_h) == [12 + len(m[0]) for m in msgs]


def test_workload_columns_are_the_workload_before_marshalling(oracle):
    """workload.generate_columns() vs the oracle's decode of workload.generate(), configs 0 and 1; and marshalling the
    columns gives the payload arena back (this is what the encoder bench and the full-shape GPU test rely on)."""
    from __graft_entry__ import load_package
    pkg = load_package()
    for config in (0, 1):
        n_conv, n_tok = 250, 40
        n = n_conv * n_tok
        arena, off = pkg.workload.generate(config, n_conv, n_tok)
        cols = pkg.workload.generate_columns(config, n_conv, n_tok)
        for got, want in zip(cols, oracle.decode_columns(arena, off)):
            k = len(want) if want.dtype != np.uint8 or len(want) == n else len(want) - 32
            assert np.array_equal(got[:k], want[:k])
        recs, _ = oracle.encode(cols)
        assert b"".join(recs) == arena[: int(off[n])].tobytes()


def test_marshal_against_cpython_json(lib, oracle):
    """Independent witness for the encoder half: for valid UTF-8 strings without < > & U+2028 U+2029 (which only Go escapes),
    CPython's json.dumps(ensure_ascii=False, separators=(",", ":")) spells a message exactly as Go >= 1.22 does (Go 1.21 differs
    in \\b and \\f only, covered by the hand vectors)."""
    import json
    import random
    rng = random.Random(8)
    alphabet = ["a", "Z", " ", "the", " quick", '"', "\\", "\n", "\t", "\r", "\x00", "\x08", "\x0c", "\x1f", "\x7f", "é", "世", "\U0001f600", "�", "/"]
    msgs, want = [], []
    for i in range(3000):
        cid = "".join(rng.choice(alphabet) for _ in range(rng.randrange(0, 8)))
        tok = "".join(rng.choice(alphabet) for _ in range(rng.randrange(0, 40)))
        seq, done, ts = rng.choice(ec._INTS), rng.random() < 0.5, rng.choice(ec._INTS)
        msgs.append((cid.encode(), tok.encode(), seq, done, ts))
        want.append(json.dumps({"conversation_id": cid, "token": tok, "sequence": seq, "done": done, "timestamp": ts},
                               ensure_ascii=False, separators=(",", ":")).encode())
    cols = ec.columns(msgs)
    assert oracle.encode(cols, go122=True)[0] == want
    assert hc.encode(lib, cols, hc.OP_GO122)[0] == want
