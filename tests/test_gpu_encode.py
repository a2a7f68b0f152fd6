"""GPU parity tests of the producer-side encoder (row N3): k_encode_messages through the C ABI against the hand-written
known answers, the oracle and the Python model, the overflow report, and encode -> adapter chained on the device."""
import numpy as np
import pytest

import encode_case as ec
from __graft_entry__ import load_package
from gpu_util import normalise

pytestmark = pytest.mark.gpu
OP_BRIDGE, OP_GO122, OP_RESP = 1, 2, 256


@pytest.fixture(scope="module")
def pkg():
    return load_package()


@pytest.fixture(scope="module")
def eng(pkg):
    e = pkg.SseGpu(device=0, max_records=1 << 20, max_in_bytes=64 << 20, ring_slots=1)
    yield e
    e.close()


def gpu_encode(eng, cols, op=0, cap=None, keep=False):
    """-> (records as bytes, chan_len, rc).  keep=True returns the device tensors too."""
    import torch
    cid, cid_off, tok, tok_off, seq, done, ts = cols
    n = len(seq)
    if cap is None:
        cap = (int(cid_off[n]) * 8 + int(tok_off[n]) * 6) + 256 * n + 64
    dev = lambda a, t: torch.from_numpy(a.view(t)).cuda()  # noqa: E731
    d = dict(cid=dev(cid, np.uint8), cid_off=dev(cid_off, np.int32), tok=dev(tok, np.uint8), tok_off=dev(tok_off, np.int32),
             seq=dev(seq, np.int64), done=dev(done, np.uint8), ts=dev(ts, np.int64),
             out=torch.full((cap + 16,), 0xAA, dtype=torch.uint8, device="cuda"), out_off=torch.zeros(n + 1, dtype=torch.int64, device="cuda"),
             out_len=torch.zeros(max(n, 1), dtype=torch.int32, device="cuda"), chan=torch.zeros(max(n, 1), dtype=torch.int32, device="cuda"))
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    eng.encode_launch(d["cid"].data_ptr(), d["cid_off"].data_ptr(), d["tok"].data_ptr(), d["tok_off"].data_ptr(), d["seq"].data_ptr(),
                      d["done"].data_ptr(), d["ts"].data_ptr(), n, op, d["out"].data_ptr(), cap, d["out_off"].data_ptr(),
                      d["out_len"].data_ptr(), d["chan"].data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    rc, used = eng.encode_result()
    out = d["out"].cpu().numpy()
    assert (out[cap:] == 0xAA).all(), "wrote past out_cap"
    if rc != 0:
        return None, None, rc
    out_off = d["out_off"].cpu().numpy().astype(np.uint64)
    out_len = d["out_len"].cpu().numpy().astype(np.uint32)[:n]
    assert used == int(out_off[n])
    arena, want_off = normalise(out, out_off, out_len, False)
    recs = [arena[int(want_off[i]):int(want_off[i + 1])].tobytes() for i in range(n)]
    res = (recs, d["chan"].cpu().numpy().astype(np.uint32)[:n], rc)
    return res + (d,) if keep else res


def test_hand_vectors(eng):
    got, _, rc = gpu_encode(eng, ec.columns([m for m, _ in ec.HAND]))
    assert rc == 0 and got == [w for _, w in ec.HAND]
    got, _, rc = gpu_encode(eng, ec.columns([m for m, _ in ec.HAND_GO122]), OP_GO122)
    assert rc == 0 and got == [w for _, w in ec.HAND_GO122]
    got, chan, rc = gpu_encode(eng, ec.columns([m for m, _, _ in ec.HAND_CHANNEL]), OP_BRIDGE)
    assert rc == 0 and got == [c + w for _, c, w in ec.HAND_CHANNEL]
    assert list(chan) == [len(c) for _, c, _ in ec.HAND_CHANNEL]


@pytest.mark.parametrize("n", [0, 1, 127, 128, 129, 5000])
@pytest.mark.parametrize("op", [0, OP_GO122, OP_BRIDGE])
def test_random_messages_vs_oracle_and_model(eng, oracle, n, op):
    msgs = ec.random_messages(n, seed=100 + n + op)
    cols = ec.columns(msgs)
    got, chan, rc = gpu_encode(eng, cols, op)
    assert rc == 0
    want, want_chan = oracle.encode(cols, go122=bool(op & OP_GO122), with_channel=bool(op & OP_BRIDGE))
    assert got == want
    assert np.array_equal(chan, want_chan)
    model = [(ec.model_channel(m) if op & OP_BRIDGE else b"") + ec.model_encode(m, bool(op & OP_GO122)) for m in msgs]
    assert got == model


@pytest.mark.parametrize("op", [OP_RESP, OP_RESP | OP_GO122, OP_RESP | OP_BRIDGE])
def test_resp_publish_framing(eng, oracle, op):
    """SSEGPU_OP_RESP: records are the RESP2 PUBLISH commands go-redis writes (llm-stream-proxy/main.go:244)."""
    got, chan, rc = gpu_encode(eng, ec.columns([m for m, _ in ec.HAND_RESP]), OP_RESP)
    assert rc == 0 and got == [w for _, w in ec.HAND_RESP]
    msgs = ec.random_messages(5000, seed=31 + op)
    cols = ec.columns(msgs)
    got, chan, rc = gpu_encode(eng, cols, op)
    want, want_chan = oracle.encode_resp(cols, go122=bool(op & OP_GO122))
    assert rc == 0 and got == want and np.array_equal(chan, want_chan)
    assert got == [ec.model_resp(m, bool(op & OP_GO122)) for m in msgs]


def test_overflow_is_reported_and_nothing_is_written_past_the_cap(eng):
    msgs = ec.random_messages(3000, seed=9)
    cols = ec.columns(msgs)
    need = sum(len(ec.model_encode(m)) for m in msgs)
    _, _, rc = gpu_encode(eng, cols, 0, cap=need // 2)
    assert rc == -6                                  # SSEGPU_E_OVERFLOW
    _, _, rc = gpu_encode(eng, cols, 0, cap=need + 16 * (len(msgs) // 128 + 1))
    assert rc == 0


def test_argument_checks(eng, pkg):
    with pytest.raises(pkg.SseGpuError) as err:
        eng.encode_launch(0, 0, 0, 0, 0, 0, 0, 5, 0, 0, 0, 0, 0)
    assert err.value.code == -1
    with pytest.raises(pkg.SseGpuError) as err:       # only BRIDGE and GO122 mean anything here
        eng.encode_launch(4, 4, 4, 4, 4, 4, 4, 5, 8, 16, 100, 4, 4)
    assert err.value.code == -1
    with pytest.raises(pkg.SseGpuError) as err:
        eng.encode_launch(4, 4, 4, 4, 4, 4, 4, (1 << 20) + 1, 0, 16, 100, 4, 4)
    assert err.value.code == -5


def test_encode_then_adapter_on_the_device(pkg, oracle):
    """producer -> consumer without leaving HBM: 2000 conversations x 100 tokens encoded on the GPU, the encoded arena
    framed by the adapter kernel in ordered mode; every frame == header + payload, all on the canonical fast lane."""
    import torch
    n_conv, n_tok = 2000, 100
    n = n_conv * n_tok
    words = [b" the", b" of", b"Hello", b",", b" world", b".", b"\n", b" caf\xc3\xa9", b" <b>", b"\"", b" token"]
    rng = np.random.default_rng(4)
    pick = rng.integers(0, len(words), n)
    msgs = [(b"conv-%06d" % (i % n_conv), words[pick[i]], i // n_conv + 1, i // n_conv == n_tok - 1, 1700000000000000000 + i) for i in range(n)]
    cols = ec.columns(msgs)
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=64 << 20, ring_slots=1)
    try:
        recs, _, rc, d = gpu_encode(eng, cols, 0, keep=True)
        assert rc == 0
        want, _ = oracle.encode(cols)
        assert recs == want
        # the encoder's out_off/out_len are not a prefix array (tiles land in completion order), so compact on the host
        # side of the test only to build the adapter's input offsets; the bytes stay the device's.
        arena, off = oracle.pack(recs)
        d_in = torch.from_numpy(arena).cuda()
        d_off = torch.from_numpy(off.view(np.int32)).cuda()
        cap = int(off[n]) + 48 * n
        d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        d_out_off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        d_out_len = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_status = torch.zeros(n, dtype=torch.int32, device="cuda")
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        eng.frame_launch(d_in.data_ptr(), d_off.data_ptr(), 0, n, 8, d_out.data_ptr(), cap, d_out_off.data_ptr(), d_out_len.data_ptr(), 0,
                         d_status.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        rc, c = eng.launch_result()
        assert rc == 0 and c["frames"] == n and c["fast_path"] == n and c["done"] == n_conv
        out = d_out.cpu().numpy().tobytes()
        oo = d_out_off.cpu().numpy()
        for i in list(range(0, n, 997)) + [n - 1]:
            assert out[int(oo[i]):int(oo[i + 1])] == b"event: token\nid: %d\ndata: " % msgs[i][2] + recs[i] + b"\n\n"
    finally:
        eng.close()


def test_encode_reproduces_the_baseline_workload_byte_for_byte(pkg, oracle):
    """Full-shape property: the config-1 workload (10 000 conversations x 200 tokens, UUID ids, 1.5 % escapes, 1 % multi-byte
    UTF-8) as columns -> GPU encoder -> must be the workload's payload arena itself; the columns are also what the oracle
    decodes from that arena."""
    n_conv, n_tok = 10000, 200
    arena, off = pkg.workload.generate(1, n_conv, n_tok)
    n = n_conv * n_tok
    cols = pkg.workload.generate_columns(1, n_conv, n_tok)
    for got, want in zip(cols, oracle.decode_columns(arena, off)):
        k = len(want) if want.dtype != np.uint8 or len(want) == n else len(want) - 32          # byte arenas: ignore the padding
        assert np.array_equal(got[:k], want[:k])
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=64 << 20, ring_slots=1)
    try:
        recs, _, rc, d = gpu_encode(eng, cols, 0, keep=True)
        assert rc == 0
        assert b"".join(recs) == arena[: int(off[n])].tobytes()
        assert [len(r) for r in recs[:1000]] == list(np.diff(off[:1001].astype(np.int64)))
    finally:
        eng.close()


def test_long_and_hostile_strings_take_every_path(eng, oracle):
    """ids and tokens longer than one warp trip (32 positions), multi-byte sequences across the trip boundary, tiles that
    overflow the staging tile (emitted serially into the arena) and tiles whose strings overflow the input buffers."""
    import random
    rng = random.Random(77)
    pieces = [b"a", b"\n", b"\"", "é".encode(), "世".encode(), "\U0001f600".encode(), b"\xff", b"\xe4\xb8", b"<", " ".encode()]
    msgs = []
    for i in range(1500):
        k = rng.choice([0, 1, 5, 30, 31, 32, 33, 34, 63, 64, 65, 200, 700])
        tok = b"".join(rng.choice(pieces) for _ in range(k))
        cid = b"".join(rng.choice(pieces) for _ in range(rng.choice([0, 3, 36, 40, 100])))
        msgs.append((cid, tok, i, i % 7 == 0, 1700000000000000000 + i))
    cols = ec.columns(msgs)
    for op in (0, OP_GO122, OP_BRIDGE, OP_RESP):
        got, chan, rc = gpu_encode(eng, cols, op)
        assert rc == 0
        if op & OP_RESP:
            want = [ec.model_resp(m) for m in msgs]
        else:
            want = [(ec.model_channel(m) if op & OP_BRIDGE else b"") + ec.model_encode(m, bool(op & OP_GO122)) for m in msgs]
        assert got == want
