"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against the hand-typed
vectors, the committed fixtures and the C oracle on the same seeded inputs.  Bit-exact: bytes, offsets, status words."""
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hand_vectors as hv  # noqa: E402
import make_golden as mg  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    return load_package()


@pytest.fixture(scope="module")
def eng(pkg):
    e = pkg.SseGpu(device=0, max_records=1 << 20, max_in_bytes=200 << 20, ring_slots=2)
    yield e
    e.close()


from gpu_util import run_batch, slices, normalise  # noqa: E402


def test_hand_vectors(pkg, eng, oracle):
    arena, off = oracle.pack([v[1] for v in hv.ADAPTER])
    start = np.array([v[2] for v in hv.ADAPTER], dtype=np.int64)
    for op in (0, pkg.OP_NO_FAST):
        out, out_off, status, _ = run_batch(eng, arena, off, start, op)
        for (name, _, _, st, frame), g, s in zip(hv.ADAPTER, slices(out, out_off), status):
            assert int(s) == st, name
            assert g == frame, name
    g5 = [v[1] for v in hv.ADAPTER if v[0] == "G5_escapes_go121"]
    arena, off = oracle.pack(g5)
    out, out_off, _, _ = run_batch(eng, arena, off, None, pkg.OP_GO122)
    assert slices(out, out_off)[0] == hv.G5_GO122_FRAME
    arena, off = oracle.pack([b[0] for b in hv.BRIDGE])
    out, out_off, status, subj = run_batch(eng, arena, off, None, pkg.OP_BRIDGE)
    for (payload, st, sj, js), g, sl, s in zip(hv.BRIDGE, slices(out, out_off), subj, status):
        assert int(s) == st and g[: int(sl)] == sj and g[int(sl):] == js


@pytest.mark.parametrize("op", [0, 4, 2, 1, 1 | 4, 8, 8 | 1])
def test_golden_fixtures(pkg, eng, oracle, golden, op):
    arena, off = oracle.pack([bytes.fromhex(c["in"]) for c in golden])
    start = np.array([c["start"] for c in golden], dtype=np.int64)
    out, out_off, status, subj = run_batch(eng, arena, off, None if op & 1 else start, op)
    for c, g, sl, s in zip(golden, slices(out, out_off), subj, status):
        if op & pkg.OP_BRIDGE:
            assert int(s) == c["bstatus"], c["name"]
            assert g[: int(sl)].hex() == c["subject"] and g[int(sl):].hex() == c["json"], c["name"]
        else:
            assert int(s) == c["status"], c["name"]
            assert g.hex() == (c.get("frame122", c["frame"]) if op & pkg.OP_GO122 else c["frame"]), c["name"]


def test_differential_fuzz_vs_oracle(pkg, eng, oracle):
    rng = random.Random(99)
    base = mg.canon("7f3c2b1a-1111-4222-8333-944455556666", " hello", 42, False, 1700000000000000000)
    payloads = []
    for k in range(30000):
        r = rng.random()
        if r < 0.4:
            payloads.append(mg.canon(mg.rand_text(rng, rng.randrange(0, 40), 0.15), mg.rand_text(rng, rng.randrange(0, 20), 0.3),
                                     rng.randrange(-5, 5000), rng.random() < 0.05, rng.randrange(-(1 << 63), 1 << 63)))
        elif r < 0.7:
            payloads.append(mg.structured(rng))
        else:
            payloads.append(mg.mutate(rng, rng.choice([base, payloads[-1] if payloads else base])))
    arena, off = oracle.pack(payloads)
    start = np.array([rng.choice([0, 0, 3, 100]) for _ in payloads], dtype=np.int64)
    for op, go122 in ((0, False), (pkg.OP_GO122, True), (pkg.OP_NO_FAST, False), (pkg.OP_ORDERED, False)):
        want_out, want_off, want_status = oracle.adapter(arena, off, start, go122=go122)
        out, out_off, status, _ = run_batch(eng, arena, off, start, op)
        assert np.array_equal(status, want_status)
        assert np.array_equal(out_off, want_off)
        assert np.array_equal(out, want_out)
    want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
    out, out_off, status, subj = run_batch(eng, arena, off, None, pkg.OP_BRIDGE)
    assert np.array_equal(status, want_status) and np.array_equal(out_off, want_off)
    assert np.array_equal(subj, want_subj) and np.array_equal(out, want_out)


@pytest.mark.parametrize("op", [1 | 128, 1 | 128 | 4, 1 | 128 | 2, 1 | 128 | 8])
def test_nats_pub_framing(pkg, eng, oracle, golden, op):
    """SSEGPU_OP_NATS_PUB: bridge records as complete NATS PUB messages; in ordered mode the arena IS the byte stream
    nats.go would have written for the same sequence of nc.Publish calls (redis-nats-bridge/main.go:164)."""
    arena, off = oracle.pack([v[0] for v in hv.BRIDGE_PUB])
    out, out_off, status, subj = run_batch(eng, arena, off, None, op)
    assert slices(out, out_off) == [v[2] for v in hv.BRIDGE_PUB]
    assert [int(x) for x in status] == [v[1] for v in hv.BRIDGE_PUB]
    rng = random.Random(5)
    payloads = [bytes.fromhex(c["in"]) for c in golden]
    base = mg.canon("7f3c2b1a-1111-4222-8333-944455556666", " hello", 42, False, 1700000000000000000)
    for k in range(8000):
        r = rng.random()
        payloads.append(mg.canon(mg.rand_text(rng, rng.randrange(0, 40), 0.15), mg.rand_text(rng, rng.randrange(0, 200), 0.3),
                                 rng.randrange(-5, 5000), rng.random() < 0.05, rng.randrange(-(1 << 63), 1 << 63)) if r < 0.6
                        else mg.mutate(rng, rng.choice([base, payloads[-1]])))
    arena, off = oracle.pack(payloads)
    want_out, want_off, want_subj, want_status = oracle.bridge_pub(arena, off, go122=bool(op & 2))
    out, out_off, status, subj = run_batch(eng, arena, off, None, op)
    assert np.array_equal(status, want_status) and np.array_equal(subj, want_subj)
    assert np.array_equal(out_off, want_off) and np.array_equal(out, want_out)


def test_nats_pub_needs_bridge_mode(pkg, eng, oracle):
    arena, off = oracle.pack([b"{}"])
    b = eng.acquire()
    try:
        b.fill(arena, off, None)
        for op in (128, 256, 1 | 256):
            with pytest.raises(pkg.SseGpuError) as err:
                b.submit(1, op, use_filter=False)
            assert err.value.code == -1
    finally:
        b.release()


@pytest.mark.parametrize("pct", [0.05, 1.0, 20.0])
@pytest.mark.parametrize("op", [0, 1, 1 | 128, 8, 8 | 1, 8 | 1 | 128])
def test_sparse_non_canonical_records_in_canonical_traffic(pkg, eng, oracle, pct, op):
    """Mostly Go-marshalled traffic with a few records that need the general decoder (a "Token" key, a payload with a
    space after the brace, a syntax error, a type error): in arrival mode the tile kernel leaves them to its tail phase
    when they are few in their warp, in ordered mode both passes do (the count pass's tail adds their lengths to the tile
    totals, the emit pass leaves holes that its tail fills) -- statuses, frames, counters and the arena invariants must not notice."""
    n_conv, n_tok = 2000, 30
    arena, off = pkg.workload.generate(1, n_conv, n_tok)
    n = n_conv * n_tok
    payloads = [arena[int(off[i]):int(off[i + 1])].tobytes() for i in range(n)]
    rng = random.Random(int(pct * 100) + op)
    odd = 0
    for i in range(n):
        if rng.random() < pct / 100.0:
            p, k = payloads[i], rng.randrange(4)
            payloads[i] = (p.replace(b'"token"', b'"Token"') if k == 0 else b"{ " + p[1:] if k == 1 else p[:-1] if k == 2
                           else p.replace(b'"sequence":', b'"sequence":"x",'  b'"q":'))
            odd += 1
    arena, off = oracle.pack(payloads)
    start = None if op & 1 else np.array([rng.choice([0, 0, 0, 7]) for _ in range(n)], dtype=np.int64)
    if op & 128:
        want_out, want_off, want_subj, want_status = oracle.bridge_pub(arena, off)
    elif op & 1:
        want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
    else:
        want_out, want_off, want_status = oracle.adapter(arena, off, start)
        want_subj = None
    before = eng.counters()
    out, out_off, status, subj = run_batch(eng, arena, off, start, op)
    after = eng.counters()
    assert np.array_equal(status, want_status)
    assert np.array_equal(out_off, want_off) and np.array_equal(out, want_out)
    if want_subj is not None:
        assert np.array_equal(subj, want_subj)
    code = want_status & 0xF
    assert after["frames"] - before["frames"] == int((code == 0).sum())
    assert after["err_syntax"] - before["err_syntax"] == int((code == 1).sum())
    assert after["err_type"] - before["err_type"] == int((code == 2).sum())
    assert after["filtered"] - before["filtered"] == int((code == 3).sum())
    assert after["done"] - before["done"] == int(((want_status & 0x100) != 0).sum())
    assert after["bytes_out"] - before["bytes_out"] == int(want_off[n])
    assert after["fast_path"] - before["fast_path"] <= n - odd + int((code == 3).sum())


@pytest.mark.parametrize("op", [0, 8])
@pytest.mark.parametrize("n", [0, 1, 2, 127, 128, 129, 255, 256, 257, 1000])
def test_tile_boundaries(pkg, eng, oracle, n, op):
    arena, off = pkg.workload.generate(1, max(n, 1), 1)
    off = off[: n + 1].copy()
    want_out, want_off, want_status = oracle.adapter(arena, off)
    out, out_off, status, _ = run_batch(eng, arena, off, None, op)
    assert np.array_equal(status, want_status) and np.array_equal(out_off, want_off) and np.array_equal(out, want_out)


@pytest.mark.parametrize("op", [0, 8, 1])
@pytest.mark.parametrize("shift", [0, 1, 2, 3])
def test_offset_array_alignment(pkg, oracle, op, shift):
    """The kernel brings a tile's record offsets on chip with a bulk copy when the offset array is 16-byte aligned and loads
    them per thread otherwise: a device-resident launch with in_off at every word alignment, several tiles and a ragged last one."""
    import torch
    n = 128 * 9 + 77
    arena, off = pkg.workload.generate(1, 50, (n + 49) // 50)
    off = off[: n + 1].copy()
    if op & 1:
        want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
    else:
        want_out, want_off, want_status = oracle.adapter(arena, off)
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=int(off[n]) + 64)
    try:
        d_in = torch.from_numpy(arena).cuda()
        d_off_buf = torch.zeros(n + 1 + 8, dtype=torch.int32, device="cuda")
        d_off = d_off_buf[shift: shift + n + 1]
        d_off.copy_(torch.from_numpy(off.view(np.int32)))
        assert d_off.data_ptr() % 16 == 4 * shift
        cap = int(off[n]) + 64 * n
        d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        d_out_off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        d_out_len = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_subj = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_status = torch.zeros(n, dtype=torch.int32, device="cuda")
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        eng.frame_launch(d_in.data_ptr(), d_off.data_ptr(), 0, n, op, d_out.data_ptr(), cap, d_out_off.data_ptr(), d_out_len.data_ptr(),
                         d_subj.data_ptr(), d_status.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        rc, c = eng.launch_result()
        assert rc == 0 and c["frames"] == n
        out, out_off = normalise(d_out.cpu().numpy(), d_out_off.cpu().numpy().astype(np.uint64), d_out_len.cpu().numpy().astype(np.uint32), bool(op & 8))
        assert np.array_equal(d_status.cpu().numpy().astype(np.uint32), want_status)
        assert np.array_equal(out_off, want_off) and np.array_equal(out[: int(want_off[n])], want_out)
    finally:
        eng.close()


def test_empty_and_ragged_records(pkg, eng, oracle):
    payloads = [b"", b"", b"null", b"", b"x" * 5, b"", mg.canon("c", "t", 1, False, 2), b""] * 40 + [b""] * 300
    arena, off = oracle.pack(payloads)
    want_out, want_off, want_status = oracle.adapter(arena, off)
    out, out_off, status, _ = run_batch(eng, arena, off)
    assert np.array_equal(status, want_status) and np.array_equal(out_off, want_off) and np.array_equal(out, want_out)


def test_oversize_tiles_use_global_path(pkg, eng, oracle):
    """Tiles whose bytes exceed the shared-memory caps (input and/or output side) stay bit-exact."""
    rng = random.Random(5)
    payloads = []
    for k in range(600):
        r = rng.random()
        if r < 0.05:
            payloads.append(mg.canon("c", "x" * rng.randrange(5000, 40000), k, False, 2))
        elif r < 0.1:
            payloads.append(mg.canon("c", "<\n&" * rng.randrange(1000, 3000), k, False, 2))      # output 4x input
        elif r < 0.15:
            payloads.append(b'{"pad":"' + b"p" * rng.randrange(1000, 30000) + b'","token":"t"}')  # big in, tiny out
        else:
            payloads.append(mg.canon("conv-%d" % k, " tok", k, False, 2))
    arena, off = oracle.pack(payloads)
    for op in (0, pkg.OP_BRIDGE, pkg.OP_ORDERED):
        if not op & pkg.OP_BRIDGE:
            want_out, want_off, want_status = oracle.adapter(arena, off)
        else:
            want_out, want_off, _, want_status = oracle.bridge(arena, off)
        out, out_off, status, _ = run_batch(eng, arena, off, None, op)
        assert np.array_equal(status, want_status) and np.array_equal(out_off, want_off) and np.array_equal(out, want_out)


@pytest.mark.parametrize("op", [8, 8 | 1, 8 | 1 | 128, 0])
def test_ordered_plan_word_field_limits(pkg, eng, oracle, op):
    """Ordered mode's counting pass packs a canonical record's plan into 32 bits (start of the sequence literal < 8192, end of
    the id < 2048).  Go-marshalled records beyond those widths that still fit the shared-memory tile must take the other
    route (decoded again by the emit pass or its tail) and come out the same."""
    payloads = []
    for k in range(128 * 6):
        if k % 128 == 5:
            payloads.append(mg.canon("i" * 2100, "t", k, False, 2))            # id ends beyond 2047
        elif k % 128 == 77:
            payloads.append(mg.canon("c", "y" * 8200, k, k % 256 == 77, 2))    # sequence literal starts beyond 8191
        elif k % 128 == 100:
            payloads.append(mg.canon("j" * 2020, "t", k, False, 2))            # just inside / outside the id limit
        else:
            payloads.append(mg.canon("c", "t", k, False, 2))
    arena, off = oracle.pack(payloads)
    if op & 128:
        want_out, want_off, want_subj, want_status = oracle.bridge_pub(arena, off)
    elif op & 1:
        want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
    else:
        want_out, want_off, want_status = oracle.adapter(arena, off)
        want_subj = None
    out, out_off, status, subj = run_batch(eng, arena, off, None, op)
    assert np.array_equal(status, want_status) and np.array_equal(out_off, want_off) and np.array_equal(out, want_out)
    if want_subj is not None:
        assert np.array_equal(subj, want_subj)


@pytest.mark.parametrize("op", [0, 8])
@pytest.mark.parametrize("config,n_conv,n_tok", [(0, 1000, 512), (1, 10000, 100)])
def test_baseline_workloads_bit_exact(pkg, eng, oracle, config, n_conv, n_tok, op):
    arena, off = pkg.workload.generate(config, n_conv, n_tok)
    want_out, want_off, want_status = oracle.adapter(arena, off)
    before = eng.counters()
    out, out_off, status, _ = run_batch(eng, arena, off, None, op)
    after = eng.counters()
    assert np.array_equal(status, want_status)
    assert np.array_equal(out_off, want_off)
    assert np.array_equal(out, want_out)
    n = n_conv * n_tok
    assert after["frames"] - before["frames"] == n and after["fast_path"] - before["fast_path"] == n
    assert after["done"] - before["done"] == n_conv
    assert after["bytes_in"] - before["bytes_in"] == int(off[n]) and after["bytes_out"] - before["bytes_out"] == int(want_off[n])


def test_output_overflow_is_reported_not_truncated(pkg, oracle):
    arena, off = pkg.workload.generate(1, 1000, 4)
    e = pkg.SseGpu(device=0, max_records=4096, max_in_bytes=1 << 20, max_out_bytes=100000, ring_slots=1)
    try:
        b = e.acquire()
        b.fill(arena, off)
        b.submit(len(off) - 1, 0)
        with pytest.raises(pkg.SseGpuError) as err:
            b.wait()
        assert err.value.code == -6
        b.release()
    finally:
        e.close()


@pytest.mark.parametrize("op", [0, 8])
def test_device_resident_launch_full_size_properties(pkg, oracle, op):
    """BASELINE config-1 shape at a size the GPU does in milliseconds: checked by size-independent properties
    (offsets == prefix sum of per-record lengths; every frame == header + payload + trailer) and against the
    oracle on a sampled window."""
    import torch
    n_conv, n_tok = 10000, 200
    arena, off = pkg.workload.generate(1, n_conv, n_tok)
    n = n_conv * n_tok
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=int(off[n]) + 64)
    try:
        d_in = torch.from_numpy(arena).cuda()
        d_off = torch.from_numpy(off.view(np.int32)).cuda()
        cap = int(off[n]) + 48 * n
        d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        d_out_off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        d_out_len = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_status = torch.zeros(n, dtype=torch.int32, device="cuda")
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        st = stream.cuda_stream
        for _ in range(2):          # second launch reuses the zeroed scratch: must be idempotent
            eng.frame_launch(d_in.data_ptr(), d_off.data_ptr(), 0, n, op, d_out.data_ptr(), cap, d_out_off.data_ptr(),
                             d_out_len.data_ptr(), 0, d_status.data_ptr(), st)
        torch.cuda.synchronize()
        rc, c = eng.launch_result()
        assert rc == 0 and c["frames"] == n and c["fast_path"] == n and c["done"] == n_conv
        status = d_status.cpu().numpy().astype(np.uint32)
        seq_digits = np.char.str_len((np.arange(n) // n_conv + 1).astype(str))
        lens = np.diff(off.astype(np.int64)) + 26 + seq_digits
        assert np.array_equal(d_out_len.cpu().numpy().astype(np.int64), lens)
        out, out_off = normalise(d_out.cpu().numpy(), d_out_off.cpu().numpy().astype(np.uint64),
                                 d_out_len.cpu().numpy().astype(np.uint32), bool(op & 8))
        assert c["bytes_out"] == int(out_off[n])
        assert ((status & 0xF) == 0).all() and int((status & 0x100 != 0).sum()) == n_conv
        lo, hi = n // 2 - 5000, n // 2 + 5000
        sub_off = (off[lo:hi + 1] - off[lo]).astype(np.uint32)
        sub_arena = np.concatenate([arena[off[lo]:off[hi]], np.zeros(16, np.uint8)])
        want_out, want_off, _ = oracle.adapter(sub_arena, sub_off)
        assert np.array_equal(out[int(out_off[lo]):int(out_off[hi])], want_out)
        # checksum of checksums: the payload bytes survive verbatim inside the frames
        assert int(out[: int(out_off[n])].astype(np.uint64).sum()) == int(arena[: int(off[n])].astype(np.uint64).sum()) + \
            sum(b"event: token\nid: \ndata: \n\n") * n + int(sum((ord(ch) for s in (np.arange(n_tok) + 1).astype(str) for ch in s))) * n_conv
    finally:
        eng.close()


@pytest.mark.parametrize("op", [0, 8])
def test_baseline_config1_full_size_bit_exact(pkg, oracle, op):
    """The BASELINE workload at its full size -- 10 000 conversations x 1 000 tokens = 10 M records in one launch -- against the
    oracle over EVERY record: statuses, lengths and all 1.6 GB of frames (arrival mode after the placement check, ordered mode
    as the arena lies)."""
    import torch
    n_conv, n_tok = 10000, 1000
    arena, off = pkg.workload.generate(1, n_conv, n_tok)
    n = n_conv * n_tok
    want_out, want_off, want_status = oracle.adapter(arena, off)
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=int(off[n]) + 64, ring_slots=1)
    try:
        d_in = torch.from_numpy(arena).cuda()
        d_off = torch.from_numpy(off.view(np.int32)).cuda()
        cap = int(off[n]) + 48 * n
        d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        d_out_off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        d_out_len = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_status = torch.zeros(n, dtype=torch.int32, device="cuda")
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        eng.frame_launch(d_in.data_ptr(), d_off.data_ptr(), 0, n, op, d_out.data_ptr(), cap, d_out_off.data_ptr(), d_out_len.data_ptr(), 0,
                         d_status.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        rc, c = eng.launch_result()
        assert rc == 0 and c["frames"] == n and c["fast_path"] == n and c["bytes_out"] == int(want_off[n])
        assert np.array_equal(d_status.cpu().numpy().astype(np.uint32), want_status)
        out, out_off = normalise(d_out.cpu().numpy(), d_out_off.cpu().numpy().astype(np.uint64), d_out_len.cpu().numpy().astype(np.uint32), bool(op & 8))
        assert np.array_equal(out_off, want_off)
        assert np.array_equal(out[: int(want_off[n])], want_out)
    finally:
        eng.close()


def test_concurrent_submitters_on_one_context(pkg, oracle):
    """The ABI's thread model: every entry point may be called from any OS thread, a batch belongs to one thread between
    acquire and release.  Four threads hammer one context (ctypes drops the GIL inside the calls), each on its own slots,
    mixing modes; every result must be the oracle's and the context totals must add up."""
    import threading
    n_conv, n_tok = 400, 8
    arena, off = pkg.workload.generate(1, n_conv, n_tok)
    n = n_conv * n_tok
    want = {0: oracle.adapter(arena, off), 8: oracle.adapter(arena, off), 1: oracle.bridge(arena, off)}
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=int(off[n]) + 64, ring_slots=4)
    errors = []

    def worker(k):
        try:
            for it in range(40):
                op = (0, 8, 1)[(k + it) % 3]
                while True:
                    try:
                        b = eng.acquire()
                        break
                    except pkg.SseGpuError as e:          # ring momentarily empty
                        assert e.code == -4
                b.fill(arena, off, None)
                b.submit(n, op)
                b.wait()
                out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), bool(op & 8))
                assert np.array_equal(b.status, want[op][-1])
                assert np.array_equal(out_off, want[op][1]) and np.array_equal(out, want[op][0])
                b.release()
        except BaseException as e:                        # noqa: BLE001 -- reported by the main thread
            errors.append((k, repr(e)))

    try:
        before = eng.counters()
        threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors
        after = eng.counters()
        assert after["records"] - before["records"] == 4 * 40 * n and after["frames"] - before["frames"] == 4 * 40 * n
        assert after["batches"] - before["batches"] == 160
    finally:
        eng.close()


def test_malformed_offsets_are_refused_not_followed(pkg, oracle):
    """in_off that is not monotonic or runs past in_off[n] must neither crash nor read outside the arena."""
    arena, off = pkg.workload.generate(1, 300, 2)
    n = len(off) - 1
    e = pkg.SseGpu(device=0, max_records=4096, max_in_bytes=1 << 20, ring_slots=1)
    try:
        for mutate in (lambda o: o.__setitem__(100, o[100] + 5000000), lambda o: o.__setitem__(200, 3),
                       lambda o: o.__setitem__(slice(50, 60), o[50:60][::-1].copy())):
            bad = off.copy()
            mutate(bad)
            b = e.acquire()
            b.fill(arena, off)
            b.in_off[: n + 1] = bad
            b.submit(n, 0)
            with pytest.raises(pkg.SseGpuError) as err:
                b.wait()
            assert err.value.code == -1
            b.release()
        # and a good batch afterwards is processed normally
        out, out_off, status, _ = run_batch(e, arena, off)
        want_out, want_off, want_status = oracle.adapter(arena, off)
        assert np.array_equal(out, want_out) and np.array_equal(status, want_status)
    finally:
        e.close()


def test_api_error_paths(pkg, oracle):
    """The ABI reports misuse with error codes instead of undefined behaviour."""
    arena, off = pkg.workload.generate(1, 50, 2)
    n = len(off) - 1
    e = pkg.SseGpu(device=0, max_records=64, max_in_bytes=1 << 16, ring_slots=1)
    try:
        b = e.acquire()
        with pytest.raises(pkg.SseGpuError) as err:          # ring of one slot is exhausted
            e.acquire()
        assert err.value.code == -4
        with pytest.raises(pkg.SseGpuError) as err:          # wait before submit
            b.wait()
        assert err.value.code == -7
        b.fill(arena, off[:61].copy())
        with pytest.raises(pkg.SseGpuError) as err:          # more records than the slot holds
            b.submit(n, 0)
        assert err.value.code == -5
        with pytest.raises(pkg.SseGpuError) as err:          # inspection requested without a loaded pattern set
            b.submit(32, pkg.OP_INSPECT)
        assert err.value.code == -7
        with pytest.raises(pkg.SseGpuError) as err:          # interest lookup requested without a table
            b.submit(32, pkg.OP_MATCH)
        assert err.value.code == -7
        b.submit(32, 0)                                      # and the slot is still usable afterwards
        b.wait()
        want_out, want_off, want_status = oracle.adapter(arena, off[:33].copy())
        assert np.array_equal(b.status, want_status) and int(b.out_off[32]) >= int(want_off[32])
        b.release()
        b2 = e.acquire()                                     # released slot can be acquired again
        b2.release()
    finally:
        e.close()
    with pytest.raises(pkg.SseGpuError) as err:              # no such device
        pkg.SseGpu(device=63, max_records=16, max_in_bytes=4096)
    assert err.value.code == -2


@pytest.mark.parametrize("op", [0, 8, 1, 1 | 8, 1 | 128, 1 | 128 | 8])
def test_header_only_egress_reassembles_to_the_same_bytes(pkg, eng, oracle, golden, op):
    """SSEGPU_OP_HEADER_ONLY: a record that is exactly json.Marshal's output comes back as its header only (status bit
    SSEGPU_ST_HEADER_ONLY); header + payload + trailer must be the frame the full mode and the oracle produce."""
    rng = random.Random(5)
    payloads = [bytes.fromhex(c["in"]) for c in golden]
    arena0, off0 = pkg.workload.generate(1, 300, 7)
    payloads += [arena0[off0[i]:off0[i + 1]].tobytes() for i in range(len(off0) - 1)]
    rng.shuffle(payloads)
    arena, off = oracle.pack(payloads)
    n = len(payloads)
    if op & 128:
        want_out, want_off, want_subj, want_status = oracle.bridge_pub(arena, off)
    elif op & 1:
        want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
    else:
        want_out, want_off, want_status = oracle.adapter(arena, off)
    want = slices(want_out, want_off)
    out, out_off, status, subj = run_batch(eng, arena, off, None, op | pkg.OP_HEADER_ONLY)
    got = slices(out, out_off)
    trailer = b"\r\n" if op & 128 else (b"" if op & 1 else b"\n\n")
    n_hdr = 0
    for i in range(n):
        st = int(status[i])
        assert (st & ~0x400) == int(want_status[i]), i
        if st & 0x400:
            n_hdr += 1
            assert got[i] + payloads[i] + trailer == want[i], i
        else:
            assert got[i] == want[i], i
        if op & 1 and (st & 0xF) == 0:
            k = 4 if op & 128 else 0
            assert want[i][k:k + int(subj[i])].startswith(b"chat.") and int(subj[i]) == int(want_subj[i])
    assert n_hdr >= 1500        # (Go-marshalled records that share a tile with a giant record go the general way and come back whole)
