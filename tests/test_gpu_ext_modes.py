"""GPU parity tests of the optional stages (content inspection, interest lookup) in the arena layouts and entry points
round 1 left untested: SSEGPU_OP_ORDERED with inspection (verdict-only and inline, per-token and streaming) and with
the interest table; the device-resident entry (ssegpu_inspect_bind / ssegpu_interest_bind + ssegpu_frame_launch); the
interest table at BASELINE config 3's shape (100 000 subjects, Zipf fan-out, wildcard consumers, 10 % unsubscribed).
Expectations come from the brute-force inspection oracle, the Python model of the Go path and the NATS subject model;
the hook these flags stand for is the comment block at src/sse-adapter/sse_handler.go:193-208."""
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import interest_case  # noqa: E402
import make_golden as mg  # noqa: E402
import nats_model  # noqa: E402
import pymodel  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import normalise, slices  # noqa: E402
from inspect_sets import SET_A, SET_B  # noqa: E402
from test_gpu_inspect import odd_payload  # noqa: E402
from test_inspect_cpu import rand_token  # noqa: E402

pytestmark = pytest.mark.gpu
ORD = 8


@pytest.fixture(scope="module")
def pkg():
    return load_package()


def run(eng, arena, off, op, conv=None):
    b = eng.acquire()
    b.fill(arena, off, None, conv)
    b.submit(len(off) - 1, op)
    b.wait()
    out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), bool(op & ORD))
    res = dict(out=out, off=out_off, status=b.status.copy(), verdict=b.verdict.copy(), match=b.match.copy(),
               hit=b.sub_span.copy(), subj=b.subj_len.copy())
    b.release()
    return res


def expected_inspect(insp, payloads, inline, bridge=False, stream_conv=None, hist=None):
    """-> (status list, frame list, (verdict, begin, end) list) the way sse_handler.go:200-207 would act on the oracle's verdicts.
    stream_conv: per-record conversation index (streaming mode; `hist` carries each conversation's text so far)."""
    sts, frames, verdicts = [], [], []
    for i, p in enumerate(payloads):
        st, m = pymodel.unmarshal(p)
        if st != 0:
            sts.append(st); frames.append(b""); verdicts.append((0, 0, 0))
            continue
        text = m["tok"].encode("utf-8")
        if stream_conv is None:
            v = insp.token(text)
        else:
            c = stream_conv[i]
            s = len(hist[c])
            hist[c] += text
            v = insp.stream(hist[c], s, len(hist[c]), m["done"])
            if m["done"]:
                hist[c] = b""
        verdicts.append(v)
        act = v[0] & 0xFF
        flags = 0x100 if m["done"] else 0
        if inline and act == 2:
            sts.append(4); frames.append(b"")
            continue
        if inline and act == 1:
            m = dict(m, tok="[REDACTED]")
            flags |= 0x200
        sts.append(flags)
        frames.append((b"chat." + m["cid"].encode("utf-8") + b".tokens" + pymodel.marshal(m)) if bridge else pymodel.frame(m))
    return sts, frames, verdicts


def mixed_payloads(rng, n, n_conv=None, odd=0.06):
    """mostly Go-marshalled records with planted pattern fragments, a sprinkling of odd spellings and a few drops"""
    convs, payloads = [], []
    for i in range(n):
        c = rng.randrange(n_conv) if n_conv else i
        t = rand_token(rng).decode("utf-8", "replace")
        done = rng.random() < 0.03
        r = rng.random()
        if r < odd:
            p = odd_payload(rng, "conv-%d" % c, t, i, done, 5)
        elif r < odd + 0.01:
            p = rng.choice([b'{"token":1}', b"{}x", b"null", b'{"sequence":1.5}'])
        else:
            p = mg.canon("conv-%d" % c, t, i, done, 1700000000000000000 + i)
        convs.append(c)
        payloads.append(p)
    return convs, payloads


@pytest.mark.parametrize("inline", [False, True])
@pytest.mark.parametrize("ordered", [False, True])
@pytest.mark.parametrize("bridge", [False, True])
def test_per_token_inspection_in_every_layout(pkg, oracle, inline, ordered, bridge):
    rng = random.Random(1000 + 4 * inline + 2 * ordered + bridge)
    pats = SET_A + SET_B
    insp = oracle.inspector(pats)
    _, payloads = mixed_payloads(rng, 6000)
    arena, off = oracle.pack(payloads)
    eng = pkg.SseGpu(device=0, max_records=1 << 14, max_in_bytes=4 << 20, ring_slots=1)
    try:
        eng.load_patterns(pats, pkg.INSPECT_PER_TOKEN)
        op = (pkg.OP_INSPECT_INLINE if inline else pkg.OP_INSPECT) | (ORD if ordered else 0) | (pkg.OP_BRIDGE if bridge else 0)
        r = run(eng, arena, off, op)
        sts, frames, verdicts = expected_inspect(insp, payloads, inline, bridge)
        assert [int(s) for s in r["status"]] == sts
        assert slices(r["out"], r["off"]) == frames
        got = [(int(v), int(b), int(e)) for v, (b, e) in zip(r["verdict"], r["match"])]
        assert got == verdicts
        acts = [v[0] & 0xFF for v in verdicts]
        assert acts.count(1) > 30 and acts.count(2) > 30 and acts.count(3) > 5
    finally:
        eng.close()


@pytest.mark.parametrize("inline", [False, True])
@pytest.mark.parametrize("ordered", [False, True])
def test_streaming_inspection_in_every_layout(pkg, oracle, inline, ordered):
    """state carried per conversation slot across batches; within a batch a conversation occurs any number of times"""
    rng = random.Random(2000 + 2 * inline + ordered)
    pats = SET_A + SET_B
    insp = oracle.inspector(pats)
    n_conv = 300
    hist = [b""] * n_conv
    eng = pkg.SseGpu(device=0, max_records=1 << 13, max_in_bytes=2 << 20, ring_slots=1, conv_slots=n_conv)
    try:
        eng.load_patterns(pats, pkg.INSPECT_STREAMING)
        op = (pkg.OP_INSPECT_INLINE if inline else pkg.OP_INSPECT) | (ORD if ordered else 0)
        hits = 0
        for batch in range(6):
            convs, payloads = mixed_payloads(rng, 2500, n_conv)
            arena, off = oracle.pack(payloads)
            r = run(eng, arena, off, op, np.array(convs, dtype=np.uint32))
            sts, frames, verdicts = expected_inspect(insp, payloads, inline, False, convs, hist)
            got = [(int(v), int(b), int(e)) for v, (b, e) in zip(r["verdict"], r["match"])]
            bad = [i for i in range(len(payloads)) if got[i] != verdicts[i]]
            assert not bad, (batch, bad[:5], [(got[i], verdicts[i]) for i in bad[:5]])
            assert [int(s) for s in r["status"]] == sts
            assert slices(r["out"], r["off"]) == frames
            hits += sum(1 for v in verdicts if v[0])
        assert hits > 300
    finally:
        eng.close()


@pytest.mark.parametrize("generic", [True, False])
def test_interest_lookup_ordered_arena(pkg, oracle, generic):
    deltas, live, payloads, subjects = interest_case.build(generic_wildcard=generic)
    arena, off = oracle.pack(payloads)
    n = len(off) - 1
    eng = pkg.SseGpu(device=0, max_records=1 << 13, max_in_bytes=4 << 20, ring_slots=1)
    try:
        eng.interest_apply(deltas)
        r = run(eng, arena, off, pkg.OP_BRIDGE | pkg.OP_MATCH | ORD)
        assert interest_case.check(r["hit"], r["status"], eng.interest_subs(), eng.interest_wild_ids(), live, subjects) > 1000
        want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
        got, want = slices(r["out"], r["off"]), slices(want_out, want_off)
        for i in range(n):
            assert got[i] == (want[i] if int(r["status"][i]) & 0xF == 0 else b"")
    finally:
        eng.close()


class DevBuffers:
    def __init__(self, torch, arena, off, extra=128):
        n = len(off) - 1
        self.n = n
        self.d_in = torch.from_numpy(arena).cuda()
        self.d_off = torch.from_numpy(off.view(np.int32)).cuda()
        self.cap = int(off[n]) * 2 + extra * n + 4096
        z = lambda m, dt: torch.zeros(max(m, 1), dtype=dt, device="cuda")                               # noqa: E731
        self.d_out, self.d_out_off, self.d_out_len = z(self.cap, torch.uint8), z(n + 1, torch.int64), z(n, torch.int32)
        self.d_status, self.d_subj, self.d_verdict = z(n, torch.int32), z(n, torch.int32), z(n, torch.int32)
        self.d_match, self.d_hit = z(2 * n, torch.int32), z(3 * n, torch.int32)

    def launch(self, eng, op, d_conv=0, stream=0):
        eng.inspect_bind(d_conv, self.d_verdict.data_ptr(), self.d_match.data_ptr())
        eng.interest_bind(self.d_hit.data_ptr())
        eng.frame_launch(self.d_in.data_ptr(), self.d_off.data_ptr(), 0, self.n, op, self.d_out.data_ptr(), self.cap,
                         self.d_out_off.data_ptr(), self.d_out_len.data_ptr(), self.d_subj.data_ptr(), self.d_status.data_ptr(), stream)

    def results(self, ordered):
        import torch
        torch.cuda.synchronize()
        n = self.n
        out, out_off = normalise(self.d_out.cpu().numpy(), self.d_out_off.cpu().numpy().astype(np.uint64),
                                 self.d_out_len.cpu().numpy().astype(np.uint32)[:n], ordered)
        return dict(out=out, off=out_off, status=self.d_status.cpu().numpy().astype(np.uint32)[:n],
                    verdict=self.d_verdict.cpu().numpy().astype(np.uint32)[:n], match=self.d_match.cpu().numpy()[:2 * n].reshape(n, 2),
                    hit=self.d_hit.cpu().numpy().astype(np.uint32)[:3 * n].reshape(n, 3), subj=self.d_subj.cpu().numpy()[:n])


@pytest.mark.parametrize("ordered", [False, True])
def test_device_resident_entry_with_bound_buffers(pkg, oracle, ordered):
    """ssegpu_inspect_bind / ssegpu_interest_bind + ssegpu_frame_launch: the kernel-level entry the bench uses"""
    import torch
    rng = random.Random(77 + ordered)
    pats = SET_A + SET_B
    insp = oracle.inspector(pats)
    n_conv = 200
    convs, payloads = mixed_payloads(rng, 5000, n_conv)
    arena, off = oracle.pack(payloads)
    o = ORD if ordered else 0
    eng = pkg.SseGpu(device=0, max_records=1 << 13, max_in_bytes=2 << 20, conv_slots=n_conv)
    try:
        db = DevBuffers(torch, arena, off)
        # per-token, inline
        eng.load_patterns(pats, pkg.INSPECT_PER_TOKEN)
        db.launch(eng, pkg.OP_INSPECT_INLINE | o)
        r = db.results(ordered)
        rc, c = eng.launch_result()
        sts, frames, verdicts = expected_inspect(insp, payloads, True)
        assert rc == 0 and [int(s) for s in r["status"]] == sts and slices(r["out"], r["off"]) == frames
        assert [(int(v), int(b), int(e)) for v, (b, e) in zip(r["verdict"], r["match"])] == verdicts
        assert c["frames"] == sum(1 for s in sts if s & 0xF == 0)
        # streaming, verdicts only, the same batch twice: the second launch continues every conversation's text
        eng.load_patterns(pats, pkg.INSPECT_STREAMING)
        d_conv = torch.from_numpy(np.array(convs, dtype=np.int32)).cuda()
        hist = [b""] * n_conv
        for _ in range(2):
            db.launch(eng, pkg.OP_INSPECT | o, d_conv.data_ptr())
            r = db.results(ordered)
            rc, _ = eng.launch_result()
            sts, frames, verdicts = expected_inspect(insp, payloads, False, False, convs, hist)
            assert rc == 0 and [int(s) for s in r["status"]] == sts and slices(r["out"], r["off"]) == frames
            assert [(int(v), int(b), int(e)) for v, (b, e) in zip(r["verdict"], r["match"])] == verdicts
        # interest table through the bound span buffer
        deltas, live, ipay, subjects = interest_case.build(generic_wildcard=False)
        ia, ioff = oracle.pack(ipay)
        eng.interest_apply(deltas)
        db2 = DevBuffers(torch, ia, ioff)
        db2.launch(eng, pkg.OP_BRIDGE | pkg.OP_MATCH | o)
        r = db2.results(ordered)
        assert interest_case.check(r["hit"], r["status"], eng.interest_subs(), eng.interest_wild_ids(), live, subjects) > 1000
    finally:
        eng.close()


def test_interest_table_at_config3_shape(pkg, oracle):
    """100 000 subjects, Zipf(1.1) fan-out 1..1024, the two wildcard consumers of docs/security-inspection-patterns.md:110,268,
    10 % of the subjects without a subscriber; every record's subscriber set against the NATS subject model."""
    import torch
    n_subj, n_tok = 100_000, 2
    rng = np.random.default_rng(0x5EED0004)
    ids = [pkg.workload.conversation_id(1, i) for i in range(n_subj)]
    fan = np.maximum(1, (1024.0 / np.arange(1, n_subj + 1) ** 1.1)).astype(np.int64)
    rng.shuffle(fan)
    listened = rng.random(n_subj) >= 0.10
    deltas, want_ids, sid = [], {}, 1
    for i in range(n_subj):
        if listened[i]:
            subj = b"chat." + ids[i] + b".tokens"
            want_ids[i] = list(range(sid, sid + int(fan[i])))
            for _ in range(int(fan[i])):
                deltas.append((0, sid, subj))
                sid += 1
    wild = [(b"chat.*.tokens", sid), (b"chat.*.control", sid + 1)]
    arena, off = pkg.workload.generate(1, n_subj, n_tok)
    eng = pkg.SseGpu(device=0, max_records=n_subj * n_tok, max_in_bytes=1 << 20)
    try:
        for with_wild in (False, True):
            eng.interest_apply(deltas if not with_wild else [(0, i, s) for s, i in wild])
            db = DevBuffers(torch, arena, off)
            db.launch(eng, pkg.OP_BRIDGE | pkg.OP_MATCH)
            r = db.results(False)
            subs, wid = eng.interest_subs(), eng.interest_wild_ids()
            hit = r["hit"]
            wmask = 0
            for k, w in enumerate(wid):
                pat = [s for s, i in wild if i == int(w)][0]
                if nats_model.subject_matches(pat, b"chat." + ids[0] + b".tokens"):
                    wmask |= 1 << k
            assert (wmask != 0) == with_wild
            for j in range(n_tok):
                h = hit[j * n_subj:(j + 1) * n_subj]
                assert np.array_equal(h[:, 1], np.where(listened, fan, 0).astype(np.uint32))
                assert (h[:, 2] == wmask).all()
                st = r["status"][j * n_subj:(j + 1) * n_subj] & 0xF
                assert np.array_equal(st == 0, listened | with_wild) and np.array_equal(st == 5, ~(listened | with_wild))
            # the id lists themselves, on a sample (every 37th subject) plus the most popular ones
            for i in list(range(0, n_subj, 37)) + [int(x) for x in np.argsort(-fan)[:50]]:
                o, c = int(hit[i, 0]), int(hit[i, 1])
                assert sorted(subs[o:o + c].tolist()) == want_ids.get(i, []), i
            want_out, want_off, _, _ = oracle.bridge(arena, off)
            got, want = slices(r["out"], r["off"]), slices(want_out, want_off)
            for i in range(0, n_subj * n_tok, 101):
                assert got[i] == (want[i] if int(r["status"][i]) & 0xF == 0 else b"")
    finally:
        eng.close()


def test_streaming_batches_on_different_ring_slots_run_in_submit_order(pkg, oracle):
    """A conversation with tokens in two batches that are in flight together: the second batch must start from the state the
    first one commits (streaming batches of a context execute in submit order), so a pattern split across them is found."""
    insp = oracle.inspector(SET_B)
    n_conv = 2000
    eng = pkg.SseGpu(device=0, max_records=1 << 13, max_in_bytes=2 << 20, ring_slots=4, conv_slots=n_conv)
    try:
        eng.load_patterns(SET_B, pkg.INSPECT_STREAMING)
        halves = [("my pass", "word is"), ("ignore pre", "vious now"), ("sec", "ret"), ("plain ", "text")]
        for rep in range(5):
            first = [mg.canon("conv-%d" % c, halves[c % 4][0], 2 * rep + 1, False, 5) for c in range(n_conv)]
            second = [mg.canon("conv-%d" % c, halves[c % 4][1], 2 * rep + 2, True, 5) for c in range(n_conv)]
            conv = np.arange(n_conv, dtype=np.uint32)
            slots = []
            for payloads in (first, second):                      # both submitted before either is waited for
                arena, off = oracle.pack(payloads)
                b = eng.acquire()
                b.fill(arena, off, None, conv)
                b.submit(n_conv, pkg.OP_INSPECT)
                slots.append(b)
            v1 = slots[0].wait().verdict.copy()
            v2 = slots[1].wait().verdict.copy()
            for b in slots:
                b.release()
            assert not v1.any()
            for c in range(n_conv):
                a, z = halves[c % 4]
                text = (a + z).encode()
                assert int(v2[c]) == insp.stream(text, len(a), len(text), True)[0], (rep, c)
            assert (v2[0::4] != 0).all() and (v2[3::4] == 0).all()
    finally:
        eng.close()


def test_fanout_expansion_of_match_spans(pkg, oracle):
    """ssegpu_fanout_launch: (subscriber, record) pairs from the spans of an OP_MATCH launch, literal subscribers in list order
    then the matching wildcard subscriptions; also after in-place table updates (the device copy of the id arena grows with them)."""
    import torch
    rng = random.Random(91)
    n_subj = 3000
    subs_of = {j: [] for j in range(n_subj)}
    deltas, sid = [], 1
    for j in range(n_subj):
        fan = 0 if rng.random() < 0.1 else (1 if rng.random() < 0.7 else rng.randrange(2, 600))
        for _ in range(fan):
            deltas.append((0, sid, b"chat.subj-%d.tokens" % j))
            subs_of[j].append(sid)
            sid += 1
    deltas += [(0, 900001, b"chat.*.tokens"), (0, 900002, b"chat.*.control"), (0, 900003, b"chat.>")]
    eng = pkg.SseGpu(device=0, max_records=1 << 15, max_in_bytes=8 << 20)
    try:
        eng.interest_apply(deltas)
        for step in range(3):
            if step:                                    # a few in-place updates between launches
                for _ in range(40):
                    j = rng.randrange(n_subj)
                    if subs_of[j] and rng.random() < 0.4:
                        x = subs_of[j].pop(rng.randrange(len(subs_of[j])))
                        eng.interest_apply([(1, x, b"chat.subj-%d.tokens" % j)])
                    else:
                        eng.interest_apply([(0, sid, b"chat.subj-%d.tokens" % j)])
                        subs_of[j].append(sid)
                        sid += 1
            picks = [rng.randrange(n_subj + 50) for _ in range(20000)]
            arena, off = oracle.pack([mg.canon("subj-%d" % j, " t", k, False, 5) for k, j in enumerate(picks)])
            db = DevBuffers(torch, arena, off)
            db.launch(eng, pkg.OP_BRIDGE | pkg.OP_MATCH)
            n = db.n
            cap = 4_000_000
            d_sub, d_rec = torch.zeros(cap, dtype=torch.int32, device="cuda"), torch.zeros(cap, dtype=torch.int32, device="cuda")
            d_roff = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
            eng.fanout_launch(db.d_hit.data_ptr(), n, d_sub.data_ptr(), d_rec.data_ptr(), cap, d_roff.data_ptr())
            torch.cuda.synchronize()
            rc, total = eng.fanout_result()
            assert rc == 0
            roff, ps, pr = d_roff.cpu().numpy(), d_sub.cpu().numpy()[:total], d_rec.cpu().numpy()[:total]
            assert int(roff[n]) == total and np.array_equal(pr, np.repeat(np.arange(n), np.diff(roff)))
            wild = eng.interest_wild_ids().tolist()
            for i in list(range(0, n, 97)) + [int(x) for x in np.argsort(-np.diff(roff))[:20]]:
                j = picks[i]
                got = ps[int(roff[i]):int(roff[i + 1])].tolist()
                lit = subs_of.get(j, [])
                assert sorted(got[: len(lit)]) == sorted(lit) and sorted(got[len(lit):]) == sorted(w for w in wild if w in (900001, 900003)), (step, i, j)
            assert total > 200000
            rc, _ = (eng.fanout_launch(db.d_hit.data_ptr(), n, d_sub.data_ptr(), d_rec.data_ptr(), 1000, d_roff.data_ptr()), torch.cuda.synchronize(), eng.fanout_result())[2]
            assert rc == -6
    finally:
        eng.close()
