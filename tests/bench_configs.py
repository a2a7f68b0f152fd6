#!/usr/bin/env python3
"""Measurements for BASELINE.json configs 0, 2, 3 and 4 (bench.py covers config 1, the headline).

  python tests/bench_configs.py [--config 0|2|3|4|all] [--scale S]

One JSON line per measurement.  Device-resident inputs, CUDA events on the launching stream, >= 3 warm-up launches.
config 2  full inspection pattern set (Set A and Set A+B) on 100 000 conversations: per-token mode in one launch and
          streaming mode as one launch per arrival round (a conversation occurs once per batch); matches are planted
          in 0.1 % of the conversations, half of them split across token boundaries, and must all be found
config 3  bridge pipeline decode -> subject match -> reframe at 100 000 subjects, Zipf(1.1) fan-out 1..1024,
          two wildcard consumers, 10 % of the messages addressed to subjects nobody subscribed to
config 4  bursty arrivals: batches cut every 1 ms from an on/off process (on 200 ms at 10x rate, off 800 ms) over the
          conversations one GPU owns out of 1 M on 8 GPUs; p50/p99 submit->wait latency through the pinned ring
"""
import argparse
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from __graft_entry__ import load_package  # noqa: E402
from inspect_sets import SET_A, SET_B  # noqa: E402

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
PLANTS = [b"123-45-6789", b"4111111111111111", b"bob@example.com", b"ignore previous instructions", b"disregard the above",
          b"new instructions:", b"system prompt:"]


def canon(cid, tok, seq, done, ts):
    import pymodel
    return pymodel.marshal({"cid": cid, "tok": tok, "seq": seq, "done": done, "ts": ts})


def time_launches(fn, stream, reps):
    for _ in range(3):
        fn()
    stream.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        fn()
    b.record(stream)
    stream.synchronize()
    return a.elapsed_time(b) / reps


class DevBatch:
    def __init__(self, arena, off, extra_cap=64):
        self.n = len(off) - 1
        self.in_bytes = int(off[self.n])
        self.d_in = torch.from_numpy(arena).cuda()
        self.d_off = torch.from_numpy(off.view(np.int32)).cuda()
        self.cap = self.in_bytes + extra_cap * self.n + 4096
        self.d_out = torch.empty(self.cap, dtype=torch.uint8, device="cuda")
        self.d_out_off = torch.empty(self.n + 1, dtype=torch.int64, device="cuda")
        self.d_out_len = torch.empty(self.n, dtype=torch.int32, device="cuda")
        self.d_status = torch.empty(self.n, dtype=torch.int32, device="cuda")
        self.d_subj = torch.empty(self.n, dtype=torch.int32, device="cuda")
        self.d_verdict = torch.zeros(self.n, dtype=torch.int32, device="cuda")
        self.d_match = torch.zeros(2 * self.n, dtype=torch.int32, device="cuda")
        self.d_hit = torch.zeros(3 * self.n, dtype=torch.int32, device="cuda")

    def launch(self, eng, op, stream, d_conv=0):
        eng.inspect_bind(d_conv, self.d_verdict.data_ptr(), self.d_match.data_ptr())
        eng.interest_bind(self.d_hit.data_ptr())
        eng.frame_launch(self.d_in.data_ptr(), self.d_off.data_ptr(), 0, self.n, op, self.d_out.data_ptr(), self.cap,
                         self.d_out_off.data_ptr(), self.d_out_len.data_ptr(), self.d_subj.data_ptr(), self.d_status.data_ptr(),
                         stream.cuda_stream)


def config0(pkg, scale):
    """configs[0]: the reference's own CPU-runnable case -- 1 000 conversations x 512 load-generator tokens."""
    import oracle_lib
    oracle = oracle_lib.load()
    arena, off = pkg.workload.generate(0, int(1000 * scale), 512)
    n = len(off) - 1
    stream = torch.cuda.Stream()
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=1 << 20)
    db = DevBatch(arena, off)
    ms = time_launches(lambda: db.launch(eng, 0, stream), stream, 10)
    rc, c = eng.launch_result()
    threads = os.cpu_count() or 1
    sec, frames, out_bytes = oracle.bench_adapter(arena, off, threads, reps=3)
    sec1, _, _ = oracle.bench_adapter(arena, off, 1, reps=2)
    # bit-exact against the oracle on the whole config
    want_out, want_off, want_status = oracle.adapter(arena, off)
    from gpu_util import normalise
    out, out_off = normalise(db.d_out.cpu().numpy(), db.d_out_off.cpu().numpy().astype(np.uint64),
                             db.d_out_len.cpu().numpy().astype(np.uint32), False)
    assert np.array_equal(out, want_out) and np.array_equal(db.d_status.cpu().numpy().astype(np.uint32), want_status)
    eng.close()
    return [{"config": 0, "records": n, "payload_bytes": db.in_bytes, "frame_bytes": int(c["bytes_out"]), "gpu_ms_per_launch": ms,
             "gpu_frames_per_sec": n / ms * 1e3, "gpu_alg_GBps": (db.in_bytes + c["bytes_out"] + 16 * n) / ms / 1e6,
             "cpu_oracle_frames_per_sec_all_cores": frames / sec, "cpu_cores": threads, "cpu_oracle_frames_per_sec_1_core": frames / sec1,
             "bit_exact_vs_oracle": True,
             "note": "CPU arm = C restatement of the Go path (no Go toolchain on the box); 66 MB of input fits in L2, so this small config is not an HBM measurement"}]


def config2(pkg, scale):
    n_conv, n_tok = int(100_000 * scale), 32
    n_plant = max(7, n_conv // 1000)
    rng = random.Random(0x5EED0003)
    stream = torch.cuda.Stream()
    # planted conversations: each carries one pattern, whole in one token or split over two/three consecutive tokens
    plant_rows = [[] for _ in range(n_tok)]
    for p in range(n_plant):
        frag = PLANTS[p % len(PLANTS)]
        split = p % 2 == 1
        at = rng.randrange(2, n_tok - 4)
        cut = rng.randrange(1, len(frag)) if split else len(frag)
        for j in range(n_tok):
            tok = b" filler"
            if j == at:
                tok = b" " + frag[:cut]
            elif split and j == at + 1:
                tok = frag[cut:] + b" "
            elif j == at + 1:
                tok = b" after"
            plant_rows[j].append(canon("planted-%06d" % p, tok.decode(), j + 1, j == n_tok - 1, 1700000000000000000 + j))
    rounds = []
    g_arena, g_off = pkg.workload.generate(1, n_conv, n_tok)        # arrival order: round j = records [j*n_conv, (j+1)*n_conv)
    tok_bytes = int(g_off[-1]) - n_conv * n_tok * (66 + 36 + 19 + 5) - n_conv * sum(len(str(j + 1)) for j in range(n_tok))
    for j in range(n_tok):
        lo, hi = j * n_conv, (j + 1) * n_conv
        base = int(g_off[lo])
        off = (g_off[lo:hi + 1] - g_off[lo]).astype(np.uint32)
        rows = b"".join(plant_rows[j])
        lens = np.array([len(r) for r in plant_rows[j]], dtype=np.uint32)
        total = int(off[-1])
        arena2 = np.concatenate([g_arena[base:base + total], np.frombuffer(rows, dtype=np.uint8), np.zeros(64, np.uint8)])
        off2 = np.concatenate([off, total + np.cumsum(lens, dtype=np.uint64).astype(np.uint32)])
        rounds.append((arena2, off2))
    per_round = len(rounds[0][1]) - 1
    conv = torch.arange(per_round, dtype=torch.int32, device="cuda")
    out = []
    for name, pats in (("SetA", SET_A), ("SetA+B", SET_A + SET_B)):
        # --- per-token mode: everything in one launch
        eng = pkg.SseGpu(device=0, max_records=per_round * n_tok, max_in_bytes=1 << 20, conv_slots=per_round)
        eng.load_patterns(pats, pkg.INSPECT_PER_TOKEN)
        big_arena = np.concatenate([r[0][: int(r[1][-1])] for r in rounds] + [np.zeros(64, np.uint8)])
        sizes = np.concatenate([np.diff(r[1].astype(np.int64)) for r in rounds])
        big_off = np.zeros(len(sizes) + 1, dtype=np.uint32)
        big_off[1:] = np.cumsum(sizes).astype(np.uint32)
        db = DevBatch(big_arena, big_off)
        ms = time_launches(lambda: db.launch(eng, pkg.OP_INSPECT, stream), stream, 5)
        ms_plain = time_launches(lambda: db.launch(eng, 0, stream), stream, 5)
        v = db.d_verdict.cpu().numpy()
        rc, c = eng.launch_result()
        out.append({"config": 2, "mode": "per_token", "patterns": name, "records": db.n, "ms_per_launch": ms, "frames_per_sec": db.n / ms * 1e3,
                    "ms_without_inspection": ms_plain, "hits": int((v != 0).sum()), "planted_whole_in_one_token": (n_plant + 1) // 2, "inspected_token_bytes_per_sec": tok_bytes / ms * 1e3,
                    "alg_GBps": (db.in_bytes + c["bytes_out"] + 32 * db.n) / ms / 1e6, "frac_of_measured_hbm": (db.in_bytes + c["bytes_out"] + 32 * db.n) / ms / 1e6 / PEAK})
        eng.close()
        # --- streaming mode: the whole batch in ONE call (every conversation occurs n_tok times in it); the library sorts the
        # records by conversation and walks each conversation's tokens in arrival order with the carried DFA state
        eng = pkg.SseGpu(device=0, max_records=per_round * n_tok, max_in_bytes=1 << 20, conv_slots=per_round)
        eng.load_patterns(pats, pkg.INSPECT_STREAMING)
        conv_all = torch.arange(per_round, dtype=torch.int32, device="cuda").repeat(n_tok)
        ms = time_launches(lambda: db.launch(eng, pkg.OP_INSPECT, stream, conv_all.data_ptr()), stream, 5)
        eng.conv_reset()
        db.launch(eng, pkg.OP_INSPECT, stream, conv_all.data_ptr())
        stream.synchronize()
        rc, c = eng.launch_result()
        v = db.d_verdict.cpu().numpy().reshape(n_tok, per_round)
        hits = (v != 0).any(axis=0)
        found = int(hits[n_conv:].sum())                # planted conversations sit at the end of every round
        out.append({"config": 2, "mode": "streaming", "patterns": name, "records": per_round * n_tok, "kernel_launches_per_batch": eng.launch_count() // 9,
                    "ms_per_batch": ms, "frames_per_sec": per_round * n_tok / ms * 1e3, "inspected_token_bytes_per_sec": tok_bytes / ms * 1e3,
                    "alg_GBps": (db.in_bytes + c["bytes_out"] + 32 * db.n) / ms / 1e6, "frac_of_measured_hbm": (db.in_bytes + c["bytes_out"] + 32 * db.n) / ms / 1e6 / PEAK,
                    "planted": n_plant, "planted_found": found, "false_hits_in_generated_traffic": int(hits[:n_conv].sum())})
        assert rc == 0 and found == n_plant, (rc, found, n_plant)
        eng.close()
    return out


def config3(pkg, scale):
    n_subj, n_tok = int(100_000 * scale), 16
    rng = np.random.default_rng(0x5EED0004)
    stream = torch.cuda.Stream()
    eng = pkg.SseGpu(device=0, max_records=n_subj * n_tok, max_in_bytes=1 << 20)
    ids = [pkg.workload.conversation_id(1, i) for i in range(n_subj)]
    # Zipf(1.1) popularity: the r-th most popular subject has max(1, 1024 / r^1.1) subscribers (1 .. 1024)
    fan = np.maximum(1, (1024.0 / np.arange(1, n_subj + 1) ** 1.1)).astype(np.int64)
    rng.shuffle(fan)
    listened = rng.random(n_subj) >= 0.10                       # 10 % of the subjects have no subscriber
    deltas, sid = [], 1
    for i in range(n_subj):
        if listened[i]:
            subj = b"chat." + ids[i] + b".tokens"
            for _ in range(int(fan[i])):
                deltas.append((0, sid, subj))
                sid += 1
    t0 = time.perf_counter()
    eng.interest_apply(deltas)
    t_build = time.perf_counter() - t0
    arena, off = pkg.workload.generate(1, n_subj, n_tok)
    db = DevBatch(arena, off, extra_cap=96)
    op = pkg.OP_BRIDGE | pkg.OP_MATCH
    ms = time_launches(lambda: db.launch(eng, op, stream), stream, 5)
    rc, c = eng.launch_result()
    status = db.d_status.cpu().numpy()
    hit = db.d_hit.cpu().numpy().reshape(-1, 3)
    want_cnt = np.where(listened, fan, 0)
    assert np.array_equal(hit[:, 1], np.tile(want_cnt, n_tok).astype(hit.dtype))
    assert np.array_equal((status & 0xF) == 0, np.tile(listened, n_tok))
    res = [{"config": 3, "mode": "bridge+match, literal subscriptions only", "subjects": n_subj, "subscriptions": len(deltas), "records": db.n,
            "ms_per_launch": ms, "frames_per_sec": db.n / ms * 1e3, "published": int(c["frames"]), "table_build_s": t_build,
            "alg_GBps": (db.in_bytes + c["bytes_out"] + 32 * db.n) / ms / 1e6}]
    eng.interest_apply([(0, sid, b"chat.*.tokens"), (0, sid + 1, b"chat.*.control")])   # the two wildcard consumers
    ms = time_launches(lambda: db.launch(eng, op, stream), stream, 5)
    rc, c = eng.launch_result()
    hit = db.d_hit.cpu().numpy().reshape(-1, 3)
    assert (hit[:, 2] == 1).all() and c["frames"] == db.n
    res.append({"config": 3, "mode": "bridge+match, + wildcard consumers chat.*.tokens, chat.*.control", "records": db.n,
                "ms_per_launch": ms, "frames_per_sec": db.n / ms * 1e3, "published": int(c["frames"]),
                "alg_GBps": (db.in_bytes + c["bytes_out"] + 32 * db.n) / ms / 1e6})
    ms = time_launches(lambda: db.launch(eng, pkg.OP_BRIDGE, stream), stream, 5)
    res.append({"config": 3, "mode": "bridge only (no table)", "records": db.n, "ms_per_launch": ms, "frames_per_sec": db.n / ms * 1e3})
    eng.close()
    return res


def config4(pkg, scale):
    n_conv = int(125_000 * scale)                  # one GPU's share of 1 M conversations on 8 GPUs
    rng = np.random.default_rng(0x5EED0005)
    base_rate = n_conv / 0.0625                    # tokens/s: one token per 50 ms + U(0,25) ms per conversation
    cap = 65536
    eng = pkg.SseGpu(device=0, max_records=cap, max_in_bytes=cap * 160, ring_slots=4)
    arena, off = pkg.workload.generate(1, n_conv, 2)
    lat, sizes = [], []
    slot_q = []
    t_ms = 0
    for k in range(3000):
        on = (t_ms % 1000) < 200
        lam = base_rate * (10 if on else 1) * 1e-3
        n = int(min(cap, max(1, rng.poisson(lam))))
        lo = int(rng.integers(0, len(off) - 1 - n))
        b = eng.acquire()
        m = int(off[lo + n]) - int(off[lo])
        b.in_bytes[:m] = arena[int(off[lo]):int(off[lo + n])]
        b.in_off[: n + 1] = off[lo:lo + n + 1] - off[lo]
        b.n = n
        t0 = time.perf_counter()                   # the two C-ABI calls themselves (ctypes overhead included, numpy views not)
        rc1 = eng.lib.ssegpu_batch_submit(eng.ctx, b.h, n, 0, 0)
        rc2 = eng.lib.ssegpu_batch_wait(eng.ctx, b.h, 0)
        lat.append((time.perf_counter() - t0) * 1e6)
        assert rc1 == 0 and rc2 == 0
        if k % 500 == 0:
            b.wait()                                # already done: builds the views; spot check
            assert int(b.status[: n].max()) & 0xF == 0 and int(b.out_len[0]) == int(b.in_off[1]) + 27
        sizes.append(n)
        b.release()
        t_ms += 1
    lat = np.array(lat[200:])
    sizes = np.array(sizes[200:])
    eng.close()
    return [{"config": 4, "conversations_on_this_gpu": n_conv, "batches": len(lat), "batch_records_p50": int(np.median(sizes)),
             "batch_records_max": int(sizes.max()), "submit_to_wait_us_p50": float(np.percentile(lat, 50)),
             "submit_to_wait_us_p99": float(np.percentile(lat, 99)), "submit_to_wait_us_max": float(lat.max()),
             "note": "1 GPU measured; the 8-GPU deployment runs eight independent copies of this (no cross-GPU step)"}]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="all")
    ap.add_argument("--scale", type=float, default=1.0)
    args = ap.parse_args()
    pkg = load_package()
    for name, fn in (("0", config0), ("2", config2), ("3", config3), ("4", config4)):
        if args.config in ("all", name):
            for line in fn(pkg, args.scale):
                print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
