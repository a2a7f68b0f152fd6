#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on BASELINE.json's configs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--no-extras]

Metric: SSE frames/s (and bytes/s, inspected bytes/s) through the hot path decode -> filter -> (inspect / match) -> encode
-> frame.  Headline workload = BASELINE.json configs[1]: 10 000 concurrent conversations x 1 000 tokens, UUID ids,
Mistral-7B-like token lengths, arrival order, synthetic.  One step = one pass over the whole 10 M-record batch.

  value     frames/s with the batch already resident in HBM (CUDA events around K launches, max over ranks)
  e2e       frames/s through the C-ABI batch ring with HOST pinned buffers: H2D copies, kernel, D2H copies of
            frames + offsets + status all inside the timed region (wall clock between device syncs)
  roofline  algorithmic bytes (payload + frame + 16 B/record) / kernel time vs the measured HBM copy peak
  cpu_baseline  the C oracle (port of the reference Go path; no Go toolchain here) on the box's host cores, bounded sample
  inspect   configs[2]: the full inspection pattern set on 100 000 conversations x 256 tokens per GPU (25.6 M records, planted
            matches, half of them split across tokens): per-token mode (Set A, Set A+B) and streaming mode (Set A+B, the whole
            batch in one call), frames/s + inspected token bytes/s + roofline at 328.7 algorithmic B/record
  bridge_match  configs[3]: bridge decode -> subject match -> reframe at 100 000 subjects, Zipf(1.1) fan-out, two wildcard
            consumers, 10 % unsubscribed
  latency   configs[4]: this GPU's share of 1 M conversations, bursty arrivals, 1 ms batches through the pinned ring,
            p50/p99 of submit -> wait

N > 1 (torchrun): conversations are sharded by fnv1a64(conversation_id) % N exactly as the host router does; each
rank owns ~10 000 (config 1) / ~100 000 (config 2, 3) conversations (weak scaling), no data-path collective; only the
timing/counter gather crosses ranks.  Rates are sums over ranks of (units / max-over-ranks time).
"""
import argparse
import json
import os
import random
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_CONV, N_TOK = 10000, 1000
METRIC, UNIT = "sse_frames_per_sec", "frames/s"
E2E_BATCH = 1_000_000          # records per pinned ring slot on the e2e path
WORKLOAD = "configs[1]: 10k conversations x 1k tokens per GPU, UUID ids, Mistral-7B-like token lengths, arrival order"
C2_CONV, C2_TOK = 100_000, 256
C3_SUBJ, C3_TOK = 100_000, 16
PLANTS = [b"123-45-6789", b"4111111111111111", b"bob@example.com", b"ignore previous instructions", b"disregard the above",
          b"new instructions:", b"system prompt:"]


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        # under load = samples at or above half the maximum clock seen (idle samples between phases are dropped)
        sm = [int(r[0]) for r in self.rows if r and r[0].isdigit()]
        sm = sorted(x for x in sm if 2 * x >= max(sm)) if sm else []
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def measured_traffic(kernel, n_records):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the committed ncu --set full captures
    (profiles/traffic.json), scaled to this launch's record count."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        t = t.get("kernels", {}).get(kernel) or (t if t.get("kernel") == kernel else None)
        return (t["dram_bytes_read"] + t["dram_bytes_write"]) * n_records / t["records_per_launch"]
    except Exception:
        return None


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def make_workload(pkg, rank, world):
    return pkg.sharding.rank_workload(1, N_CONV * world, N_TOK, rank, world)


def cpu_baseline(arena, off, threads, budget_s=12.0):
    """Times the oracle (CPU port of the reference path) on a bounded prefix of the same workload, warm buffers."""
    import oracle_lib
    oracle = oracle_lib.load()
    n = len(off) - 1
    probe = min(n, 200_000)
    sec, frames, _ = oracle.bench_adapter(arena, off[: probe + 1], threads, reps=1)
    rate = frames / sec
    sample = int(min(n, max(probe, rate * budget_s / 4)))
    prep = oracle.bench_prepared(arena, off[: sample + 1], threads)
    runs = [prep.run() for _ in range(4)][1:]
    prep.close()
    sec, frames, out_bytes = min(runs)
    return {"value": frames / sec, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "first %d records of the workload (arrival order), best of 3 warm passes, %d pthreads, C oracle "
                      "(restatement of the Go path; Go toolchain absent)" % (sample, threads),
            "bytes_per_sec": (int(off[sample]) + out_bytes) / sec,
            "inspection": "no CPU arm: the reference's inspection hook is a comment (sse_handler.go:193-208)"}, sample


def run_reference(args):
    """The reference's own CPU implementation of the path on all host cores: the C port (Go is absent), the
    same workload and record count per step as the GPU arm, output buffers allocated and touched once before timing."""
    rank, local, world = dist_env()
    if rank != 0:
        return
    from __graft_entry__ import load_package
    import oracle_lib
    pkg = load_package()
    oracle = oracle_lib.load()
    arena, off = make_workload(pkg, 0, world)
    n = len(off) - 1
    threads = os.cpu_count() or 1
    prep = oracle.bench_prepared(arena, off, threads)
    for _ in range(max(1, args.warmup)):
        prep.run()
    secs, frames_total = [], 0
    for _ in range(args.steps):
        sec, frames, _ = prep.run()
        secs.append(sec)
        frames_total += frames
    prep.close()
    t_total = sum(secs)
    value = frames_total / t_total
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": max(1, args.warmup), "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "records_per_step": n,
                   "note": "each step = one pass over rank 0's whole batch (%d records) on the host cores; warm output buffers" % n},
        "best_step_value": n / min(secs),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "all %d records per step, %d pthreads, C oracle (Go toolchain absent, reference unbuildable here)" % (n, threads)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------ helpers (GPU arm) ----
class DevBatch:
    def __init__(self, torch, np, arena, off, extra_cap=64):
        self.n = len(off) - 1
        self.in_bytes = int(off[self.n])
        self.d_in = torch.from_numpy(arena).cuda()
        self.d_off = torch.from_numpy(off.view(np.int32)).cuda()
        self.cap = self.in_bytes + extra_cap * self.n + 4096
        e = lambda m, dt: torch.empty(max(m, 1), dtype=dt, device="cuda")                  # noqa: E731
        self.d_out, self.d_out_off, self.d_out_len = e(self.cap, torch.uint8), e(self.n + 1, torch.int64), e(self.n, torch.int32)
        self.d_status, self.d_subj = e(self.n, torch.int32), e(self.n, torch.int32)
        self.d_verdict = self.d_match = self.d_hit = None

    def bind(self, torch, eng, d_conv=0, inspect=False, match=False):
        if inspect and self.d_verdict is None:
            self.d_verdict = torch.zeros(self.n, dtype=torch.int32, device="cuda")
            self.d_match = torch.zeros(2 * self.n, dtype=torch.int32, device="cuda")
        if match and self.d_hit is None:
            self.d_hit = torch.zeros(3 * self.n, dtype=torch.int32, device="cuda")
        if inspect:
            eng.inspect_bind(d_conv, self.d_verdict.data_ptr(), self.d_match.data_ptr())
        if match:
            eng.interest_bind(self.d_hit.data_ptr())

    def launch(self, eng, op, stream):
        eng.frame_launch(self.d_in.data_ptr(), self.d_off.data_ptr(), 0, self.n, op, self.d_out.data_ptr(), self.cap,
                         self.d_out_off.data_ptr(), self.d_out_len.data_ptr(), self.d_subj.data_ptr(), self.d_status.data_ptr(),
                         stream.cuda_stream)


def time_launches(torch, fn, stream, reps, warm=3):
    for _ in range(warm):
        fn()
    stream.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        fn()
    b.record(stream)
    stream.synchronize()
    return a.elapsed_time(b) / reps


def planted_rows(n_plant, n_tok, seed):
    """Conversations that carry one pattern each, whole in one token or split over two consecutive tokens."""
    import pymodel
    rng = random.Random(seed)
    rows, text_bytes = [], 0
    for p in range(n_plant):
        frag = PLANTS[p % len(PLANTS)]
        split = p % 2 == 1
        at = rng.randrange(2, n_tok - 4)
        cut = rng.randrange(1, len(frag)) if split else len(frag)
        for j in range(n_tok):
            tok = b" filler"
            if j == at:
                tok = b" " + frag[:cut]
            elif j == at + 1:
                tok = (frag[cut:] + b" ") if split else b" after"
            text_bytes += len(tok)
            rows.append(pymodel.marshal({"cid": "planted-%06d" % p, "tok": tok.decode(), "seq": j + 1, "done": j == n_tok - 1,
                                         "ts": 1700000000000000000 + j}))
    return rows, text_bytes


def inspect_section(pkg, torch, np, dist, rank, world, local, peak, reps):
    """configs[2] at SURVEY 8(d)'s size: 100 000 conversations x 256 tokens per GPU, Set A and Set A+B."""
    from inspect_sets import SET_A, SET_B
    stream = torch.cuda.Stream()
    conv_map = None if world == 1 else pkg.workload.shard_members(1, C2_CONV * world, rank, world)
    n_conv = C2_CONV if conv_map is None else len(conv_map)
    arena, off = pkg.workload.generate(1, n_conv, C2_TOK, conv_map=conv_map)
    cols = pkg.workload.generate_columns(1, n_conv, C2_TOK, conv_map=conv_map)
    tok_bytes = int(cols[3][-1])
    del cols
    n_plant = 100
    rows, planted_text = planted_rows(n_plant, C2_TOK, 0x5EED0003 + rank)
    n_gen = n_conv * C2_TOK
    lens = np.array([len(r) for r in rows], dtype=np.uint64)
    total = int(off[n_gen])
    arena = np.concatenate([arena[:total], np.frombuffer(b"".join(rows), dtype=np.uint8), np.zeros(64, np.uint8)])
    off = np.concatenate([off[: n_gen + 1], (total + np.cumsum(lens)).astype(np.uint32)])
    tok_bytes += planted_text
    n = len(off) - 1
    conv = np.concatenate([np.tile(np.arange(n_conv, dtype=np.int32), C2_TOK),
                           np.repeat(np.arange(n_conv, n_conv + n_plant, dtype=np.int32), C2_TOK)])
    d_conv = torch.from_numpy(conv).cuda()
    db = DevBatch(torch, np, arena, off)
    out = {"workload": "configs[2]: %d conversations x %d tokens per GPU (+%d planted conversations, half of the planted patterns split "
                       "across two tokens), one batch of %d records" % (n_conv, C2_TOK, n_plant, n),
           "records": n, "token_text_bytes": tok_bytes, "alg_bytes_per_record": "payload + frame + 32", "modes": []}
    for name, pats, mode in (("SetA per_token", SET_A, pkg.INSPECT_PER_TOKEN), ("SetA+B per_token", SET_A + SET_B, pkg.INSPECT_PER_TOKEN),
                             ("SetA+B streaming", SET_A + SET_B, pkg.INSPECT_STREAMING)):
        eng = pkg.SseGpu(device=local, max_records=n, max_in_bytes=1 << 20, conv_slots=n_conv + n_plant)
        eng.load_patterns(pats, mode)
        streaming = mode == pkg.INSPECT_STREAMING
        db.bind(torch, eng, d_conv.data_ptr() if streaming else 0, inspect=True)
        l0 = eng.launch_count()
        ms = time_launches(torch, lambda: db.launch(eng, pkg.OP_INSPECT, stream), stream, reps)
        launches = (eng.launch_count() - l0) // (reps + 3)
        if streaming:
            eng.conv_reset()
            db.launch(eng, pkg.OP_INSPECT, stream)
            stream.synchronize()
        rc, c = eng.launch_result()
        assert rc == 0, rc
        hit = (db.d_verdict[n_gen:] != 0).view(n_plant, C2_TOK).any(dim=1)
        found = int(hit.sum().item())
        false_hits = int((db.d_verdict[:n_gen] != 0).sum().item())
        alg = db.in_bytes + c["bytes_out"] + 32 * n
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        s = torch.tensor([float(n), float(tok_bytes), float(alg), float(found)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(s, op=dist.ReduceOp.SUM)
        ms_max = t.item()
        n_all, tb_all, alg_all, found_all = s.tolist()
        out["modes"].append({
            "mode": name, "ms_per_batch": ms_max, "kernel_launches_per_batch": int(launches),
            "frames_per_sec": n_all / ms_max * 1e3, "inspected_bytes_per_sec": tb_all / ms_max * 1e3,
            "roofline": {"bound": "hbm", "achieved": alg / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": alg / ms / 1e6 / peak,
                         "alg_bytes_per_batch": int(alg), "traffic": measured_traffic("k_frame_tiles<inspect>", n) if not streaming else None},
            "planted": n_plant * world, "planted_found": int(found_all),
            "planted_expected_at_least": (n_plant * world) if streaming else (n_plant * world + 1) // 2,
            "hits_in_generated_traffic": false_hits})
        eng.close()
    return out


def bridge_match_section(pkg, torch, np, dist, rank, world, local, peak, reps):
    """configs[3]: bridge decode -> subject match -> reframe at 100 000 subjects with a high-fanout interest table."""
    stream = torch.cuda.Stream()
    conv_map = None if world == 1 else pkg.workload.shard_members(1, C3_SUBJ * world, rank, world)
    n_subj = C3_SUBJ if conv_map is None else len(conv_map)
    rng = np.random.default_rng(0x5EED0004 + rank)
    ids = [pkg.workload.conversation_id(1, int(i)) for i in (range(n_subj) if conv_map is None else conv_map)]
    fan = np.maximum(1, (1024.0 / np.arange(1, n_subj + 1) ** 1.1)).astype(np.int64)
    rng.shuffle(fan)
    listened = rng.random(n_subj) >= 0.10
    deltas, sid = [], 1
    for i in range(n_subj):
        if listened[i]:
            subj = b"chat." + ids[i] + b".tokens"
            for _ in range(int(fan[i])):
                deltas.append((0, sid, subj))
                sid += 1
    eng = pkg.SseGpu(device=local, max_records=n_subj * C3_TOK, max_in_bytes=1 << 20)
    t0 = time.perf_counter()
    eng.interest_apply(deltas)
    t_build = time.perf_counter() - t0
    arena, off = pkg.workload.generate(1, n_subj, C3_TOK, conv_map=conv_map)
    db = DevBatch(torch, np, arena, off, extra_cap=96)
    db.bind(torch, eng, match=True)
    op = pkg.OP_BRIDGE | pkg.OP_MATCH
    ms = time_launches(torch, lambda: db.launch(eng, op, stream), stream, reps)
    rc, c = eng.launch_result()
    hit = db.d_hit.view(-1, 3)
    want = torch.from_numpy(np.tile(np.where(listened, fan, 0), C3_TOK).astype(np.int32)).cuda()
    ok = bool((hit[:, 1] == want).all().item())
    eng.interest_apply([(0, sid, b"chat.*.tokens"), (0, sid + 1, b"chat.*.control")])     # the two wildcard consumers
    ms_w = time_launches(torch, lambda: db.launch(eng, op, stream), stream, reps)
    rc2, c2 = eng.launch_result()
    ms_b = time_launches(torch, lambda: db.launch(eng, pkg.OP_BRIDGE, stream), stream, reps)
    eng.close()
    alg = db.in_bytes + c["bytes_out"] + 32 * db.n
    t = torch.tensor([ms, ms_w, ms_b], dtype=torch.float64, device="cuda")
    s = torch.tensor([float(db.n), float(c["frames"]), float(ok)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    tm, sm = t.tolist(), s.tolist()
    return {"workload": "configs[3]: %d subjects x %d tokens per GPU, %d literal subscriptions (Zipf(1.1) fan-out 1..1024), 10 %% of the subjects "
                        "unsubscribed; second figure with the wildcard consumers chat.*.tokens, chat.*.control" % (n_subj, C3_TOK, len(deltas)),
            "records": int(sm[0]), "published": int(sm[1]), "subscriber_counts_match_model": int(sm[2]) == world, "table_build_s": t_build,
            "ms_per_batch": tm[0], "frames_per_sec": sm[0] / tm[0] * 1e3, "ms_per_batch_with_wildcards": tm[1],
            "ms_per_batch_bridge_only": tm[2],
            "roofline": {"bound": "hbm", "achieved": alg / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": alg / ms / 1e6 / peak,
                         "alg_bytes_per_batch": int(alg), "traffic": measured_traffic("k_frame_tiles<interest>", db.n)}}


def latency_section(pkg, torch, np, dist, rank, world, local):
    """configs[4]: this GPU's share of 1 M conversations, on/off bursts, batches cut every 1 ms, pinned ring, submit -> wait."""
    n_conv = 1_000_000 // max(world, 8)            # 125 000 per GPU (the config is quoted on 8 GPUs)
    rng = np.random.default_rng(0x5EED0005 + rank)
    base_rate = n_conv / 0.0625                    # tokens/s: one token per 50 ms + U(0,25) ms per conversation
    cap = 65536
    eng = pkg.SseGpu(device=local, max_records=cap, max_in_bytes=cap * 160, ring_slots=4)
    arena, off = pkg.workload.generate(1, min(n_conv, 50_000), 2)
    lat, sizes = [], []
    t_ms = 0
    for k in range(2200):
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
        sizes.append(n)
        b.release()
        t_ms += 1
    lat, sizes = np.array(lat[200:]), np.array(sizes[200:])
    eng.close()
    p = torch.tensor([float(np.percentile(lat, 50)), float(np.percentile(lat, 99)), float(lat.max())], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(p, op=dist.ReduceOp.MAX)
    p50, p99, mx = p.tolist()
    return {"workload": "configs[4]: %d conversations on each GPU (1 M over 8), on 200 ms at 10x rate / off 800 ms, batches cut every 1 ms, "
                        "pinned ring" % n_conv,
            "batches": int(len(lat)), "batch_records_p50": int(np.median(sizes)), "batch_records_max": int(sizes.max()),
            "submit_to_wait_us_p50": p50, "submit_to_wait_us_p99": p99, "submit_to_wait_us_max": mx,
            "note": "worst rank's percentiles; ranks are independent (no cross-GPU step)"}


def run_gpu(args):
    import numpy as np
    import torch
    from __graft_entry__ import load_package
    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    use_dist = world > 1
    dist = None
    if use_dist:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = load_package()
    arena, off = make_workload(pkg, rank, world)
    n = len(off) - 1
    in_bytes = int(off[n])

    # ---------------- device-resident arm ("value", roofline) ----------------
    eng = pkg.SseGpu(device=local, max_records=n, max_in_bytes=in_bytes + 64)
    d_in = torch.from_numpy(arena).cuda()
    d_off = torch.from_numpy(off.view(np.int32)).cuda()
    cap = in_bytes + 48 * n
    d_out = torch.empty(cap, dtype=torch.uint8, device="cuda")
    d_out_off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    d_out_len = torch.empty(n, dtype=torch.int32, device="cuda")
    d_status = torch.empty(n, dtype=torch.int32, device="cuda")
    stream = torch.cuda.Stream()          # explicit non-default stream: launches and timing events share it
    torch.cuda.synchronize()

    op = pkg.OP_ADAPTER | (pkg.OP_ORDERED if args.ordered else 0)

    def launch():
        eng.frame_launch(d_in.data_ptr(), d_off.data_ptr(), 0, n, op, d_out.data_ptr(), cap, d_out_off.data_ptr(),
                         d_out_len.data_ptr(), 0, d_status.data_ptr(), stream.cuda_stream)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                         # samples through warm-up, the device-timed region and the e2e region
    for _ in range(max(args.warmup, 3)):
        launch()
    torch.cuda.synchronize()
    rc, c0 = eng.launch_result()
    assert rc == 0 and c0["frames"] == n, (rc, c0)
    out_bytes = c0["bytes_out"]
    if use_dist:
        dist.barrier()
    launches0 = eng.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    torch.cuda.synchronize()
    t_all0 = torch.cuda.Event(enable_timing=True)
    t_all1 = torch.cuda.Event(enable_timing=True)
    t_all0.record(stream)
    for a, b in evs:
        a.record(stream)
        launch()
        b.record(stream)
    t_all1.record(stream)
    torch.cuda.synchronize()
    if use_dist:
        dist.barrier()
    launches = eng.launch_count() - launches0
    dev_ms = t_all0.elapsed_time(t_all1)
    kern_ms = sum(a.elapsed_time(b) for a, b in evs) / args.steps

    # ---------------- end-to-end arm: pinned ring, H2D + kernel + D2H ----------------
    nb = (n + E2E_BATCH - 1) // E2E_BATCH
    bounds = [(k * E2E_BATCH, min(n, (k + 1) * E2E_BATCH)) for k in range(nb)]
    max_in = max(int(off[hi]) - int(off[lo]) for lo, hi in bounds)
    e2e_eng = pkg.SseGpu(device=local, max_records=E2E_BATCH, max_in_bytes=max_in + 64,
                         max_out_bytes=max_in + 48 * E2E_BATCH, ring_slots=nb)
    slots = []
    h2d = d2h = 0
    for lo, hi in bounds:                      # the "NATS callback" side: payloads appended to pinned slots up front
        b = e2e_eng.acquire()
        m = hi - lo
        b.in_bytes[: int(off[hi]) - int(off[lo])] = arena[int(off[lo]):int(off[hi])]
        b.in_off[: m + 1] = off[lo:hi + 1] - off[lo]
        b.n = m
        slots.append(b)
        nbytes = int(off[hi]) - int(off[lo])
        a16 = lambda x: (x + 15) & ~15                                                      # noqa: E731
        # exactly what ssegpu_batch_submit copies: payload (+32 pad) and offsets in; the packed result block (header + counters,
        # out_off, out_len, status) and the speculative frame copy out
        h2d += a16(nbytes) + 32 + 4 * (m + 1)
        d2h += 4160 + a16(8 * (m + 1)) + 2 * a16(4 * m + 4) + nbytes + 32 * m + 16 * ((m + 127) // 128)

    def e2e_step(flags=0):
        for b in slots:
            b.submit(b.n, op | flags)
        frames = 0
        for b in slots:
            b.wait()
            frames += b.n
        return frames

    def e2e_timed(flags=0):
        for _ in range(2):
            e2e_step(flags)
        torch.cuda.synchronize()
        if use_dist:
            dist.barrier()
        t0 = time.perf_counter()
        frames = 0
        for _ in range(args.steps):
            frames += e2e_step(flags)
        torch.cuda.synchronize()
        return frames, time.perf_counter() - t0

    # SSEGPU_OP_HEADER_ONLY egress first (the default mode runs last so that the spot check below sees its output): the GPU
    # returns only the header of every Go-marshalled frame; the host still holds the payload in the pinned input slot and
    # appends it when it writes to the connection (host/adapter_host.cpp does, tests/test_gpu_adapter_host.py checks the bytes)
    _, e2e_hdr_s = e2e_timed(pkg.OP_HEADER_ONLY)
    hdr = slots[0]
    assert int(hdr.status[0]) & pkg.ST_HEADER_ONLY
    hdr_frame0 = hdr.out[int(hdr.out_off[0]):int(hdr.out_off[0]) + int(hdr.out_len[0])].tobytes() + \
        arena[int(off[0]):int(off[1])].tobytes() + b"\n\n"
    d2h_hdr = sum(4160 + ((8 * (b.n + 1) + 15) & ~15) + 2 * ((4 * b.n + 4 + 15) & ~15) + 32 * b.n + 16 * ((b.n + 127) // 128) for b in slots)
    e2e_frames, e2e_s = e2e_timed()
    first = slots[0]
    assert hdr_frame0 == first.out[int(first.out_off[0]):int(first.out_off[0]) + int(first.out_len[0])].tobytes()
    clocks = sampler.stop() if rank == 0 else None
    # spot check the e2e output of the last step against the device-resident arm
    last = slots[-1]
    lo, hi = bounds[-1]
    assert np.array_equal(last.out_len, d_out_len[lo:hi].cpu().numpy().astype(np.uint32))
    k0, g0, ln = int(last.out_off[0]), int(d_out_off[lo].item()), int(last.out_len[0])
    assert np.array_equal(last.out[k0:k0 + ln], d_out[g0:g0 + ln].cpu().numpy())
    e2e_launches = e2e_eng.launch_count()
    eng.close()
    e2e_eng.close()
    del d_in, d_off, d_out, d_out_off, d_out_len, d_status, slots
    torch.cuda.empty_cache()

    # ---------------- the rest of the metric: configs 2, 3, 4 (every rank, its own shard) ----------------
    peak, which = measured_peak()
    extras = {}
    if not args.no_extras:
        extras["inspect"] = inspect_section(pkg, torch, np, dist, rank, world, local, peak, reps=5)
        torch.cuda.empty_cache()
        extras["bridge_match"] = bridge_match_section(pkg, torch, np, dist, rank, world, local, peak, reps=10)
        torch.cuda.empty_cache()
        extras["latency"] = latency_section(pkg, torch, np, dist, rank, world, local)

    # ---------------- gather over ranks ----------------
    tot = torch.tensor([float(n), float(in_bytes), float(out_bytes), float(e2e_frames)], dtype=torch.float64, device="cuda")
    mx = torch.tensor([dev_ms, kern_ms, e2e_s, e2e_hdr_s], dtype=torch.float64, device="cuda")
    if use_dist:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    tot, mx = tot.tolist(), mx.tolist()
    if rank == 0:
        frames_all, in_all, out_all, e2e_all = tot
        dev_s = mx[0] / 1e3
        value = frames_all * args.steps / dev_s
        alg_bytes = in_bytes + out_bytes + 16 * n                 # per launch, this rank's kernel
        achieved = alg_bytes / (kern_ms / 1e3) / 1e9
        cpu, sample = cpu_baseline(arena, off, os.cpu_count() or 1)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": mx[0] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "records_per_step": int(frames_all), "payload_bytes_per_step": int(in_all), "frame_bytes_per_step": int(out_all),
                       "sharding": "fnv1a64(conversation_id) % n_gpus, no data-path collective",
                       "arena_layout": "record order (count pass + tile scan + emit pass)" if args.ordered else "tile completion order (atomic cursor)",
                       "l2": "inputs (%.2f GB per GPU) larger than L2, no flush needed" % (in_bytes / 1e9),
                       "bytes_per_sec": (in_all + out_all) * args.steps / dev_s},
            "e2e": {"value": e2e_all / mx[2], "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "batches_per_step": nb, "records_per_batch": E2E_BATCH, "ms_per_step": 1e3 * mx[2] / args.steps,
                    "bytes_per_sec": (in_all + out_all) * args.steps / mx[2]},
            "e2e_header_only": {"value": e2e_all / mx[3], "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h_hdr),
                                "ms_per_step": 1e3 * mx[3] / args.steps,
                                "note": "SSEGPU_OP_HEADER_ONLY: only frame headers + offsets + status cross back; the host appends the payload it "
                                        "still holds in the pinned input slot when it writes to the connection"},
            "gpu_launches": int(launches), "gpu_launches_e2e": int(e2e_launches),
            "roofline": {"bound": "hbm", "kernel": "k_frame_tiles", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": which, "frac_of_8000_datasheet": achieved / 8000.0, "traffic": measured_traffic("k_frame_tiles", n),
                         "alg_bytes_per_launch": int(alg_bytes), "kernel_ms": kern_ms,
                         "note": "kernel_ms = CUDA events around memset(scratch)+k_frame_tiles on the launching stream, mean of %d" % args.steps},
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        line.update(extras)
        print(json.dumps(line))
    if use_dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--ordered", action="store_true", help="SSEGPU_OP_ORDERED: record-ordered arena (two decode passes)")
    ap.add_argument("--no-extras", action="store_true", help="config 1 only (skip the inspect / bridge_match / latency sections)")
    args = ap.parse_args()
    # Exactly one JSON line may reach stdout.  Libraries write there behind Python's back (NCCL prints its version banner
    # from C when the first communicator is created, whatever NCCL_DEBUG says short of NONE), so the process's fd 1 is
    # pointed at stderr for the whole run and the result line goes to the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
