"""Kernel-level timing of the producer-side encoder (row N3) on the BASELINE config-1 shape: 10k conversations x
1k tokens, columns resident in HBM.  Prints one JSON line.  python tools/bench_encode.py [--steps K] [--warmup W]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--conversations", type=int, default=10000)
    ap.add_argument("--tokens", type=int, default=1000)
    a = ap.parse_args()
    import torch
    pkg = load_package()
    n_conv, n_tok = a.conversations, a.tokens
    n = n_conv * n_tok
    rng = np.random.default_rng(1)
    vocab = [b" the", b" of", b" and", b"Hello", b",", b" world", b".", b"\n", b" caf\xc3\xa9", b" token", b" stream", b" a", b" is", b" response"]
    ids = np.array([b"%08x-%04x-4%03x-a%03x-%012x" % (i, i & 0xFFFF, i & 0xFFF, i & 0xFFF, i * 2654435761 & 0xFFFFFFFFFFFF) for i in range(n_conv)])
    conv = np.tile(np.arange(n_conv), n_tok)                       # message i belongs to conversation i % n_conv (interleaved arrival)
    cid = np.frombuffer(b"".join(ids[conv].tolist()) + b"\0" * 16, dtype=np.uint8)
    cid_off = (np.arange(n + 1, dtype=np.uint64) * 36).astype(np.uint32)
    pick = rng.integers(0, len(vocab), n)
    lens = np.array([len(v) for v in vocab])[pick]
    tok_off = np.zeros(n + 1, dtype=np.uint32)
    tok_off[1:] = np.cumsum(lens)
    tok = np.frombuffer(b"".join(vocab[k] for k in pick) + b"\0" * 16, dtype=np.uint8)
    seq = (np.arange(n) // n_conv + 1).astype(np.int64)
    done = (seq == n_tok).astype(np.uint8)
    ts = 1700000000000000000 + np.arange(n, dtype=np.int64) * 1000
    eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=64 << 20, ring_slots=1)
    dev = lambda x, t: torch.from_numpy(np.ascontiguousarray(x).view(t)).cuda()  # noqa: E731
    d = [dev(cid, np.uint8), dev(cid_off, np.int32), dev(tok, np.uint8), dev(tok_off, np.int32), dev(seq, np.int64), dev(done, np.uint8), dev(ts, np.int64)]
    cap = n * 160
    out = torch.empty(cap, dtype=torch.uint8, device="cuda")
    out_off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
    out_len = torch.zeros(n, dtype=torch.int32, device="cuda")
    stream = torch.cuda.Stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    evs = []
    with torch.cuda.stream(stream):
        for it in range(a.warmup + a.steps):
            flush.fill_(it & 0xFF)                                  # L2 flush between iterations (256 MB > 126 MB L2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            eng.encode_launch(*[t.data_ptr() for t in d], n, 0, out.data_ptr(), cap, out_off.data_ptr(), out_len.data_ptr(), 0, stream.cuda_stream)
            e1.record(stream)
            evs.append((e0, e1))
    torch.cuda.synchronize()
    rc, used = eng.encode_result()
    assert rc == 0
    ms = float(np.mean([e0.elapsed_time(e1) for e0, e1 in evs[a.warmup:]]))
    in_bytes = int(cid_off[n]) + int(tok_off[n]) + n * (4 + 4 + 8 + 1 + 8)
    out_bytes = int(out_len.sum().item()) + n * (8 + 4)
    print(json.dumps({"metric": "messages_encoded_per_s", "value": n / ms * 1e3, "ms": ms, "n": n, "algorithmic_bytes": in_bytes + out_bytes,
                      "GBps": (in_bytes + out_bytes) / ms / 1e6, "out_arena_bytes": used, "steps": a.steps, "warmup": a.warmup,
                      "workload": "config-1 shape, %d conversations x %d tokens, columns in HBM, L2 flushed between iterations" % (n_conv, n_tok)}))
    eng.close()


if __name__ == "__main__":
    main()
