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
    ap.add_argument("--plain", action="store_true", help="experiment: replace every token byte that needs escaping or is not ASCII by 'x'")
    a = ap.parse_args()
    import torch
    pkg = load_package()
    n_conv, n_tok = a.conversations, a.tokens
    n = n_conv * n_tok
    # columns = the BASELINE config-1 workload before marshalling (UUID ids, Mistral-7B-like token lengths, 1.5 % tokens with an
    # escape, 1 % with multi-byte UTF-8); the encoder must reproduce the workload's payloads, checked below
    arena, off = pkg.workload.generate(1, n_conv, n_tok)
    cid, cid_off, tok, tok_off, seq, done, ts = pkg.workload.generate_columns(1, n_conv, n_tok)
    if a.plain:
        bad = (tok < 0x20) | (tok >= 0x80) | np.isin(tok, np.frombuffer(b'"\\<>&', dtype=np.uint8))
        tok = np.where(bad, np.uint8(ord("x")), tok)
        tok[int(tok_off[n]):] = 0
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
    got_len = out_len.cpu().numpy().astype(np.int64)
    assert os.environ.get("SSEGPU_LIB") or a.plain or np.array_equal(got_len, np.diff(off.astype(np.int64))), "encoded lengths differ from the workload's payloads"
    k0 = int(out_off[12345].item())
    assert os.environ.get("SSEGPU_LIB") or a.plain or out[k0:k0 + int(got_len[12345])].cpu().numpy().tobytes() == arena[int(off[12345]):int(off[12346])].tobytes()
    ms = float(np.mean([e0.elapsed_time(e1) for e0, e1 in evs[a.warmup:]]))
    in_bytes = int(cid_off[n]) + int(tok_off[n]) + n * (4 + 4 + 8 + 1 + 8)
    out_bytes = int(out_len.sum().item()) + n * (8 + 4)
    print(json.dumps({"metric": "messages_encoded_per_s", "value": n / ms * 1e3, "ms": ms, "n": n, "algorithmic_bytes": in_bytes + out_bytes,
                      "GBps": (in_bytes + out_bytes) / ms / 1e6, "out_arena_bytes": used, "steps": a.steps, "warmup": a.warmup,
                      "workload": "config-1 shape, %d conversations x %d tokens, columns in HBM, L2 flushed between iterations" % (n_conv, n_tok)}))
    eng.close()


if __name__ == "__main__":
    main()
