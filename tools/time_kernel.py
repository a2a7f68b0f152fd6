"""Kernel-only timing of ssegpu_frame_launch on the config-1 workload (no checks: for experimental builds selected with
SSEGPU_LIB=...).  python tools/time_kernel.py [--op N] [--steps K]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--op", type=int, default=0)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--conversations", type=int, default=10000)
ap.add_argument("--tokens", type=int, default=1000)
ap.add_argument("--noncanon", type=float, default=0.0, help="percent of records made non-canonical (same length: \"Token\" for \"token\")")
a = ap.parse_args()
import torch  # noqa: E402
pkg = load_package()
arena, off = pkg.workload.generate(1, a.conversations, a.tokens)
n = a.conversations * a.tokens
if a.noncanon > 0:
    pick = np.nonzero(np.random.default_rng(3).random(n) < a.noncanon / 100.0)[0]
    pos = off[pick].astype(np.int64) + 59                       # the t of "token" behind a 36-byte id
    assert (arena[pos] == ord("t")).all()
    arena[pos] = ord("T")
eng = pkg.SseGpu(device=0, max_records=n, max_in_bytes=int(off[n]) + 64, ring_slots=1)
d_in = torch.from_numpy(arena).cuda()
d_off = torch.from_numpy(off.view(np.int32)).cuda()
cap = int(off[n]) + 64 * n
d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
d_out_off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
d_out_len = torch.zeros(n, dtype=torch.int32, device="cuda")
d_subj = torch.zeros(n, dtype=torch.int32, device="cuda")
d_status = torch.zeros(n, dtype=torch.int32, device="cuda")
stream = torch.cuda.Stream()
evs = []
with torch.cuda.stream(stream):
    for it in range(3 + a.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        eng.frame_launch(d_in.data_ptr(), d_off.data_ptr(), 0, n, a.op, d_out.data_ptr(), cap, d_out_off.data_ptr(), d_out_len.data_ptr(),
                         d_subj.data_ptr(), d_status.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        evs.append((e0, e1))
torch.cuda.synchronize()
rc, c = eng.launch_result()
ms = [e0.elapsed_time(e1) for e0, e1 in evs[3:]]
print("noncanon=%.2f%% " % a.noncanon, end="")
print("lib=%s op=%d rc=%d frames=%d fast=%d bytes_out=%d arena=%d  ms mean %.4f min %.4f" % (
    os.environ.get("SSEGPU_LIB", "default"), a.op, rc, c["frames"], c["fast_path"], c["bytes_out"], int(d_out_off[n].item()), float(np.mean(ms)), min(ms)))
