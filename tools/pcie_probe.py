"""Host<->device copy rates of this box (pinned memory, 1 GiB per direction per GPU): H2D alone, D2H alone, both at once.
The e2e number of bench.py cannot exceed what this prints.

  python tools/pcie_probe.py                 one GPU (cuda:0)
  python tools/pcie_probe.py --gpus N        N processes, one per GPU, measuring AT THE SAME TIME (a barrier file keeps the
                                             phases aligned), as bench.py's ranks do: shows whether the box's host fabric,
                                             not the GPUs, caps the aggregate.  Also prints nvidia-smi topo -m and the NUMA layout.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time


def worker(dev, world, sync_dir):
    import torch
    torch.cuda.set_device(dev)
    n = 1 << 30
    h_a = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_b = torch.empty(n, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier(tag):
        if world == 1:
            return
        open(os.path.join(sync_dir, "%s.%d" % (tag, dev)), "w").close()
        t0 = time.time()
        while len([f for f in os.listdir(sync_dir) if f.startswith(tag + ".")]) < world:
            if time.time() - t0 > 120:
                raise SystemExit("pcie_probe: barrier timeout")
            time.sleep(0.0005)

    def run(h2d, d2h, tag, reps=6):
        rates = []
        for r in range(reps):
            torch.cuda.synchronize()
            barrier("%s%d" % (tag, r))
            t0 = time.perf_counter()
            if h2d:
                with torch.cuda.stream(s1):
                    d_a.copy_(h_a, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_b.copy_(d_b, non_blocking=True)
            torch.cuda.synchronize()
            rates.append(n / (time.perf_counter() - t0) / 1e9)
        rates.sort()
        return rates[len(rates) // 2]                      # median: under contention the best rep is not the sustained rate

    out = {"gpu": dev, "h2d_GBps": run(True, False, "a"), "d2h_GBps": run(False, True, "b"), "both_each_GBps": run(True, True, "c")}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--worker", type=int, default=-1)
    ap.add_argument("--sync-dir", default="")
    a = ap.parse_args()
    if a.worker >= 0:
        worker(a.worker, a.gpus, a.sync_dir)
        return
    if a.gpus == 1:
        worker(0, 1, "")
        return
    sync_dir = tempfile.mkdtemp()
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--gpus", str(a.gpus), "--worker", str(d), "--sync-dir", sync_dir],
                              stdout=subprocess.PIPE, text=True) for d in range(a.gpus)]
    rows = []
    for p in procs:
        o, _ = p.communicate()
        rows += [json.loads(line) for line in o.splitlines() if line.startswith("{")]
    agg = {k: sum(r[k] for r in rows) for k in ("h2d_GBps", "d2h_GBps", "both_each_GBps")}
    print(json.dumps({"gpus": a.gpus, "per_gpu": rows, "aggregate_h2d_GBps": agg["h2d_GBps"], "aggregate_d2h_GBps": agg["d2h_GBps"],
                      "aggregate_both_each_direction_GBps": agg["both_each_GBps"]}))
    for cmd in (["nvidia-smi", "topo", "-m"], ["numactl", "-H"], ["lscpu"]):
        try:
            print("$ " + " ".join(cmd))
            print(subprocess.run(cmd, capture_output=True, text=True, timeout=20).stdout[:3000])
        except Exception as e:                                                   # noqa: BLE001
            print("(unavailable: %s)" % e)


if __name__ == "__main__":
    main()
