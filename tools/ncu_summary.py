#!/usr/bin/env python3
"""profiles/<tag>_summary.md (+ an entry in profiles/traffic.json) from one `ncu --set full` capture.

usage: tools/ncu_summary.py <report.ncu-rep> <tag> <traffic key> <records per launch> <algorithmic bytes per launch> "<what was profiled>" [--mangled M] [--unit U --kernel K]
One row of the report per profiled launch (ordered mode: two rows); the traffic entry is the sum over the rows."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag, key, n_rec, alg, what = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), float(sys.argv[5]), sys.argv[6]
extra = sys.argv[7:]
rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__warps_active.avg.per_cycle_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "nsecond": 1e-9, "msecond": 1e-3}
L = ["# %s — ncu --set full --clock-control none, B200" % tag, "", what, ""]
rd = wr = dur = 0.0
for vals in rows[2:]:
    if len(vals) < len(hdr):
        continue
    d = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
    L += ["## `%s`" % d["Kernel Name"][0], "", "| metric | value | unit |", "|---|---|---|"]
    L += ["| %s | %s | %s |" % (k, d[k][0], d[k][1]) for k in want if k in d]
    st = sorted(((float(v[0].replace(",", "")), k) for k, v in d.items() if 'issue_stalled' in k and 'ratio' in k and v[0]), reverse=True)
    L += ["", "Warp stall reasons (cycles per issued instruction): " +
          ", ".join("%s %.2f" % (k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v) for v, k in st if v > 0.1), ""]

    def num(name):
        v, u = d[name]
        return float(v.replace(",", "")) * scale.get(u, 1.0)
    rd += num('dram__bytes_read.sum'); wr += num('dram__bytes_write.sum'); dur += num('gpu__time_duration.sum')
L += ["DRAM traffic per launch (all rows): %.3f GB read + %.3f GB written = %.3f GB; algorithmic bytes per launch %.3f GB -> traffic / algorithmic = %.2f; "
      "under ncu (cold caches, serialised) the launch took %.1f us -> %.0f GB/s algorithmic." % (rd / 1e9, wr / 1e9, (rd + wr) / 1e9, alg / 1e9, (rd + wr) / alg, dur * 1e6, alg / dur / 1e9),
      "No tensor pipe activity (nothing on this path is a contraction).", ""]
if "--mangled" in extra or "--kernel" in extra:
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), rep, "--top", "16", "--lost"] + extra, capture_output=True, text=True).stdout
    L += ["## Where the warp-instructions go (tools/ncu_lines.py: ncu SASS page joined with nvdisasm -gi line info)", "", "```", out.strip(), "```"]
open(os.path.join(ROOT, "profiles", "%s_summary.md" % tag), "w").write("\n".join(L) + "\n")
tp = os.path.join(ROOT, "profiles", "traffic.json")
t = json.load(open(tp)) if os.path.exists(tp) else {}
t.setdefault("kernels", {})[key] = {"dram_bytes_read": rd, "dram_bytes_write": wr, "records_per_launch": n_rec, "alg_bytes_per_launch": alg,
                                   "source": "profiles/%s_summary.md (ncu --set full, one launch)" % tag}
json.dump(t, open(tp, "w"), indent=1)
print("\n".join(L[:12]))
