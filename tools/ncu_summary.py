#!/usr/bin/env python3
"""profiles/<tag>_summary.md + profiles/traffic.json from one `ncu --set full` capture of k_frame_tiles.
usage: tools/ncu_summary.py <report.ncu-rep> <tag> "<command that was profiled>" """
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__warps_active.avg.per_cycle_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']
d = {h: (vals[i], units[i]) for i, h in enumerate(hdr) if h in want or ('issue_stalled' in h and 'ratio' in h)}
L = ["# %s — ncu --set full, k_frame_tiles, 10 M records per launch, B200" % tag, "", "Command: `%s`" % cmd, "",
     "| metric | value | unit |", "|---|---|---|"]
L += ["| %s | %s | %s |" % (k, d[k][0], d[k][1]) for k in want if k in d]
L += ["", "Warp stall reasons (cycles per issued instruction):", "", "| reason | value |", "|---|---|"]
st = sorted(((float(v[0]), k) for k, v in d.items() if 'issue_stalled' in k), reverse=True)
L += ["| %s | %.3f |" % (k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v) for v, k in st if v > 0.05]


def gb(name):
    v, u = d[name]
    return float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u]


rd, wr = gb('dram__bytes_read.sum'), gb('dram__bytes_write.sum')
L += ["", "DRAM traffic per launch: %.3f GB read + %.3f GB written = %.3f GB; algorithmic bytes per launch 3.128 GB (payload 1.339 + "
      "frames 1.628 + 16 B/record) -> traffic/algorithmic = %.2f (the kernel also writes 4 B out_len per record and <=15 pad bytes per tile)."
      % (rd / 1e9, wr / 1e9, (rd + wr) / 1e9, (rd + wr) / 3.12784e9),
      "No tensor pipe activity (nothing on this path is a contraction). SASS: `UBLKCP.S.G` (TMA bulk load), `UBLKCP.G.S` (TMA bulk store).", ""]
out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), rep, "--top", "20", "--lost"], capture_output=True, text=True).stdout
L += ["## Where the warp-instructions go (tools/ncu_lines.py: ncu SASS page joined with nvdisasm -gi line info)", "", "```", out.strip(), "```"]
open(os.path.join(ROOT, "profiles", "%s_summary.md" % tag), "w").write("\n".join(L) + "\n")
json.dump({"kernel": "k_frame_tiles", "version": tag, "dram_bytes_read": rd, "dram_bytes_write": wr, "records_per_launch": 10000000,
           "source": "profiles/%s_summary.md (ncu --set full, one launch)" % tag}, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print("\n".join(L[:40]))
