#!/usr/bin/env python3
"""Per-source-line instruction and stall-sample totals for the frame kernel.

usage: tools/ncu_lines.py <report.ncu-rep> [libssegpu.so] [--top N] [--func regex] [--unit sse_encode --kernel k_encode_messages]
Joins `ncu --page source --csv` (SASS rows, in order) with `nvdisasm -g` line info of the kernel cubin
extracted from the library that was profiled (must be the same build)."""
import csv
import os
import re
import subprocess
import sys
import tempfile

rep = sys.argv[1]
lib = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else os.path.join(
    os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "distributed-sse-for-llm-response_b200", "libssegpu.so")
top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 45

tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", lib], cwd=tmp, stdout=subprocess.DEVNULL)
unit = sys.argv[sys.argv.index("--unit") + 1] if "--unit" in sys.argv else "sse_kernels"
kern = sys.argv[sys.argv.index("--kernel") + 1] if "--kernel" in sys.argv else "k_frame_tiles"
MANGLED = sys.argv[sys.argv.index("--mangled") + 1] if "--mangled" in sys.argv else ""      # e.g. Lb1ELi0 = k_frame_tiles<true, 0>
cubin = [f for f in os.listdir(tmp) if f.startswith(unit + ".")][0]
dis = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout

# instruction order -> (innermost file:line, chain of frames innermost..outermost)
lines = []
chains = []
cur_chain = []
pending = []
in_kernel = False
for ln in dis.splitlines():
    if ln.startswith(".text."):
        in_kernel = kern in ln and (MANGLED in ln if MANGLED else (kern != "k_frame_tiles" or "Lb0" in ln or "ILb0" in ln))
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        pending.append((os.path.basename(m.group(1)), int(m.group(2))))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        if pending:
            cur_chain = pending
            pending = []
        if in_kernel:
            lines.append((cur_chain[0] if cur_chain else None, ln.split("*/", 1)[1].strip()[:60]))
            chains.append(cur_chain)

rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()))
hi = [i for i, r in enumerate(rows) if "Instructions Executed" in r][0]
hdr = rows[hi]
ie, smp, srcc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
body = [r for r in rows[hi + 1:] if len(r) > ie and r[ie].isdigit()]
print("sass rows: ncu %d, nvdisasm %d" % (len(body), len(lines)))
n = min(len(body), len(lines))
agg = {}
tot_i = tot_s = 0
for k in range(n):
    key = lines[k][0]
    a = agg.setdefault(key, [0, 0])
    a[0] += int(body[k][ie])
    a[1] += int(body[k][smp])
    tot_i += int(body[k][ie])
    tot_s += int(body[k][smp])
print("total warp-instructions %d, samples %d" % (tot_i, tot_s))
src_cache = {}


def src_line(key):
    if key is None:
        return ""
    f, l = key
    for d in ("distributed-sse-for-llm-response_b200/csrc",):
        p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), d, f)
        if os.path.exists(p):
            if p not in src_cache:
                src_cache[p] = open(p).read().splitlines()
            if l - 1 < len(src_cache[p]):
                return src_cache[p][l - 1].strip()[:100]
    return ""


# ---- per function (any frame of the inline chain), and per kernel-level call site ----
CS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "distributed-sse-for-llm-response_b200", "csrc")
fn_cache = {}


def func_of(key):
    f, l = key
    p = os.path.join(CS, f)
    if not os.path.exists(p):
        return f
    if p not in fn_cache:
        fn_cache[p] = open(p).read().splitlines()
    src = fn_cache[p]
    for k in range(min(l, len(src)) - 1, -1, -1):
        t = src[k]
        m = re.match(r"^(?:    )?(?:template\s*<[^>]*>\s*)?(?:SSE_HD(?:_NOINLINE)?|__device__|__global__|static|inline)\b.*?(\w+)\s*\(", t)
        if m and not t.strip().startswith("//"):
            return m.group(1)
    return f


fn_incl, site, fn_static = {}, {}, {}
for k in range(n):
    ie_k, s_k = int(body[k][ie]), int(body[k][smp])
    seen = set()
    for fr in chains[k]:
        fn = func_of(fr)
        if fn not in seen:
            seen.add(fn)
            a = fn_incl.setdefault(fn, [0, 0])
            a[0] += ie_k
            a[1] += s_k
            fn_static[fn] = fn_static.get(fn, 0) + 1
if "--static" in sys.argv:
    print("---- static SASS instructions per function (inclusive of what is inlined into it), of %d ----" % n)
    for fn, c in sorted(fn_static.items(), key=lambda kv: -kv[1])[:30]:
        print("%6d  %s" % (c, fn))
    if chains[k]:
        outer = chains[k][-1]
        a = site.setdefault(outer, [0, 0])
        a[0] += ie_k
        a[1] += s_k
print("---- inclusive per function ----")
for fn, (i, s_) in sorted(fn_incl.items(), key=lambda kv: -kv[1][0])[:28]:
    print("%5.1f%% inst %5.1f%% smp  %s" % (100.0 * i / tot_i, 100.0 * s_ / max(tot_s, 1), fn))
print("---- per kernel-level line (outermost frame) ----")
for key, (i, s_) in sorted(site.items(), key=lambda kv: -kv[1][0])[:22]:
    print("%5.1f%% inst %5.1f%% smp  %-22s %s" % (100.0 * i / tot_i, 100.0 * s_ / max(tot_s, 1), "%s:%d" % key, src_line(key)))
if "--lost" in sys.argv:
    te = hdr.index("Thread Instructions Executed")
    fn_eff, line_eff = {}, {}
    for k in range(n):
        i_k, t_k = int(body[k][ie]), int(body[k][te])
        seen = set()
        for fr in chains[k]:
            fn = func_of(fr)
            if fn not in seen:
                seen.add(fn)
                a = fn_eff.setdefault(fn, [0, 0])
                a[0] += i_k
                a[1] += t_k
        a = line_eff.setdefault(lines[k][0], [0, 0])
        a[0] += i_k
        a[1] += t_k
    tot_lost = sum(32 * a[0] - a[1] for a in line_eff.values())
    print("---- divergence: share of warp-instructions, average active threads, share of all lost lane-slots ----")
    for fn, (i, t) in sorted(fn_eff.items(), key=lambda kv: -(32 * kv[1][0] - kv[1][1]))[:16]:
        print("%-24s %5.1f%% inst  %4.1f thr  %5.1f%% of lost" % (fn, 100.0 * i / tot_i, t / max(i, 1), 100.0 * (32 * i - t) / tot_lost))
    for key, (i, t) in sorted(line_eff.items(), key=lambda kv: -(32 * kv[1][0] - kv[1][1]))[:24]:
        print("%-26s %5.1f%% inst  %4.1f thr  %5.1f%% of lost  %s" % ("%s:%d" % key if key else "?", 100.0 * i / tot_i, t / max(i, 1),
                                                                     100.0 * (32 * i - t) / tot_lost, src_line(key)[:70]))
print("---- per innermost line ----")
for key, (i, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% inst %5.1f%% smp  %-28s %s" % (100.0 * i / tot_i, 100.0 * s / max(tot_s, 1), "%s:%d" % key if key else "?", src_line(key)))
if "--sass" in sys.argv:          # --sass <function>: the SASS rows inlined from that function, with executed counts per record
    want = sys.argv[sys.argv.index("--sass") + 1]
    nrec = float(sys.argv[sys.argv.index("--records") + 1]) if "--records" in sys.argv else 1.0
    print("---- SASS of %s (warp-instructions executed / record count %g) ----" % (want, nrec))
    for k in range(n):
        if any(func_of(fr) == want for fr in chains[k]):
            print("%8.3f  %-12s %s" % (int(body[k][ie]) / nrec, "%s:%d" % lines[k][0] if lines[k][0] else "?", lines[k][1]))
