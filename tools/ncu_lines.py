#!/usr/bin/env python
"""Summarise an .ncu-rep (needs -lineinfo + --import-source on): headline metrics, executed SASS opcode
histogram, executed instructions per source line.  usage: ncu_lines.py rep.ncu-rep [configs_per_launch] [top]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ncfg = float(sys.argv[2]) if len(sys.argv) > 2 else 1048576.0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
W = ncfg / 32.0

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum", "lts__t_sector_hit_rate.pct",
        "launch__grid_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]
for h, u, v in zip(hdr, units, vals):
    if h in want or h.startswith("smsp__average_warps_issue_stalled") and float(v or 0) > 0.15:
        print("%-80s %-12s %s" % (h, u, v))

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True,
                     text=True).stdout
cur_file = None
agg = collections.Counter()
smp = collections.Counter()
opagg = collections.defaultdict(collections.Counter)
ops = collections.Counter()
text = {}
line = None
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if len(r) < 8 or r[0] == "Line No":
        continue
    if r[0] != "":
        line = (cur_file, int(r[0]))
        text[line] = r[1].strip()
        continue
    try:
        e = int(r[7])
    except ValueError:
        continue
    agg[line] += e
    try:
        smp[line] += int(r[6])
    except ValueError:
        pass
    s = r[3].strip()
    if s.startswith("@"):
        s = s.split(None, 1)[1]
    op = s.split()[0].split(".")[0]
    opagg[line][op] += e
    ops[op] += e
# kernels compiled at run time carry only file names (embedded headers): take the text from the source tree
import os
_src_cache = {}
def _text(f, l):
    t = text.get((f, l), "")
    if t and "Unable to open" not in t:
        return t
    if f not in _src_cache:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tinyfk_b200", "csrc", f)
        _src_cache[f] = open(path).read().splitlines() if os.path.exists(path) else []
    lines = _src_cache[f]
    return lines[l - 1].strip() if 0 < l <= len(lines) else t
for k in list(text):
    text[k] = _text(*k)
tot = sum(agg.values())
print("\nexecuted warp instructions per 32 configs: %.0f" % (tot / W))
print(" ".join("%s=%.0f" % (k, v / W) for k, v in ops.most_common(24)))
byfile = collections.Counter()
for (f, l), c in agg.items():
    byfile[f] += c
print({k: round(v / W) for k, v in byfile.most_common()})
for (f, l), c in agg.most_common(top):
    print("%-16s %4d %6.0f  %-64s %s" % (f, l, c / W, text[(f, l)][:64], dict((k, round(v / W)) for k, v in opagg[(f, l)].most_common(4))))

# warp-stall samples per source line: where the warps' TIME goes (samples ~ warp-cycles), next to the instruction share
ts = float(sum(smp.values())) or 1.0
print("\nstall samples: %d; share of samples vs share of executed instructions, per file" % ts)
sf = collections.Counter()
for (f, l), c in smp.items():
    sf[f] += c
print({k: "%.1f%% / %.1f%%" % (100 * v / ts, 100.0 * byfile[k] / tot) for k, v in sf.most_common()})
for (f, l), c in smp.most_common(top):
    print("%-16s %4d %5.1f%% (instr %4.1f%%)  %-64s" % (f, l, 100 * c / ts, 100.0 * agg[(f, l)] / tot, text[(f, l)][:64]))
