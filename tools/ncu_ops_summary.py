#!/usr/bin/env python
"""Summarise the raw-page CSV of an `ncu --set full` capture (tools/gpu_run.sh ncu_ops) per kernel: duration, DRAM bytes and
share of the peak, the busiest unit (sm__throughput), issue-slot, FMA / ALU / XU (MUFU) / tensor pipe activity, occupancy.
usage: ncu_ops_summary.py raw.csv > summary.txt"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
def f(r, name):
    v = r[col[name]].replace(",", "")
    try: return float(v)
    except ValueError: return float("nan")
def mbytes(r, name):
    v, u = f(r, name), units[col[name]]
    return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}[u]
def usec(r, name):
    v, u = f(r, name), units[col[name]]
    return v * {"ns": 1e-3, "us": 1.0, "ms": 1e3}[u]
agg = collections.OrderedDict()
for r in data:
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void ", "").replace("b200asr::", "")
    key = (name[:60], r[col["Grid Size"]], r[col["Block Size"]])
    agg.setdefault(key, []).append(r)
M = ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
     "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
     "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
     "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
     "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
print("ncu --set full --clock-control none, tools/op_one.py at the sizes of BASELINE configs[1] (8 x 10 s: 3992 tokens, T = 499, 16 heads);")
print("per kernel (mean over its captured launches): duration, DRAM MB read + written (GB/s), then % of peak: DRAM, busiest SM unit,")
print("issue slots, FMA pipe, ALU pipe, XU (MUFU) pipe, tensor pipe, warps active (occupancy), L1/TEX, L2\n")
print(f"{'kernel':62s} {'grid':>10s} {'n':>2s} {'us':>7s} {'MB rd+wr':>14s} {'GB/s':>6s} | dram   sm  issue  fma  alu   xu  tens  occ   l1   l2")
for (name, grid, block), rs in agg.items():
    n = len(rs)
    t = sum(usec(r, "gpu__time_duration.sum") for r in rs) / n
    rd = sum(mbytes(r, "dram__bytes_read.sum") for r in rs) / n
    wr = sum(mbytes(r, "dram__bytes_write.sum") for r in rs) / n
    vals = [sum(f(r, m) for r in rs) / n for m in M]
    print(f"{name:62s} {grid.replace(' ', ''):>10s} {n:2d} {t:7.1f} {rd:6.1f}+{wr:6.1f} {(rd + wr) / t * 1e3:7.0f} | " + " ".join(f"{v:4.0f}" for v in vals))
