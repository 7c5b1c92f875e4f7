#!/usr/bin/env python
"""Summarise an ncu per-launch capture of gemm_tc_kernel (gpu__time_duration, dram bytes read/write, tensor-pipe
active %) into a table + a JSON blob bench.py reads for roofline.traffic.
usage: gemm_traffic_summary.py capture.csv out.txt out.json"""
import collections, csv, hashlib, json, os, re, sys
rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
per = collections.OrderedDict()
for r in csv.DictReader(rows):
    d = per.setdefault(r["ID"], {"name": re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", ""), "grid": r.get("Grid Size")})
    d[r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
ns = lambda v, u: v * {"ns": 1, "us": 1e3, "ms": 1e6}[u]
by = lambda v, u: v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
T = Bt = 0.0
for d in per.values():
    t = ns(*d["gpu__time_duration.sum"])
    b = by(*d["dram__bytes_read.sum"]) + by(*d["dram__bytes_write.sum"])
    tp = d["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"][0]
    a = agg[(d["name"][:48], d["grid"], round(t / 4e3) * 4)]
    a[0] += 1; a[1] += t; a[2] += b; a[3] += tp * t
    T += t; Bt += b
n = len(per)
tw = sum(a[3] for a in agg.values()) / T
out = [f"{n} gemm_tc_kernel launches (one training step, eager, ncu --clock-control none; cold-cache serialised timings)",
       f"total {T/1e6:.3f} ms; mean DRAM traffic {Bt/n/1e6:.2f} MB/launch; time-weighted tensor-pipe active {tw:.1f}%", ""]
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"{a[1]/1e6:7.3f} ms n={a[0]:3d} avg {a[1]/a[0]/1e3:6.1f} us  dram {a[2]/a[0]/1e6:6.1f} MB/launch {a[2]/a[1]:7.1f} GB/s  "
               f"tensor {a[3]/a[1]:5.1f}%  {key[0]} grid {key[1]}")
open(sys.argv[2], "w").write("\n".join(out) + "\n")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sha = hashlib.sha256(open(os.path.join(ROOT, "privacy-preserve-federated-asr_b200", "csrc", "gemm_tcgen05.cu"), "rb").read()).hexdigest()[:16]
json.dump({"kernel": "gemm_tc_kernel", "launches": n, "dram_bytes_per_launch": Bt / n, "tensor_pipe_active_pct_time_weighted": tw,
           "source": sys.argv[1], "gemm_source_sha16": sha}, open(sys.argv[3], "w"))
print(out[1])
