#!/usr/bin/env python
"""Summarise an `ncu --set full` report (exported with `ncu -i X.ncu-rep --page raw --csv`) per kernel:
launches captured, mean duration, DRAM bytes read + written per launch (the `roofline.traffic` figure of bench.py),
tensor-pipe and memory-pipe activity.  usage: ncu_traffic.py raw.csv out.json"""
import csv, json, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
def val(r, name):
    i = col.get(name)
    if i is None or r[i] in ("", "n/a"):
        return None
    v = float(r[i].replace(",", ""))
    u = units[i]
    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ns": 1e-3, "ms": 1e3}.get(u, 1.0)
    return v * scale
out = {}
for r in rows[2:]:
    if len(r) != len(hdr):
        continue
    name = re.sub(r"^void ", "", r[col["Kernel Name"]]).split("(")[0]
    short = name.split("<")[0].split("::")[-1]
    e = out.setdefault(short, {"launches": 0, "duration_us": 0.0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0,
                               "tensor_pipe_active_pct": 0.0, "lsu_mem_pct": 0.0, "variants": set()})
    e["variants"].add(name)
    e["launches"] += 1
    e["duration_us"] += val(r, "gpu__time_duration.sum") or 0.0
    e["dram_read_bytes"] += val(r, "dram__bytes_read.sum") or 0.0
    e["dram_write_bytes"] += val(r, "dram__bytes_write.sum") or 0.0
    e["tensor_pipe_active_pct"] += val(r, "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active") or 0.0
    e["lsu_mem_pct"] += val(r, "l1tex__throughput.avg.pct_of_peak_sustained_active") or 0.0
for e in out.values():
    n = e["launches"]
    for k in ("duration_us", "dram_read_bytes", "dram_write_bytes", "tensor_pipe_active_pct", "lsu_mem_pct"):
        e[k] = round(e[k] / n, 2)
    e["dram_bytes_per_launch"] = round(e["dram_read_bytes"] + e["dram_write_bytes"], 1)
    e["variants"] = sorted(e["variants"])
json.dump(out, open(sys.argv[2], "w"), indent=1)
for k, e in sorted(out.items(), key=lambda kv: -kv[1]["duration_us"] * kv[1]["launches"]):
    print(f"{k:28s} n={e['launches']:3d} {e['duration_us']:8.1f} us  dram {e['dram_bytes_per_launch'] / 1e6:8.1f} MB/launch  tensor {e['tensor_pipe_active_pct']:5.1f}%")
