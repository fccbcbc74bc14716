"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel launches, total and mean duration,
share.  Usage: python tools/launch_summary.py gpurun_out/launches.csv [last_n]  (last_n: only the last n launches)."""
import csv
import re
import sys
from collections import OrderedDict


def main(path, last=None):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
        name = re.sub(r"<.*", "", r["Kernel Name"]).replace("mdgen::", "")
        name = re.sub(r"\(.*", "", name)
        rows.append((name, us))
    if last:
        rows = rows[-last:]
    agg = OrderedDict()
    for n, us in rows:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print(f"{len(rows)} launches, total {tot:.1f} us")
    for n, (k, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:46s} {k:5d} launches {us:10.1f} us {us / k:9.1f} us avg {100 * us / tot:6.1f}%")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else None)
