#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: stall-reason totals and the hottest instructions."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = next(r for r in rows if '# Samples' in r)
si, src = hdr.index('# Samples'), hdr.index('Source')
stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = [r for r in rows if len(r) == len(hdr) and r[si].isdigit()]
tot = sum(int(r[si]) for r in data)
print('total samples', tot)
agg = {hdr[i]: sum(int(r[i]) for r in data) for i in stalls}
for k, v in sorted(agg.items(), key=lambda x: -x[1])[:10]:
    print(f"  {k:28s} {v:8d} {v / tot * 100:5.1f}%")
for r in sorted(data, key=lambda r: -int(r[si]))[:n]:
    st = sorted([(int(r[i]), hdr[i][6:]) for i in stalls], reverse=True)[:2]
    print(f"{int(r[si]):7d} {r[src].strip()[:64]:64s} {st}")
