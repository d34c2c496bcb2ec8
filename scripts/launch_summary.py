"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name, launches and total time.
usage: python scripts/launch_summary.py launches.csv [first_id last_id]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None
d = collections.OrderedDict()
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 10**9
tot = 0.0
for r in rows:
    if len(r) > 10 and r[0] == 'ID':
        hdr = r; continue
    if hdr and len(r) == len(hdr):
        i = int(r[0])
        if i < lo or i > hi: continue
        k = r[hdr.index('Kernel Name')][:70] + ' g=' + r[hdr.index('Grid Size')]
        v = float(r[hdr.index('Metric Value')].replace(',', ''))
        if r[hdr.index('Metric Unit')] == 'ns': v /= 1e3
        d.setdefault(k, [0, 0]); d[k][0] += v; d[k][1] += 1; tot += v
for k, (v, n) in sorted(d.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:100s} {n:4d} {v:10.1f} us {100*v/tot:5.1f}%")
print("total", round(tot, 1), "us")
