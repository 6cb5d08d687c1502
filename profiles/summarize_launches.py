"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total time and share per kernel.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/<name>.txt"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
gi, bi = hdr.index('Grid Size'), hdr.index('Block Size')
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(',', ''))
    except ValueError:
        continue
    a = agg.setdefault(r[ki], [0, 0.0, r[gi], r[bi]])
    a[0] += 1
    a[1] += v
tot = sum(v[1] for v in agg.values())
print(f'# {sys.argv[1]}: {sum(v[0] for v in agg.values())} launches, {tot / 1e6:.3f} ms total (ncu, cold-cache, serialised)')
print(f'{"kernel":70s} {"launches":>8s} {"total_us":>12s} {"avg_us":>10s} {"share":>6s}  last grid/block')
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f'{k[:70]:70s} {v[0]:8d} {v[1] / 1e3:12.1f} {v[1] / 1e3 / v[0]:10.1f} {v[1] / tot:6.3f}  {v[2]} {v[3]}')
