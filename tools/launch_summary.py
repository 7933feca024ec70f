"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = None
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if hdr is None:
        if 'Kernel Name' in r:
            hdr = r
        continue
    if len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    if d.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    v = float(d['Metric Value'].replace(',', ''))
    u = d['Metric Unit']
    v = v / 1000 if u in ('ns', 'nsecond') else v * 1000 if u in ('ms', 'msecond') else v
    k = re.sub(r'\(.*', '', d['Kernel Name'])[:90]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
    print(f"{v[1]/1000:9.2f} ms {v[0]:5d} x {v[1]/v[0]:8.1f} us  {100*v[1]/tot:5.1f}%  {k}")
print(f"total {tot/1000:.2f} ms")
