#!/usr/bin/env python
"""Last-launch device time per distinct kernel from an `ncu --metrics gpu__time_duration.sum --csv` log.
  tools/launch_list.py gpurun_out/x.csv [more.csv ...]"""
import csv
import sys

for f in sys.argv[1:]:
    rows = list(csv.reader(open(f)))
    hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
    h = rows[hi]
    kn, mv, mn = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Name')
    seen = {}
    for r in rows[hi + 1:]:
        if len(r) > mv and r[mn] == 'gpu__time_duration.sum':
            seen.setdefault(r[kn].split('(')[0][:70], []).append(float(r[mv].replace(',', '')) / 1000)
    print(f)
    for k, v in seen.items():
        print(f"{v[-1]:9.1f} us (last of {len(v)}, mean {sum(v)/len(v):7.1f})  {k}")
