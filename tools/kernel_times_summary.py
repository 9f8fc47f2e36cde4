#!/usr/bin/env python
"""Summarise tools/kernel_times.sh output: last launch of each kernel name -> us, instructions."""
import csv, sys, re, collections
rows = []
with open(sys.argv[1]) as f:
    for line in f:
        if line.startswith('"ID"'):
            break
    for r in csv.reader(f):
        if len(r) >= 15:
            rows.append((r[4], r[12], float(r[14].replace(",", ""))))
last = collections.OrderedDict()
for k, m, v in rows:
    k = re.sub(r"\(.*", "", k).replace("void ", "")
    last.setdefault(k, {})[m] = v
for k, d in last.items():
    print(f"{k[:60]:60s} {d.get('gpu__time_duration.sum', 0)/1e3:10.1f} us {d.get('smsp__inst_executed.sum', 0)/1e6:10.1f} Minst")
