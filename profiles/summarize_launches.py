#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: the launches of the LAST forward
(between the last two norm/mean markers is fragile, so: the last N launches where N = launches per forward,
found from the period of vfi_b200 kernels) grouped by kernel name, with shares.
usage: summarize_launches.py launches.csv [n_forwards_in_file]"""
import csv, sys, re, collections
rows = []
with open(sys.argv[1]) as f:
    for line in f:
        if line.startswith('"ID"'):
            break
    rd = csv.reader(f)
    for r in rd:
        if len(r) >= 15 and r[12] == "gpu__time_duration.sum":
            rows.append((r[4], float(r[14])))
nfw = int(sys.argv[2]) if len(sys.argv) > 2 else 4
# index of dattn launches: 3 per forward
idx = [i for i, (k, _) in enumerate(rows) if "dattn" in k]
assert len(idx) >= 6, len(idx)
# one forward = from after the lv1 dattn of forward n-1 ... shift to start at first kernel after prev forward's end:
# use the period between consecutive lv3 dattn launches (same position in each forward)
start, end = idx[-6], idx[-3]
fw = rows[start:end]
tot = sum(t for _, t in fw)
def short(k):
    k = re.sub(r"\(.*", "", k); k = re.sub(r"<.*", "", k); return k.replace("void ", "")[:70]
g = collections.OrderedDict()
for k, t in fw:
    s = short(k); a = g.setdefault(s, [0, 0.0]); a[0] += 1; a[1] += t
print(f"total launches in file: {len(rows)}; one steady-state forward = {len(fw)} launches, {tot/1e6:.2f} ms (sum of serialised, cold-cache durations)")
print(f"{'kernel':72s} {'n':>5s} {'ms':>9s} {'share':>7s}")
for k, (n, t) in sorted(g.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:72s} {n:5d} {t/1e6:9.3f} {100*t/tot:6.1f}%")
