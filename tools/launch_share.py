#!/usr/bin/env python3
"""Kernel shares from an ncu launch list (--metrics gpu__time_duration.sum --csv).  usage: launch_share.py list.csv"""
import collections
import csv
import re
import sys

tot = collections.defaultdict(lambda: [0, 0.0])
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
h = rows[0]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("rtb::", "").replace("<unnamed>::", "")
    name = re.sub(r"<.*", "", name) if name.startswith("at::") else name
    tot[name][0] += 1
    tot[name][1] += float(r[vi].replace(",", ""))
ours = sum(v[1] for k, v in tot.items() if k.startswith("k_"))
print("launches %d, ours %.3f ms (per-launch times under ncu are serialised and cold-cache: compare SHARES)" % (len(rows) - 1, ours * 1e-6))
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-46s n=%5d  %10.3f ms  %5.1f %% of our kernels" % (k[:46], v[0], v[1] * 1e-6, 100 * v[1] / ours if k.startswith("k_") else 0.0))
