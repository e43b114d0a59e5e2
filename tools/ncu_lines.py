#!/usr/bin/env python3
"""Per-source-line summary of an .ncu-rep (needs -lineinfo + --import-source on): share of executed warp
instructions, average active threads per instruction, share of stall samples.  usage: ncu_lines.py rep [top]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
hdr, cur, out = None, None, []
for r in csv.reader(txt.splitlines()):
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].isdigit():
        g = lambda name: int(float(r[hdr.index(name)] or 0)) if name in hdr else 0
        out.append((g("Instructions Executed"), g("Thread Instructions Executed"), g("# Samples"), cur, r[0], r[1].strip()[:100]))
tot = sum(o[0] for o in out) or 1
ts = sum(o[2] for o in out) or 1
thr = sum(o[1] for o in out)
print("warp instructions %d, avg active threads %.2f, samples %d" % (tot, thr / tot, ts))
for o in sorted(out, reverse=True)[:top]:
    print("%5.1f%% inst  thr/inst %4.1f  samp %4.1f%%  %s:%s  %s" % (100 * o[0] / tot, o[1] / max(1, o[0]), 100 * o[2] / ts, o[3], o[4], o[5]))
