#!/usr/bin/env python
"""Instruction / sample share of source-line regions of one kernel (ncu --import-source on).
usage: python scripts/ncu_regions.py REP KERNEL_REGEX FILE name:lo-hi [name:lo-hi ...]"""
import collections
import csv
import subprocess
import sys

rep, kern, fname = sys.argv[1:4]
regs = []
for a in sys.argv[4:]:
    n, r = a.split(":")
    lo, hi = r.split("-")
    regs.append((n, int(lo), int(hi)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k",
                      "regex:" + kern, "--launch-count", "1"], capture_output=True, text=True).stdout
rows, cur = [], None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] in ("Function Name", "Line No") or len(r) < 10:
        continue
    if r[2] == "-":
        try:
            rows.append((cur, int(r[0]), int(r[6] or 0), int(r[7] or 0), int(r[8] or 0)))
        except ValueError:
            pass


def region(f, l):
    if f != fname:
        return "other:" + f
    for n, lo, hi in regs:
        if lo <= l <= hi:
            return n
    return fname + ":other"


ti, ts = sum(r[3] for r in rows) or 1, sum(r[2] for r in rows) or 1
ai, as_, at = collections.Counter(), collections.Counter(), collections.Counter()
for f, l, s, i, t in rows:
    k = region(f, l)
    ai[k] += i
    as_[k] += s
    at[k] += t
print("# %s: %d warp instructions, %d samples" % (kern, ti, ts))
for k, v in ai.most_common():
    print("%-32s inst %5.1f%%  samples %5.1f%%  threads/inst %4.1f" % (k, 100.0 * v / ti, 100.0 * as_[k] / ts, at[k] / max(v, 1)))
