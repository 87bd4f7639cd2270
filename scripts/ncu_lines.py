#!/usr/bin/env python
"""Per-source-line hot spots of one kernel from an .ncu-rep captured with --import-source on.
usage: python scripts/ncu_lines.py REP KERNEL_REGEX [top_n]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k",
                      "regex:" + kern, "--launch-count", "1"], capture_output=True, text=True).stdout
rows, cur_file = [], None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] in ("Function Name", "Line No") or len(r) < 10:
        continue
    if r[2] == "-":     # aggregate row of a source line
        try:
            rows.append((cur_file, int(r[0]), r[1].strip(), int(r[6] or 0), int(r[7] or 0), int(r[8] or 0)))
        except ValueError:
            pass
tot_s = sum(r[3] for r in rows) or 1
tot_i = sum(r[4] for r in rows) or 1
tot_t = sum(r[5] for r in rows) or 1
print("# %s: samples %d, warp inst %d, thread inst %d (avg %.1f threads/inst)" % (kern, tot_s, tot_i, tot_t, tot_t / tot_i))
print("%-14s %5s %7s %7s %6s  %s" % ("file", "line", "samp%", "inst%", "thr/in", "source"))
for f, ln, src, s, i, t in sorted(rows, key=lambda r: -r[4])[:top]:
    print("%-14s %5d %6.2f%% %6.2f%% %6.1f  %s" % (f, ln, 100.0 * s / tot_s, 100.0 * i / tot_i, t / max(i, 1), src[:110]))
