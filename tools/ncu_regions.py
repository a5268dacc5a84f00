#!/usr/bin/env python
"""Per-region instruction / stall breakdown of one kernel launch from `ncu -i X.ncu-rep --page source --csv` output.
usage: ncu_regions.py <source.csv> <first line of the launch's section (1-based)> <last line>  [--list]
Regions are delimited by landmark instructions (LDG.E.128 groups, WARPSYNC, VOTE) printed with --list."""
import csv, sys
path, a, b = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
rows = list(csv.reader(open(path)))[a - 1:b]
hdr = rows[1]
ia, isrc, ist, iex, ith = (hdr.index(k) for k in ("Address", "Source", "Warp Stall Sampling (All Samples)", "Instructions Executed", "Avg. Threads Executed"))
data = []
for r in rows[2:]:
    try:
        data.append((int(r[ia], 16), r[isrc].strip(), int(r[ist]), int(r[iex]), float(r[ith])))
    except Exception:
        pass
tot_ex, tot_st = sum(d[3] for d in data), sum(d[2] for d in data)
print("SASS instructions %d, warp instructions executed %.1f M, stall samples %d" % (len(data), tot_ex / 1e6, tot_st))
if "--list" in sys.argv:
    for i, d in enumerate(data):
        if any(k in d[1] for k in ("LDG", "WARPSYNC", "VOTE", "EXIT", "ATOM", "BAR", "LDS", "STS", "RET", "CALL")):
            print(i, d[1][:64], "stall", d[2], "exec %.2fM" % (d[3] / 1e6), "thr %.1f" % d[4])
    sys.exit(0)
marks = [int(x) for x in sys.argv[4].split(",")]
names = sys.argv[5].split(",")
for (lo, hi), nm in zip(zip(marks[:-1], marks[1:]), names):
    ex = sum(d[3] for d in data[lo:hi]); st = sum(d[2] for d in data[lo:hi]); thr = sum(d[3] * d[4] for d in data[lo:hi]) / max(1, ex)
    print("%-34s idx %4d-%4d  exec %7.1fM (%4.1f%%)  avg threads %4.1f  stall %5d (%4.1f%%)" % (nm, lo, hi, ex / 1e6, 100 * ex / tot_ex, thr, st, 100 * st / max(1, tot_st)))
top = sorted(range(len(data)), key=lambda i: -data[i][2])[:12]
for i in top:
    print("  top stall", i, data[i][1][:60], data[i][2], "exec %.2fM" % (data[i][3] / 1e6), "thr %.1f" % data[i][4])
