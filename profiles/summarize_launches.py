#!/usr/bin/env python
"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/r01_capture.sh).
usage: python profiles/summarize_launches.py launches.csv[.gz] > summary.txt"""
import csv, gzip, re, sys, collections
fn = sys.argv[1]
f = gzip.open(fn, "rt") if fn.endswith(".gz") else open(fn)
rows = [r for r in csv.reader(l for l in f if l.startswith('"'))]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); gi = hdr.index("Grid Size"); bi = hdr.index("Block Size")
tot = collections.OrderedDict()
for r in rows[1:]:
    name = re.match(r"(?:void )?([A-Za-z_0-9]+)", r[ki]).group(1)
    if name.startswith("k_amg_") and not name.startswith("k_amg_enter"):      # one launch per multigrid level: keep the levels apart by grid size
        name += "[%s]" % r[gi].split(",")[0].strip("( ")
    t = tot.setdefault(name, [0, 0.0, r[gi], r[bi]])
    t[0] += 1; t[1] += float(r[vi].replace(",", "")) * 1e-6
alltime = sum(t[1] for t in tot.values())
print("%d launches, %.1f ms of kernel time (cold-cache, serialised: compare shares)" % (len(rows) - 1, alltime))
print("%-28s %8s %12s %10s %7s  %s" % ("kernel", "launches", "total ms", "avg ms", "share", "grid x block (first launch)"))
for name, t in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-28s %8d %12.3f %10.4f %6.1f%%  %s x %s" % (name[:28], t[0], t[1], t[1] / t[0], 100 * t[1] / alltime, t[2], t[3]))
