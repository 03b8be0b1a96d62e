#!/usr/bin/env python
"""Per-CUDA-line warp-stall samples of one kernel from an `ncu --set full --import-source on` report.
usage: python profiles/ncu_source_lines.py <report.ncu-rep> <kernel regex> [min share, default 0.02]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.02
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
hdr = rows[hi]
si = hdr.index("# Samples")
names = ["stall_long_sb", "stall_short_sb", "stall_barrier", "stall_math", "stall_wait", "stall_lg", "stall_mio", "stall_not_selected"]
ci = [hdr.index(n) for n in names]
body = [r for r in rows[hi + 1:] if len(r) > max(ci) and r[si].isdigit() and r[0].strip()]
tot = sum(int(r[si]) for r in body)
print("kernel %s: %d warp-stall samples; lines with >= %.0f%% of them" % (kern, tot, 100 * thr))
print("share  line  " + " ".join(n.replace("stall_", "") for n in names))
for r in body:
    s = int(r[si])
    if s >= tot * thr:
        print("%5.1f%% L%-4s %s | %s" % (100.0 * s / tot, r[0], " ".join("%6s" % r[c] for c in ci), r[1].strip()[:110]))
