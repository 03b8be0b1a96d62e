#!/usr/bin/env python
"""Summarise `ncu --set full` reports (profiles/r01_capture.sh) into a text table and a per-kernel DRAM-traffic file.

usage: python profiles/make_ncu_summary.py <report.ncu-rep> <units> <out.txt> [traffic.json]
  units = what one launch of these kernels processes at the captured size: wet (node, layer) pairs for the tracer
          kernels, matrix rows for the SSH kernels (bench.py prints both in its JSON line)
The traffic file maps kernel name -> DRAM bytes per unit (dram__bytes_read.sum + dram__bytes_write.sum of one launch /
units); bench.py multiplies it by the units of its own run to fill roofline.traffic."""
import csv, json, re, subprocess, sys, collections, io

METRICS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
           ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"), ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
           ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %"),
           ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue act %"), ("launch__registers_per_thread", "regs"),
           ("smsp__inst_executed.sum", "warp insts"), ("l1tex__data_pipe_lsu_wavefronts.sum", "L1 wavefronts"),
           ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64 pipe %")]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def main():
    rep, units, out = sys.argv[1], float(sys.argv[2]), sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, unit_row = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    per = collections.OrderedDict()
    for r in rows[2:]:
        name = re.match(r"(?:void )?([A-Za-z_0-9]+)", r[ki]).group(1)      # template arguments dropped: k_lo<0, 1> -> k_lo
        if name.startswith("k_amg_") and "Grid Size" in hdr:               # one kernel per multigrid level: keep the levels apart
            name += "[grid %s]" % r[hdr.index("Grid Size")].replace(" ", "")
        d = per.setdefault(name, collections.defaultdict(list))
        for m, _ in METRICS:
            if m in hdr:
                i = hdr.index(m)
                try:
                    d[m].append(float(r[i].replace(",", "")) * SCALE.get(unit_row[i], 1.0))
                except ValueError:
                    pass
    lines = ["source: %s   (ncu --set full --clock-control none; per-launch values, mean over the captured launches)" % rep,
             "units per launch at the captured size: %d" % units, ""]
    traffic = {}
    for name, d in per.items():
        n = len(d["gpu__time_duration.sum"])
        mean = {m: (sum(v) / len(v) if v else float("nan")) for m, v in d.items()}
        t = mean["gpu__time_duration.sum"]
        by = mean.get("dram__bytes_read.sum", 0.0) + mean.get("dram__bytes_write.sum", 0.0)
        traffic[name] = {"dram_bytes_per_unit": by / units, "dram_bytes_per_launch": by, "units": units, "launches": n}
        lines.append("%s   (%d launches)" % (name, n))
        lines.append("    time %.4f ms   DRAM read %.3f GB + write %.3f GB = %.1f B/unit -> %.0f GB/s under ncu" % (
            t * 1e3, mean.get("dram__bytes_read.sum", 0) / 1e9, mean.get("dram__bytes_write.sum", 0) / 1e9, by / units, by / t / 1e9))
        lines.append("    " + "   ".join("%s %.4g" % (lab, mean[m]) for m, lab in METRICS[3:] if m in mean))
        lines.append("")
    open(out, "w").write("\n".join(lines))
    if len(sys.argv) > 4:
        try:
            old = json.load(open(sys.argv[4]))
        except (OSError, ValueError):
            old = {}
        old.update(traffic)
        json.dump(old, open(sys.argv[4], "w"), indent=1, sort_keys=True)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
