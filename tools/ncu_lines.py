#!/usr/bin/env python3
"""Per-source-line summary of an ncu capture (warp instructions, shared wavefronts, stall samples
per unit of work).  usage: ncu_lines.py prof.ncu-rep units [min_inst_per_unit]"""
import csv, subprocess, sys
rep, units = sys.argv[1], float(sys.argv[2])
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 5.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname = ""
h = None
tot = 0.0
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        h = r
        ie, iw, ism = h.index("Instructions Executed"), h.index("L1 Wavefronts Shared"), h.index("# Samples")
    elif h and r and r[0].isdigit() and len(r) > ie:
        try:
            n = float(r[ie])
        except ValueError:
            continue
        tot += n
        if n / units >= thr:
            print("%-22s %4s %8.1f inst %7.1f wf %7s smp  %s" % (fname, r[0], n / units, float(r[iw] or 0) / units, r[ism], r[1].strip()[:80]))
print("total inst/unit", tot / units)
