#!/usr/bin/env python3
"""Dynamic SASS opcode mix of one ncu capture, per unit of work (executed warp-instructions, shared-memory
wavefronts and stall samples per opcode).  usage: ncu_opcodes.py prof.ncu-rep units"""
import collections, csv, subprocess, sys
rep, units = sys.argv[1], float(sys.argv[2])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[1]
ie, iw, ism = h.index("Instructions Executed"), h.index("L1 Wavefronts Shared"), h.index("# Samples")
cnt, wf, smp = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows[2:]:
    try:
        n = float(r[ie])
    except (ValueError, IndexError):
        continue
    s = r[1].split()
    op = (s[1] if s[0].startswith("@") else s[0]).rstrip(";")
    base = op.split(".")[0]
    if base in ("LDS", "STS", "MUFU", "SHFL"):
        base = op
    cnt[base] += n
    wf[base] += float(r[iw] or 0)
    smp[base] += float(r[ism] or 0)
print(rows[0][1])
print("total %.1f warp-instructions, %.1f shared wavefronts per unit (%g units), %d samples"
      % (sum(cnt.values()) / units, sum(wf.values()) / units, units, sum(smp.values())))
for k, v in cnt.most_common():
    if v / units >= 1.0:
        print("%-12s %7.1f inst %7.1f wf %7d smp" % (k, v / units, wf[k] / units, smp[k]))
