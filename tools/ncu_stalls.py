#!/usr/bin/env python3
"""Top warp-stall reasons per kernel from an ncu report (warps stalled per issue-active cycle, first launch of each kernel).
  python tools/ncu_stalls.py gpurun_out/X.ncu-rep"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
ki = hdr.index("Kernel Name")
pre, suf = "smsp__average_warps_issue_stalled_", "_per_issue_active.ratio"
cols = [i for i, h in enumerate(hdr) if h.startswith(pre) and h.endswith(suf)]
seen = set()
for r in rows[2:]:
    n = r[ki].split("(")[0]
    if n in seen:
        continue
    seen.add(n)
    vals = sorted(((float(r[i]), hdr[i][len(pre):-len(suf)]) for i in cols if r[i] not in ("", "n/a")), reverse=True)[:6]
    print("%-26s" % n, "  ".join("%s=%.2f" % (b, a) for a, b in vals))
