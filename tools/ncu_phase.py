#!/usr/bin/env python3
"""Bucket the executed warp instructions of one kernel (ncu --page source csv) by CUDA source line ranges.
  python tools/ncu_phase.py src.csv file.cu 218-229:load 231-263:phase1 ...   (helper-header lines are listed apart)"""
import csv, sys
path, fname = sys.argv[1], sys.argv[2]
rng = []
for a in sys.argv[3:]:
    r, n = a.split(":"); lo, hi = r.split("-"); rng.append((int(lo), int(hi), n))
cur, hdr = "", None
tot = {}; other = {}
for r in csv.reader(open(path, errors="replace")):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < 8 or not r[0].strip().isdigit(): continue
    try: v = float(r[hdr.index("Instructions Executed")])
    except ValueError: continue
    ln = int(r[0])
    if cur == fname:
        for lo, hi, n in rng:
            if lo <= ln <= hi: tot[n] = tot.get(n, 0) + v; break
        else: tot["?%d" % ln] = tot.get("?%d" % ln, 0) + v
    else: other["%s:%d" % (cur, ln)] = other.get("%s:%d" % (cur, ln), 0) + v
s = sum(tot.values()) + sum(other.values())
for k, v in sorted(tot.items(), key=lambda x: -x[1]): print("%6.2f%%  %s" % (v / s * 100, k))
print("-- inlined helpers")
for k, v in sorted(other.items(), key=lambda x: -x[1])[:25]: print("%6.2f%%  %s" % (v / s * 100, k))
