#!/usr/bin/env python3
"""Executed warp instructions by SASS opcode for one kernel (ncu --page source --print-source sass --csv)."""
import csv, sys, collections
hdr = None; agg = collections.Counter()
for r in csv.reader(open(sys.argv[1], errors="replace")):
    if len(r) > 5 and ("Source" in r and "Instructions Executed" in r): hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    try: v = float(r[hdr.index("Instructions Executed")])
    except ValueError: continue
    src = r[hdr.index("Source")].strip()
    t = src.split()
    if not t: continue
    op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    agg[op.split(".")[0] if len(sys.argv) < 3 else op] += v
tot = sum(agg.values())
print("total %.4g" % tot)
for k, v in agg.most_common(30): print("%6.2f%%  %s" % (v / tot * 100, k))
