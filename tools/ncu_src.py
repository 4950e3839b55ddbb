#!/usr/bin/env python3
"""Summarise `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv -k regex:<kernel>` output:
top CUDA source lines by executed warp instructions, with stall samples."""
import csv, sys
path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
out, cur_file, hdr = [], "", None
for r in csv.reader(open(path, errors="replace")):
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < 8 or not r[0].strip().isdigit():
        continue
    ie = hdr.index("Instructions Executed"); sm = hdr.index("# Samples")
    try: v = float(r[ie])
    except ValueError: continue
    out.append((v, cur_file, r[0], r[1], r[sm]))
agg = {}
for v, f, ln, s, smp in out:  # aggregate SASS rows (and repeated launches) by CUDA source line
    k = (f, ln, s)
    a = agg.setdefault(k, [0.0, 0])
    a[0] += v
    try: a[1] += int(smp)
    except ValueError: pass
out = [(a[0], k[0], k[1], k[2], a[1]) for k, a in agg.items()]
tot = sum(o[0] for o in out)
print("total warp-instructions: %.4g" % tot)
out.sort(reverse=True)
for v, f, ln, s, smp in out[:top]:
    print("%5.1f%% smp=%6s %-13s:%-4s %s" % (v / tot * 100, smp, f, ln, s.strip()[:120]))
