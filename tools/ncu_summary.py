#!/usr/bin/env python3
"""Turn gpurun_out ncu artefacts into the tracked summaries under profiles/.
  python tools/ncu_summary.py <tag> [launches.csv] [prof.ncu-rep]
writes profiles/<tag>_launches.md (per-kernel share of the step from the gpu__time_duration pass) and
profiles/<tag>_kernels.md (per-kernel metrics of the --set full capture) + profiles/ncu_traffic.json."""
import collections, csv, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
launches = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "launches.csv")
rep = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "prof1.ncu-rep")
out_dir = os.path.join(ROOT, "profiles")

if os.path.exists(launches):
    rows = [r for r in csv.reader(open(launches, errors="replace")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    rows = rows[rows.index(hdr) + 1:]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows:
        agg.setdefault(r[ki].split("(")[0], []).append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    with open(os.path.join(out_dir, tag + "_launches.md"), "w") as f:
        f.write("# %s — ncu launch list (gpu__time_duration.sum, --clock-control none)\n\n" % tag)
        f.write("Cold-cache, serialised launches: compare SHARES, not absolutes.\n\n| kernel | launches | avg us | share of captured time |\n|---|---|---|---|\n")
        for n, v in agg.items():
            f.write("| %s | %d | %.1f | %.1f %% |\n" % (n, len(v), sum(v) / len(v) / 1e3, sum(v) / tot * 100))
    print(open(os.path.join(out_dir, tag + "_launches.md")).read())

if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "lts__t_sector_hit_rate.pct"]
    idx = [(w, hdr.index(w)) for w in want if w in hdr]
    ki = hdr.index("Kernel Name")
    per = collections.OrderedDict()
    for r in rows[2:]:
        per.setdefault(r[ki].split("(")[0], []).append(r)
    traffic = {}
    issue = {}
    with open(os.path.join(out_dir, tag + "_kernels.md"), "w") as f:
        f.write("# %s — ncu --set full --clock-control none (per launch, first captured launch of each kernel)\n\n" % tag)
        f.write("| kernel | " + " | ".join("%s [%s]" % (w, units[i]) for w, i in idx) + " |\n|---|" + "---|" * len(idx) + "\n")
        for n, rs in per.items():
            r = rs[0]
            f.write("| %s | " % n + " | ".join(r[i] for _, i in idx) + " |\n")
            try:   # the binding utilisations of an issue-bound kernel: issue slots and the busiest pipes (POPC issues on XU)
                short = n.replace("void ", "").split("<")[0].strip()
                issue[short] = {k2: round(float(r[hdr.index(m)]), 1) for k2, m in (
                    ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    ("pipe_alu_pct", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                    ("pipe_xu_pct", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
                    ("pipe_fma_pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                    ("pipe_lsu_pct", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                    ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")) if m in hdr}
                issue[short]["source"] = tag
            except Exception:
                pass
            try:
                rd, wr = float(r[hdr.index("dram__bytes_read.sum")]), float(r[hdr.index("dram__bytes_write.sum")])
                mul = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(units[hdr.index("dram__bytes_read.sum")], 1)
                mulw = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(units[hdr.index("dram__bytes_write.sum")], 1)
                traffic[n] = int(rd * mul + wr * mulw)
                if n in ("k_resize", "k_resize_tma") and len(rs) >= 7:   # the pyramid stage = the 7 per-level launches of one batch
                    traffic[n] = int(sum(float(q[hdr.index("dram__bytes_read.sum")]) * mul + float(q[hdr.index("dram__bytes_write.sum")]) * mulw for q in rs[:7]))
                    traffic[n + "_launches_summed"] = 7
            except Exception:
                pass
    traffic["frames_per_launch"] = int(os.environ.get("KFS_NCU_FRAMES", "128"))
    traffic["source"] = "profiles/%s_kernels.md (ncu --set full of that build, not of the bench run that quotes it)" % tag
    json.dump(traffic, open(os.path.join(out_dir, "ncu_traffic.json"), "w"), indent=1)
    json.dump(issue, open(os.path.join(out_dir, "ncu_issue.json"), "w"), indent=1)
    print(open(os.path.join(out_dir, tag + "_kernels.md")).read()[:3000])
