#!/usr/bin/env python
"""Summarises an .ncu-rep (run here, no GPU needed): key raw metrics of the first profiled kernel, and the
source lines with the most stall samples.  python tools/ncu_summary.py gpurun_out/x.ncu-rep [--source]"""
import csv, io, subprocess, sys, collections, re
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_bytes.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
d = dict(zip(hdr, zip(units, vals)))
print("kernel:", d.get("Kernel Name", ("", ""))[1][:100])
for w in want:
    if w in d:
        print("%-85s %-10s %s" % (w, d[w][0], d[w][1]))
for h in hdr:
    if "issue_stalled" in h and "per_issue_active.ratio" in h and "not_issued" not in h:
        v = float(d[h][1] or 0)
        if v > 0.05:
            print("  stall %-40s %.3f" % (h.split("issue_stalled_")[1].split("_per_issue")[0], v))
if "--source" in sys.argv:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    # find header
    for i, r in enumerate(rows):
        if "Source" in r and any("Samples" in c for c in r):
            h = r; body = rows[i + 1:]; break
    else:
        print("no source page"); sys.exit(0)
    si = h.index("Source"); col = [i for i, c in enumerate(h) if c.strip() == "# Samples" or c.strip() == "Warp Stall Sampling (All Samples)"]
    ci = col[0] if col else None
    ii = [i for i, c in enumerate(h) if "Instructions Executed" == c.strip()]
    agg = []
    for r in body:
        if len(r) <= max(si, ci or 0): continue
        try: s = float(r[ci])
        except: continue
        agg.append((s, r[si].strip()[:110], r[ii[0]] if ii else ""))
    tot = sum(a[0] for a in agg) or 1
    for s, t, n in sorted(agg, reverse=True)[:28]:
        print("%6.2f%%  inst=%-12s %s" % (100 * s / tot, n, t))
