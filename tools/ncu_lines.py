"""Summarise an ncu report: headline metrics + instructions per source line."""
import csv, subprocess, sys
rep = sys.argv[1]; units = float(sys.argv[2]) if len(sys.argv) > 2 else 81920.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, vals = rows[0], rows[2]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__shared_mem_per_block_dynamic", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]
for h, v in zip(hdr, vals):
    if h in want or ("issue_stalled" in h and "per_issue_active" in h and float(v.replace(",", "") or 0) > 0.2):
        print(f"{h:90s} {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass,cuda", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
agg, hd = [], None
for r in rows:
    if len(r) > 8 and r[0] == "Line No": hd = r; continue
    if hd and len(r) > 8 and r[0] != "":
        try: agg.append((int(r[7]), int(r[6]), r[0], r[1][:100]))
        except ValueError: pass
tot = sum(a[0] for a in agg)
print("total inst", tot, "per unit", tot / units)
agg.sort(reverse=True)
for a in agg[: int(sys.argv[3]) if len(sys.argv) > 3 else 30]:
    print(f"{a[0]/units:8.1f} inst/unit  samples {a[1]:6d}  L{a[2]:>5}  {a[3]}")
