"""One-paragraph summary of an `ncu --page raw --csv` capture: duration, IPC, issue utilisation, occupancy, active lanes, cache hit
rates, DRAM bytes and the warp-stall breakdown.  usage: python tools/ncu_summary.py <name> <raw.csv> [...]"""
import csv, sys
def summary(name, path):
    rows = list(csv.reader(open(path))); h, u, v = rows[0], rows[1], rows[2]
    g = {n: (v[i], u[i]) for i, n in enumerate(h)}
    f = lambda n: float(g[n][0].replace(",", "")) if n in g and g[n][0] not in ("", "n/a") else float("nan")
    unit = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1}
    b = lambda n: f(n) * unit.get(g[n][1], 1) if n in g else float("nan")
    tms = f("gpu__time_duration.sum") * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(g["gpu__time_duration.sum"][1], 1)
    print(f"{name}: {g['Kernel Name'][0] if 'Kernel Name' in g else ''} grid {g.get('Grid Size', ('?',))[0]} block {g.get('Block Size', ('?',))[0]}")
    print(f"  duration {tms:.3f} ms, registers/thread {g.get('launch__registers_per_thread', ('?',))[0]}, warps active {f('sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} % of 64/SM, "
          f"IPC {f('sm__inst_executed.avg.per_cycle_active'):.2f}/SM, issue slots busy {f('smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f} %, "
          f"active lanes {f('smsp__thread_inst_executed_per_inst_executed.ratio'):.1f}/32")
    print(f"  icc hit {f('sm__icc_request_hit_rate.pct'):.0f} %, L1 global-load hit {f('l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct'):.0f} %, local-load requests "
          f"{f('l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum'):.3g} (hit {f('l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct'):.0f} %), DRAM read {b('dram__bytes_read.sum')/1e6:.1f} MB + write {b('dram__bytes_write.sum')/1e6:.1f} MB")
    st = sorted(((f(n), n.split("issue_stalled_")[1].split("_per_")[0]) for n in h if n.startswith("smsp__average_warps_issue_stalled") and n.endswith("_per_issue_active.ratio")), reverse=True)
    tot = sum(x for x, _ in st)
    print("  warp stalls: " + ", ".join(f"{n} {100 * x / tot:.0f} %" for x, n in st[:8]))
for k in range(1, len(sys.argv) - 1, 2):
    summary(sys.argv[k], sys.argv[k + 1])
