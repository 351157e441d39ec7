"""Per-opcode view of an `ncu --page source --csv` SASS page (tools/ncu_bench.sh -> prof_<P>_source.csv): dynamic warp
instructions and stall samples per opcode class and stall reason, and the instructions where long-scoreboard samples sit.
usage: python tools/ncu_sass_stalls.py gpurun_out/prof_f64_source.csv <units>"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1]))); units = float(sys.argv[2])
hd = rows[1]; ci = {n: i for i, n in enumerate(hd)}
reasons = [n for n in hd if n.startswith("stall_") and "Not Issued" not in n]
ops = collections.defaultdict(lambda: [0, 0, collections.Counter()]); tot_ex = tot_s = 0; rs = collections.Counter()
waits = collections.Counter()
for r in rows[2:]:
    if len(r) < len(hd): continue
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[ci["Source"]])
    if not m: continue
    op = m.group(2); ex = int(r[ci["Instructions Executed"]] or 0); s = int(r[ci["# Samples"]] or 0)
    o = ops[op]; o[0] += ex; o[1] += s; tot_ex += ex; tot_s += s
    for n in reasons:
        v = int(r[ci[n]] or 0); o[2][n] += v; rs[n] += v
print(f"warp instructions {tot_ex} = {tot_ex/units:.0f} per unit; samples {tot_s}")
print("stall reasons:", ", ".join(f"{n[6:]} {100*v/tot_s:.1f}%" for n, v in rs.most_common(10)))
print(f"{'opcode':12s} {'inst/unit':>9s} {'inst%':>6s} {'samples%':>8s}  top stall reasons at this opcode")
for op, (ex, s, c) in sorted(ops.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f"{op:12s} {ex/units:9.1f} {100*ex/tot_ex:6.1f} {100*s/tot_s:8.1f}  " + ", ".join(f"{n[6:]} {100*v/max(s,1):.0f}%" for n, v in c.most_common(3)))
