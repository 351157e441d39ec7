"""Aggregate `ncu --page source --print-source cuda,sass --csv` per CUDA source line and per engine stage.
usage: ncu -i rep --page source --print-source cuda,sass --csv | python tools/ncu_lines2.py <units> [top]"""
import csv, sys, collections
units = float(sys.argv[1]); top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(sys.stdin))
STAGES = [(167, 237, "kinematics"), (243, 381, "dynamics(CRB/RNE/act)"), (386, 599, "collide"), (609, 736, "make_rows"), (739, 770, "contact_eval"),
          (772, 788, "contact_apply"), (790, 800, "jt_force"), (805, 905, "block_direction"), (908, 939, "dense_direction"), (942, 1096, "solve(body)"),
          (1098, 1117, "touch"), (1121, 1190, "integrate"), (1225, 1330, "ur_ik"), (1370, 1700, "task/io")]
hd = None; per_line = collections.defaultdict(lambda: [0, 0, ""]); cur_file = ""
for r in rows:
    if "Instructions Executed" in r and "Source" in r:
        hd = r; ie = r.index("Instructions Executed"); isamp = r.index("# Samples"); isrc = r.index("Source"); il = 0
        continue
    if hd is None or len(r) <= max(ie, isamp): continue
    try: line = int(r[il]); ex = int(r[ie] or 0); sm = int(r[isamp] or 0)
    except ValueError: continue
    p = per_line[line]; p[0] += ex; p[1] += sm; p[2] = r[isrc][:110]
tot = sum(v[0] for v in per_line.values()); ts = sum(v[1] for v in per_line.values())
print(f"total warp-instructions {tot}  per unit {tot/units:.0f}  samples {ts}")
st = collections.OrderedDict((n, [0, 0]) for _, _, n in STAGES); st["other(helpers/headers)"] = [0, 0]
for line, (ex, sm, _) in per_line.items():
    for lo, hi, n in STAGES:
        if lo <= line <= hi: st[n][0] += ex; st[n][1] += sm; break
    else: st["other(helpers/headers)"][0] += ex; st["other(helpers/headers)"][1] += sm
print("NOTE: line numbers of inlined helper headers (mjs_common.cuh, cuda headers) alias into the ranges; treat as approximate")
for n, (ex, sm) in st.items(): print(f"  {n:26s} inst/unit {ex/units:8.0f} ({100*ex/max(tot,1):5.1f}%)  samples {100*sm/max(ts,1):5.1f}%")
for line, (ex, sm, src) in sorted(per_line.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"  L{line:5d} inst/unit {ex/units:7.1f} samples {100*sm/max(ts,1):5.2f}%  {src}")
