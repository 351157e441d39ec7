"""Aggregate `ncu --page source --print-source cuda,sass --csv` per CUDA source line and per engine stage.
usage: ncu -i rep --page source --print-source cuda,sass --csv | python tools/ncu_lines2.py <units> [top]"""
import csv, sys, collections
sys.path.insert(0, __file__.rsplit("/", 1)[0])
units = float(sys.argv[1]); top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(sys.stdin))
# line ranges of csrc/arm_kernel.cuh at the commit the profile was taken from (update when the file moves; LINE_SHIFT maps an older profile)
from ncu_lines2_stages import STAGES
import os
SHIFT = [tuple(map(int, x.split(":"))) for x in os.environ.get("LINE_SHIFT", "").split(",") if x]  # "from:delta,..." for a profile of an older source
def shifted(line):
    d = 0
    for lo, dd in SHIFT:
        if line >= lo: d = dd
    return line + d
hd = None; per_line = collections.defaultdict(lambda: [0, 0, "", 0]); cur_file = ""
for r in rows:
    if "Instructions Executed" in r and "Source" in r:
        hd = r; ie = r.index("Instructions Executed"); isamp = r.index("# Samples"); isrc = r.index("Source"); il = 0
        continue
    if hd is None or len(r) <= max(ie, isamp): continue
    try: line = shifted(int(r[il])); ex = int(r[ie] or 0); sm = int(r[isamp] or 0)
    except ValueError: continue
    p = per_line[line]; p[0] += ex; p[1] += sm; p[2] = r[isrc][:110]; p[3] += 1 if ex >= 0.9 * units else 0
tot = sum(v[0] for v in per_line.values()); ts = sum(v[1] for v in per_line.values())
print(f"total warp-instructions {tot}  per unit {tot/units:.0f}  samples {ts}")
st = collections.OrderedDict((n, [0, 0, 0]) for _, _, n in STAGES); st["other(helpers/headers)"] = [0, 0, 0]
for line, (ex, sm, _, hot) in per_line.items():
    for lo, hi, n in STAGES:
        if lo <= line <= hi: st[n][0] += ex; st[n][1] += sm; st[n][2] += hot; break
    else: st["other(helpers/headers)"][0] += ex; st["other(helpers/headers)"][1] += sm; st["other(helpers/headers)"][2] += hot
print("NOTE: line numbers of inlined helper headers (mjs_common.cuh, cuda headers) alias into the ranges; treat as approximate")
for n, (ex, sm, hot) in st.items(): print(f"  {n:26s} inst/unit {ex/units:8.0f} ({100*ex/max(tot,1):5.1f}%)  samples {100*sm/max(ts,1):5.1f}%  hot code {hot*16/1024:6.1f} KB")
print("largest hot-code lines (SASS instructions executed >= 0.9 x per env-substep):")
for line, (ex, sm, src, hot) in sorted(per_line.items(), key=lambda kv: -kv[1][3])[:top]:
    print(f"  L{line:5d} hot {hot*16/1024:6.2f} KB inst/unit {ex/units:7.1f}  {src}")
print("most sampled lines:")
for line, (ex, sm, src, hot) in sorted(per_line.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"  L{line:5d} inst/unit {ex/units:7.1f} samples {100*sm/max(ts,1):5.2f}%  {src}")
