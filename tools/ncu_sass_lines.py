"""Per source line / per engine stage view of an `ncu --page source --csv` SASS page, with the line of every SASS instruction taken
from `nvdisasm -g -c` of the same cubin (the two listings are matched instruction by instruction).
usage: cuobjdump -xelf all _build/arm_f64.o; nvdisasm -g -c arm_f64.sm_100a.cubin > all.sass
       python tools/ncu_sass_lines.py prof_f64_source.csv all.sass <mangled kernel name> <units> [top]"""
import collections, csv, re, sys
sys.path.insert(0, __file__.rsplit("/", 1)[0])
page_csv, sass, kname, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4]); top = int(sys.argv[5]) if len(sys.argv) > 5 else 30
rows = list(csv.reader(open(page_csv))); hd = rows[1]; ci = {n: i for i, n in enumerate(hd)}
page = [r for r in rows[2:] if len(r) >= len(hd)]
cur = None; dis = []; on = False
for l in open(sass):
    if l.startswith("//--------------------- .text."): on = (".text." + kname + " ") in l or l.rstrip().endswith(kname + " --------------------------")
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: dis.append((m.group(2).strip(), cur))
assert len(dis) == len(page), (len(dis), len(page))
from ncu_lines2_stages import STAGES
per = collections.defaultdict(lambda: [0, 0, collections.Counter()])
reasons = [n for n in hd if n.startswith("stall_") and "Not Issued" not in n]
for k, r in enumerate(page):
    f, ln = dis[k][1] if dis[k][1] else ("?", 0)
    p = per[(f, ln)]; p[0] += int(r[ci["Instructions Executed"]] or 0); p[1] += int(r[ci["# Samples"]] or 0)
    for n in reasons: p[2][n[6:]] += int(r[ci[n]] or 0)
tot = sum(v[0] for v in per.values()); ts = sum(v[1] for v in per.values())
st = collections.OrderedDict((n, [0, 0, collections.Counter()]) for _, _, n in STAGES); st["other files (cuda headers, mjs_common, mjs_convex)"] = [0, 0, collections.Counter()]
for (f, ln), (ex, sm, c) in per.items():
    key = "other files (cuda headers, mjs_common, mjs_convex)"
    if f == "arm_kernel.cuh":
        for lo, hi, n in STAGES:
            if lo <= ln <= hi: key = n; break
    st[key][0] += ex; st[key][1] += sm; st[key][2].update(c)
print(f"warp instructions {tot/units:.0f} per unit, samples {ts}")
for n, (ex, sm, c) in st.items():
    print(f"  {n:48s} inst/unit {ex/units:7.0f} ({100*ex/tot:5.1f}%)  samples {100*sm/ts:5.1f}%  cycles/inst x{(sm/ts)/(ex/tot) if ex else 0:4.2f}  " + ", ".join(f"{a} {100*b/max(sm,1):.0f}%" for a, b in c.most_common(3)))
print("most sampled source lines:")
src = {}
try:
    for i, l in enumerate(open("mujoco_sim_b200/csrc/arm_kernel.cuh"), 1): src[("arm_kernel.cuh", i)] = l.strip()[:100]
except OSError: pass
for (f, ln), (ex, sm, c) in sorted(per.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"  {f}:{ln:<5d} inst/unit {ex/units:6.1f} samples {100*sm/ts:5.2f}%  {c.most_common(1)[0][0]:12s} {src.get((f, ln), '')}")
