#!/bin/bash
# round-2 final single-GPU measurements: full GPU suite (with the numbers the tests print), bench lines of all workloads, reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/gputest_final.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/gputest_final.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2_f64.json 2> gpurun_out/bench_r2_f64.err; echo "bench rc=$?"
python bench.py --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/bench_r2_f64_100steps.json 2>/dev/null
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_r2_reference.json 2>/dev/null
for w in point_mass_reach robot_reach push_button; do python bench.py --workload $w --steps 50 --warmup 5 > gpurun_out/bench_r2_$w.json 2>/dev/null; done
python bench.py --precision mixed --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_r2_mixed.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_r2_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], round(d["value"]), d.get("ms_per_step"), "e2e", round(d["e2e"]["value"]), "frac", d.get("roofline", {}).get("frac"))
    except Exception as e:
        print(f, "ERR", e)
PY
