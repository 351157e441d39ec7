#!/bin/bash
# round-2 multi-GPU lines: $1 = number of GPUs.  Planar Push (headline) and Button Push (BASELINE config 5), plus the reference arm under torchrun.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_r2_f64_${N}gpu.json 2> gpurun_out/bench_r2_f64_${N}gpu.err; echo "push rc=$?"
$TR bench.py --gpus $N --steps 50 --warmup 5 --workload push_button > gpurun_out/bench_r2_push_button_${N}gpu.json 2> gpurun_out/bench_r2_push_button_${N}gpu.err; echo "button rc=$?"
$TR bench.py --gpus $N --impl reference --steps 5 --warmup 3 > gpurun_out/bench_r2_reference_${N}gpu.json 2>/dev/null; echo "ref rc=$?"
python - <<PY
import json
for f in ("f64", "push_button", "reference"):
    try:
        d = json.loads(open(f"gpurun_out/bench_r2_{f}_${N}gpu.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], round(d["value"]), d.get("ms_per_step"), "e2e", round(d["e2e"]["value"]), d.get("stats_allreduce", {}).get("count_in_timed_region"))
    except Exception as e:
        print(f, "ERR", e)
PY
