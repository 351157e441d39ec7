#!/bin/bash
# Round-end validation on one B200: GPU parity tests, smoke, the bench lines of all workloads, the reference (CPU) arm, ncu evidence.
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5 ) > gpurun_out/final_pytest.log 2>&1
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 ) > gpurun_out/final_smoke.log 2>&1
timeout 900 python bench.py > gpurun_out/final_bench_planar_push.json 2> gpurun_out/final_bench_planar_push.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err
for W in robot_reach push_button point_mass_reach; do
  timeout 600 python bench.py --workload $W --no-cpu-baseline --no-alt > gpurun_out/final_bench_$W.json 2> gpurun_out/final_bench_$W.err
done
timeout 300 python bench.py --envs 4096 --no-cpu-baseline --no-alt > gpurun_out/final_bench_planar_push_4096.json 2> gpurun_out/final_bench_4096.err
UNITS=852480 timeout 900 bash tools/ncu_bench.sh f64 > gpurun_out/final_ncu.log 2>&1
tail -3 gpurun_out/final_pytest.log gpurun_out/final_smoke.log; cat gpurun_out/final_bench_*.json | cut -c1-700; tail -5 gpurun_out/final_ncu.log
