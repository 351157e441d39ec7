#!/bin/bash
# One full-set ncu capture of the step kernel of every other workload / precision (the Planar Push f64 kernel has its own: ncu_bench.sh).
mkdir -p gpurun_out
cap() {  # name, kernel regex, bench args
  local T=$1 K=$2; shift 2
  CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-alt --burn-in 20 $*"
  $CMD > gpurun_out/plain_$T.log 2>&1 || { echo "plain run failed: $T"; return; }
  ncu --set full --clock-control none -k regex:$K -s 6 -c 1 -o /tmp/o_$T -f $CMD > gpurun_out/ncu_$T.log 2>&1
  ncu -i /tmp/o_$T.ncu-rep --page raw --csv > gpurun_out/prof_${T}_raw.csv 2>/dev/null
}
cap pointmass_f64 pm_step_kernel --workload point_mass_reach
ncu -i /tmp/o_pointmass_f64.ncu-rep --page source --csv > gpurun_out/prof_pointmass_f64_source.csv 2>/dev/null
cap reach_f64 arm_step_kernel --workload robot_reach
cap button_f64 arm_step_kernel --workload push_button
cap push_mixed arm_step_kernel --workload planar_push --precision mixed
ls -la gpurun_out/prof_*_raw.csv
