#!/bin/bash
# A/B of library variants on one box: bash tools/gpu_ab_libs.sh libA.so libB.so ...  (files under mujoco_sim_b200/; the shipped library runs first and last)
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-alt --no-small"
for lib in libmjsim_b200.so "$@" libmjsim_b200.so; do
  MJS_LIB=mujoco_sim_b200/$lib $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib', round(d['value']), round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms'],3))"
done
