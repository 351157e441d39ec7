#!/bin/bash
# One full-set ncu capture of arm_step_kernel (bench command, $1 = precision, $2 = envs, $3 = MJS_SYNC_LEVEL); brings back the raw
# page and the gzip'd SASS-level source page (per-instruction executed counts and stall samples) for local analysis.
P=${1:-f64}; N=${2:-5328}; S=${3:-2}
export MJS_SYNC_LEVEL=$S
CMD="python bench.py --steps 4 --warmup 3 --precision $P --envs $N --no-cpu-baseline --no-alt"
T=s${S}_$P
$CMD > gpurun_out/plain_$T.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:arm_step_kernel -s 4 -c 1 -o /tmp/prof_$T -f $CMD > gpurun_out/ncu_full_$T.log 2>&1
ncu -i /tmp/prof_$T.ncu-rep --page raw --csv > gpurun_out/prof_${T}_raw.csv 2>/dev/null
ncu -i /tmp/prof_$T.ncu-rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/prof_${T}_sass.csv.gz
ncu -i /tmp/prof_$T.ncu-rep --page source --print-source cuda,sass --csv 2>/dev/null | gzip -9 > gpurun_out/prof_${T}_cuda.csv.gz
ls -la /tmp/prof_$T.ncu-rep gpurun_out/prof_${T}_*
