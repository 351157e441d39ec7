#!/bin/bash
# ncu evidence for the bench command (run under gpurun).  $1 = precision.  SKIP = step-kernel launches to skip before the
# full-set capture (default 554: past the 550-step burn-in, i.e. a launch of the stationary state mix; SKIP=4 = fresh scenes).  Writes CSV pages (the
# .ncu-rep is too large to travel back: gpurun_out is capped at 64 MiB).
P=${1:-f64}
CMD="python bench.py --steps 4 --warmup 3 --precision $P --no-cpu-baseline --no-alt --no-small $BENCH_EXTRA"
$CMD > gpurun_out/plain_$P.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_$P.csv $CMD > gpurun_out/ncu_list_$P.log 2>&1
$CMD > gpurun_out/plain2_$P.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:arm_step_kernel -s ${SKIP:-554} -c 1 -o /tmp/prof_$P -f $CMD > gpurun_out/ncu_full_$P.log 2>&1
ncu -i /tmp/prof_$P.ncu-rep --page raw --csv > gpurun_out/prof_${P}_raw.csv 2>/dev/null
ncu -i /tmp/prof_$P.ncu-rep --page source --csv > gpurun_out/prof_${P}_source.csv 2>/dev/null
ncu -i /tmp/prof_$P.ncu-rep --page source --print-source cuda,sass --csv 2>/dev/null > /tmp/prof_${P}_cuda.csv; head -c 3000 /tmp/prof_${P}_cuda.csv > gpurun_out/prof_${P}_cuda_head.csv; python tools/ncu_lines2.py ${UNITS:-852480} 45 < /tmp/prof_${P}_cuda.csv > gpurun_out/prof_${P}_lines.txt 2>&1
ls -la /tmp/prof_$P.ncu-rep gpurun_out/prof_${P}_*.csv
