#!/bin/bash
# Light ncu capture for profiles/r2/traffic.json (DRAM bytes of one stationary-mix launch of the bench command; a few passes),
# then -- time permitting -- the full-set capture of tools/ncu_bench.sh.  Every ncu call runs under its own timeout.
P=${1:-f64}
CMD="python bench.py --steps 4 --warmup 3 --precision $P --no-cpu-baseline --no-alt --no-small"
M=dram__bytes_read.sum,dram__bytes_write.sum,launch__grid_size,launch__registers_per_thread,gpu__time_duration.sum,sm__inst_executed.avg.per_cycle_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__icc_request_hit_rate.pct,smsp__inst_executed.sum
timeout 300 ncu --metrics $M --clock-control none -k regex:arm_step_kernel -s ${SKIP:-554} -c 1 -o /tmp/light_$P -f $CMD > gpurun_out/ncu_light_$P.log 2>&1; echo "light rc=$?"
ncu -i /tmp/light_$P.ncu-rep --page raw --csv > gpurun_out/prof_${P}_light_raw.csv 2>/dev/null; ls -la gpurun_out/prof_${P}_light_raw.csv
if [ "${FULL:-1}" = "1" ]; then
  timeout ${FULL_TIMEOUT:-420} ncu --set full --clock-control none --import-source on -k regex:arm_step_kernel -s ${SKIP:-554} -c 1 -o /tmp/prof_$P -f $CMD > gpurun_out/ncu_full_$P.log 2>&1; echo "full rc=$?"
  ncu -i /tmp/prof_$P.ncu-rep --page raw --csv > gpurun_out/prof_${P}_raw.csv 2>/dev/null
  ncu -i /tmp/prof_$P.ncu-rep --page source --csv > gpurun_out/prof_${P}_source.csv 2>/dev/null
  ls -la gpurun_out/prof_${P}_raw.csv gpurun_out/prof_${P}_source.csv
fi
