#!/bin/bash
# round-2 first GPU call: baseline tests, FMA peaks, sanitizer logs, baseline bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
tools/_bin/fma_peak > gpurun_out/fma_peak.json 2> gpurun_out/fma_peak.err
cat gpurun_out/fma_peak.json
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gputest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/gputest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_base.json 2> gpurun_out/bench_base.err; tail -1 gpurun_out/bench_base.json | cut -c1-400
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/sanitizer_$tool.log python tools/sanitize_rollout.py 96 5 > gpurun_out/sanitizer_$tool.out 2>&1
  echo "$tool rc=$?"; tail -2 gpurun_out/sanitizer_$tool.log
done
