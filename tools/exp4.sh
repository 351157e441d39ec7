#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp4.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
( MJS_SYNC_LEVEL=11 timeout 600 python -m pytest tests/test_arm_gpu.py tests/test_edge_cases_gpu.py tests/test_golden_gpu.py -m gpu -x -q 2>&1 | tail -3 ) >> $O 2>&1
run() { echo "## sync=$1 $2 $3" >> $O; MJS_SYNC_LEVEL=$1 timeout 300 python tools/bench_variant.py - $2 $3 >> $O 2>&1; }
for S in 1 2 10 11; do run $S f64 42624; done
for S in 2 10 11; do run $S mixed 37888; done
for S in 1 2 10 11; do run $S f64 5328; done
for S in 2 10 11; do run $S mixed 7104; done
cat $O
