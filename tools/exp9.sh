#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp9.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
run() { echo "## sync=$1 $2 $3" >> $O; MJS_SYNC_LEVEL=$1 timeout 300 python tools/bench_variant.py - $2 $3 >> $O 2>&1; }
for S in 0 1 2 10; do run $S f64 42624; done
for S in 2 10 1; do run $S mixed 37888; done
for S in 0 1 2; do run $S f64 5328; done
cat $O
