#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp11.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
run() { echo "## $1 $2 $3" >> $O; timeout 300 python tools/bench_variant.py $1 $2 $3 >> $O 2>&1; }
for L in - $PWD/libmjsim_v_f8.so $PWD/libmjsim_v_f4.so; do run $L mixed 37888; done
cat $O
