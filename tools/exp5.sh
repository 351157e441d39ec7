#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp5.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
( MJS_SYNC_LEVEL=49 timeout 600 python -m pytest tests/test_arm_gpu.py tests/test_golden_gpu.py -m gpu -x -q 2>&1 | tail -3 ) >> $O 2>&1
run() { echo "## $1 sync=$2 $3 $4" >> $O; MJS_SYNC_LEVEL=$2 timeout 300 python tools/bench_variant.py $1 $3 $4 >> $O 2>&1; }
for S in 1 17 49 65 145; do run - $S f64 42624; done
for G in 2 3 4 12; do run $PWD/libmjsim_v_g$G.so 1 f64 42624; done
run $PWD/libmjsim_v_g3.so 49 f64 42624
run $PWD/libmjsim_v_g12.so 49 f64 42624
for S in 10 26 58; do run - $S mixed 37888; done
run $PWD/libmjsim_v_f8.so 10 mixed 37888
for S in 1 49; do run - $S f64 5328; done
cat $O
