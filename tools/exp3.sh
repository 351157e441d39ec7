#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp3.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
REP=$PWD/libmjsim_v_rep.so
MJS_LIB=$REP MJS_SYNC_LEVEL=0 python tools/gpu_stage_repeat.py 19536 f64 >> $O 2>&1
MJS_LIB=$REP MJS_SYNC_LEVEL=2 python tools/gpu_stage_repeat.py 19536 f64 >> $O 2>&1
MJS_LIB=$REP MJS_SYNC_LEVEL=0 python tools/gpu_stage_repeat.py 1628 f64 >> $O 2>&1
run() { echo "## $1 sync=$2 $3 $4" >> $O; MJS_SYNC_LEVEL=$2 python tools/bench_variant.py $1 $3 $4 >> $O 2>&1; }
for L in $PWD/libmjsim_v_r1.so $PWD/libmjsim_v_scr1.so; do
  for S in 2 0; do run $L $S f64 42624; done
  run $L 2 mixed 37888
done
run - 1 f64 42624
cat $O
