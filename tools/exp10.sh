#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp10.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) >> $O 2>&1
run() { echo "## $1 $2 $3" >> $O; timeout 300 python tools/bench_variant.py $1 $2 $3 >> $O 2>&1; }
V=$PWD/libmjsim_v_hessv1.so
for L in $V - $V -; do run $L f64 42624; done
for L in $V -; do run $L mixed 37888; done
cat $O
