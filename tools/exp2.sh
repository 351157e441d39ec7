#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp2.txt
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/exp2_pytest.log 2>&1
SC=$PWD/libmjsim_v_sc.so
( MJS_LIB=$SC timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/exp2_pytest_sc.log 2>&1
run() { echo "## $1 sync=$2 $3 $4" >> $O; MJS_SYNC_LEVEL=$2 python tools/bench_variant.py $1 $3 $4 >> $O 2>&1; }
for S in 2 1 0; do run - $S f64 42624; done
for S in 2 0; do run $SC $S f64 42624; done
for S in 2 0; do run - $S f64 5328; run $SC $S f64 5328; done
for S in 2 0; do run - $S mixed 37888; run $SC $S mixed 37888; done
for S in 2 0; do run - $S mixed 7104; done
cat gpurun_out/exp2_pytest.log gpurun_out/exp2_pytest_sc.log $O
