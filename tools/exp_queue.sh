#!/bin/bash
# A/B on one B200: static lock-step blocks vs the persistent work-queue grid, base library vs the new one.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/exp_queue.txt
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/exp_pytest_static.log 2>&1
( MJS_QUEUE=1 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/exp_pytest_queue.log 2>&1
B=$PWD/libmjsim_v_base.so
for N in 42624 4736; do
  MJS_QUEUE=0 python tools/bench_variant.py $B f64 $N >> gpurun_out/exp_queue.txt 2>&1
  echo "new static sync2:" >> gpurun_out/exp_queue.txt
  MJS_QUEUE=0 python tools/bench_variant.py - f64 $N >> gpurun_out/exp_queue.txt 2>&1
  echo "new static sync0:" >> gpurun_out/exp_queue.txt
  MJS_QUEUE=0 MJS_SYNC_LEVEL=0 python tools/bench_variant.py - f64 $N >> gpurun_out/exp_queue.txt 2>&1
  echo "new queue:" >> gpurun_out/exp_queue.txt
  MJS_QUEUE=1 python tools/bench_variant.py - f64 $N >> gpurun_out/exp_queue.txt 2>&1
done
echo "mixed:" >> gpurun_out/exp_queue.txt
MJS_QUEUE=0 python tools/bench_variant.py $B mixed 37888 >> gpurun_out/exp_queue.txt 2>&1
MJS_QUEUE=0 python tools/bench_variant.py - mixed 37888 >> gpurun_out/exp_queue.txt 2>&1
MJS_QUEUE=1 python tools/bench_variant.py - mixed 37888 >> gpurun_out/exp_queue.txt 2>&1
cat gpurun_out/exp_pytest_static.log gpurun_out/exp_pytest_queue.log gpurun_out/exp_queue.txt
