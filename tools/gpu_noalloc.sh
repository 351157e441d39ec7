#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo alone rc=$?; tail -15 gpurun_out/noalloc_alone.log
timeout 800 python -m pytest tests -m gpu -x -q > gpurun_out/gputest_all.log 2>&1; echo all rc=$?; tail -5 gpurun_out/gputest_all.log
