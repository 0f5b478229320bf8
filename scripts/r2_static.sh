#!/bin/bash
# the static-scratch entry point: alone in a fresh process, then after other backward tests in one process
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/zz_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/zz_alone.log
tail -4 gpurun_out/zz_alone.log
timeout 80 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q \
  -k "backward_sgd_vs_reference_step or static_scratch" > gpurun_out/zz_after.log 2>&1
echo "after rc=$?" >> gpurun_out/zz_after.log
tail -4 gpurun_out/zz_after.log
