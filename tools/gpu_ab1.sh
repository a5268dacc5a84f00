#!/bin/bash
# GPU experiment 1: parity tests on the working library, then stage times of the traversal variants.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -8 > gpurun_out/pytest_gpu_ab1.log
tail -3 gpurun_out/pytest_gpu_ab1.log
for v in A B C D; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:64 cfg1_1m:64 2>&1 | tail -2
done
for r in 20 30; do
  MTSB_REFILL=$r MTSB200_LIB=$PWD/_exp/C.so timeout 300 python tools/trace_ab.py cfg2_10m:64 2>&1 | tail -1 | sed "s/^/refill$r /"
done
