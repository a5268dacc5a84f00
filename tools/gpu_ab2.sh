#!/bin/bash
# GPU experiment 2 (2 GPUs): parity tests of the new wavefront loop / multi-GPU path, then stage times.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -12 > gpurun_out/pytest_gpu_ab2.log
tail -6 gpurun_out/pytest_gpu_ab2.log
for v in A N N9 NP; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:64 cfg1_1m:64 2>&1 | tail -2
done
for v in A N; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:512 2>&1 | tail -1
done
