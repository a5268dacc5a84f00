#!/bin/bash
mkdir -p gpurun_out
for v in H V H V; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:64 cfg1_1m:64 2>&1 | tail -2
done
MTSB200_LIB=$PWD/_exp/V.so timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_trace_big.py tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -2
