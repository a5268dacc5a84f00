#!/bin/bash
# GPU experiment 4: trace kernel with hit records in shared memory (U) and shading-kernel occupancy variants (S5/S6/S8 = U + launch bounds)
mkdir -p gpurun_out
for v in T U S5 S6 S8; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:64 cfg1_1m:64 2>&1 | tail -2
done
MTSB200_LIB=$PWD/_exp/U.so timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
