#!/bin/bash
# GPU experiment 3 (2 GPUs): parity tests with the cooperative-triangle kernel, then stage times N (lock-step per-lane triangles) vs T.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider --deselect tests/test_gpu_render_full.py 2>&1 | tail -15 > gpurun_out/pytest_gpu_ab3.log
tail -6 gpurun_out/pytest_gpu_ab3.log
for v in N T T24 TS4; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:64 cfg1_1m:64 2>&1 | tail -2
done
for r in 22 30; do
  MTSB_REFILL=$r MTSB200_LIB=$PWD/_exp/T.so timeout 300 python tools/trace_ab.py cfg2_10m:64 2>&1 | tail -1 | sed "s/^/refill$r /"
done
