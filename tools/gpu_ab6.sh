#!/bin/bash
mkdir -p gpurun_out
for v in H F16 F20 F28 A8 A16; do
  MTSB200_LIB=$PWD/_exp/$v.so timeout 300 python tools/trace_ab.py cfg2_10m:64 2>&1 | tail -1
done
for r in 22 24 28; do MTSB_REFILL=$r MTSB200_LIB=$PWD/_exp/H.so timeout 300 python tools/trace_ab.py cfg2_10m:64 2>&1 | tail -1 | sed "s/^/refill$r /"; done
