#!/bin/bash
TAG=${1:-t}
mkdir -p gpurun_out
timeout 2400 python -X faulthandler -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest_gpu_$TAG.full.log 2>&1
grep -v site-packages gpurun_out/pytest_gpu_$TAG.full.log | tail -15 | tee gpurun_out/pytest_gpu_$TAG.log
