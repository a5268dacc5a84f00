#!/bin/bash
# Runs the small sanitize case and the GPU parity tests under the CHECKED build (_exp/CHK.so = -DMTSB_CHECKED: device-side bounds
# checks on queue writes, node / triangle fetches, the warp queue and film writes; see trace.cuh).  compute-sanitizer is closed on this pool.
mkdir -p gpurun_out
export MTSB200_LIB=$PWD/_exp/CHK.so
python tools/sanitize_case.py 2>&1 | tail -4 | tee gpurun_out/checked_case.log
timeout 2400 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_multi.py tests/test_gpu_records.py tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4 | tee gpurun_out/checked_pytest.log
