#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3 | tee gpurun_out/pytest_gpu_last.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -2 | tee gpurun_out/pytest_zz_alone_last.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
