#!/bin/bash
# ncu capture of the traversal kernels on the 10M-triangle workload (cfg2) at 4 spp.  usage: tools/gpu_ncu_cfg2.sh <tag>
TAG=${1:-c2}
mkdir -p gpurun_out
P="python bench.py --workload cfg2_10m --steps 1 --warmup 0 --spp 4 --no-cpu-baseline"
$P > gpurun_out/plain_cfg2_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_trace" -c 6 -o gpurun_out/prof_cfg2_$TAG $P > gpurun_out/ncu_cfg2_$TAG.log 2>&1
tail -2 gpurun_out/ncu_cfg2_$TAG.log
