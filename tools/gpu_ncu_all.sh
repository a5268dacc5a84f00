#!/bin/bash
# ncu counters for EVERY kernel of the render path on cfg2 (10M triangles) at 4 spp: duration, DRAM / L2 bytes, warp occupancy,
# issue utilisation, threads per instruction, branch efficiency.  Raw CSV -> gpurun_out/ncu_all_<tag>.csv   usage: tools/gpu_ncu_all.sh <tag>
TAG=${1:-all}
mkdir -p gpurun_out
P="python bench.py --workload cfg2_10m --steps 1 --warmup 0 --spp 4 --no-cpu-baseline"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,smsp__sass_average_branch_targets_threads_uniform.pct,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,l1tex__throughput.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__grid_size,launch__block_size
$P > gpurun_out/plain_all_$TAG.log 2>&1 && ncu --metrics $M --clock-control none -c 260 --csv --log-file gpurun_out/ncu_all_$TAG.csv $P > gpurun_out/ncu_all_$TAG.log 2>&1
tail -2 gpurun_out/ncu_all_$TAG.log; wc -l gpurun_out/ncu_all_$TAG.csv
