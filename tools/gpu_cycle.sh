#!/bin/bash
# One measurement cycle on the GPU box: parity tests, bench lines, ncu launch list + full capture of the traversal kernel.
# usage: tools/gpu_cycle.sh <tag> [ncu]
TAG=${1:-run}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/pytest_gpu_$TAG.log
tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_cfg1_$TAG.json 2> gpurun_out/bench_cfg1_$TAG.err
python - <<PY
import json
for w in ("cfg1",):
    try:
        d=json.loads(open("gpurun_out/bench_%s_$TAG.json"%w).read().strip().splitlines()[-1])
        print(w, "value %.1fM e2e %.1fM Mrays %.0f stages %s roof %.3f"%(d["value"]/1e6, d["e2e"]["value"]/1e6, d["mrays_per_s"], {k:round(v,1) for k,v in d["stage_ms_last_step"].items()}, d["roofline"]["frac"]))
    except Exception as e: print(w, "failed", e)
PY
timeout 400 python bench.py --no-cpu-baseline --workload cfg2_10m --steps 2 > gpurun_out/bench_cfg2_$TAG.json 2> gpurun_out/bench_cfg2_$TAG.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_cfg2_$TAG.json").read().strip().splitlines()[-1])
    print("cfg2", "value %.1fM e2e %.1fM Mrays %.0f stages %s roof %.3f"%(d["value"]/1e6, d["e2e"]["value"]/1e6, d["mrays_per_s"], {k:round(v,1) for k,v in d["stage_ms_last_step"].items()}, d["roofline"]["frac"]))
except Exception as e: print("cfg2 failed", e)
PY
if [ "$2" == "ncu" ]; then
  P="python bench.py --steps 1 --warmup 0 --spp 4 --no-cpu-baseline"
  $P > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $P > gpurun_out/ncu1_$TAG.log 2>&1
  $P > gpurun_out/plain2_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_trace|k_shade|k_splat|k_sort" -s 2 -c 12 -o gpurun_out/prof_$TAG $P > gpurun_out/ncu2_$TAG.log 2>&1
  tail -2 gpurun_out/ncu2_$TAG.log
fi
