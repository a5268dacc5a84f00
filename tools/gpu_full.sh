#!/bin/bash
# Full GPU check: every -m gpu test (the no-alloc test is collected last), that test alone in a fresh process, a short bench line,
# then an ncu --set full capture (with source counters) of the traversal kernels on cfg2 at 4 spp.   usage: tools/gpu_full.sh <tag> [ncu]
TAG=${1:-full}
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3 | tee gpurun_out/pytest_zz_alone_$TAG.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_cfg2_$TAG.json 2> gpurun_out/bench_cfg2_$TAG.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_cfg2_$TAG.json").read().strip().splitlines()[-1])
    print("cfg2 value %.1fM e2e %.1fM Mrays %.0f stages %s roof %.3f cpu %s"%(d["value"]/1e6, d["e2e"]["value"]/1e6, d["mrays_per_s"], {k:round(v,1) for k,v in d["stage_ms_last_step"].items()}, d["roofline"]["frac"], d.get("cpu_baseline",{}).get("value")))
except Exception as e: print("cfg2 failed", e); print(open("gpurun_out/bench_cfg2_$TAG.err").read()[-2000:])
PY
if [ "$2" == "ncu" ]; then
  P="python bench.py --workload cfg2_10m --steps 1 --warmup 0 --spp 4 --no-cpu-baseline"
  $P > gpurun_out/plain_cfg2_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_trace" -c 6 -o gpurun_out/prof_cfg2_$TAG $P > gpurun_out/ncu_cfg2_$TAG.log 2>&1
  tail -2 gpurun_out/ncu_cfg2_$TAG.log
fi
