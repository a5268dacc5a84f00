#!/bin/bash
# full GPU suite + bench + cfg3 ray microbenchmark   usage: tools/gpu_full2.sh <tag>
TAG=${1:-f2}
mkdir -p gpurun_out
timeout 2400 python -X faulthandler -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest_gpu_$TAG.full.log 2>&1
grep -v site-packages gpurun_out/pytest_gpu_$TAG.full.log | tail -15 | tee gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_cfg2_$TAG.json 2> gpurun_out/bench_cfg2_$TAG.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_cfg2_$TAG.json").read().strip().splitlines()[-1])
    print("cfg2 value %.1fM e2e %.1fM Mrays %.0f stages %s roof %.3f cpu %s"%(d["value"]/1e6, d["e2e"]["value"]/1e6, d["mrays_per_s"], {k:round(v,1) for k,v in d["stage_ms_last_step"].items()}, d["roofline"]["frac"], d.get("cpu_baseline",{}).get("value")))
except Exception as e: print("cfg2 failed", e); print(open("gpurun_out/bench_cfg2_$TAG.err").read()[-2000:])
PY
timeout 600 python tools/bench_rays.py > gpurun_out/rays_cfg3_$TAG.json 2> gpurun_out/rays_cfg3_$TAG.err
python - <<PY
import json
for l in open("gpurun_out/rays_cfg3_$TAG.json"):
    try:
        d=json.loads(l); print(d["config"]["workload"], "closest %.0f any %.0f Mrays/s; e2e pageable %.0f pinned %.0f"%(d["closest"]["mrays_per_s"], d["any"]["mrays_per_s"], d["e2e"]["mrays_per_s"], d["e2e_pinned"]["mrays_per_s"]))
    except Exception as e: print("rays line failed", e)
PY
tail -3 gpurun_out/rays_cfg3_$TAG.err
