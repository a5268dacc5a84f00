#!/bin/bash
# Quick GPU check: parity tests + the cfg1 bench line (no cfg2, no ncu).  usage: tools/gpu_quick.sh <tag>
TAG=${1:-q}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/pytest_gpu_$TAG.log
tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_cfg1_$TAG.json 2> gpurun_out/bench_cfg1_$TAG.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_cfg1_$TAG.json").read().strip().splitlines()[-1])
print("cfg1 value %.1fM e2e %.1fM stages %s launches %d"%(d["value"]/1e6, d["e2e"]["value"]/1e6, {k:round(v,1) for k,v in d["stage_ms_last_step"].items()}, d["gpu_launches"]))
print(d["e2e"]["wall_ms_per_step"])
PY
