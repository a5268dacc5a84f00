#!/bin/bash
TAG=${1:-tb}
mkdir -p gpurun_out
timeout 2400 python -X faulthandler -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest_gpu_$TAG.full.log 2>&1
grep -v site-packages gpurun_out/pytest_gpu_$TAG.full.log | tail -8 | tee gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg2_$TAG.json 2> gpurun_out/bench_cfg2_$TAG.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg2_n2_$TAG.json 2> gpurun_out/bench_cfg2_n2_$TAG.err
python - <<PY
import json
for f in ("gpurun_out/bench_cfg2_$TAG.json", "gpurun_out/bench_cfg2_n2_$TAG.json"):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(d["n_gpus"], "value %.1fM e2e %.1fM"%(d["value"]/1e6, d["e2e"]["value"]/1e6), d["e2e"]["wall_ms_per_step"], d["detail"])
    except Exception as e: print(f, "failed", e)
PY
