#!/bin/bash
# strong scaling of ONE frame (bench.py default workload) over 1/2/4/8 GPUs of one box.  usage: tools/gpu_scale.sh <tag> [workload]
TAG=${1:-s}; WL=${2:-cfg2_10m}
mkdir -p gpurun_out
for N in 8 4 2 1; do
  if [ $N == 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps ${STEPS:-3} --warmup ${WARM:-3} --workload $WL --no-cpu-baseline > gpurun_out/scale_${WL}_n${N}_$TAG.json 2> gpurun_out/scale_${WL}_n${N}_$TAG.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps ${STEPS:-3} --warmup ${WARM:-3} --workload $WL --no-cpu-baseline > gpurun_out/scale_${WL}_n${N}_$TAG.json 2> gpurun_out/scale_${WL}_n${N}_$TAG.err
  fi
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/scale_${WL}_n${N}_$TAG.json") if l.startswith("{")][-1])
    print("N=$N", d["scaling"], "value %.1fM e2e %.1fM ms/step %.1f"%(d["value"]/1e6, d["e2e"]["value"]/1e6, d["ms_per_step"]), {k:round(v,1) for k,v in d["stage_ms_last_step"].items()}, d["config"]["spp_total"])
except Exception as e:
    print("N=$N failed", e); print(open("gpurun_out/scale_${WL}_n${N}_$TAG.err").read()[-1500:])
PY
done
