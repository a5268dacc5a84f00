#!/bin/bash
# Final validation on ONE GPU, as the driver runs it: -m gpu tests, smoke(), default bench (+cpu_baseline), reference arm, other configs, ncu for roofline.traffic
TAG=${1:-fin}
mkdir -p gpurun_out
timeout 2400 python -X faulthandler -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu_$TAG.full.log 2>&1
grep -v site-packages gpurun_out/pytest_gpu_$TAG.full.log | tail -5 | tee gpurun_out/pytest_gpu_$TAG.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke_$TAG.log
timeout 1200 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_default_$TAG.json 2> gpurun_out/bench_default_$TAG.err
timeout 1200 python bench.py --impl reference --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_reference_$TAG.json 2> gpurun_out/bench_reference_$TAG.err
timeout 600 python bench.py --workload cfg1_1m --steps 5 --warmup 3 > gpurun_out/bench_cfg1_$TAG.json 2> gpurun_out/bench_cfg1_$TAG.err
timeout 600 python bench.py --workload cfg0_cbox --steps 5 --warmup 3 > gpurun_out/bench_cfg0_$TAG.json 2> gpurun_out/bench_cfg0_$TAG.err
python - <<PY
import json
for f in ("default","reference","cfg1","cfg0"):
    try:
        d=json.loads([l for l in open("gpurun_out/bench_%s_$TAG.json"%f) if l.startswith("{")][-1])
        print(f, d["config"]["workload"], "value %.2fM e2e %.2fM ms/step %.1f"%(d["value"]/1e6, d["e2e"]["value"]/1e6, d["ms_per_step"]), "cpu", d.get("cpu_baseline",{}).get("value"), d.get("cpu_baseline",{}).get("cores"), "roof", d.get("roofline",{}).get("frac"))
    except Exception as e: print(f, "failed", e); print(open("gpurun_out/bench_%s_$TAG.err"%f).read()[-800:])
PY
P="python bench.py --workload cfg2_10m --steps 1 --warmup 0 --spp 4 --no-cpu-baseline"
$P > gpurun_out/plain_cfg2_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_trace" -c 6 -o gpurun_out/prof_cfg2_$TAG $P > gpurun_out/ncu_cfg2_$TAG.log 2>&1
tail -1 gpurun_out/ncu_cfg2_$TAG.log
