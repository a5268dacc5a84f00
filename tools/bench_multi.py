#!/usr/bin/env python
"""One frame of bench.py's default workload through the IN-LIBRARY multi-GPU path (mtsb_multi_create / mtsb_multi_render:
one host thread per GPU, ncclReduce of the film; no torch.distributed, one process).  usage: bench_multi.py [--gpus 1,2,4,8]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mitsuba_b200 import lib, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", default="1,2,4,8")
ap.add_argument("--steps", type=int, default=3)
a = ap.parse_args()
b = scenes.scene_10m()
desc = b.desc()
have = lib.device_count()
for n in [int(x) for x in a.gpus.split(",")]:
    if n > have:
        continue
    t0 = time.time()
    ms = lib.MultiScene(desc, n)
    t_create = time.time() - t0
    ms.render(spp=512, seed=1)                      # warm-up (path buffers, NCCL channels)
    best = None
    for k in range(a.steps):
        t0 = time.time()
        st = ms.render(spp=512, seed=10 + k)
        img = ms.film_develop()
        wall = time.time() - t0
        if best is None or st.ms_total < best[0].ms_total:
            best = (st, wall)
    st, wall = best
    print(json.dumps({"path": "mtsb_multi_render (in-library, one process)", "n_gpus": n, "samples": int(st.samples), "ms_total": st.ms_total,
                      "ms_reduce": st.ms_reduce, "samples_per_s": st.samples / st.ms_total * 1e3, "wall_ms_render_plus_develop": wall * 1e3,
                      "create_s_all_gpus": t_create, "launches": int(st.kernel_launches), "mean": float(img.mean())}), flush=True)
    ms.close()
