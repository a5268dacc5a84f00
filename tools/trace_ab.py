#!/usr/bin/env python
"""Stage times of mtsb_render for the library named by MTSB200_LIB (kernel A/B experiments).
usage: MTSB200_LIB=_exp/X.so python tools/trace_ab.py [cfg2_10m:64] [cfg1_1m:64] ..."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mitsuba_b200 import lib, scenes

lib.init(0)
from mitsuba_b200 import abi
abi.ABI_VERSION = lib.lib().mtsb_abi_version()      # older variant libraries check the description against their own version
jobs = sys.argv[1:] or ["cfg2_10m:64", "cfg1_1m:64"]
tag = os.path.basename(os.environ.get("MTSB200_LIB", "default"))
for j in jobs:
    name, spp = j.split(":")
    spp = int(spp)
    b = scenes.scene_10m(spp=spp) if name == "cfg2_10m" else scenes.scene_1m(spp=spp)
    sc = lib.Scene(b)
    sc.render(spp=spp, seed=1)
    best = None
    for k in range(2):
        st = sc.render(spp=spp, seed=2 + k)
        if best is None or st.ms_total < best.ms_total:
            best = st
    st = best
    print("%-12s %-10s spp %3d: total %8.1f ms  closest %7.1f any %7.1f shade %6.1f splat %5.1f raygen %5.1f | %.0f Msamples/s, %d launches, rays c %.1fM s %.1fM"
          % (tag, name, spp, st.ms_total, st.ms_trace_closest, st.ms_trace_any, st.ms_shade, st.ms_splat, st.ms_raygen,
             st.samples / st.ms_total / 1e3, st.kernel_launches, st.closest_rays / 1e6, st.shadow_rays / 1e6), flush=True)
    sc.close()
