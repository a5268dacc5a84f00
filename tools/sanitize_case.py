#!/usr/bin/env python
"""The small case compute-sanitizer runs (memcheck / racecheck): scene creation + BVH build, host-buffer ray queries (chunked
path, no-alloc path), one wavefront render of the Cornell box with every stage (raygen, trace, sort, shade, shadow rays,
splat, develop), a render with textures and an environment map.  usage: compute-sanitizer --tool X python tools/sanitize_case.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from mitsuba_b200 import abi, lib, scenes

lib.init(0)
b = scenes.cbox(width=48, height=48, spp=4)
sc = lib.Scene(b)
rays = scenes.chord_rays([278, 273, 280], 600.0, 5000, seed=3)
h = sc.trace(rays); a = sc.trace(rays, abi.TRACE_ANY); hs = sc.trace_static(rays)
assert np.array_equal(h["t"], hs["t"])
st = sc.render(spp=4, seed=1)
img = sc.film_develop()
print("cbox: %d samples, %d launches, mean %.4f, hits %.3f" % (st.samples, st.kernel_launches, float(img.mean()), float((h["prim"] != abi.NO_HIT).mean())))
sc.close()
b = scenes.cbox_textured(width=40, height=30, spp=2)
sc = lib.Scene(b); st = sc.render(spp=2, seed=2); print("textured: mean %.4f" % float(sc.film_develop().mean())); sc.close()
b = scenes.scene_10m(nu=16, nv=8, n_obj=6, width=40, height=24, spp=2)
b.emitters = [e for e in b.emitters if e.type != abi.EMITTER_CONSTANT]
b.add_envmap(scenes.synthetic_sky(256, 128), scale=0.6)
sc = lib.Scene(b); st = sc.render(spp=2, seed=3); print("envmap + mixed BSDFs: mean %.4f" % float(sc.film_develop().mean())); sc.close()
