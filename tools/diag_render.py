"""Diagnostic (GPU box): GPU film vs the oracle rendered with the SAME counter-based random numbers."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from mitsuba_b200 import lib, scenes, abi
from mitsuba_b200.scenedesc import look_at

lib.init(0)


def cmp(b, label, spp=1, seed=9):
    sc, orc = lib.Scene(b), oracle.Oracle(b)
    st = sc.render(spp=spp, seed=seed)
    fg = sc.film_download().astype(np.float64)
    ost = orc.render(spp=spp, seed=seed, counter_rng=True)
    fo = orc.film_download().astype(np.float64)
    d = np.abs(fg - fo)
    tol = 1e-3 * np.maximum(np.abs(fo), 1e-3)
    badpix = (d > tol)[..., :3].any(-1)
    print("%-28s spp=%d bad px %6d / %d (%.4f)  mean g=%.6f o=%.6f  | gpu closest %d shadow %d shade %d plen %d | cpu closest %d shadow %d shade %d plen %d"
          % (label, spp, badpix.sum(), badpix.size, badpix.mean(), fg[..., :3].mean(), fo[..., :3].mean(), st.closest_rays, st.shadow_rays,
             st.shade_events, st.path_length_sum, ost.closest_rays, ost.shadow_rays, ost.shade_events, ost.path_length_sum))
    ys, xs = np.nonzero(badpix)
    for y, x in list(zip(ys, xs))[:4]:
        print("      px", x, y, "gpu", fg[y, x], "cpu", fo[y, x])
    sc.close()
    return badpix.mean()


def boxed(b, w, h, cam):
    b.set_sensor(cam[0], cam[1], w, h, rfilter="box", near_clip=cam[2], far_clip=cam[3])
    return b


cb_cam = (look_at([278, 273, -800], [278, 273, -799], [0, 1, 0]), 39.3077, 10.0, 2800.0)
for md in (1, 2, 3, 8, -1):
    cmp(boxed(scenes.cbox(width=64, height=64, spp=1, max_depth=md), 64, 64, cb_cam), "cbox box maxDepth=%d" % md)
cmp(boxed(scenes.cbox(width=64, height=64, spp=4, max_depth=8), 64, 64, cb_cam), "cbox box spp4", spp=4)
cmp(scenes.cbox(width=64, height=64, spp=4, max_depth=8), "cbox gaussian spp4", spp=4)
cmp(scenes.cbox(width=64, height=64, spp=4, max_depth=8, twosided=True), "cbox twosided", spp=4)
for md in (2, 3, 10):
    cmp(scenes.scene_10m(width=96, height=54, spp=2, nu=24, nv=12, n_obj=12, max_depth=md), "mixed maxDepth=%d" % md, spp=2)
cmp(scenes.scene_1m(width=96, height=54, spp=2, nu=24, nv=12, n_obj=8), "1m-small", spp=2)
b = scenes.cbox(width=48, height=48, spp=2)
b.add_point_light([278, 400, 279], [4e5, 4e5, 4e5])
b.set_integrator(max_depth=5, rr_depth=3, hide_emitters=True, sample_count=2)
cmp(b, "point+hide", spp=2)
