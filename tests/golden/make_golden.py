"""Generates tests/golden/*.npz from oracle/_ref/libmtsref.so -- the library compiled from the
reference tree itself (TriAccel, SAH kd-tree + Havran traversal, MicrofacetDistribution, warp,
erf/erfinv).  Run in the container that has /root/reference:

    make -C oracle all ref && python tests/golden/make_golden.py

The fixtures pin (a) the CPU restatement in oracle/ and (b) the CUDA path, on the GPU box where
the reference tree does not exist.  Scenes are regenerated from mitsuba_b200.scenes (seeded);
a checksum of the vertex data is stored to prove that the inputs are the same.
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from mitsuba_b200 import abi, scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def golden_scenes():
    return {
        "cbox": scenes.cbox(width=64, height=64, spp=4),
        "mesh": scenes.scene_1m(width=96, height=54, spp=4, nu=48, nv=24, n_obj=4),
    }


def scene_digest(b):
    pos, nrm, uv, idx = b.arrays()
    h = hashlib.sha256()
    h.update(pos.tobytes())
    h.update(idx.tobytes())
    return h.hexdigest()


def ray_sets(b, orc, n=6000, seed=7):
    """chord rays (test_kd.cpp:112-117 construction), camera rays, and secondary rays with
    mint == Epsilon (the adaptive-epsilon branch of skdtree.cpp:127-130)."""
    tight, _, _ = orc.bounds()
    c = (tight[:3] + tight[3:]) / 2
    r = float(np.linalg.norm(tight[3:] - c))
    chords = scenes.chord_rays(c, r, n, seed=seed)
    rng = np.random.default_rng(seed + 1)
    w, h = b.sensor.crop_width, b.sensor.crop_height
    cam = orc.sensor_rays(np.stack([rng.random(n) * w, rng.random(n) * h], -1).astype(np.float32))
    # secondary rays: start on the surfaces the camera rays hit, cosine-ish random directions
    hits = orc.trace(cam, abi.TRACE_CLOSEST)
    ok = hits["prim"] != abi.NO_HIT
    sec = np.zeros(int(ok.sum()), dtype=abi.RAY_DTYPE)
    sec["o"] = cam["o"][ok] + cam["d"][ok] * hits["t"][ok][:, None]
    d = rng.normal(size=(sec.shape[0], 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec["d"] = d.astype(np.float32)
    sec["mint"] = np.float32(1e-4)
    sec["maxt"] = np.inf
    return {"chord": chords, "camera": cam, "secondary": sec}


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def main():
    R = oracle.ref()
    if R is None:
        raise SystemExit("oracle/_ref/libmtsref.so missing: run `make -C oracle all ref` where /root/reference exists")
    R.ref_set_log_level(5)
    # ---- hit records of the reference kd-tree --------------------------------------------
    for name, b in golden_scenes().items():
        orc = oracle.Oracle(b)
        kd = oracle.RefKDTree(b)
        out = {"digest": np.frombuffer(scene_digest(b).encode(), dtype=np.uint8)}
        t, e = kd.bounds()
        out["tight"], out["enlarged"] = t, e
        for rname, rays in ray_sets(b, orc).items():
            out["rays_" + rname] = rays
            out["closest_" + rname] = kd.trace(rays, abi.TRACE_CLOSEST)
            out["any_" + rname] = (kd.trace(rays, abi.TRACE_ANY)["prim"] != abi.NO_HIT)
        np.savez_compressed(os.path.join(HERE, "hits_%s.npz" % name), **out)
        print(name, {k: v.shape for k, v in out.items()})
    # ---- TriAccel records -----------------------------------------------------------------------
    rng = np.random.default_rng(11)
    tri = (rng.normal(size=(512, 3, 3)) * rng.choice([0.01, 1.0, 100.0], size=(512, 1, 1))).astype(np.float32)
    tri[:8, 2] = tri[:8, 1]                      # degenerate: denom == 0 -> k = 3
    tri[8:16, :, 1] = 0.0                        # axis-aligned flats
    rec = np.zeros((512, 12), np.uint32)
    for i in range(512):
        R.ref_triaccel_load(vp(tri[i, 0].copy()), vp(tri[i, 1].copy()), vp(tri[i, 2].copy()), vp(rec[i]))
    np.savez_compressed(os.path.join(HERE, "triaccel.npz"), tri=tri, rec=rec)
    # ---- microfacet / warp / erf ------------------------------------------------------------------
    cases, outs = [], []
    for typ in (0, 1):
        for au, av in ((0.1, 0.1), (0.3, 0.3), (0.5, 0.3), (0.05, 0.05)):
            for vis in (0, 1):
                for _ in range(64):
                    wi = rng.normal(size=3)
                    wi[2] = abs(wi[2]) + 1e-3
                    wi /= np.linalg.norm(wi)
                    u = rng.random(2)
                    o = np.zeros(6, np.float32)
                    wi32, u32 = wi.astype(np.float32), u.astype(np.float32)
                    R.ref_microfacet(typ, au, av, vis, vp(wi32), vp(u32), vp(o))
                    cases.append([typ, au, av, vis, *wi32, *u32])
                    outs.append(o)
    np.savez_compressed(os.path.join(HERE, "microfacet.npz"), cases=np.array(cases, np.float32), out=np.array(outs, np.float32))
    u = rng.random((256, 2)).astype(np.float32)
    warp = np.zeros((4, 256, 3), np.float32)
    for k in range(4):
        for i in range(256):
            R.ref_warp(k, vp(u[i].copy()), vp(warp[k, i]))
    x = np.linspace(-0.999, 0.999, 401).astype(np.float32)
    xe = np.linspace(-4, 4, 401).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "scalar.npz"), u=u, warp=warp, x=x, erfinv=np.array([R.ref_erfinv(float(v)) for v in x], np.float32),
                        xe=xe, erf=np.array([R.ref_erf(float(v)) for v in xe], np.float32))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
