"""Generates tests/golden/texture.npz from oracle/_ref/libmtsref.so: the reference's own TMIPMap (include/mitsuba/render/
mipmap.h, the class behind `bitmap` textures, src/textures/bitmap.cpp:172-177) instantiated on small random images --
the Lanczos MIP pyramid (Bitmap::resample / Resampler), box and bilinear level-0 lookups and the filtered lookup
(trilinear / EWA incl. the anisotropy clamp) for every wrap mode.  Glue: oracle/ref_texture_glue.cpp.

    make -C oracle all ref && python tests/golden/make_golden_texture.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
# (width, height, wrapU, wrapV, filterType, maxAnisotropy): wrap 0 clamp 1 repeat 2 mirror 3 zero 4 one; filter 0 nearest 1 bilinear 2 trilinear 3 ewa
CASES = [(37, 21, 1, 2, 3, 20.0), (64, 64, 0, 1, 3, 4.0), (5, 130, 3, 4, 2, 20.0), (16, 9, 1, 1, 1, 20.0), (9, 9, 2, 0, 0, 20.0),
         (32, 16, 1, 1, 3, 20.0), (1, 7, 1, 1, 3, 20.0)]


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def main():
    oracle.build(ref=True)
    R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libmtsref.so"))
    R.ref_static_init()
    R.ref_mipmap_create.restype = C.c_void_p
    R.ref_mipmap_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float]
    for f in ("ref_mipmap_levels", "ref_mipmap_level_size", "ref_mipmap_level", "ref_mipmap_eval", "ref_mipmap_stats", "ref_mipmap_destroy"):
        getattr(R, f).argtypes = {"ref_mipmap_levels": [C.c_void_p], "ref_mipmap_level_size": [C.c_void_p, C.c_int, C.c_void_p],
                                  "ref_mipmap_level": [C.c_void_p, C.c_int, C.c_void_p],
                                  "ref_mipmap_eval": [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p],
                                  "ref_mipmap_stats": [C.c_void_p, C.c_void_p], "ref_mipmap_destroy": [C.c_void_p]}[f]
    rng = np.random.default_rng(20261108)
    out = {"cases": np.array(CASES, np.float32)}
    for ci, (w, h, wu, wv, filt, aniso) in enumerate(CASES):
        img = (rng.random((h, w, 3), dtype=np.float32) ** 2 * 1.3).astype(np.float32)        # some texels above 1: the pyramid clamps
        img[rng.random((h, w)) < 0.05] = 0.0
        hnd = R.ref_mipmap_create(w, h, vp(img), wu, wv, filt, aniso)
        n = R.ref_mipmap_levels(hnd)
        out["img%d" % ci] = img
        sizes = []
        for lv in range(n):
            wh = (C.c_int * 2)()
            R.ref_mipmap_level_size(hnd, lv, wh)
            a = np.zeros((wh[1], wh[0], 3), np.float32)
            R.ref_mipmap_level(hnd, lv, vp(a))
            sizes.append((wh[0], wh[1]))
            if lv > 0:
                out["lvl%d_%d" % (ci, lv)] = a.astype(np.float16)      # the values ARE half precision: lossless
                assert np.array_equal(a.astype(np.float16).astype(np.float32), a)
        out["sizes%d" % ci] = np.array(sizes, np.int32)
        N = 1500
        q = np.zeros((N, 6), np.float32)
        q[:, :2] = rng.random((N, 2)) * 3 - 1
        s = 10 ** rng.uniform(-3, -0.3, (N, 1))
        q[:, 2:] = (rng.standard_normal((N, 4)) * s).astype(np.float32)
        q[:40, 4:] = q[:40, 2:4] * np.float32(0.01)          # very skinny ellipses: the maxAnisotropy clamp
        q[40:50, 2:] = 0                                      # degenerate footprint
        out["q%d" % ci] = q
        for mode, name in ((0, "box"), (1, "bilinear"), (2, "filtered")):
            o = np.zeros((N, 3), np.float32)
            R.ref_mipmap_eval(hnd, mode, N, vp(q), vp(o))
            out["%s%d" % (name, ci)] = o
        st = np.zeros(9, np.float32)
        R.ref_mipmap_stats(hnd, vp(st))
        out["stats%d" % ci] = st
        R.ref_mipmap_destroy(hnd)
    np.savez_compressed(os.path.join(HERE, "texture.npz"), **out)
    print("wrote texture.npz:", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "texture.npz")), "bytes")


if __name__ == "__main__":
    main()
