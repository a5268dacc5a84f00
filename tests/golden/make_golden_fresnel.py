"""Generates tests/golden/fresnel.npz from oracle/_ref/libmtsref.so (compiled from the reference tree): the Fresnel
terms of src/libcore/util.cpp:651-860 that the smooth and rough BSDFs share, the diffuse Fresnel reflectance integral
(plastic.cpp:195-197) and the SPD -> RGB conversion of three data/ior presets (conductor `material=` names).

    make -C oracle all ref && python tests/golden/make_golden_fresnel.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def main():
    oracle.build(ref=True)
    R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libmtsref.so"))
    R.ref_static_init()
    R.ref_fresnel_dielectric_ext.restype = C.c_float
    R.ref_fresnel_dielectric_ext.argtypes = [C.c_float, C.c_float, C.POINTER(C.c_float)]
    R.ref_fresnel_conductor_exact.argtypes = [C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    R.ref_fresnel_diffuse_reflectance.restype = C.c_float
    R.ref_fresnel_diffuse_reflectance.argtypes = [C.c_float, C.c_int]
    rng = np.random.default_rng(20261104)
    # dielectric: cos in [-1, 1] incl. grazing / normal incidence, eta on both sides of one and the index-matched case
    cos_i = np.concatenate([rng.uniform(-1, 1, 500), [1.0, -1.0, 0.0, 1e-4, -1e-4, 0.74535599, -0.74535599]]).astype(np.float32)
    etas = np.array([1.5, 1 / 1.5, 1.5046 / 1.000277, 1.000277 / 1.5046, 1.33, 2.4, 1.0], np.float32)
    F = np.zeros((etas.size, cos_i.size), np.float32)
    cos_t = np.zeros_like(F)
    t = C.c_float()
    for a, eta in enumerate(etas):
        for b, c in enumerate(cos_i):
            F[a, b] = R.ref_fresnel_dielectric_ext(float(c), float(eta), C.byref(t))
            cos_t[a, b] = t.value
    # conductor: cos in [0, 1], a few (eta, k) triples
    cos_c = np.concatenate([rng.uniform(0, 1, 300), [0.0, 1.0]]).astype(np.float32)
    ek = np.array([[[0.2, 0.92, 1.1], [3.9, 2.45, 2.14]], [[1.66, 0.88, 0.52], [9.2, 6.3, 4.8]], [[0.0, 0.0, 0.0], [1.0, 1.0, 1.0]],
                   [[1.5, 2.5, 3.5], [0.0, 0.1, 0.01]]], np.float32)
    Fc = np.zeros((ek.shape[0], cos_c.size, 3), np.float32)
    for a in range(ek.shape[0]):
        for b, c in enumerate(cos_c):
            R.ref_fresnel_conductor_exact(float(c), vp(ek[a, 0].copy()), vp(ek[a, 1].copy()), vp(Fc[a, b]))
    # diffuse Fresnel reflectance, exact (adaptive quadrature) mode
    eta_d = np.array([1.05, 1.2, 1.33, 1.4895874, 1.5, 1.9, 2.5, 1 / 1.05, 1 / 1.2, 1 / 1.33, 1 / 1.4895874, 1 / 1.5, 1 / 1.9, 0.4], np.float32)
    fdr = np.array([R.ref_fresnel_diffuse_reflectance(float(e), 0) for e in eta_d], np.float32)
    # SPD -> RGB for three presets
    names = ["Cu", "Au", "Al"]
    rgb = np.zeros((3, 2, 3), np.float32)
    for a, n in enumerate(names):
        for b, w in enumerate(("eta", "k")):
            assert R.ref_spd_to_rgb(("/root/reference/data/ior/%s.%s.spd" % (n, w)).encode(), vp(rgb[a, b])) == 0
    np.savez_compressed(os.path.join(HERE, "fresnel.npz"), cos_i=cos_i, etas=etas, F=F, cos_t=cos_t, cos_c=cos_c, ek=ek, Fc=Fc,
                        eta_d=eta_d, fdr=fdr, ior_names=np.array(names), ior_rgb=rgb)
    print("wrote fresnel.npz")


if __name__ == "__main__":
    main()
