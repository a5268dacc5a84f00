"""Generates tests/golden/envmap.npz from oracle/_ref/libmtsref.so: float -> half -> float conversions of the
reference's src/libcore/half.cpp (the storage type of the environment map, envmap.cpp:101-103) and the tent warp of
src/libcore/warp.cpp:157-176 used by EnvironmentMap::internalSampleDirection.

    make -C oracle all ref && python tests/golden/make_golden_envmap.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    oracle.build(ref=True)
    R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libmtsref.so"))
    R.ref_static_init()
    R.ref_float_to_half.restype = C.c_uint16
    R.ref_float_to_half.argtypes = [C.c_float]
    R.ref_half_to_float.restype = C.c_float
    R.ref_half_to_float.argtypes = [C.c_uint16]
    R.ref_square_to_tent.argtypes = [C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(20261107)
    x = np.concatenate([rng.uniform(0, 5, 2000), 10 ** rng.uniform(-9, 4.8, 3000), -(10 ** rng.uniform(-6, 3, 200)),
                        [0, 65504, 65519.9, 65520, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-08, 6.1e-5, 6.1035156e-05, 1, 0.1]]).astype(np.float32)
    bits = np.array([R.ref_float_to_half(float(v)) for v in x], np.uint16)
    all_bits = np.arange(0, 65536, 7, dtype=np.uint16)
    all_floats = np.array([R.ref_half_to_float(int(v)) for v in all_bits], np.float32)
    u = np.concatenate([rng.uniform(0, 1, (500, 2)), [[0, 0], [0.5, 0.5], [0.49999997, 0.99999994], [0.25, 0.75]]]).astype(np.float32)
    tent = np.zeros_like(u)
    for i in range(u.shape[0]):
        o = np.zeros(2, np.float32)
        R.ref_square_to_tent(u[i].ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
        tent[i] = o
    np.savez_compressed(os.path.join(HERE, "envmap.npz"), x=x, half_bits=bits, all_bits=all_bits, all_floats=all_floats, u=u, tent=tent)
    print("wrote envmap.npz")


if __name__ == "__main__":
    main()
