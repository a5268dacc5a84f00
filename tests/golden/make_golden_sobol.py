"""Generates tests/golden/qmc_sobol.npz from oracle/_ref/libmtsref.so (the reference's src/samplers/sobolseq.cpp and
sobolseq.h, compiled where they lie): the complete 32-bit generator matrices, the pixel tables behind sobol::look_up
for every resolution, and sample / look_up values.

    make -C oracle all ref && python tests/golden/make_golden_sobol.py
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
    R.ref_sobol_matrix32.restype = C.c_uint32
    R.ref_sobol_matrix32.argtypes = [C.c_uint32, C.c_uint32]
    R.ref_sobol_vdc.restype = C.c_uint64
    R.ref_sobol_vdc.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
    R.ref_sobol_sample.restype = C.c_float
    R.ref_sobol_sample.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64]
    R.ref_sobol_look_up.restype = C.c_uint64
    R.ref_sobol_look_up.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64]
    rng = np.random.default_rng(20261106)
    mat = np.array([[R.ref_sobol_matrix32(d, c) for c in range(52)] for d in range(1024)], np.uint32)
    n_m = 25           # rows of Matrices::vdc_sobol_matrices (m = 1..25); the inverse table has one more
    vdc = np.array([[R.ref_sobol_vdc(m, c, 0) for c in range(52)] for m in range(1, n_m + 1)], np.uint64)
    inv = np.array([[R.ref_sobol_vdc(m, c, 1) for c in range(52)] for m in range(1, n_m + 1)], np.uint64)
    dims = np.concatenate([np.arange(0, 24), [100, 511, 1023]]).astype(np.uint32)
    idx = np.concatenate([np.arange(0, 64), rng.integers(64, 1 << 32, 200), rng.integers(1 << 32, 1 << 50, 100)]).astype(np.uint64)
    scr = np.array([0, 0x9E3779B97F4A7C15, 12345], np.uint64)
    val = np.array([[[R.ref_sobol_sample(int(i), int(d), int(s)) for i in idx] for d in dims] for s in scr], np.float32)
    cases = []
    for m in (2, 5, 6, 9, 11, 12):
        for _ in range(40):
            cases.append((m, int(rng.integers(0, 4096)), int(rng.integers(0, 1 << m)), int(rng.integers(0, 1 << m)), int(scr[rng.integers(0, 3)])))
    look = np.array([R.ref_sobol_look_up(*c) for c in cases], np.uint64)
    np.savez_compressed(os.path.join(HERE, "qmc_sobol.npz"), matrices32=mat, vdc=vdc, vdc_inv=inv, dims=dims, idx=idx, scramble=scr,
                        values=val, look_cases=np.array(cases, np.uint64), look=look)
    print("wrote qmc_sobol.npz")


if __name__ == "__main__":
    main()
