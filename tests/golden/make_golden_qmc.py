"""Generates tests/golden/qmc.npz from oracle/_ref/libmtsref.so (the reference's own src/libcore/qmc.cpp and
src/samplers/faure.cpp, compiled where they lie): radical inverses, digit permutations, (0,2)-sequence points and
the TEA generator -- the building blocks of the `halton` and `ldsampler` plugins.

    make -C oracle all ref && python tests/golden/make_golden_qmc.py
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
    R.ref_radical_inverse_fast.restype = C.c_float
    R.ref_radical_inverse_fast.argtypes = [C.c_int, C.c_uint64]
    R.ref_scrambled_radical_inverse_fast.restype = C.c_float
    R.ref_scrambled_radical_inverse_fast.argtypes = [C.c_int, C.c_uint64, C.c_void_p]
    R.ref_radical_inverse.restype = C.c_float
    R.ref_radical_inverse.argtypes = [C.c_int, C.c_uint64]
    R.ref_radical_inverse_incremental.restype = C.c_float
    R.ref_radical_inverse_incremental.argtypes = [C.c_int, C.c_float]
    R.ref_permutation.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
    R.ref_sample02.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
    R.ref_sample_tea.restype = C.c_uint64
    R.ref_sample_tea.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
    rng = np.random.default_rng(20261105)
    primes = np.array([R.ref_prime(i) for i in range(1024)], np.int32)
    # radical inverses: small, medium and > 2^32 indices over low and high dimensions
    dims = np.concatenate([np.arange(0, 40), [63, 100, 255, 511, 1000, 1023]]).astype(np.int32)
    idx = np.concatenate([np.arange(0, 64), rng.integers(64, 1 << 20, 200), rng.integers(1 << 20, 1 << 32, 200),
                          rng.integers(1 << 32, 1 << 45, 100)]).astype(np.uint64)
    rinv = np.array([[R.ref_radical_inverse_fast(int(d), int(i)) for i in idx] for d in dims], np.float32)
    perm_bases = [0, 1, 2, 3, 4, 5, 10, 25, 100, 500, 1023]
    perms = {}
    for scr in (-1, 7):
        for b in perm_bases:
            buf = (C.c_uint16 * int(primes[b]))()
            n = R.ref_permutation(scr, b, 0, buf)
            perms["perm_%s_%d" % ("faure" if scr == -1 else "s7", b)] = np.array(buf[:n], np.uint16)
    inv = {}
    for scr in (-1, 7):
        for b in (0, 1):
            buf = (C.c_uint16 * int(primes[b]))()
            R.ref_permutation(scr, b, 1, buf)
            inv["invperm_%s_%d" % ("faure" if scr == -1 else "s7", b)] = np.array(buf[:], np.uint16)
    sdims = np.array(perm_bases, np.int32)
    srinv = {}
    for scr in (-1, 7):
        out = np.zeros((sdims.size, idx.size), np.float32)
        for a, b in enumerate(perm_bases):
            p = perms["perm_%s_%d" % ("faure" if scr == -1 else "s7", b)].copy()
            for c, i in enumerate(idx):
                out[a, c] = R.ref_scrambled_radical_inverse_fast(b, int(i), p.ctypes.data_as(C.c_void_p))
        srinv["srinv_%s" % ("faure" if scr == -1 else "s7")] = out
    # generic + incremental radical inverse (test_samplers.cpp:79-86)
    gen = np.array([[R.ref_radical_inverse(int(b), int(i)) for i in range(40)] for b in (2, 3, 5, 7)], np.float32)
    x = np.float32(0.0)
    incr = []
    for i in range(20):
        incr.append(x)
        x = np.float32(R.ref_radical_inverse_incremental(2, float(x)))
    # (0,2)-sequence
    scr2 = rng.integers(0, 1 << 32, (6, 2)).astype(np.uint32)
    scr2[0] = 0
    n02 = np.concatenate([np.arange(0, 256), rng.integers(256, 1 << 32, 64)]).astype(np.uint32)
    s02 = np.zeros((scr2.shape[0], n02.size, 2), np.float32)
    for a in range(scr2.shape[0]):
        s = scr2[a].copy()
        for c, n in enumerate(n02):
            o = np.zeros(2, np.float32)
            R.ref_sample02(int(n), s.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
            s02[a, c] = o
    tea_in = rng.integers(0, 1 << 32, (64, 2)).astype(np.uint32)
    tea = np.array([[R.ref_sample_tea(int(a), int(b), r) for a, b in tea_in] for r in (4, 8)], np.uint64)
    np.savez_compressed(os.path.join(HERE, "qmc.npz"), primes=primes, dims=dims, idx=idx, rinv=rinv, sdims=sdims,
                        generic=gen, incremental=np.array(incr, np.float32), scr2=scr2, n02=n02, s02=s02,
                        tea_in=tea_in, tea=tea, **perms, **inv, **srinv)
    print("wrote qmc.npz")


if __name__ == "__main__":
    main()
