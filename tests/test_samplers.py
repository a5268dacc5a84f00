"""CPU: the `halton` and `ldsampler` sample generators.

Three layers are checked against each other and against the reference:
  * tests/golden/qmc.npz -- values produced by the reference's own src/libcore/qmc.cpp and src/samplers/faure.cpp
    (compiled into oracle/_ref by tests/golden/make_golden_qmc.py);
  * oracle/orc_sampler.h -- the stateful restatement of HaltonSampler / LowDiscrepancySampler;
  * mitsuba_b200/csrc/qmc.h -- the random-access form the kernels use, compiled with g++ by tests/emu.
Known-answer table: src/tests/test_samplers.cpp:37-56 (MATLAB haltonset) and :79-86 (incremental radical inverse)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle
from tests import util

EMU_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu")


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-C", EMU_DIR], stdout=subprocess.DEVNULL)
    E = C.CDLL(os.path.join(EMU_DIR, "libemu.so"))
    E.emu_halton_sequence.argtypes = [C.c_int] * 9 + [C.c_void_p]
    E.emu_radical_inverse.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p]
    E.emu_radical_inverse.restype = C.c_float
    E.emu_permutation.argtypes = [C.c_int, C.c_int, C.c_void_p]
    E.emu_sample02.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
    E.emu_permute_pow2.argtypes = [C.c_uint32] * 4
    E.emu_permute_pow2.restype = C.c_uint32
    E.emu_ldsampler_sets.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    E.emu_sobol_sequence.argtypes = [C.c_int] * 9 + [C.c_void_p]
    E.emu_sobol_matrix32.argtypes = [C.c_uint32, C.c_uint32]
    E.emu_sobol_matrix32.restype = C.c_uint32
    E.emu_sobol_vdc.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
    return E


@pytest.fixture(scope="module")
def gold():
    return util.load_golden("qmc.npz")


@pytest.fixture(scope="module")
def L():
    return oracle.lib()


# MATLAB: p = haltonset(5); net(p,5)   (src/tests/test_samplers.cpp:37-44)
HALTON_KAT = np.array([
    [0, 0, 0, 0, 0],
    [0.5, 1 / 3, 0.2, 1 / 7, 1 / 11],
    [0.25, 2 / 3, 0.4, 2 / 7, 2 / 11],
    [0.75, 1 / 9, 0.6, 3 / 7, 3 / 11],
    [0.125, 4 / 9, 0.8, 4 / 7, 4 / 11]])


def halton(fn, res, scramble, px, py, first, n, n2d, n1d):
    out = np.zeros((n, 2 * n2d + n1d), np.float32)
    assert fn(res[0], res[1], scramble, px, py, first, n, n2d, n1d, vp(out)) == 0
    return out


def test_halton_known_answers(L, emu):
    """test01_Halton: a freshly constructed sampler (no permutations, no space partition), five next1D per sample."""
    for fn in (L.orc_halton_sequence, emu.emu_halton_sequence):
        got = halton(fn, (0, 0), 0, 0, 0, 0, 5, 0, 5)
        assert np.abs(got - HALTON_KAT).max() <= 1e-7


def test_radical_inverse_incremental(L):
    """test03_radicalInverseIncr: exact agreement of the incremental and the direct form in base 2."""
    x = np.float32(0.0)
    for i in range(20):
        assert x == np.float32(L.orc_radical_inverse(2, i))
        x = np.float32(L.orc_radical_inverse_incremental(2, float(x)))


def test_prime_table(gold, emu):
    emu.emu_prime.restype = C.c_int
    assert [emu.emu_prime(i) for i in range(0, 1024, 37)] == [int(gold["primes"][i]) for i in range(0, 1024, 37)]
    assert int(gold["primes"][1023]) == 8161


def test_radical_inverse_against_reference(gold, L, emu):
    """radicalInverseFast (qmc.cpp:198-): bit-exact in the oracle and in the kernel form, including indices > 2^32."""
    for a, d in enumerate(gold["dims"]):
        base = int(gold["primes"][d])
        o = np.array([L.orc_radical_inverse_fast(int(d), int(i)) for i in gold["idx"]], np.float32)
        e = np.array([emu.emu_radical_inverse(base, int(i), None) for i in gold["idx"]], np.float32)
        assert np.array_equal(o, gold["rinv"][a]), "oracle, dimension %d" % d
        assert np.array_equal(e, gold["rinv"][a]), "kernel form, dimension %d" % d
    g = np.array([[L.orc_radical_inverse(b, i) for i in range(40)] for b in (2, 3, 5, 7)], np.float32)
    assert np.array_equal(g, gold["generic"])
    assert np.array_equal(gold["incremental"], np.array([L.orc_radical_inverse(2, i) for i in range(20)], np.float32))


@pytest.mark.parametrize("name,scramble", [("faure", -1), ("s7", 7)])
def test_permutations_against_reference(gold, L, emu, name, scramble):
    """PermutationStorage (faure.cpp): Faure permutations and the TEA-shuffled ones, and the scrambled radical inverse."""
    for a, b in enumerate(gold["sdims"]):
        want = gold["perm_%s_%d" % (name, b)]
        o = np.zeros(want.size, np.uint16)
        e = np.zeros(want.size, np.uint16)
        assert L.orc_permutation(scramble, int(b), vp(o)) == want.size
        assert emu.emu_permutation(scramble, int(b), vp(e)) == want.size
        assert np.array_equal(o, want) and np.array_equal(e, want), "base index %d" % b
        assert sorted(want.tolist()) == list(range(want.size))
        so = np.array([L.orc_scrambled_radical_inverse_fast(int(b), int(i), vp(o)) for i in gold["idx"]], np.float32)
        se = np.array([emu.emu_radical_inverse(int(gold["primes"][b]), int(i), vp(e)) for i in gold["idx"]], np.float32)
        assert np.array_equal(so, gold["srinv_%s" % name][a]) and np.array_equal(se, gold["srinv_%s" % name][a])
    for b in (0, 1):
        p = gold["perm_%s_%d" % (name, b)]
        inv = np.zeros_like(p)
        inv[p] = np.arange(p.size, dtype=np.uint16)
        assert np.array_equal(inv, gold["invperm_%s_%d" % (name, b)])


def test_sample02_and_tea_against_reference(gold, L, emu):
    for a in range(gold["scr2"].shape[0]):
        s = gold["scr2"][a].copy()
        for c, n in enumerate(gold["n02"]):
            o, e = np.zeros(2, np.float32), np.zeros(2, np.float32)
            L.orc_sample02(int(n), vp(s), vp(o))
            emu.emu_sample02(int(n), vp(s), vp(e))
            assert np.array_equal(o, gold["s02"][a, c]) and np.array_equal(e, gold["s02"][a, c])
    for r, rounds in enumerate((4, 8)):
        got = np.array([L.orc_sample_tea(int(a), int(b), rounds) for a, b in gold["tea_in"]], np.uint64)
        assert np.array_equal(got, gold["tea"][r])


@pytest.mark.parametrize("res,scramble", [((512, 512), -1), ((1920, 1080), -1), ((64, 48), 0), ((300, 200), 7)])
def test_halton_kernel_form_equals_stateful_sampler(L, emu, res, scramble):
    """The random-access form used by k_raygen / k_shade returns the stateful sampler's numbers bit for bit, for
    pixels all over the film, late sample indices and deep dimensions."""
    rng = np.random.default_rng(5)
    for _ in range(12):
        px, py = int(rng.integers(0, res[0])), int(rng.integers(0, res[1]))
        first = int(rng.integers(0, 5000))
        a = halton(L.orc_halton_sequence, res, scramble, px, py, first, 6, 9, 5)
        b = halton(emu.emu_halton_sequence, res, scramble, px, py, first, 6, 9, 5)
        assert np.array_equal(a, b)
        # the first two dimensions address the pixel: the offset stays inside it (up to rounding of the product)
        assert (a[:, :2] > -1e-4).all() and (a[:, :2] < 1 + 1e-4).all()


def test_halton_enumeration_covers_pixels(L):
    """Gruenschloss enumeration (halton.cpp:277-296): sample j of pixel (x, y) is element offset + j*stride of the
    global sequence, and that element falls into the pixel (modulo the 128-pixel period)."""
    res = (200, 100)
    for px, py in [(0, 0), (5, 77), (127, 99), (130, 3), (199, 50)]:
        s = halton(L.orc_halton_sequence, res, 0, px, py, 0, 4, 1, 0)
        assert (np.floor(s) == 0).all()


def test_permute_pow2_is_a_bijection(emu):
    rng = np.random.default_rng(9)
    for bits in (0, 1, 2, 3, 4, 6, 8, 10, 12):
        n = 1 << bits
        for _ in range(20):
            ka, kb = (int(v) for v in rng.integers(0, 1 << 32, 2))
            img = sorted(emu.emu_permute_pow2(i, bits, ka, kb) for i in range(n))
            assert img == list(range(n))
    # positions are uniform over keys: chi-square of where index 3 of 16 lands over 8000 keys
    hits = np.zeros(16)
    for _ in range(8000):
        ka, kb = (int(v) for v in rng.integers(0, 1 << 32, 2))
        hits[emu.emu_permute_pow2(3, 4, ka, kb)] += 1
    chi2 = ((hits - 500.0) ** 2 / 500.0).sum()
    assert chi2 < 45.0, chi2          # 15 degrees of freedom, p ~ 1e-4


def stratification_1d(v):
    n = v.size
    return np.array_equal(np.sort(np.floor(v * n).astype(int)), np.arange(n))


def test_ldsampler_point_sets(L, emu):
    """ldsampler.cpp:152-197: every dimension of a pixel is a scrambled, shuffled (0,2)-sequence prefix -- a
    permutation of the n strata in 1D, and one point per elementary interval in 2D.  Holds for the oracle's
    literal restatement and for the keyed random-access form of the kernels alike."""
    n, dim = 64, 4
    o1, o2 = np.zeros((dim, n), np.float32), np.zeros((dim, n, 2), np.float32)
    e1, e2 = np.zeros((dim, n), np.float32), np.zeros((dim, n, 2), np.float32)
    assert L.orc_ldsampler_sets(n, dim, 11, vp(o1), vp(o2)) == n
    assert emu.emu_ldsampler_sets(n, dim, 11, 17, 5, vp(e1), vp(e2)) == 0
    for s1, s2 in ((o1, o2), (e1, e2)):
        for d in range(dim):
            assert stratification_1d(s1[d])
            assert stratification_1d(s2[d, :, 0]) and stratification_1d(s2[d, :, 1])
            for a, b in ((8, 8), (16, 4), (4, 16), (2, 32)):          # elementary intervals of the (0,2) net
                cells = np.floor(s2[d, :, 0] * a).astype(int) * b + np.floor(s2[d, :, 1] * b).astype(int)
                assert np.array_equal(np.sort(cells), np.arange(n))
        # different dimensions are differently scrambled and shuffled
        assert not np.array_equal(s1[0], s1[1]) and not np.array_equal(s2[0], s2[1])
    # neighbouring pixels get different sets in the kernel form
    f1, f2 = np.zeros((dim, n), np.float32), np.zeros((dim, n, 2), np.float32)
    emu.emu_ldsampler_sets(n, dim, 11, 18, 5, vp(f1), vp(f2))
    assert not np.array_equal(e1, f1)


def test_ldsampler_order_is_decorrelated(emu):
    """The shuffle must break the correlation between dimensions: with identity order, sample i of every dimension
    would be the same radical-inverse digit pattern.  Rank correlation between two dimensions stays small."""
    n, dim = 256, 4
    cors = []
    for px in range(40):
        e1, e2 = np.zeros((dim, n), np.float32), np.zeros((dim, n, 2), np.float32)
        emu.emu_ldsampler_sets(n, dim, 3, px, 9, vp(e1), vp(e2))
        cors.append(np.corrcoef(e1[0], e1[1])[0, 1])
        cors.append(np.corrcoef(e2[0, :, 0], e2[1, :, 1])[0, 1])
    cors = np.array(cors)
    assert np.abs(cors.mean()) < 0.03 and np.abs(cors).max() < 0.3, (cors.mean(), np.abs(cors).max())


# ---- sobol (src/samplers/sobol.cpp, sobolseq.h) ------------------------------------------------------------------

@pytest.fixture(scope="module")
def sgold():
    return util.load_golden("qmc_sobol.npz")


def test_sobol_tables_against_reference(sgold, L, emu):
    """The generator matrices expanded from the Joe-Kuo seeds equal the reference's literal table entry for entry
    (1024 dimensions x 52 columns), and so do the pixel tables of sobol::look_up for every resolution 2^1 .. 2^25 --
    in the oracle (direction-number recurrence, row elimination) and in the product's host code (integer recurrence,
    column elimination), two independent derivations."""
    mo = np.array([[L.orc_sobol_matrix32(d, c) for c in range(52)] for d in range(1024)], np.uint32)
    me = np.array([[emu.emu_sobol_matrix32(d, c) for c in range(52)] for d in range(1024)], np.uint32)
    assert np.array_equal(mo, sgold["matrices32"]) and np.array_equal(me, sgold["matrices32"])
    for m in range(1, sgold["vdc"].shape[0] + 1):
        a, b, c, d = (np.zeros(52, np.uint64) for _ in range(4))
        L.orc_sobol_vdc(m, vp(a), vp(b))
        assert emu.emu_sobol_vdc(m, vp(c), vp(d)) == 0
        n = 52 - 2 * m
        for got_vdc, got_inv in ((a, b), (c, d)):
            assert np.array_equal(got_vdc[:n], sgold["vdc"][m - 1][:n]), m
            assert np.array_equal(got_inv[:2 * m], sgold["vdc_inv"][m - 1][:2 * m]), m


def test_sobol_values_against_reference(sgold, L):
    for a, scr in enumerate(sgold["scramble"]):
        for b, d in enumerate(sgold["dims"]):
            got = np.array([L.orc_sobol_sample(int(i), int(d), int(scr) & 0xFFFFFFFF) for i in sgold["idx"]], np.float32)
            assert np.array_equal(got, sgold["values"][a, b]), (scr, d)


@pytest.mark.parametrize("res,scramble", [((512, 512), 0), ((1920, 1080), 0), ((64, 48), 3), ((300, 200), 77)])
def test_sobol_kernel_form_equals_stateful_sampler(L, emu, res, scramble):
    rng = np.random.default_rng(6)
    for _ in range(12):
        px, py = int(rng.integers(0, res[0])), int(rng.integers(0, res[1]))
        first = int(rng.integers(0, 3000))
        a = halton(L.orc_sobol_sequence, res, scramble, px, py, first, 6, 9, 5)
        b = halton(emu.emu_sobol_sequence, res, scramble, px, py, first, 6, 9, 5)
        assert np.array_equal(a, b)
        assert (a[:, :2] > -1e-4).all() and (a[:, :2] < 1 + 1e-4).all()       # the enumeration stays inside the pixel


def test_sobol_look_up_against_reference(sgold, L):
    """sobol::look_up through the stateful sampler: sample `frame` of pixel (px, py) is element look_up(...) of the
    sequence; dimension 2 of that element is compared with the reference's sample(look_up(...), 2)."""
    R = None
    for case, want in zip(sgold["look_cases"], sgold["look"]):
        m, frame, px, py, scr = (int(v) for v in case)
        if scr > 0xFFFFFFFF or scr == 12345:       # the sampler derives its mask from the property by TEA; raw masks are covered above
            continue
        res = 1 << m
        out = halton(L.orc_sobol_sequence, (res, res), 0, px, py, frame, 1, 1, 1)
        assert out[0, 2] == np.float32(L.orc_sobol_sample(int(want), 2, 0))
        if int(want) != frame:
            assert np.floor(out[0, :2]).tolist() == [0.0, 0.0]
