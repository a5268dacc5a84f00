"""CPU: the oracle restatement against fixtures produced by the reference's own code
(tests/golden/make_golden.py, oracle/_ref) and against the reference's known-answer test
src/tests/test_kd.cpp:34-84."""
import ctypes as C
import os

import numpy as np
import pytest

from mitsuba_b200 import abi
from tests import util


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("name", ["cbox", "mesh"])
def test_scene_inputs_match_fixture(name):
    g = util.load_golden("hits_%s.npz" % name)
    assert bytes(g["digest"]).decode() == util.scene_digest(util.golden_scenes()[name])


@pytest.mark.parametrize("name", ["cbox", "mesh"])
def test_oracle_hits_vs_reference_kdtree(oracle_mod, name):
    g = util.load_golden("hits_%s.npz" % name)
    orc = oracle_mod.Oracle(util.golden_scenes()[name])
    tight, enlarged, _ = orc.bounds()
    assert np.array_equal(tight, g["tight"]) and np.array_equal(enlarged, g["enlarged"])
    for rs in ("chord", "camera", "secondary"):
        rays = g["rays_" + rs]
        st = util.compare_hits(orc.trace(rays, abi.TRACE_CLOSEST), g["closest_" + rs], label=name + "/" + rs)
        assert st["bit_exact"] >= 0.995 * st["n"], st
        occl = orc.trace(rays, abi.TRACE_ANY)["prim"] != abi.NO_HIT
        assert (occl != g["any_" + rs]).sum() <= max(3, 2e-3 * len(rays))


def test_triaccel_records_bit_exact(oracle_mod):
    g = util.load_golden("triaccel.npz")
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    b = SceneBuilder()
    m = b.add_material(diffuse(0.5))
    tri = g["tri"]
    b.add_mesh(tri.reshape(-1, 3), np.arange(tri.shape[0] * 3, dtype=np.uint32).reshape(-1, 3), m)
    orc = oracle_mod.Oracle(b)
    for i in range(tri.shape[0]):
        rec = orc.triaccel(i)
        ref = g["rec"][i]
        assert rec[0] == ref[0]
        if ref[0] != 3:
            assert np.array_equal(rec[:10], ref[:10]), i


def test_microfacet_vs_reference(oracle_mod):
    g = util.load_golden("microfacet.npz")
    L = oracle_mod.lib()
    worst = 0.0
    for c, ref in zip(g["cases"], g["out"]):
        o = np.zeros(6, np.float32)
        wi, u = c[4:7].copy(), c[7:9].copy()
        L.orc_microfacet(int(c[0]), float(c[1]), float(c[2]), int(c[3]), vp(wi), vp(u), vp(o))
        assert np.all(util.rel_close(o, ref, 2e-5, 1e-7)), (c, o, ref)
        worst = max(worst, float(np.max(np.abs(o - ref) / np.maximum(np.abs(ref), 1e-3))))
    assert worst < 1e-4


def test_scalar_functions_vs_reference(oracle_mod):
    g = util.load_golden("scalar.npz")
    L = oracle_mod.lib()
    assert np.array_equal(np.array([L.orc_erfinv(float(v)) for v in g["x"]], np.float32), g["erfinv"])
    assert np.array_equal(np.array([L.orc_erf(float(v)) for v in g["xe"]], np.float32), g["erf"])
    for k in range(4):
        for i in range(g["u"].shape[0]):
            o = np.zeros(3, np.float32)
            L.orc_warp(k, vp(g["u"][i].copy()), vp(o))
            assert np.all(util.rel_close(o, g["warp"][k, i], 1e-6, 1e-7))


@pytest.mark.skipif(not os.path.isdir("/root/reference/include/mitsuba"), reason="reference tree absent (GPU box)")
def test_reference_clipped_aabb_known_answers(oracle_mod):
    """src/tests/test_kd.cpp:34-84 -- pins the library compiled from the reference tree."""
    R = oracle_mod.ref()
    assert R is not None
    tri = np.array([0, 0, 0, 1, 0, 0, 1, 1, 0], np.float32)
    cases = [
        ([0, .5, -1, 1, 1, 1], True, [.5, .5, 0], [1, 1, 0]),
        ([2, 2, 2, 3, 3, 3], False, None, None),
        ([-1, -1, -1, 1, 1, 1], True, [0, 0, 0], [1, 1, 0]),
        ([-100, -100, 0, 100, 100, 0], True, [0, 0, 0], [1, 1, 0]),
        ([0, 1, 0, 1, 2, 0], True, [1, 1, 0], [1, 1, 0]),
    ]
    for box, valid, lo, hi in cases:
        out = np.zeros(6, np.float32)
        ok = R.ref_clipped_aabb(vp(tri), vp(np.array(box, np.float32)), vp(out))
        assert bool(ok) == valid
        if valid:
            assert np.allclose(out[:3], lo, atol=1e-6) and np.allclose(out[3:], hi, atol=1e-6)
