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


def test_fresnel_terms_vs_reference(oracle_mod):
    """util.cpp:651-761 as compiled from the reference (tests/golden/make_golden_fresnel.py): the oracle's Fresnel terms
    are the same arithmetic, so they must agree bit for bit (strict IEEE on both sides)."""
    g = util.load_golden("fresnel.npz")
    L = oracle_mod.lib()
    t = C.c_float()
    for a, eta in enumerate(g["etas"]):
        got = np.zeros((2, g["cos_i"].size), np.float32)
        for b, c in enumerate(g["cos_i"]):
            got[0, b] = L.orc_fresnel_dielectric_ext(float(c), float(eta), C.byref(t))
            got[1, b] = t.value
        assert np.array_equal(got[0], g["F"][a]) and np.array_equal(got[1], g["cos_t"][a]), float(eta)
    for a in range(g["ek"].shape[0]):
        o = np.zeros(3, np.float32)
        for b, c in enumerate(g["cos_c"]):
            L.orc_fresnel_conductor_exact(float(c), vp(g["ek"][a, 0].copy()), vp(g["ek"][a, 1].copy()), vp(o))
            assert np.all(util.rel_close(o, g["Fc"][a, b], 1e-6, 1e-7)), (a, float(c), o, g["Fc"][a, b])


def test_diffuse_fresnel_reflectance_vs_reference():
    """plastic.cpp:195-197 -> fresnelDiffuseReflectance(eta, false): the reference integrates adaptively to a relative
    tolerance of 1e-5 (util.cpp:856); the host-side restatement must sit inside that tolerance."""
    from mitsuba_b200.scenedesc import fresnel_diffuse_reflectance
    g = util.load_golden("fresnel.npz")
    for eta, want in zip(g["eta_d"], g["fdr"]):
        assert float(fresnel_diffuse_reflectance(eta)) == pytest.approx(float(want), rel=1e-5)


def test_conductor_presets_vs_reference():
    """`material=` names: mitsuba_b200/host/ior_rgb.inc and scenes.METALS hold what the reference's own SPD -> RGB
    conversion produces (spectrum.cpp:172-191)."""
    from mitsuba_b200 import scenes
    g = util.load_golden("fresnel.npz")
    table = {}
    with open(os.path.join(os.path.dirname(__file__), "..", "mitsuba_b200", "host", "ior_rgb.inc")) as f:
        for line in f:
            if line.startswith("{"):
                name = line.split('"')[1]
                nums = [float(x.rstrip("f")) for x in line.replace("{", " ").replace("}", " ").replace(",", " ").split()[1:]]
                table[name] = np.array(nums, np.float32).reshape(2, 3)
    assert len(table) == 71 and "Ag" in table and "W" in table
    for name, rgb in zip(g["ior_names"], g["ior_rgb"]):
        assert np.array_equal(table[str(name)], rgb), name
        assert np.array_equal(np.array(scenes.METALS[str(name)], np.float32), rgb), name


@pytest.mark.skipif(not os.path.isdir("/root/reference/data/ior"), reason="reference tree absent (GPU box)")
def test_conductor_preset_table_is_current(oracle_mod, tmp_path):
    """Every row of ior_rgb.inc equals a fresh conversion by the reference's code."""
    R = oracle_mod.ref()
    R.ref_static_init()
    out = np.zeros(3, np.float32)
    with open(os.path.join(os.path.dirname(__file__), "..", "mitsuba_b200", "host", "ior_rgb.inc")) as f:
        rows = [l for l in f if l.startswith("{")]
    for line in rows:
        name = line.split('"')[1]
        nums = np.array([float(x.rstrip("f")) for x in line.replace("{", " ").replace("}", " ").replace(",", " ").split()[1:]], np.float32)
        for j, w in enumerate(("eta", "k")):
            assert R.ref_spd_to_rgb(("/root/reference/data/ior/%s.%s.spd" % (name, w)).encode(), vp(out)) == 0
            assert np.array_equal(out, nums[3 * j:3 * j + 3]), (name, w)
