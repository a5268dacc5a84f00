"""GPU: BSDF / emitter / intersection / sensor / film records through the C ABI against the oracle."""
import numpy as np
import pytest

from mitsuba_b200 import abi, scenes
from mitsuba_b200.scenedesc import SceneBuilder, conductor, dielectric, diffuse, look_at, plastic, roughconductor, roughdielectric
from tests import util

pytestmark = pytest.mark.gpu

# fp32 with FMA contraction on the device vs strict IEEE on the host: a few ulp per operation
RTOL, ATOL = 5e-4, 2e-6


def material_scene():
    b = SceneBuilder()
    eta, k = scenes.METALS["Cu"]
    mats = {
        "diffuse": b.add_material(diffuse([0.5, 0.3, 0.7])),
        "diffuse2s": b.add_material(diffuse([0.5, 0.3, 0.7], twosided=True)),
        "rc_beck": b.add_material(roughconductor(eta, k, alpha=0.3, distribution="beckmann")),
        "rc_ggx": b.add_material(roughconductor(eta, k, alpha=0.2, distribution="ggx")),
        "rc_ggx_aniso_all": b.add_material(roughconductor(eta, k, distribution="ggx", alpha_u=0.5, alpha_v=0.3, sample_visible=False)),
        "rc_beck_all": b.add_material(roughconductor(eta, k, alpha=0.3, distribution="beckmann", sample_visible=False)),
        "rd_beck": b.add_material(roughdielectric(1.5, 1.0, alpha=0.3, distribution="beckmann")),
        "rd_ggx": b.add_material(roughdielectric(1.5, 1.0, alpha=0.3, distribution="ggx")),
        "rd_ggx_all": b.add_material(roughdielectric(1.5, 1.0, alpha=0.2, distribution="ggx", sample_visible=False)),
    }
    # smooth (Dirac) models, SURVEY 8(f) N3
    mats["glass"] = b.add_material(dielectric(1.5046, 1.000277, [0.9, 0.8, 0.7], [0.6, 0.5, 0.4]))
    mats["glass_inv"] = b.add_material(dielectric(1.0, 1.33))
    mats["mirror"] = b.add_material(conductor(eta, k, specular_reflectance=[1.0, 0.5, 0.25]))
    mats["mirror2s"] = b.add_material(conductor(0.0, 1.0, twosided=True))
    mats["mirror_over_diffuse"] = b.add_twosided(conductor(eta, k), diffuse([0.2, 0.4, 0.6]))
    mats["plastic"] = b.add_material(plastic(1.49, 1.000277, [0.6, 0.3, 0.1], [1.0, 0.9, 0.8]))
    mats["plastic_nonlinear2s"] = b.add_material(plastic(1.9, 1.33, [0.1, 0.2, 0.9], 0.5, nonlinear=True, twosided=True))
    mats["plastic_over_mirror"] = b.add_twosided(plastic(), conductor(eta, k))
    mats["mixed2s"] = b.add_twosided(roughconductor(eta, k, alpha=0.2, distribution="ggx"), diffuse([0.2, 0.4, 0.6]))
    b.add_quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), mats["diffuse"])
    b.add_quad((-0.3, 2, -0.3), (0.3, 2, -0.3), (0.3, 2, 0.3), (-0.3, 2, 0.3), mats["diffuse"], area_radiance=[5, 4, 3])
    b.add_point_light([0.5, 1.5, 0.2], [3, 3, 3], sampling_weight=0.5)
    b.add_constant_env([0.3, 0.4, 0.5], sampling_weight=2.0)
    b.set_sensor(look_at([0, 1, -4], [0, 0.5, 0], [0, 1, 0]), 40.0, 48, 32)
    return b, mats


def queries(n, seed):
    rng = np.random.default_rng(seed)
    q = np.zeros(n, dtype=abi.BSDF_QUERY_DTYPE)
    for f in ("wi", "wo"):
        v = rng.normal(size=(n, 3))
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        q[f] = v.astype(np.float32)
    q["wi"][: n // 2, 2] = np.abs(q["wi"][: n // 2, 2])       # half of them on the front side
    q["u"] = rng.random((n, 3)).astype(np.float32)
    return q


def test_bsdf_eval_pdf_sample_vs_oracle(gpu_lib, oracle_mod):
    b, mats = material_scene()
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    q = queries(20000, 1)
    for name, mid in mats.items():
        g, o = sc.bsdf_query(mid, q), orc.bsdf_query(mid, q)
        # the discrete decisions (sample succeeded, reflection vs transmission) must agree except where
        # a comparison sits within rounding of its threshold
        agree = (g["spdf"] > 0) == (o["spdf"] > 0)
        agree &= np.sign(g["swo"][:, 2]) == np.sign(o["swo"][:, 2])
        assert agree.mean() > 0.999, (name, agree.mean())
        for f in ("eval", "pdf"):
            assert util.frac_close(g[f], o[f], RTOL, ATOL) > 0.999, (name, f)
        for f in ("weight", "spdf", "swo", "seta"):
            fr = util.frac_close(g[f][agree], o[f][agree], 2e-3, 1e-5)
            assert fr > 0.995, (name, f, fr)
    sc.close()


def test_emitter_sampling_vs_oracle(gpu_lib, oracle_mod):
    b, _ = material_scene()
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    rng = np.random.default_rng(2)
    n = 20000
    q = np.zeros(n, dtype=abi.DIRECT_QUERY_DTYPE)
    q["ref"] = (rng.random((n, 3)) * [2, 1.5, 2] - [1, 0, 1]).astype(np.float32)
    nn = rng.normal(size=(n, 3))
    q["ref_n"] = (nn / np.linalg.norm(nn, axis=1, keepdims=True)).astype(np.float32)
    q["ref_n"][::4] = 0            # refN == 0: BSDFs with transmission (records.inl:175-179)
    q["u"] = rng.random((n, 2)).astype(np.float32)
    g, o = sc.emitter_sample(q), orc.emitter_sample(q)
    assert (g["emitter"] == o["emitter"]).mean() > 0.9995
    ok = (g["emitter"] == o["emitter"]) & (o["emitter"] >= 0)
    assert set(np.unique(o["emitter"][ok])) == {0, 1, 2}
    assert np.array_equal(g["measure"][ok], o["measure"][ok])
    for f in ("value", "pdf", "d", "dist", "p", "n"):
        assert util.frac_close(g[f][ok], o[f][ok], RTOL, 1e-5) > 0.999, f
    sc.close()


def test_intersection_records_and_sensor_rays_vs_oracle(gpu_lib, oracle_mod):
    b = scenes.scene_1m(width=160, height=90, spp=1, nu=32, nv=16, n_obj=6)     # vertex normals + flat quads
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    rng = np.random.default_rng(4)
    pos = np.stack([rng.random(20000) * 160, rng.random(20000) * 90], -1).astype(np.float32)
    rg, ro = sc.sensor_rays(pos), orc.sensor_rays(pos)
    for f in ("o", "d", "mint", "maxt"):
        assert np.array_equal(rg[f], ro[f]), f          # bit-exact: same operation order, no contraction
    hits = orc.trace(ro)
    assert (hits["prim"] != abi.NO_HIT).mean() > 0.5
    ig, io = sc.fill_intersection(ro, hits), orc.fill_intersection(ro, hits)
    assert util.frac_close(ig, io, 1e-4, 1e-5) > 0.9995
    sc.close()


@pytest.mark.parametrize("rfilter", ["gaussian", "box"])
def test_film_put_and_develop_vs_oracle(gpu_lib, oracle_mod, rfilter):
    b = scenes.cbox(width=50, height=37, spp=1)           # not a multiple of the tile size
    b.set_sensor(look_at([278, 273, -800], [278, 273, -799], [0, 1, 0]), 39.3, 50, 37, rfilter=rfilter, near_clip=10, far_clip=2800)
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    assert sc.film_size() == orc.film_size()
    rng = np.random.default_rng(6)
    n = 5000
    s = np.zeros((n, 6), np.float32)
    s[:, 0] = rng.random(n) * 54 - 2           # also positions outside the crop window (clipped footprints)
    s[:, 1] = rng.random(n) * 41 - 2
    s[:, 2:5] = rng.random((n, 3)) * 3
    s[:, 5] = rng.integers(0, 2, n)
    s[7, 2] = np.nan                           # rejected samples (imageblock.h:147-153)
    s[9, 3] = np.inf
    sc.film_clear(); orc.film_clear()
    sc.film_put(s); orc.film_put(s)
    fg, fo = sc.film_download(), orc.film_download()
    assert fg.shape == fo.shape
    assert np.allclose(fg, fo, rtol=1e-4, atol=1e-5)            # fp32 add order differs (atomics)
    assert np.allclose(sc.film_develop(), orc.film_develop(), rtol=1e-4, atol=1e-5)
    sc.film_clear()
    assert not sc.film_download().any()
    sc.close()


def test_envmap_sampling_vs_oracle(gpu_lib, oracle_mod):
    """Scene::sampleEmitterDirect with an `envmap` emitter (envmap.cpp:516-543, 567-600) next to an area
    light; the CDF search is exact (same float tables), directions and values agree to fp32 tolerance."""
    b, _ = material_scene()
    rot = np.array([[0.36, 0.48, -0.8, 0], [-0.8, 0.6, 0, 0], [0.48, 0.64, 0.6, 0], [0, 0, 0, 1]], np.float64)
    b.emitters = [e for e in b.emitters if e.type != abi.EMITTER_CONSTANT]
    b.add_envmap(scenes.synthetic_sky(64, 32), to_world=rot, scale=0.75)
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    rng = np.random.default_rng(21)
    n = 30000
    q = np.zeros(n, dtype=abi.DIRECT_QUERY_DTYPE)
    q["ref"] = (rng.random((n, 3)) * [2, 1.5, 2] - [1, 0, 1]).astype(np.float32)
    nn = rng.normal(size=(n, 3))
    q["ref_n"] = (nn / np.linalg.norm(nn, axis=1, keepdims=True)).astype(np.float32)
    q["ref_n"][::4] = 0
    q["u"] = rng.random((n, 2)).astype(np.float32)
    g, o = sc.emitter_sample(q), orc.emitter_sample(q)
    env = max(int(e) for e in np.unique(o["emitter"]))
    assert (g["emitter"] == o["emitter"]).mean() > 0.9995
    ok = (g["emitter"] == o["emitter"]) & (o["emitter"] == env)
    assert ok.sum() > 5000
    for f in ("value", "pdf", "d", "dist", "p", "n"):
        assert util.frac_close(g[f][ok], o[f][ok], RTOL, 1e-5) > 0.999, f
    sc.close()
