"""CPU: the oracle's smooth `dielectric` / `conductor` restatement against closed forms.

The reference keeps no fixtures for these plugins (its chi-square test needs the full plugin stack), so the
restatement is pinned on what the published model fixes: Snell's law, the Fresnel reflectance at normal
incidence, total internal reflection, the radiance scale eta^2 across the interface (dielectric.cpp:289-295)
and the conductor reflectance ((n-1)^2+k^2)/((n+1)^2+k^2)."""
import numpy as np
import pytest

from mitsuba_b200 import abi
from mitsuba_b200.scenedesc import SceneBuilder, conductor, dielectric, diffuse, fresnel_diffuse_reflectance, look_at, plastic


def _scene():
    b = SceneBuilder()
    mats = {
        "glass": b.add_material(dielectric(1.5, 1.0, specular_reflectance=[0.9, 0.8, 0.7], specular_transmittance=[0.6, 0.5, 0.4])),
        "mirror": b.add_material(conductor([0.2, 0.9, 1.1], [3.9, 2.4, 2.1], specular_reflectance=[1.0, 0.5, 0.25], ext_eta=1.0)),
        "mirror2s": b.add_material(conductor(0.0, 1.0, ext_eta=1.0, twosided=True)),
    }
    mats["mirror_over_diffuse"] = b.add_twosided(conductor(0.0, 1.0, ext_eta=1.0), diffuse([0.2, 0.4, 0.6]))
    b.add_quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), mats["glass"])
    b.add_quad((-0.3, 2, -0.3), (0.3, 2, -0.3), (0.3, 2, 0.3), (-0.3, 2, 0.3), mats["glass"], area_radiance=[5, 4, 3])
    b.set_sensor(look_at([0, 1, -4], [0, 0.5, 0], [0, 1, 0]), 40.0, 16, 16)
    return b, mats


def _q(wi, u0):
    wi = np.asarray(wi, np.float32)
    q = np.zeros(wi.shape[0], dtype=abi.BSDF_QUERY_DTYPE)
    q["wi"] = wi
    q["wo"] = [0, 0, 1]
    q["u"][:, 0] = u0
    q["u"][:, 1] = 0.5
    return q


def _dirs(n, seed, front=None):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    if front is True:
        v[:, 2] = np.abs(v[:, 2])
    if front is False:
        v[:, 2] = -np.abs(v[:, 2])
    return v.astype(np.float32)


def test_dielectric_snell_fresnel_and_radiance_scale(oracle_mod):
    b, mats = _scene()
    orc = oracle_mod.Oracle(b)
    eta = np.float32(1.5)
    # eval / pdf are zero under the solid-angle measure of `path` (dielectric.cpp:224-226, 256-259)
    wi = _dirs(4000, 3)
    r = orc.bsdf_query(mats["glass"], _q(wi, 0.999999))
    assert not r["eval"].any() and not r["pdf"].any()
    # u.x close to 1: transmission unless totally reflected
    cos_i = wi[:, 2].astype(np.float64)
    sin2_t = (1 - cos_i ** 2) * np.where(cos_i > 0, 1 / 1.5 ** 2, 1.5 ** 2)
    tir = sin2_t >= 1
    assert tir.any() and (~tir).any()
    trans = ~tir & (np.abs(sin2_t - 1) > 1e-4)
    wo = r["swo"].astype(np.float64)
    # Snell: sin(theta_t) = sin(theta_i) / eta (outside -> inside), opposite hemisphere, same plane of incidence
    assert np.all(np.sign(wo[trans, 2]) == -np.sign(cos_i[trans]))
    assert np.allclose(1 - wo[trans, 2] ** 2, sin2_t[trans], atol=2e-6)
    assert np.allclose(np.linalg.norm(wo[trans], axis=1), 1.0, atol=2e-6)
    cross = wo[trans, 0] * wi[trans, 1] - wo[trans, 1] * wi[trans, 0]
    assert np.allclose(cross, 0, atol=1e-6)
    # bRec.eta and the radiance scale (1/eta^2 going in, eta^2 going out)
    inward = trans & (cos_i > 0)
    outward = trans & (cos_i < 0)
    assert np.allclose(r["seta"][inward], eta) and np.allclose(r["seta"][outward], 1 / eta)
    assert np.allclose(r["weight"][inward], np.array([0.6, 0.5, 0.4]) / 1.5 ** 2, rtol=1e-6)
    assert np.allclose(r["weight"][outward], np.array([0.6, 0.5, 0.4]) * 1.5 ** 2, rtol=1e-6)
    # total internal reflection: the reflection lobe with probability one
    t = tir & (np.abs(sin2_t - 1) > 1e-4)
    assert np.allclose(r["spdf"][t], 1.0) and np.allclose(r["swo"][t], wi[t] * np.array([-1, -1, 1], np.float32))
    assert np.allclose(r["weight"][t], [0.9, 0.8, 0.7]) and np.allclose(r["seta"][t], 1.0)


def test_dielectric_lobe_probabilities(oracle_mod):
    b, mats = _scene()
    orc = oracle_mod.Oracle(b)
    # normal incidence: F = ((eta-1)/(eta+1))^2 = 0.04 from either side
    for z in (1.0, -1.0):
        refl = orc.bsdf_query(mats["glass"], _q([[0, 0, z]], 0.0399))
        tran = orc.bsdf_query(mats["glass"], _q([[0, 0, z]], 0.0401))
        assert refl["spdf"][0] == pytest.approx(0.04, rel=1e-5) and refl["swo"][0].tolist() == [0, 0, z]
        assert tran["spdf"][0] == pytest.approx(0.96, rel=1e-5) and tran["swo"][0, 2] == -z
        assert refl["weight"][0].tolist() == pytest.approx([0.9, 0.8, 0.7])
    # the two lobe probabilities sum to one for every direction, and F rises towards grazing incidence
    wi = _dirs(2000, 5, front=True)
    a = orc.bsdf_query(mats["glass"], _q(wi, 0.0))              # u.x = 0 <= F: reflection
    c = orc.bsdf_query(mats["glass"], _q(wi, 0.9999999))
    assert np.allclose(a["swo"], wi * np.array([-1, -1, 1], np.float32))
    assert np.allclose(a["spdf"] + c["spdf"], 1.0, atol=1e-6)
    order = np.argsort(wi[:, 2])
    assert np.all(np.diff(a["spdf"][order]) <= 1e-6)            # monotone in cos(theta_i) for eta = 1.5


def test_conductor_mirror_direction_and_fresnel(oracle_mod):
    b, mats = _scene()
    orc = oracle_mod.Oracle(b)
    wi = _dirs(2000, 7)
    r = orc.bsdf_query(mats["mirror"], _q(wi, 0.3))
    front = wi[:, 2] > 0
    assert not r["eval"].any() and not r["pdf"].any()
    assert not r["spdf"][~front].any() and not r["weight"][~front].any()            # EFrontSide only (conductor.cpp:271)
    assert np.all(r["spdf"][front] == 1.0) and np.all(r["seta"][front] == 1.0)
    assert np.array_equal(r["swo"][front], wi[front] * np.array([-1, -1, 1], np.float32))
    # normal incidence: R = ((n-1)^2 + k^2) / ((n+1)^2 + k^2), times specularReflectance
    n, k = np.array([0.2, 0.9, 1.1]), np.array([3.9, 2.4, 2.1])
    r0 = orc.bsdf_query(mats["mirror"], _q([[0, 0, 1]], 0.3))["weight"][0]
    assert r0 == pytest.approx((((n - 1) ** 2 + k ** 2) / ((n + 1) ** 2 + k ** 2)) * [1.0, 0.5, 0.25], rel=2e-5)
    # material "none" (eta = 0, k = 1) is a perfect mirror; twosided mirrors the frame for wi.z < 0
    r2 = orc.bsdf_query(mats["mirror2s"], _q(wi, 0.3))
    assert np.allclose(r2["weight"], 1.0, atol=1e-6) and np.all(r2["spdf"] == 1.0)
    assert np.array_equal(r2["swo"], wi * np.array([-1, -1, 1], np.float32))
    # mixed wrapper: mirror in front, diffuse behind
    r3 = orc.bsdf_query(mats["mirror_over_diffuse"], _q(wi, 0.3))
    assert np.all(r3["spdf"][front] == 1.0) and np.all(r3["swo"][~front, 2] < 0)
    assert np.allclose(r3["weight"][~front], [0.2, 0.4, 0.6])


def test_delta_materials_in_the_li_loop(oracle_mod):
    """A mirror facing an area light: the emitter is reached only through the delta lobe, with MIS weight 1
    (path.cpp:263-264), and no shadow rays leave the mirror (path.cpp:176-177)."""
    b = SceneBuilder()
    mirror = b.add_material(conductor(0.0, 1.0, ext_eta=1.0))
    black = b.add_material(diffuse(0.0))
    b.add_quad((-50, 0, -50), (-50, 0, 50), (50, 0, 50), (50, 0, -50), mirror)                            # facing +y
    b.add_quad((-50, 10, -50), (50, 10, -50), (50, 10, 50), (-50, 10, 50), black, area_radiance=[2, 3, 4])   # facing -y
    b.set_integrator(max_depth=3, rr_depth=10, sample_count=4)
    b.set_sensor(look_at([0, 5, 0], [0, 0, 0], [0, 0, 1]), 20.0, 8, 8)
    orc = oracle_mod.Oracle(b)
    st = orc.render(spp=4, seed=1)
    img = orc.film_develop()
    assert np.allclose(img, [2, 3, 4], rtol=1e-5)
    assert st.shadow_rays == 0 and st.closest_rays == 2 * st.samples


def test_fresnel_diffuse_reflectance_known_values():
    # util.cpp:814-853 quotes fits with <= 0.1 % error for 1 <= eta <= 2: d'Eon & Irving (eta > 1), Egan & Hilgeman (eta < 1)
    for eta in (1.1, 1.33, 1.49, 1.5, 2.0):
        inv = 1.0 / eta
        fit = 0.919317 - 3.4793 * inv + 6.75335 * inv ** 2 - 7.80989 * inv ** 3 + 4.98554 * inv ** 4 - 1.36881 * inv ** 5
        assert float(fresnel_diffuse_reflectance(eta)) == pytest.approx(fit, rel=2e-3)
        fit_int = -1.4399 * inv * inv + 0.7099 * inv + 0.6681 + 0.0636 / inv
        assert float(fresnel_diffuse_reflectance(inv)) == pytest.approx(fit_int, rel=7e-3)
        # the two are tied by 1 - Fdr(1/eta) = (1 - Fdr(eta)) / eta^2 (reciprocity of the interface)
        assert 1 - float(fresnel_diffuse_reflectance(inv)) == pytest.approx((1 - float(fresnel_diffuse_reflectance(eta))) / eta ** 2, rel=1e-4)
    assert float(fresnel_diffuse_reflectance(1.0)) == 0.0


def test_plastic_lobes_energy_and_sampling(oracle_mod):
    b = SceneBuilder()
    mats = {
        "lin": b.add_material(plastic(1.5, 1.0, [0.6, 0.3, 0.1], [1.0, 0.9, 0.8])),
        "nonlin": b.add_material(plastic(1.5, 1.0, [0.6, 0.3, 0.1], 1.0, nonlinear=True)),
    }
    b.add_quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), mats["lin"])
    b.add_quad((-0.3, 2, -0.3), (0.3, 2, -0.3), (0.3, 2, 0.3), (-0.3, 2, 0.3), mats["lin"], area_radiance=[5, 4, 3])
    b.set_sensor(look_at([0, 1, -4], [0, 0.5, 0], [0, 1, 0]), 40.0, 16, 16)
    orc = oracle_mod.Oracle(b)
    m = b.materials[mats["lin"]]
    assert m.fdr_int == pytest.approx(0.59635, rel=1e-4)
    lum = lambda c: 0.212671 * c[0] + 0.715160 * c[1] + 0.072169 * c[2]
    assert m.spec_weight == pytest.approx(lum([1.0, 0.9, 0.8]) / (lum([0.6, 0.3, 0.1]) + lum([1.0, 0.9, 0.8])), rel=1e-6)

    n = 40000
    wi = np.tile(np.array([[0.6, 0.0, 0.8]], np.float32), (n, 1))
    rng = np.random.default_rng(4)
    q = _q(wi, 0.0)
    q["u"][:, :2] = rng.random((n, 2)).astype(np.float32)
    wo = _dirs(n, 8, front=True)
    q["wo"] = wo
    r = orc.bsdf_query(mats["lin"], q)
    F = 0.5 * (((0.8 - 1.5 * np.sqrt(1 - 0.36 / 2.25)) / (0.8 + 1.5 * np.sqrt(1 - 0.36 / 2.25))) ** 2 +
               ((1.5 * 0.8 - np.sqrt(1 - 0.36 / 2.25)) / (1.5 * 0.8 + np.sqrt(1 - 0.36 / 2.25))) ** 2)
    p_spec = F * m.spec_weight / (F * m.spec_weight + (1 - F) * (1 - m.spec_weight))
    spec = q["u"][:, 0] < p_spec
    assert spec.mean() == pytest.approx(p_spec, abs=0.01)
    # delta lobe: mirror direction, pdf = probSpecular, weight = specularReflectance * F / probSpecular
    assert np.array_equal(r["swo"][spec], wi[spec] * np.array([-1, -1, 1], np.float32))
    assert np.allclose(r["spdf"][spec], p_spec, rtol=1e-5)
    assert np.allclose(r["weight"][spec], np.array([1.0, 0.9, 0.8]) * F / p_spec, rtol=1e-5)
    # diffuse lobe: cosine-distributed, pdf = (1 - probSpecular) cos / pi, and sample weight * pdf == eval
    d = ~spec
    assert np.all(r["swo"][d, 2] > 0)
    assert np.allclose(r["spdf"][d], (1 - p_spec) * r["swo"][d, 2] / np.pi, rtol=1e-5)
    q2 = q.copy()
    q2["wo"] = r["swo"]
    r2 = orc.bsdf_query(mats["lin"], q2)
    assert np.allclose(r2["eval"][d], r["weight"][d] * r["spdf"][d, None], rtol=2e-5, atol=1e-8)
    assert np.allclose(r2["pdf"][d], r["spdf"][d], rtol=1e-6)
    # energy: delta albedo F * spec + Monte Carlo estimate of the diffuse albedo; never above one, and the diffuse part
    # equals rho (1-Fi)/eta^2 * E[(1-Fo)] / (1 - fdrInt) with E over the cosine density = (1-FdrExt) ... checked numerically
    albedo = (r["weight"] * 1.0).mean(0)          # E[weight] over the sampling density
    assert np.all(albedo < 1.0) and np.all(albedo > np.array([1.0, 0.9, 0.8]) * F)
    fdr_ext = float(fresnel_diffuse_reflectance(1.5))
    expect_diffuse = np.array([0.6, 0.3, 0.1]) / (1 - m.fdr_int) * (1 - F) * (1 - fdr_ext) / 2.25
    assert albedo - np.array([1.0, 0.9, 0.8]) * F == pytest.approx(expect_diffuse, rel=0.02)
    # eval for arbitrary wo is the closed form of plastic.cpp:270-279
    Fo = np.array([_fresnel(c, 1.5) for c in wo[:200, 2]])
    want = (np.array([0.6, 0.3, 0.1]) / (1 - m.fdr_int))[None, :] * (wo[:200, 2] / np.pi / 2.25 * (1 - F) * (1 - Fo))[:, None]
    assert np.allclose(r["eval"][:200], want, rtol=2e-5)
    # nonlinear: diff / (1 - diff * fdrInt) per channel
    rn = orc.bsdf_query(mats["nonlin"], q)
    ratio = rn["eval"][:200] / r["eval"][:200]
    assert np.allclose(ratio, ((1 - m.fdr_int) / (1 - np.array([0.6, 0.3, 0.1]) * m.fdr_int))[None, :], rtol=1e-5)
    # back side: nothing
    qb = _q(_dirs(100, 9, front=False), 0.5)
    rb = orc.bsdf_query(mats["lin"], qb)
    assert not rb["spdf"].any() and not rb["eval"].any() and not rb["pdf"].any()


def _fresnel(cos_i, eta):
    s2 = (1 - cos_i * cos_i) / eta ** 2
    ct = np.sqrt(1 - s2)
    rs = (cos_i - eta * ct) / (cos_i + eta * ct)
    rp = (eta * cos_i - ct) / (eta * cos_i + ct)
    return 0.5 * (rs * rs + rp * rp)
