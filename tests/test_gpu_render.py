"""GPU: the wavefront `path` integrator (mtsb_render) against the oracle's Li loop.

Two kinds of check:
 * same-stream parity -- the oracle is rendered with the counter-based random numbers the GPU uses
   (oracle.render(counter_rng=True)); films must agree pixel by pixel (fp32 rounding only) and the
   ray / shade / path-length counters must be identical up to rounding-induced decision flips;
 * independent-stream parity -- against the oracle with its sequential generator (the reference's
   model): mean luminance inside the 3-sigma Monte Carlo interval, relMSE against a high-spp oracle
   image no worse than 1.5x the CPU oracle's own at equal spp.  SURVEY.md 8(d) states 1.25x; these images are
   ~5-9 k pixels at 8-32 spp, where the squared-error estimate itself carries ~10 % sampling error, so the
   1.25x bar is applied where it is meaningful -- BASELINE configs[0] at full size against a 4096-spp image,
   tests/test_gpu_render_full.py -- and the same-stream test remains the sharp one here."""
import numpy as np
import pytest

from mitsuba_b200 import abi, scenes
from tests import util

pytestmark = pytest.mark.gpu

LUM = np.array([0.212671, 0.715160, 0.072169])


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def check_render(gpu_lib, oracle_mod, b, spp, ref_spp, label):
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    ref_stats = orc.render(spp=ref_spp, seed=11)
    ref = orc.film_develop().astype(np.float64)
    cpu_stats = orc.render(spp=spp, seed=5)
    cpu = orc.film_develop().astype(np.float64)
    st = sc.render(spp=spp, seed=7)
    gpu = sc.film_develop().astype(np.float64)
    assert st.samples == cpu_stats.samples and st.bad_samples == 0
    assert np.isfinite(gpu).all()
    # mean luminance inside the 3-sigma Monte Carlo interval of the per-sample estimator
    n = ref_stats.samples
    mean_ref = ref_stats.lum_sum / n
    var = max(ref_stats.lum_sq_sum / n - mean_ref ** 2, 0.0)
    sigma = np.sqrt(var / st.samples + var / n)
    lum_gpu = float((gpu @ LUM).mean())
    lum_ref = float((ref @ LUM).mean())
    assert abs(lum_gpu - lum_ref) <= 3 * sigma + 1e-3 * lum_ref, (label, lum_gpu, lum_ref, sigma)
    # relMSE at equal spp no worse than the CPU oracle's own.  The squared-error estimate of a ~5 k-pixel image is
    # dominated by a few firefly pixels (glass under an environment), so both sides take the MEDIAN over five seeds.
    e_gpu_all, e_cpu_all = [rel_mse(gpu, ref)], [rel_mse(cpu, ref)]
    for k in range(1, 5):
        sc.render(spp=spp, seed=7 + 100 * k)
        e_gpu_all.append(rel_mse(sc.film_develop().astype(np.float64), ref))
        orc.render(spp=spp, seed=5 + 100 * k)
        e_cpu_all.append(rel_mse(orc.film_develop().astype(np.float64), ref))
    e_gpu, e_cpu = float(np.median(e_gpu_all)), float(np.median(e_cpu_all))
    assert e_gpu <= 1.5 * e_cpu + 1e-4, (label, e_gpu_all, e_cpu_all)
    # same amount of work per sample: ray counts and path length (reference statistics counters)
    for f, tol in (("closest_rays", 0.02), ("shade_events", 0.02), ("path_length_sum", 0.02)):
        a, c = getattr(st, f) / st.samples, getattr(cpu_stats, f) / cpu_stats.samples
        assert abs(a - c) <= tol * c, (label, f, a, c)
    sc.close()
    return e_gpu, e_cpu


def check_same_stream(gpu_lib, oracle_mod, b, spp, label, seed=9, rtol=1e-3):
    """GPU film == oracle film when both draw the same random numbers."""
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    st = sc.render(spp=spp, seed=seed)
    fg = sc.film_download().astype(np.float64)
    ost = orc.render(spp=spp, seed=seed, counter_rng=True)
    fo = orc.film_download().astype(np.float64)
    bad = (np.abs(fg - fo) > rtol * np.maximum(np.abs(fo), 1e-3))[..., :3].any(-1)
    # a comparison that rounds differently on the two sides now and then sends ONE sample down another path: its 5 x 5 filter
    # footprint (25 pixels) is allowed on top of the 0.2 % fraction, which on these small images is less than one footprint
    assert bad.sum() <= max(2e-3 * bad.size, 25), (label, int(bad.sum()), bad.size)
    assert rel_mse(sc.film_develop().astype(np.float64), orc.film_develop().astype(np.float64)) < 1e-5
    assert st.samples == ost.samples
    for f in ("closest_rays", "shade_events", "path_length_sum"):
        a, c = getattr(st, f), getattr(ost, f)
        assert abs(a - c) <= 1e-3 * c + 2, (label, f, a, c)
    # the GPU skips shadow rays whose contribution is already zero (BSDF value 0); never traces more
    assert st.shadow_rays <= ost.shadow_rays, (label, st.shadow_rays, ost.shadow_rays)
    sc.close()


@pytest.mark.parametrize("max_depth", [1, 2, 3, 8, -1])
def test_same_stream_cbox_depths(gpu_lib, oracle_mod, max_depth):
    b = scenes.cbox(width=64, height=64, spp=2, max_depth=max_depth)
    check_same_stream(gpu_lib, oracle_mod, b, 2, "cbox maxDepth=%d" % max_depth)


def test_same_stream_cbox_box_filter_and_twosided(gpu_lib, oracle_mod):
    from mitsuba_b200.scenedesc import look_at
    b = scenes.cbox(width=50, height=37, spp=3, twosided=True)
    b.set_sensor(look_at([278, 273, -800], [278, 273, -799], [0, 1, 0]), 39.3077, 50, 37, rfilter="box", near_clip=10.0, far_clip=2800.0)
    check_same_stream(gpu_lib, oracle_mod, b, 3, "cbox box twosided")


def test_same_stream_mixed_materials(gpu_lib, oracle_mod):
    b = scenes.scene_10m(width=96, height=54, spp=4, nu=24, nv=12, n_obj=12)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "mixed")


def test_same_stream_ggx_conductors_1m_family(gpu_lib, oracle_mod):
    b = scenes.scene_1m(width=96, height=54, spp=4, nu=24, nv=12, n_obj=8)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "1m-small")


def _glass_mirror_cbox(**kw):
    from mitsuba_b200.scenedesc import conductor, dielectric, diffuse
    eta, k = scenes.METALS["Al"]
    return scenes.cbox(short_material=dielectric(1.5046, 1.000277), tall_material=conductor(eta, k), **kw)


@pytest.mark.parametrize("max_depth", [2, 5, -1])
def test_same_stream_glass_and_mirror_blocks(gpu_lib, oracle_mod, max_depth):
    """smooth dielectric + conductor (Dirac lobes: no emitter sampling at the hit, MIS weight 1
    on the next emitter hit -- path.cpp:176-177, 263-264)."""
    b = _glass_mirror_cbox(width=64, height=64, spp=4, max_depth=max_depth)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "cbox glass+mirror maxDepth=%d" % max_depth)


def test_same_stream_delta_lobes_in_twosided_wrappers_and_env(gpu_lib, oracle_mod):
    from mitsuba_b200.scenedesc import conductor, diffuse, roughconductor
    eta, k = scenes.METALS["Cu"]
    b = scenes.cbox(width=60, height=44, spp=4, max_depth=6,
                    short_material=(conductor(eta, k), diffuse([0.2, 0.4, 0.6])),         # mirror outside, diffuse inside
                    tall_material=(roughconductor(eta, k, alpha=0.2, distribution="ggx"), conductor(0.0, 1.0)))
    b.add_constant_env([0.3, 0.4, 0.5])
    check_same_stream(gpu_lib, oracle_mod, b, 4, "cbox twosided delta + env")


def test_same_stream_plastic_blocks(gpu_lib, oracle_mod):
    """plastic: a Dirac lobe over a diffuse base -- emitter sampling sees the diffuse part only, a sampled specular
    bounce reaches emitters with MIS weight 1."""
    from mitsuba_b200.scenedesc import conductor, plastic
    eta, k = scenes.METALS["Au"]
    b = scenes.cbox(width=64, height=64, spp=4, max_depth=6,
                    short_material=plastic(1.49, 1.000277, [0.1, 0.27, 0.36], 1.0),
                    tall_material=(plastic(1.9, 1.0, [0.6, 0.3, 0.1], 0.7, nonlinear=True), conductor(eta, k)))
    check_same_stream(gpu_lib, oracle_mod, b, 4, "cbox plastic")


def test_glass_and_mirror_blocks_independent_streams(gpu_lib, oracle_mod):
    b = _glass_mirror_cbox(width=48, height=48, spp=32, max_depth=8)
    check_render(gpu_lib, oracle_mod, b, 32, 256, "cbox glass+mirror")


def test_same_stream_point_light_hide_emitters_strict_normals(gpu_lib, oracle_mod):
    b = scenes.cbox(width=48, height=48, spp=4)
    b.add_point_light([278, 400, 279], [4e5, 4e5, 4e5])
    b.set_integrator(max_depth=5, rr_depth=3, hide_emitters=True, strict_normals=True, sample_count=4)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "point+hide+strict")


def test_cbox_diffuse_area_light(gpu_lib, oracle_mod):
    """BASELINE configs[0] at reduced resolution: cbox, maxDepth 8, diffuse, one area light."""
    check_render(gpu_lib, oracle_mod, scenes.cbox(width=96, height=96, spp=32), 32, 256, "cbox")


def test_cbox_twosided_unbounded_depth(gpu_lib, oracle_mod):
    check_render(gpu_lib, oracle_mod, scenes.cbox(width=64, height=64, spp=16, max_depth=-1, twosided=True), 16, 128, "cbox-2s")


def test_mixed_materials_env_and_area(gpu_lib, oracle_mod):
    """configs[2] family at small size: diffuse / roughdielectric / roughconductor (Beckmann + GGX),
    constant environment + area lights, maxDepth 10, rrDepth 5."""
    b = scenes.scene_10m(width=96, height=54, spp=32, nu=24, nv=12, n_obj=12)
    check_render(gpu_lib, oracle_mod, b, 32, 256, "mixed")


def test_tile_and_sample_partition_is_exact(gpu_lib):
    """Tile-interleaved and sample-split renders add up to the single-call render (8e):
    the RNG is keyed by (pixel, global sample index), so only the fp32 add order differs."""
    b = scenes.cbox(width=80, height=48, spp=8)
    sc = gpu_lib.Scene(b)
    sc.render(spp=8, seed=3)
    whole = sc.film_download().astype(np.float64)
    sc.film_clear()
    for r in range(3):
        sc.render(spp=8, seed=3, tile_index=r, tile_count=3, keep_film=True)
    tiles = sc.film_download().astype(np.float64)
    assert np.allclose(tiles, whole, rtol=1e-4, atol=1e-4)
    sc.film_clear()
    sc.render(spp=5, seed=3, spp_offset=0, keep_film=True)
    sc.render(spp=3, seed=3, spp_offset=5, keep_film=True, spp_per_pass=2)
    split = sc.film_download().astype(np.float64)
    assert np.allclose(split, whole, rtol=1e-4, atol=1e-4)
    sc.close()


def test_point_light_and_hide_emitters(gpu_lib, oracle_mod):
    b = scenes.cbox(width=48, height=48, spp=16)
    b.add_point_light([278, 400, 279], [4e5, 4e5, 4e5])
    b.set_integrator(max_depth=5, rr_depth=3, hide_emitters=True, sample_count=16)
    check_render(gpu_lib, oracle_mod, b, 16, 128, "point")


# ---- quasi-Monte Carlo samplers ---------------------------------------------------------------

def check_same_sequence(gpu_lib, oracle_mod, b, spp, label, **render_kw):
    """halton is deterministic: the GPU and the oracle (which drives the stateful sampler like renderBlock does) draw the
    same numbers as the reference itself would, so the films agree pixel by pixel."""
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    st = sc.render(spp=spp, seed=1, **render_kw)
    fg = sc.film_download().astype(np.float64)
    ost = orc.render(spp=spp, seed=99)
    fo = orc.film_download().astype(np.float64)
    bad = (np.abs(fg - fo) > 1e-3 * np.maximum(np.abs(fo), 1e-3))[..., :3].any(-1)
    # one sample whose path took a different branch after fp32 rounding touches its whole 5x5 filter footprint
    assert bad.sum() <= max(2e-3 * bad.size, 25), (label, int(bad.sum()), bad.size)
    assert st.samples == ost.samples
    for f in ("closest_rays", "shade_events", "path_length_sum"):
        a, c = getattr(st, f), getattr(ost, f)
        assert abs(a - c) <= 1e-3 * c + 2, (label, f, a, c)
    sc.close()


@pytest.mark.parametrize("scramble", [-1, 0, 5])
def test_halton_same_sequence_cbox(gpu_lib, oracle_mod, scramble):
    b = scenes.cbox(width=64, height=48, spp=6, max_depth=8)
    b.set_sampler("halton", sample_count=6, scramble=scramble)
    check_same_sequence(gpu_lib, oracle_mod, b, 6, "cbox halton scramble=%d" % scramble)


def test_halton_same_sequence_mixed_materials_deep_paths(gpu_lib, oracle_mod):
    """roughdielectric draws an extra dimension, Russian roulette another one from rrDepth on: the dimension counter
    of every path has to follow the reference's call order (path.cpp:178,211,285; roughdielectric.cpp:554)."""
    b = scenes.scene_10m(width=96, height=54, spp=4, nu=24, nv=12, n_obj=12)
    b.set_integrator(max_depth=12, rr_depth=3, sample_count=4)
    b.set_sampler("halton", sample_count=4)
    check_same_sequence(gpu_lib, oracle_mod, b, 4, "mixed halton")


def test_halton_same_sequence_delta_lobes_and_passes(gpu_lib, oracle_mod):
    """Dirac lobes skip the emitter sample (path.cpp:176-178): two dimensions fewer at those vertices.  Rendered in
    several passes: the sequence is addressed by the global sample index."""
    b = _glass_mirror_cbox(width=200, height=40, spp=8, max_depth=7)
    b.set_sampler("halton", sample_count=8)
    check_same_sequence(gpu_lib, oracle_mod, b, 8, "glass+mirror halton", spp_per_pass=3)


@pytest.mark.parametrize("scramble", [0, 9])
def test_sobol_same_sequence(gpu_lib, oracle_mod, scramble):
    """sobol is deterministic too (sobol.cpp:204-252): per-pixel enumeration of the Joe-Kuo sequence."""
    b = scenes.scene_10m(width=96, height=54, spp=5, nu=24, nv=12, n_obj=12)
    b.set_integrator(max_depth=9, rr_depth=3, sample_count=5)
    b.set_sampler("sobol", sample_count=5, scramble=scramble)
    check_same_sequence(gpu_lib, oracle_mod, b, 5, "mixed sobol scramble=%d" % scramble, spp_per_pass=2)


def test_ldsampler_statistics_and_variance(gpu_lib, oracle_mod):
    """ldsampler shuffles with the worker's random stream in the reference, so parity is statistical: mean luminance
    inside the 3-sigma interval, error no worse than the oracle's ldsampler render, and both clearly better than
    independent sampling at the same sample count (the point of the plugin, path.cpp:79-81)."""
    b = scenes.cbox(width=64, height=64, spp=16)
    sc_ind, orc_ind = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    ref_stats = orc_ind.render(spp=512, seed=11)
    ref = orc_ind.film_develop().astype(np.float64)
    sc_ind.render(spp=16, seed=3)
    e_ind = rel_mse(sc_ind.film_develop().astype(np.float64), ref)
    sc_ind.close()
    b.set_sampler("ldsampler", sample_count=16)
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    st = sc.render(spp=16, seed=7)
    gpu = sc.film_develop().astype(np.float64)
    ost = orc.render(spp=16, seed=5)
    cpu = orc.film_develop().astype(np.float64)
    n = ref_stats.samples
    mean_ref = ref_stats.lum_sum / n
    var = max(ref_stats.lum_sq_sum / n - mean_ref ** 2, 0.0)
    sigma = np.sqrt(var / st.samples + var / n)
    assert abs(float((gpu @ LUM).mean()) - float((ref @ LUM).mean())) <= 3 * sigma + 1e-3 * float((ref @ LUM).mean())
    e_gpu, e_cpu = rel_mse(gpu, ref), rel_mse(cpu, ref)
    assert e_gpu <= 1.5 * e_cpu + 1e-4, (e_gpu, e_cpu)
    assert e_gpu < 0.9 * e_ind, (e_gpu, e_ind)
    for f, tol in (("closest_rays", 0.02), ("shade_events", 0.02), ("path_length_sum", 0.02)):
        a, c = getattr(st, f) / st.samples, getattr(ost, f) / ost.samples
        assert abs(a - c) <= tol * c, (f, a, c)
    # the point sets hold sampleCount entries (ldsampler.cpp:200): asking for more is refused, as the reference asserts
    with pytest.raises(Exception):
        sc.render(spp=17, seed=7)
    sc.close()


def test_halton_partition_is_exact(gpu_lib):
    """Tile- and sample-split renders reproduce the single call: the sequence is a function of (pixel, sample index)."""
    b = scenes.cbox(width=80, height=48, spp=8)
    b.set_sampler("halton", sample_count=8)
    sc = gpu_lib.Scene(b)
    sc.render(spp=8, seed=3)
    whole = sc.film_download().astype(np.float64)
    sc.film_clear()
    for r in range(2):
        sc.render(spp=5, seed=3, tile_index=r, tile_count=2, keep_film=True)
        sc.render(spp=3, seed=3, spp_offset=5, tile_index=r, tile_count=2, keep_film=True)
    assert np.allclose(sc.film_download().astype(np.float64), whole, rtol=1e-4, atol=1e-4)
    sc.close()


# ---- envmap emitter ----------------------------------------------------------------------------

def _sky_scene(**kw):
    b = scenes.scene_10m(nu=24, nv=12, n_obj=12, **kw)
    b.emitters = [e for e in b.emitters if e.type != abi.EMITTER_CONSTANT]
    rot = np.array([[0.8, 0, 0.6, 0], [0, 1, 0, 0], [-0.6, 0, 0.8, 0], [0, 0, 0, 1]], np.float64)
    b.add_envmap(scenes.synthetic_sky(128, 64), to_world=rot, scale=0.6)
    return b


def test_same_stream_envmap(gpu_lib, oracle_mod):
    """Environment lookups on escaped rays (first segment and MIS'd BSDF rays), importance-sampled sky + sun as an
    emitter next to area lights: films equal the oracle's with the same random numbers."""
    b = _sky_scene(width=96, height=54, spp=4)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "mixed + envmap")


def test_same_stream_envmap_minified_sky_is_filtered(gpu_lib, oracle_mod):
    """A sky map much finer than the image (1024 x 512 texels behind 64 x 36 pixels): camera rays that escape carry
    differentials, so the reference filters the lookup through MIPMap::eval (envmap.cpp:393-407: EWA on the Lanczos
    pyramid).  The film equals the oracle's under the same random numbers, and differs from an unfiltered lookup."""
    def scene(res):
        b = scenes.scene_10m(nu=24, nv=12, n_obj=12, width=64, height=36, spp=4)
        b.emitters = [e for e in b.emitters if e.type != abi.EMITTER_CONSTANT]
        sky = scenes.synthetic_sky(res, res // 2, sun=False)
        yy, xx = np.mgrid[0:res // 2, 0:res]
        sky = sky * (0.25 + 1.5 * ((xx // 2 + yy // 2) % 2))[..., None].astype(np.float32)      # 2-texel checker: aliases unless filtered
        b.add_envmap(sky.astype(np.float32), scale=0.8)
        return b
    b = scene(1024)
    # per-pixel tolerance 1 %: the EWA loop indexes its 64-entry weight table with (int) q, and q (a quadratic form of the
    # texel offset) differs by an ulp between the device (FMA contraction) and the strict-IEEE oracle, so a texel now and
    # then takes the neighbouring table entry -- visible on this map because neighbouring texels differ by a factor of 7
    check_same_stream(gpu_lib, oracle_mod, b, 4, "minified envmap", rtol=1e-2)
    # the filter matters here: unfiltered lookups (4 per pixel into a checker whose texels differ by a factor of 7) would
    # make neighbouring sky pixels jump by tens of percent; through the pyramid they vary smoothly
    sc = gpu_lib.Scene(b)
    sc.render(spp=4, seed=9)
    row = sc.film_develop().astype(np.float64)[0].sum(axis=1)            # top image row: sky
    sc.close()
    row = row[row > 0.05 * row.max()]
    jump = np.abs(np.diff(row)) / (row[1:] + row[:-1])
    assert len(row) > 32 and np.median(jump) < 0.15, float(np.median(jump))      # measured 0.067 (the sky carries 10 % texel noise); unfiltered: > 0.3


def test_envmap_halton_same_sequence(gpu_lib, oracle_mod):
    b = _sky_scene(width=96, height=54, spp=4)
    b.set_sampler("halton", sample_count=4)
    check_same_sequence(gpu_lib, oracle_mod, b, 4, "mixed + envmap, halton")


def test_envmap_independent_streams(gpu_lib, oracle_mod):
    b = _sky_scene(width=64, height=36, spp=32)
    check_render(gpu_lib, oracle_mod, b, 32, 256, "mixed + envmap")


# ---- textures: checkerboard / gridtexture / bitmap (nearest, bilinear, trilinear, EWA) -------------
def test_same_stream_textured_cbox(gpu_lib, oracle_mod):
    """Every texture kind on every material slot that takes one; the back wall's bitmap is minified (EWA over several
    MIP levels on camera-ray hits, level-0 bilinear on every later hit), the left wall's is magnified."""
    b = scenes.cbox_textured(width=96, height=72, spp=4, max_depth=6)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "textured cbox")


def test_same_stream_textured_roughness(gpu_lib, oracle_mod):
    """Textured `alpha` (roughconductor.cpp:379-384, roughdielectric.cpp:300-305): the roughness of a rough conductor and of
    a rough dielectric block comes from a checkerboard / a bitmap evaluated at the hit; films equal the oracle's."""
    from mitsuba_b200.scenedesc import roughconductor, roughdielectric
    b = scenes.cbox_textured(width=72, height=54, spp=4)
    rough = scenes.pattern_image(32, 32, seed=9)[:, :, 0] * np.float32(0.5) + np.float32(0.02)
    mc = roughconductor(*scenes.METALS["Cu"], alpha=0.1, distribution="ggx")
    mc.tex_alpha = b.add_checkerboard(0.05, 0.45, uscale=3, vscale=3)
    md = roughdielectric(1.5, 1.0, alpha=0.1, distribution="beckmann")
    md.tex_alpha = b.add_bitmap(rough, "bilinear", "repeat", uscale=2, vscale=2)
    ic, idl = b.add_material(mc), b.add_material(md)
    uv4 = np.float32([[0, 0], [1, 0], [1, 1], [0, 1]])
    for (x0, z0, m) in ((80.0, 60.0, ic), (330.0, 90.0, idl)):
        pts = np.float32([(x0, 200, z0), (x0 + 150, 200, z0), (x0 + 150, 260, z0 + 150), (x0, 260, z0 + 150)])
        b.add_mesh(pts, np.array([[0, 1, 2], [0, 2, 3]], np.uint32), m, texcoords=uv4)
    check_same_stream(gpu_lib, oracle_mod, b, 4, "textured roughness")


def test_textured_cbox_halton_same_sequence_and_passes(gpu_lib, oracle_mod):
    b = scenes.cbox_textured(width=80, height=60, spp=6, max_depth=5, tex_res=128)
    b.set_sampler("halton", sample_count=6)
    check_same_sequence(gpu_lib, oracle_mod, b, 6, "textured cbox halton", spp_per_pass=4)


def test_textured_cbox_independent_streams(gpu_lib, oracle_mod):
    b = scenes.cbox_textured(width=64, height=48, spp=64, max_depth=6)
    check_render(gpu_lib, oracle_mod, b, 64, 256, "textured cbox")


def test_texture_filtering_depends_on_the_sample_count(gpu_lib, oracle_mod):
    """The camera ray's differentials are scaled by 1/sqrt(sampleCount) (integrator.cpp:200-201): the same texture seen
    with a different sampler sample count is filtered differently, on the GPU exactly as in the oracle."""
    for count in (1, 64):
        b = scenes.cbox_textured(width=64, height=48, spp=2, max_depth=2, tex_res=256)
        b.set_integrator(max_depth=2, rr_depth=5, sample_count=count)
        check_same_stream(gpu_lib, oracle_mod, b, 2, "textured cbox sampleCount=%d" % count)
