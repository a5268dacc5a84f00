"""CPU: textures on material spectra (src/textures/{bitmap,checkerboard,gridtexture,scale}.cpp,
include/mitsuba/render/mipmap.h, src/librender/intersection.cpp).

Three statements of the same algorithm are compared bit for bit:
  * the reference's own TMIPMap, compiled into oracle/_ref and recorded in tests/golden/texture.npz
    (tests/golden/make_golden_texture.py): Lanczos MIP pyramid, box / bilinear / trilinear / EWA lookups, all wrap modes;
  * oracle/orc_texture.h -- the restatement the path-tracing oracle shades with;
  * mitsuba_b200/csrc/texture.h + texture_host.h -- the code the CUDA kernels run, compiled with g++ by tests/emu."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle
from mitsuba_b200 import abi, scenes, scenedesc
from tests import util

EMU_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu")


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def emu():
    E = C.CDLL(os.path.join(EMU_DIR, "libemu.so"))
    E.emu_texture_create.restype = C.c_void_p
    E.emu_texture_create.argtypes = [C.POINTER(abi.Texture)]
    E.emu_texture_destroy.argtypes = [C.c_void_p]
    E.emu_texture_levels.argtypes = [C.c_void_p]
    E.emu_texture_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    E.emu_texture_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    E.emu_texture_eval.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    E.emu_compute_partials.argtypes = [C.c_size_t, C.c_void_p, C.c_void_p]
    E.emu_sensor_differentials.argtypes = [C.POINTER(abi.Sensor), C.c_float, C.c_float, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p]
    E.emu_float_to_half.restype = C.c_uint32
    E.emu_float_to_half.argtypes = [C.c_float]
    return E


class Impl:
    """The same five calls on the oracle and on the emulated device code."""
    def __init__(self, L, prefix):
        self.L, self.p = L, prefix

    def __getattr__(self, name):
        return getattr(self.L, self.p + name)


def bitmap_record(img, filt, wu, wv, aniso=20.0, uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0, scale=1.0):
    b = scenedesc.SceneBuilder()
    names = {0: "clamp", 1: "repeat", 2: "mirror", 3: "zero", 4: "one"}
    filters = {0: "nearest", 1: "bilinear", 2: "trilinear", 3: "ewa"}
    b.add_bitmap(img, filter_type=filters[filt], wrap_mode_u=names[wu], wrap_mode_v=names[wv], max_anisotropy=aniso,
                 uscale=uscale, vscale=vscale, uoffset=uoffset, voffset=voffset, scale=scale)
    return b.textures[0], b


def eval_tex(impl, h, uv, partials=None):
    uv = np.ascontiguousarray(uv, np.float32)
    o = np.zeros((uv.shape[0], 3), np.float32)
    pt = None if partials is None else np.ascontiguousarray(partials, np.float32)
    impl.texture_eval(h, uv.shape[0], vp(uv), None if pt is None else vp(pt), vp(o))
    return o


@pytest.mark.parametrize("which", ["oracle", "emu"])
def test_mipmap_matches_the_reference(which, emu):
    g = util.load_golden("texture.npz")
    impl = Impl(oracle.lib(), "orc_") if which == "oracle" else Impl(emu, "emu_")
    for ci, (w, h, wu, wv, filt, aniso) in enumerate(g["cases"]):
        w, h, wu, wv, filt = int(w), int(h), int(wu), int(wv), int(filt)
        img = g["img%d" % ci]
        rec, keep = bitmap_record(img, filt, wu, wv, float(aniso))
        hnd = impl.texture_create(C.byref(rec))
        sizes = g["sizes%d" % ci]
        assert impl.texture_levels(hnd) == sizes.shape[0], (which, ci)
        for lv in range(1, sizes.shape[0]):
            wh = (C.c_int * 2)()
            impl.texture_level_size(hnd, lv, wh)
            assert (wh[0], wh[1]) == tuple(sizes[lv])
            a = np.zeros((wh[1], wh[0], 3), np.float32)
            impl.texture_level(hnd, lv, vp(a))
            assert np.array_equal(a, g["lvl%d_%d" % (ci, lv)].astype(np.float32)), (which, ci, lv)
        q = g["q%d" % ci]
        got = eval_tex(impl, hnd, q[:, :2], q[:, 2:])
        assert np.array_equal(got, g["filtered%d" % ci]), (which, ci, "filtered", int((got != g["filtered%d" % ci]).any(1).sum()))
        impl.texture_destroy(hnd)
        for mode, name in ((0, "box"), (1, "bilinear")):           # BitmapTexture::eval(uv): bitmap.cpp:431-453
            rec2, keep2 = bitmap_record(img, mode, wu, wv)
            h2 = impl.texture_create(C.byref(rec2))
            assert np.array_equal(eval_tex(impl, h2, q[:, :2]), g["%s%d" % (name, ci)]), (which, ci, name)
            impl.texture_destroy(h2)


def test_uv_transform_scale_and_luminance_images(emu):
    """Texture2D's uv scale / offset (texture.cpp:112-121), the folded `scale` plugin (scale.cpp:69-71) and one-channel
    images (bitmap.cpp:262-266: replicated into a grey spectrum)."""
    rng = np.random.default_rng(5)
    img = rng.random((8, 12), dtype=np.float32)
    uv = (rng.random((500, 2)) * 2 - 0.5).astype(np.float32)
    rec, k = bitmap_record(img, 1, 1, 1, uscale=3.0, vscale=0.5, uoffset=0.25, voffset=-0.125, scale=[0.5, 1.0, 2.0])
    rec0, k0 = bitmap_record(np.repeat(img[:, :, None], 3, 2), 1, 1, 1)
    for impl in (Impl(oracle.lib(), "orc_"), Impl(emu, "emu_")):
        h, h0 = impl.texture_create(C.byref(rec)), impl.texture_create(C.byref(rec0))
        a = eval_tex(impl, h, uv)
        uv2 = np.stack([uv[:, 0] * np.float32(3.0) + np.float32(0.25), uv[:, 1] * np.float32(0.5) + np.float32(-0.125)], 1).astype(np.float32)
        b = eval_tex(impl, h0, uv2) * np.array([0.5, 1.0, 2.0], np.float32)
        assert np.array_equal(a, b)
        assert np.array_equal(a[:, 0] * 2, a[:, 1])
        impl.texture_destroy(h); impl.texture_destroy(h0)


def test_procedural_textures(emu):
    """checkerboard.cpp:61-69 and gridtexture.cpp:66-79 against a direct numpy statement."""
    rng = np.random.default_rng(6)
    uv = np.concatenate([(rng.random((4000, 2)) * 6 - 3), [[0, 0], [0.5, 0.5], [-0.5, 0.25], [0.999, -0.001], [1.0, 1.0]]]).astype(np.float32)
    b = scenedesc.SceneBuilder()
    b.add_checkerboard(color0=[0.8, 0.2, 0.1], color1=[0.1, 0.3, 0.9], uscale=2.0, vscale=3.0, uoffset=0.1)
    b.add_gridtexture(color0=[0.2, 0.2, 0.2], color1=[0.9, 0.8, 0.1], line_width=0.05, uscale=4.0, vscale=4.0, scale=0.5)
    t = uv * np.array([2.0, 3.0], np.float32) + np.array([0.1, 0.0], np.float32)
    x = 2 * np.mod(np.trunc(t[:, 0] * 2).astype(np.int64), 2) - 1
    y = 2 * np.mod(np.trunc(t[:, 1] * 2).astype(np.int64), 2) - 1
    want_c = np.where((x * y == 1)[:, None], np.float32([0.8, 0.2, 0.1]), np.float32([0.1, 0.3, 0.9]))
    t = uv * np.float32(4.0)
    fx, fy = t[:, 0] - np.floor(t[:, 0]), t[:, 1] - np.floor(t[:, 1])
    fx, fy = np.where(fx > .5, fx - 1, fx), np.where(fy > .5, fy - 1, fy)
    line = (np.abs(fx) < np.float32(0.05)) | (np.abs(fy) < np.float32(0.05))
    want_g = np.where(line[:, None], np.float32([0.9, 0.8, 0.1]), np.float32([0.2, 0.2, 0.2])) * np.float32(0.5)
    for impl in (Impl(oracle.lib(), "orc_"), Impl(emu, "emu_")):
        hc, hg = impl.texture_create(C.byref(b.textures[0])), impl.texture_create(C.byref(b.textures[1]))
        assert np.array_equal(eval_tex(impl, hc, uv), want_c)
        assert np.array_equal(eval_tex(impl, hg, uv), want_g)
        # procedural textures ignore the partials (checkerboard.cpp:71-75)
        assert np.array_equal(eval_tex(impl, hc, uv, np.ones((uv.shape[0], 4), np.float32)), want_c)
        impl.texture_destroy(hc); impl.texture_destroy(hg)


def test_half_conversion_of_the_scene_loader(emu):
    g = util.load_golden("envmap.npz")          # float -> half of the reference's half.cpp
    got = np.array([emu.emu_float_to_half(float(v)) for v in g["x"]], np.uint16)
    assert np.array_equal(got, g["half_bits"])
    rng = np.random.default_rng(7)
    x = np.concatenate([10 ** rng.uniform(-10, 5.5, 20000), -(10 ** rng.uniform(-10, 5.5, 2000)), [0.0, np.inf, -np.inf]]).astype(np.float32)
    with np.errstate(over="ignore"):
        want = x.astype(np.float16).view(np.uint16)
    got = np.array([emu.emu_float_to_half(float(v)) for v in x], np.uint16)
    assert np.array_equal(got, want)


def partial_queries(n, seed):
    rng = np.random.default_rng(seed)
    q = np.zeros((n, 21), np.float32)
    nrm = rng.standard_normal((n, 3)); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    o = rng.standard_normal((n, 3)) * 3
    p = o + rng.standard_normal((n, 3)) * 2
    d = p - o; d /= np.linalg.norm(d, axis=1, keepdims=True)
    t1 = np.cross(nrm, rng.standard_normal((n, 3))); t2 = np.cross(nrm, t1)
    q[:, 0:3], q[:, 3:6], q[:, 6:9], q[:, 9:12], q[:, 12:15] = p, nrm, t1 * rng.uniform(0.1, 3, (n, 1)), t2 * rng.uniform(0.1, 3, (n, 1)), o
    q[:, 15:18] = d + rng.standard_normal((n, 3)) * 1e-3
    q[:, 18:21] = d + rng.standard_normal((n, 3)) * 1e-3
    q[:5, 6:12] = 0            # dpdu = dpdv = 0 -> zero partials
    q[5:8, 15:18] = np.cross(q[5:8, 3:6], [1, 0, 0])      # offset ray parallel to the surface
    return q


def test_compute_partials(emu):
    """Intersection::computePartials (intersection.cpp:5-98): device code == oracle; the solution reproduces the offsets."""
    q = partial_queries(3000, 8)
    a, b = np.zeros((q.shape[0], 4), np.float32), np.zeros((q.shape[0], 4), np.float32)
    oracle.lib().orc_compute_partials(q.shape[0], vp(q), vp(a))
    emu.emu_compute_partials(q.shape[0], vp(q), vp(b))
    assert np.array_equal(a, b)
    assert (a[:5] == 0).all()
    q64 = q[100:].astype(np.float64)
    p, n, dpdu, dpdv, o, rx = q64[:, 0:3], q64[:, 3:6], q64[:, 6:9], q64[:, 9:12], q64[:, 12:15], q64[:, 15:18]
    tx = ((p - o) * n).sum(1) / (rx * n).sum(1)
    px = o + rx * tx[:, None]
    rec = p + dpdu * a[100:, 0:1] + dpdv * a[100:, 1:2]          # p + dpdu du + dpdv dv lands on the offset ray's hit point
    err = np.linalg.norm(rec - px, axis=1) / (np.linalg.norm(px - p, axis=1) + 1e-9)
    assert np.median(err) < 1e-3 and (err < 5e-2).mean() > 0.98


def test_sensor_differentials(emu):
    """perspective.cpp:160-163, 310-314 + scaleDifferential(1/sqrt(spp)): device code == oracle; the offset rays are the
    camera rays of the neighbouring pixels pulled towards the central ray by 1/sqrt(sampleCount)."""
    b = scenes.cbox(width=48, height=32, spp=16)
    orc = oracle.Oracle(b)
    rng = np.random.default_rng(9)
    pos = (rng.random((500, 2)) * [48, 32]).astype(np.float32)
    a, e = np.zeros((500, 6), np.float32), np.zeros((500, 6), np.float32)
    oracle.lib().orc_sensor_differentials(orc.h, 500, vp(pos), vp(a))
    se = b.sensor
    near_x = np.float32(se.near_clip) * np.float32(np.tan(np.float32(0.5) * np.float32(np.deg2rad(np.float32(se.fov_x_deg)))))
    rays = orc.sensor_rays(pos)
    rays_x = orc.sensor_rays(pos + np.float32([1, 0]))
    d, dx = rays["d"], rays_x["d"]
    want = d + (dx - d) * np.float32(0.25)
    assert np.allclose(a[:, :3], want, atol=2e-6)
    nx, ny = emu_near(orc, b)
    emu.emu_sensor_differentials(C.byref(se), nx, ny, 16, 500, vp(pos), vp(e))
    assert np.allclose(a, e, atol=1e-7)


def emu_near(orc, b):
    se = b.sensor
    t = np.tan(0.5 * np.deg2rad(np.float64(se.fov_x_deg)))
    nx = float(np.float32(se.near_clip * t))
    return nx, float(np.float32(nx * se.film_height / se.film_width))


def textured_quad_scene(tex_kind, spp=64, filter_type="ewa", tex_res=16):
    """A unit quad in front of the camera under a white constant environment, direct illumination only:
    the radiance of a camera sample is the texture value at its hit point (in expectation)."""
    b = scenedesc.SceneBuilder()
    b.set_integrator(max_depth=2, rr_depth=10, sample_count=spp)
    b.set_sensor(scenedesc.look_at([0, 0, 3], [0, 0, 0], [0, 1, 0]), 30.0, 32, 32)
    if tex_kind == "checkerboard":
        t = b.add_checkerboard(color0=[0.8, 0.2, 0.1], color1=[0.1, 0.3, 0.9], uscale=2.0, vscale=2.0)
    elif tex_kind == "grid":
        t = b.add_gridtexture(color0=[0.2, 0.6, 0.2], color1=[0.9, 0.8, 0.1], line_width=0.1, uscale=2.0, vscale=2.0)
    else:
        rng = np.random.default_rng(11)
        t = b.add_bitmap(rng.random((tex_res, tex_res, 3), dtype=np.float32) * 0.9, filter_type=filter_type)
    m = scenedesc.diffuse(0.5)
    m.tex_reflectance = t
    mid = b.add_material(m)
    P = np.float32([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]])
    b.add_mesh(P, np.uint32([[0, 1, 2], [0, 2, 3]]), mid, texcoords=np.float32([[0, 0], [1, 0], [1, 1], [0, 1]]))
    b.add_constant_env([1.0, 1.0, 1.0])
    return b


@pytest.mark.parametrize("kind", ["checkerboard", "grid", "bitmap"])
def test_oracle_shades_with_the_texture(kind):
    b = textured_quad_scene(kind, filter_type="nearest")
    orc = oracle.Oracle(b)
    rng = np.random.default_rng(12)
    centres = rng.uniform(8, 24, (12, 2)).astype(np.float32)
    rays = orc.sensor_rays(centres)
    its = orc.fill_intersection(rays, orc.trace(rays))
    uv = np.ascontiguousarray(its[:, 15:17])
    assert (uv > 0).all() and (uv < 1).all()
    L = oracle.lib()
    h = L.orc_texture_create(C.byref(b.textures[0]))
    want = np.zeros((12, 3), np.float32)
    L.orc_texture_eval(h, 12, vp(uv), None, vp(want))
    L.orc_texture_destroy(h)
    for c, w in zip(centres, want):
        # a white environment around a lone diffuse quad: emitter sampling and BSDF sampling add up to the albedo exactly
        li = np.mean([orc.li(c[None, :], seed=100 + s)[0, :3] for s in range(8)], axis=0)
        assert np.allclose(li, w, rtol=1e-4, atol=1e-6), (kind, c, li, w)


def test_oracle_renders_the_textured_box():
    """The textured Cornell box: deterministic across thread counts, and the textures are what is seen (the floor is a
    checkerboard: two clearly different albedos in the lower image half)."""
    b = scenes.cbox_textured(width=48, height=36, spp=4, max_depth=3)
    orc = oracle.Oracle(b)
    orc.render(spp=4, seed=3, counter_rng=True, threads=1)
    a = orc.film_develop().copy()
    orc.render(spp=4, seed=3, counter_rng=True, threads=4)
    assert np.allclose(a, orc.film_develop(), rtol=1e-5, atol=1e-6)
    assert np.isfinite(a).all() and a[..., :3].max() > 0.05
    floor = a[30:35, 8:40, :3].mean(-1)
    assert floor.max() > 1.8 * floor.min()
