"""Deterministic synthetic scenes for BASELINE.json's configs.

cfg 1  `cbox()`        Cornell box from the published Cornell measurements (the reference does not
                       ship cbox.xml / its OBJ meshes), diffuse walls, one area light.
cfg 2  `scene_1m()`    16 displaced parametric objects (65 536 tris each) = 1 048 576 tris + ground,
                       GGX roughconductor / diffuse, one area quad.
cfg 3  `scene_10m()`   160 flattened instances of a 65 536-tri object = 10 485 760 tris + ground,
                       diffuse / roughdielectric / roughconductor, constant env + 4 area quads.
All generators are pure numpy and seeded (1004, the reference's default `seed` property,
src/samplers/independent.cpp:58).
"""
import numpy as np

from .scenedesc import SceneBuilder, diffuse, look_at, plastic, roughconductor, roughdielectric

SEED = 1004

# RGB eta/k of three metals: what the reference's spectrum -> RGB conversion gives for data/ior/{Cu,Au,Al}.*.spd
# (mitsuba_b200/host/ior_rgb.inc, tools/make_ior_table.py); passed explicitly so that no SPD lookup is needed
METALS = {
    "Cu": ([0.208209336, 0.919432044, 1.10242665], [3.92252636, 2.45611167, 2.14157939]),
    "Au": ([0.143171251, 0.377493262, 1.43917966], [3.98517585, 2.38457704, 1.60435247]),
    "Al": ([1.6603291, 0.881592274, 0.521452069], [9.22673416, 6.27122402, 4.83799553]),
}


def _quad(p):
    return np.asarray(p, np.float32), np.array([[0, 1, 2], [0, 2, 3]], np.uint32)


def cbox(width=512, height=512, spp=64, max_depth=8, rr_depth=5, twosided=False, short_material=None, tall_material=None):
    """Cornell box, `path` integrator, 512x512, 64 spp, maxDepth 8 (BASELINE.json configs[0]).
    short_material / tall_material: material records for the two blocks (default: the white diffuse)."""
    b = SceneBuilder()
    b.set_integrator(max_depth=max_depth, rr_depth=rr_depth, sample_count=spp)
    b.set_sensor(look_at([278, 273, -800], [278, 273, -799], [0, 1, 0]), 39.3077, width, height,
                 near_clip=10.0, far_clip=2800.0)
    white = b.add_material(diffuse([0.725, 0.71, 0.68], twosided=twosided))
    red = b.add_material(diffuse([0.63, 0.065, 0.05], twosided=twosided))
    green = b.add_material(diffuse([0.14, 0.45, 0.091], twosided=twosided))
    light = b.add_material(diffuse(0.0))
    walls = [
        ([(552.8, 0, 0), (0, 0, 0), (0, 0, 559.2), (549.6, 0, 559.2)], white),                 # floor
        ([(556, 548.8, 0), (556, 548.8, 559.2), (0, 548.8, 559.2), (0, 548.8, 0)], white),     # ceiling
        ([(549.6, 0, 559.2), (0, 0, 559.2), (0, 548.8, 559.2), (556, 548.8, 559.2)], white),   # back
        ([(0, 0, 559.2), (0, 0, 0), (0, 548.8, 0), (0, 548.8, 559.2)], green),                 # right
        ([(552.8, 0, 0), (549.6, 0, 559.2), (556, 548.8, 559.2), (556, 548.8, 0)], red),       # left
    ]
    for pts, m in walls:
        p, i = _quad(pts)
        b.add_mesh(p, i, m)
    short = [
        [(130, 165, 65), (82, 165, 225), (240, 165, 272), (290, 165, 114)],
        [(290, 0, 114), (290, 165, 114), (240, 165, 272), (240, 0, 272)],
        [(130, 0, 65), (130, 165, 65), (290, 165, 114), (290, 0, 114)],
        [(82, 0, 225), (82, 165, 225), (130, 165, 65), (130, 0, 65)],
        [(240, 0, 272), (240, 165, 272), (82, 165, 225), (82, 0, 225)],
    ]
    tall = [
        [(423, 330, 247), (265, 330, 296), (314, 330, 456), (472, 330, 406)],
        [(423, 0, 247), (423, 330, 247), (472, 330, 406), (472, 0, 406)],
        [(472, 0, 406), (472, 330, 406), (314, 330, 456), (314, 0, 456)],
        [(314, 0, 456), (314, 330, 456), (265, 330, 296), (265, 0, 296)],
        [(265, 0, 296), (265, 330, 296), (423, 330, 247), (423, 0, 247)],
    ]
    def block_material(m):
        if m is None:
            return white
        return b.add_twosided(*m) if isinstance(m, tuple) else b.add_material(m)
    for block, m in ((short, block_material(short_material)), (tall, block_material(tall_material))):
        pos = np.concatenate([np.asarray(f, np.float32) for f in block])
        idx = np.concatenate([np.array([[0, 1, 2], [0, 2, 3]], np.uint32) + 4 * k for k in range(5)])
        b.add_mesh(pos, idx, m)
    p, i = _quad([(343, 548.75, 227), (343, 548.75, 332), (213, 548.75, 332), (213, 548.75, 227)])
    b.add_mesh(p, i, light, area_radiance=[18.4, 15.6, 8.0])
    return b


def pattern_image(width=64, height=64, seed=3):
    """A linear RGB test image with structure at every scale (stripes, blobs, noise), values in [0, 0.95]."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:height, 0:width].astype(np.float64)
    u, v = x / width, y / height
    img = np.stack([0.5 + 0.45 * np.sin(2 * np.pi * (3 * u + v)), 0.5 + 0.45 * np.cos(2 * np.pi * (5 * v - 2 * u)),
                    0.25 + 0.7 * ((np.floor(8 * u) + np.floor(8 * v)) % 2)], -1)
    img = 0.8 * img + 0.2 * rng.random((height, width, 3))
    return np.clip(img, 0.0, 0.95).astype(np.float32)


def cbox_textured(width=96, height=72, spp=4, max_depth=6, rr_depth=5, tex_res=64):
    """The Cornell box with every kind of texture this path implements (src/textures): checkerboard floor, EWA-filtered
    bitmap on the back wall (minified at this resolution), gridtexture / nearest bitmap on the side walls, a plastic block
    with a trilinear bitmap and a twosided block (textured roughconductor over a scaled checkerboard)."""
    b = SceneBuilder()
    b.set_integrator(max_depth=max_depth, rr_depth=rr_depth, sample_count=spp)
    b.set_sensor(look_at([278, 273, -800], [278, 273, -799], [0, 1, 0]), 39.3077, width, height, near_clip=10.0, far_clip=2800.0)
    img = pattern_image(tex_res, tex_res)
    grey = pattern_image(tex_res // 2, tex_res, seed=5)[:, :, 1]

    def textured(m, refl=0, trans=0):
        m.tex_reflectance, m.tex_transmittance = refl, trans
        return m
    white = b.add_material(diffuse([0.725, 0.71, 0.68]))
    floor = b.add_material(textured(diffuse(0.5), b.add_checkerboard([0.7, 0.7, 0.65], [0.15, 0.2, 0.3], uscale=4, vscale=4)))
    back = b.add_material(textured(diffuse(0.5), b.add_bitmap(img, "ewa", "repeat", uscale=3, vscale=3, max_anisotropy=8.0)))
    right = b.add_material(textured(diffuse(0.5), b.add_gridtexture([0.14, 0.45, 0.091], [0.9, 0.9, 0.2], 0.04, uscale=5, vscale=5, uoffset=0.3)))
    left = b.add_material(textured(diffuse(0.5), b.add_bitmap(img, "nearest", "mirror", uscale=1.5, vscale=1.5, scale=[0.9, 0.3, 0.25])))
    light = b.add_material(diffuse(0.0))
    uv4 = np.float32([[0, 0], [1, 0], [1, 1], [0, 1]])
    walls = [
        ([(552.8, 0, 0), (0, 0, 0), (0, 0, 559.2), (549.6, 0, 559.2)], floor),
        ([(556, 548.8, 0), (556, 548.8, 559.2), (0, 548.8, 559.2), (0, 548.8, 0)], white),
        ([(549.6, 0, 559.2), (0, 0, 559.2), (0, 548.8, 559.2), (556, 548.8, 559.2)], back),
        ([(0, 0, 559.2), (0, 0, 0), (0, 548.8, 0), (0, 548.8, 559.2)], right),
        ([(552.8, 0, 0), (549.6, 0, 559.2), (556, 548.8, 559.2), (556, 548.8, 0)], left),
    ]
    for pts, m in walls:
        p, i = _quad(pts)
        b.add_mesh(p, i, m, texcoords=uv4)
    short = [
        [(130, 165, 65), (82, 165, 225), (240, 165, 272), (290, 165, 114)],
        [(290, 0, 114), (290, 165, 114), (240, 165, 272), (240, 0, 272)],
        [(130, 0, 65), (130, 165, 65), (290, 165, 114), (290, 0, 114)],
        [(82, 0, 225), (82, 165, 225), (130, 165, 65), (130, 0, 65)],
        [(240, 0, 272), (240, 165, 272), (82, 165, 225), (82, 0, 225)],
    ]
    tall = [
        [(423, 330, 247), (265, 330, 296), (314, 330, 456), (472, 330, 406)],
        [(423, 0, 247), (423, 330, 247), (472, 330, 406), (472, 0, 406)],
        [(472, 0, 406), (472, 330, 406), (314, 330, 456), (314, 0, 456)],
        [(314, 0, 456), (314, 330, 456), (265, 330, 296), (265, 0, 296)],
        [(265, 0, 296), (265, 330, 296), (423, 330, 247), (423, 0, 247)],
    ]
    m_short = b.add_material(textured(plastic(diffuse_reflectance=0.5), b.add_bitmap(grey, "trilinear", "clamp", uscale=2, vscale=2)))
    front = textured(roughconductor([0.2, 0.92, 1.1], [3.9, 2.45, 2.14], alpha=0.15), b.add_bitmap(img, "bilinear", "repeat", scale=0.95))
    m_tall = b.add_twosided(front, textured(diffuse(0.5), b.add_checkerboard(0.8, 0.1, uscale=2, vscale=2, scale=[0.5, 1.0, 0.5])))
    for block, m in ((short, m_short), (tall, m_tall)):
        pos = np.concatenate([np.asarray(f, np.float32) for f in block])
        idx = np.concatenate([np.array([[0, 1, 2], [0, 2, 3]], np.uint32) + 4 * k for k in range(5)])
        b.add_mesh(pos, idx, m, texcoords=np.tile(uv4, (5, 1)))
    p, i = _quad([(343, 548.75, 227), (343, 548.75, 332), (213, 548.75, 332), (213, 548.75, 227)])
    b.add_mesh(p, i, light, area_radiance=[18.4, 15.6, 8.0])
    return b


def parametric_object(nu, nv, kind, rng, radius=1.0):
    """A closed displaced surface on an nu x nv grid: 2*nu*nv triangles, smooth vertex normals."""
    u = (np.arange(nu, dtype=np.float64) / nu) * 2 * np.pi
    v = (np.arange(nv + 1, dtype=np.float64) / nv)
    uu, vv = np.meshgrid(u, v, indexing="ij")          # (nu, nv+1)
    if kind == 0:      # bumpy sphere
        th = vv * np.pi
        k1, k2 = rng.integers(3, 9), rng.integers(2, 7)
        r = radius * (1 + 0.12 * np.sin(k1 * uu) * np.sin(k2 * th) * np.sin(th))
        x, y, z = r * np.sin(th) * np.cos(uu), r * np.cos(th), r * np.sin(th) * np.sin(uu)
    else:              # twisted torus
        ph = vv * 2 * np.pi
        k1 = rng.integers(2, 6)
        rr = 0.35 * radius * (1 + 0.25 * np.sin(k1 * uu + 2 * ph))
        R = 0.75 * radius
        x, y, z = (R + rr * np.cos(ph)) * np.cos(uu), rr * np.sin(ph), (R + rr * np.cos(ph)) * np.sin(uu)
    pos = np.stack([x, y, z], -1).reshape(-1, 3)
    i0 = (np.arange(nu)[:, None] * (nv + 1) + np.arange(nv)[None, :])
    i1 = (((np.arange(nu) + 1) % nu)[:, None] * (nv + 1) + np.arange(nv)[None, :])
    a, b_, c, d = i0, i1, i1 + 1, i0 + 1
    # winding chosen so that the face normal points outward for both families (the two parametrisations have opposite
    # handedness: the sphere's polar angle runs against the torus' tube angle)
    if kind == 0:
        tris = np.concatenate([np.stack([a, b_, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)])
    else:
        tris = np.concatenate([np.stack([a, c, b_], -1).reshape(-1, 3), np.stack([a, d, c], -1).reshape(-1, 3)])
    # area-weighted vertex normals
    p = pos
    fn = np.cross(p[tris[:, 1]] - p[tris[:, 0]], p[tris[:, 2]] - p[tris[:, 0]])
    nrm = np.zeros_like(p)
    for k in range(3):
        np.add.at(nrm, tris[:, k], fn)
    ln = np.linalg.norm(nrm, axis=1, keepdims=True)
    ln[ln == 0] = 1
    nrm /= ln
    return pos.astype(np.float32), tris.astype(np.uint32), nrm.astype(np.float32)


def _rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def _object_field(b, n_obj, grid, nu, nv, materials, rng, spacing=2.6, y0=1.05):
    """Places n_obj parametric objects on a jittered grid (instances are flattened: SURVEY 2.1 row 19)."""
    base = [parametric_object(nu, nv, k, rng) for k in range(2)]
    gx, gz = grid
    for i in range(n_obj):
        ix, iz = i % gx, i // gx
        pos, tris, nrm = base[i % 2]
        R = _rot_y(rng.uniform(0, 2 * np.pi)) * rng.uniform(0.8, 1.1)
        t = np.array([(ix - (gx - 1) / 2) * spacing + rng.uniform(-0.3, 0.3), y0 + rng.uniform(0, 0.4),
                      (iz - (gz - 1) / 2) * spacing + rng.uniform(-0.3, 0.3)])
        p = (pos.astype(np.float64) @ R.T + t).astype(np.float32)
        n = nrm.astype(np.float64) @ (R / np.cbrt(np.linalg.det(R))).T
        b.add_mesh(p, tris, materials[i % len(materials)], normals=n.astype(np.float32))
    return gx * spacing, gz * spacing


def scene_1m(width=1920, height=1080, spp=256, max_depth=10, rr_depth=5, nu=256, nv=128, n_obj=16):
    """BASELINE.json configs[1]: 1M-triangle scene, GGX roughconductor + diffuse, one area light."""
    rng = np.random.default_rng(SEED)
    b = SceneBuilder()
    b.set_integrator(max_depth=max_depth, rr_depth=rr_depth, sample_count=spp)
    mats = []
    for j, (name, alpha) in enumerate([("Cu", 0.05), ("Au", 0.2), ("Al", 0.4)]):
        eta, k = METALS[name]
        mats.append(b.add_material(roughconductor(eta, k, alpha=alpha, distribution="ggx")))
        mats.append(b.add_material(diffuse(rng.uniform(0.2, 0.8, 3))))
    ground = b.add_material(diffuse([0.5, 0.5, 0.5]))
    g = int(np.ceil(np.sqrt(n_obj)))
    ex, ez = _object_field(b, n_obj, (g, g), nu, nv, mats, rng)
    s = max(ex, ez)
    b.add_quad((-s, 0, -s), (-s, 0, s), (s, 0, s), (s, 0, -s), ground)
    lm = b.add_material(diffuse(0.0))
    h = 7.0
    b.add_quad((-2.5, h, -2.5), (2.5, h, -2.5), (2.5, h, 2.5), (-2.5, h, 2.5), lm, area_radiance=[30, 28, 25])
    b.set_sensor(look_at([0, 0.55 * s, -1.25 * s], [0, 0.8, 0], [0, 1, 0]), 45.0, width, height,
                 near_clip=0.05, far_clip=500.0)
    return b


def scene_10m(width=1920, height=1080, spp=512, max_depth=10, rr_depth=5, nu=256, nv=128, n_obj=160):
    """BASELINE.json configs[2]: 10M-triangle flattened-instance scene, mixed BSDFs, env + 4 area lights."""
    rng = np.random.default_rng(SEED)
    b = SceneBuilder()
    b.set_integrator(max_depth=max_depth, rr_depth=rr_depth, sample_count=spp)
    eta, k = METALS["Au"]
    eta2, k2 = METALS["Al"]
    mats = [
        b.add_material(diffuse(rng.uniform(0.2, 0.8, 3))),
        b.add_material(roughdielectric(int_ior=1.5, ext_ior=1.0, alpha=0.1, distribution="ggx")),
        b.add_material(roughconductor(eta, k, alpha=0.2, distribution="beckmann")),
        b.add_material(roughconductor(eta2, k2, alpha=0.3, distribution="ggx")),
    ]
    ground = b.add_material(diffuse([0.45, 0.45, 0.45]))
    gx = int(np.ceil(np.sqrt(n_obj * 1.6)))
    gz = int(np.ceil(n_obj / gx))
    ex, ez = _object_field(b, n_obj, (gx, gz), nu, nv, mats, rng)
    s = max(ex, ez) * 0.6
    b.add_quad((-s, 0, -s), (-s, 0, s), (s, 0, s), (s, 0, -s), ground)
    lm = b.add_material(diffuse(0.0))
    h = 8.0
    for cx, cz in [(-0.5, -0.5), (0.5, -0.5), (-0.5, 0.5), (0.5, 0.5)]:
        x, z = cx * ex * 0.5, cz * ez * 0.5
        b.add_quad((x - 2, h, z - 2), (x + 2, h, z - 2), (x + 2, h, z + 2), (x - 2, h, z + 2), lm,
                   area_radiance=[25, 24, 22])
    b.add_constant_env([0.5, 0.5, 0.5])
    b.set_sensor(look_at([0, 0.45 * s, -1.05 * s], [0, 0.8, 0], [0, 1, 0]), 45.0, width, height,
                 near_clip=0.05, far_clip=1000.0)
    return b


def chord_rays(center, radius, n, seed=SEED):
    """Chords between uniform points on a sphere -- the ray set of src/tests/test_kd.cpp:112-117
    and src/utils/kdbench.cpp ("incoherent" set of cfg 4)."""
    rng = np.random.default_rng(seed)
    from . import abi

    def sph(k):
        u = rng.random((k, 2))
        z = 1 - 2 * u[:, 1]
        r = np.sqrt(np.maximum(0, 1 - z * z))
        ph = 2 * np.pi * u[:, 0]
        return np.stack([r * np.cos(ph), r * np.sin(ph), z], -1)
    p1 = np.asarray(center) + radius * sph(n)
    p2 = np.asarray(center) + radius * sph(n)
    d = p2 - p1
    ln = np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros(n, dtype=abi.RAY_DTYPE)
    rays["o"] = p1.astype(np.float32)
    rays["d"] = (d / ln).astype(np.float32)
    rays["mint"] = 0.0
    rays["maxt"] = ln[:, 0].astype(np.float32)
    return rays


def synthetic_sky(width=64, height=32, seed=4, sun=True):
    """A small synthetic HDR latitude-longitude map (row 0 = +Y pole): blue-ish gradient sky, a warm horizon band, a
    dim ground and (optionally) a bright sun a few texels wide -- enough dynamic range to exercise the importance
    sampling tables of the `envmap` emitter.  Deterministic."""
    rng = np.random.default_rng(seed)
    theta = (np.arange(height) + 0.5) * np.pi / height
    phi = (np.arange(width) + 0.5) * 2 * np.pi / width
    T, P = np.meshgrid(theta, phi, indexing="ij")
    up = np.cos(T)
    img = np.zeros((height, width, 3), np.float32)
    sky = np.clip(up, 0, 1)[..., None] * np.array([0.25, 0.45, 0.9]) + np.exp(-8 * np.abs(up))[..., None] * np.array([0.9, 0.6, 0.3])
    ground = np.array([0.08, 0.07, 0.05]) * (1 + 0.3 * np.sin(3 * P))[..., None]
    img[:] = np.where((up > 0)[..., None], sky, ground)
    img *= (1 + 0.1 * rng.standard_normal((height, width, 1))).clip(0.5, 1.5)
    if sun:
        sd = np.array([np.sin(0.9) * np.sin(2.2), np.cos(0.9), -np.sin(0.9) * np.cos(2.2)])
        d = np.stack([np.sin(T) * np.sin(P), np.cos(T), -np.sin(T) * np.cos(P)], -1)
        img += (np.exp(-((1 - d @ sd) / 0.004)) * 400.0)[..., None] * np.array([1.0, 0.9, 0.7])
    return img.astype(np.float32)
