"""Host-side flat scene description (the POD that crosses the C ABI).

`SceneBuilder` plays the role of `Scene::addChild` + `Scene::initialize` flattening
(reference: src/librender/scene.cpp:322-384, 558-592): shapes are appended as
triangle meshes with a BSDF record and an optional area emitter, and `desc()` emits
the `MtsbSceneDesc` with concatenated vertex / index buffers (global primitive id =
shape offset + local triangle index, src/librender/skdtree.cpp:68-70).
"""
import ctypes as C
import math

import numpy as np

from . import abi


def _f3(v):
    a = np.asarray(v, dtype=np.float32).reshape(-1)
    if a.size == 1:
        a = np.repeat(a, 3)
    return (C.c_float * 3)(*[float(x) for x in a[:3]])


def diffuse(reflectance=0.5, twosided=False):
    """`diffuse` BSDF record (src/bsdfs/diffuse.cpp:82-84; reflectance clamped by
    ensureEnergyConservation, src/librender/bsdf.cpp:96-113)."""
    r = np.asarray(reflectance, dtype=np.float32).reshape(-1)
    if r.size == 1:
        r = np.repeat(r, 3)
    mx = float(r.max())
    if mx > 1.0:
        r = r * np.float32(0.99 * (1.0 / mx))
    m = abi.Material()
    m.type = abi.BSDF_DIFFUSE
    m.flags = abi.MAT_TWOSIDED if twosided else 0
    m.back = -1
    m.alpha_u = m.alpha_v = 0.1
    m.eta = 1.0
    m.reflectance = _f3(r)
    m.transmittance = _f3(1.0)
    m.cond_eta = _f3(0.0)
    m.cond_k = _f3(1.0)
    return m


def _microfacet(m, distribution, alpha, alpha_u, alpha_v, sample_visible):
    if alpha_u is None:
        alpha_u = alpha_v = alpha
    m.distribution = {"beckmann": abi.DISTR_BECKMANN, "ggx": abi.DISTR_GGX}[distribution]
    m.alpha_u = max(float(alpha_u), 1e-4)      # microfacet.h:135-136
    m.alpha_v = max(float(alpha_v), 1e-4)
    if sample_visible:
        m.flags |= abi.MAT_SAMPLE_VISIBLE


def roughconductor(eta, k, alpha=0.1, distribution="beckmann", specular_reflectance=1.0, ext_eta=1.000277,
                   sample_visible=True, twosided=False, alpha_u=None, alpha_v=None):
    """`roughconductor` record (src/bsdfs/roughconductor.cpp:168-201): eta, k are divided by extEta."""
    m = abi.Material()
    m.type = abi.BSDF_ROUGHCONDUCTOR
    m.flags = abi.MAT_TWOSIDED if twosided else 0
    m.back = -1
    _microfacet(m, distribution, alpha, alpha_u, alpha_v, sample_visible)
    e = np.asarray(eta, dtype=np.float32).reshape(-1)
    kk = np.asarray(k, dtype=np.float32).reshape(-1)
    if e.size == 1:
        e = np.repeat(e, 3)
    if kk.size == 1:
        kk = np.repeat(kk, 3)
    m.cond_eta = _f3(e / np.float32(ext_eta))
    m.cond_k = _f3(kk / np.float32(ext_eta))
    m.reflectance = _f3(specular_reflectance)
    m.transmittance = _f3(1.0)
    m.eta = 1.0
    return m


def roughdielectric(int_ior=1.5046, ext_ior=1.000277, alpha=0.1, distribution="beckmann", specular_reflectance=1.0,
                    specular_transmittance=1.0, sample_visible=True, alpha_u=None, alpha_v=None):
    """`roughdielectric` record (src/bsdfs/roughdielectric.cpp:183-211)."""
    m = abi.Material()
    m.type = abi.BSDF_ROUGHDIELECTRIC
    m.flags = 0
    m.back = -1
    _microfacet(m, distribution, alpha, alpha_u, alpha_v, sample_visible)
    m.eta = float(np.float32(int_ior) / np.float32(ext_ior))
    m.reflectance = _f3(specular_reflectance)
    m.transmittance = _f3(specular_transmittance)
    m.cond_eta = _f3(0.0)
    m.cond_k = _f3(1.0)
    return m


def _conserve(v, max_value=1.0):
    """ensureEnergyConservation on a constant texture (src/librender/bsdf.cpp:96-113)."""
    r = np.asarray(v, dtype=np.float32).reshape(-1)
    if r.size == 1:
        r = np.repeat(r, 3)
    mx = float(r.max())
    if mx > max_value:
        r = r * np.float32(0.99 * (max_value / mx))
    return r


def dielectric(int_ior=1.5046, ext_ior=1.000277, specular_reflectance=1.0, specular_transmittance=1.0):
    """Smooth `dielectric` record (src/bsdfs/dielectric.cpp:141-168, 192-197)."""
    m = abi.Material()
    m.type = abi.BSDF_DIELECTRIC
    m.flags = 0
    m.back = -1
    m.alpha_u = m.alpha_v = 0.1
    m.eta = float(np.float32(int_ior) / np.float32(ext_ior))
    m.reflectance = _f3(_conserve(specular_reflectance))
    m.transmittance = _f3(_conserve(specular_transmittance))
    m.cond_eta = _f3(0.0)
    m.cond_k = _f3(1.0)
    return m


def conductor(eta, k, specular_reflectance=1.0, ext_eta=1.000277, twosided=False):
    """Smooth `conductor` record (src/bsdfs/conductor.cpp:146-174, 189-193): eta, k are divided by extEta."""
    m = abi.Material()
    m.type = abi.BSDF_CONDUCTOR
    m.flags = abi.MAT_TWOSIDED if twosided else 0
    m.back = -1
    m.alpha_u = m.alpha_v = 0.1
    e = np.asarray(eta, dtype=np.float32).reshape(-1)
    kk = np.asarray(k, dtype=np.float32).reshape(-1)
    if e.size == 1:
        e = np.repeat(e, 3)
    if kk.size == 1:
        kk = np.repeat(kk, 3)
    m.cond_eta = _f3(e / np.float32(ext_eta))
    m.cond_k = _f3(kk / np.float32(ext_eta))
    m.reflectance = _f3(_conserve(specular_reflectance))
    m.transmittance = _f3(1.0)
    m.eta = 1.0
    return m


def fresnel_diffuse_reflectance(eta, n=65536):
    """fresnelDiffuseReflectance(eta, false) (src/libcore/util.cpp:814-860): the integral over xi in [0,1] of
    fresnelDielectricExt(sqrt(xi), eta).  The reference integrates adaptively in single precision to a relative
    tolerance of 1e-5; this is a double-precision midpoint rule, so the two agree to that tolerance."""
    eta = float(eta)
    xi = (np.arange(n, dtype=np.float64) + 0.5) / n
    cos_i = np.sqrt(xi)
    if eta == 1.0:
        return np.float32(0.0)
    scale = 1.0 / eta                                         # cos_i > 0
    cos_t_sqr = 1.0 - (1.0 - cos_i * cos_i) * (scale * scale)
    tir = cos_t_sqr <= 0.0
    cos_t = np.sqrt(np.where(tir, 0.0, cos_t_sqr))
    rs = (cos_i - eta * cos_t) / (cos_i + eta * cos_t)
    rp = (eta * cos_i - cos_t) / (eta * cos_i + cos_t)
    f = np.where(tir, 1.0, 0.5 * (rs * rs + rp * rp))
    return np.float32(f.sum() / n)


def _luminance(rgb):
    r = np.asarray(rgb, dtype=np.float32)
    return np.float32(np.float32(r[0] * np.float32(0.212671) + r[1] * np.float32(0.715160)) + r[2] * np.float32(0.072169))


def plastic(int_ior=1.49, ext_ior=1.000277, diffuse_reflectance=0.5, specular_reflectance=1.0, nonlinear=False,
            twosided=False):
    """Smooth `plastic` record (src/bsdfs/plastic.cpp:141-165, 185-206): polypropylene / air by default."""
    m = abi.Material()
    m.type = abi.BSDF_PLASTIC
    m.flags = (abi.MAT_TWOSIDED if twosided else 0) | (abi.MAT_NONLINEAR if nonlinear else 0)
    m.back = -1
    m.alpha_u = m.alpha_v = 0.1
    eta = np.float32(int_ior) / np.float32(ext_ior)
    m.eta = float(eta)
    d, sp = _conserve(diffuse_reflectance), _conserve(specular_reflectance)
    m.reflectance = _f3(d)
    m.transmittance = _f3(sp)
    m.cond_eta = _f3(0.0)
    m.cond_k = _f3(1.0)
    m.fdr_int = float(fresnel_diffuse_reflectance(np.float32(1.0) / eta))
    d_avg, s_avg = _luminance(d), _luminance(sp)
    m.spec_weight = float(s_avg / np.float32(d_avg + s_avg))
    return m


def look_at(origin, target, up):
    """Transform::lookAt (src/libcore/transform.cpp:191-214): row-major camera-to-world."""
    p = np.asarray(origin, dtype=np.float64)
    t = np.asarray(target, dtype=np.float64)
    u = np.asarray(up, dtype=np.float64)
    d = (t - p) / np.linalg.norm(t - p)
    left = np.cross(u, d)
    left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, p
    return m.astype(np.float32)


class SceneBuilder:
    def __init__(self):
        self.materials = []
        self.shapes = []          # dicts
        self.emitters = []
        self._pos, self._nrm, self._uv, self._idx = [], [], [], []
        self._nv = 0
        self._nt = 0
        self.sensor = abi.Sensor()
        self.integrator = abi.Integrator()
        self.sampler = abi.Sampler()
        self.sampler.dimension, self.sampler.scramble = 4, -1
        self.set_integrator()
        self.set_sensor(look_at([0, 0, -5], [0, 0, 0], [0, 1, 0]), 45.0, 64, 64)
        self._keep = None

    # -- plugin-like setters ------------------------------------------------
    def set_integrator(self, max_depth=-1, rr_depth=5, strict_normals=False, hide_emitters=False, sample_count=4):
        """`path` integrator + `independent` sampler properties
        (src/librender/integrator.cpp:252-287, src/samplers/independent.cpp:57)."""
        ig = self.integrator
        ig.max_depth, ig.rr_depth = int(max_depth), int(rr_depth)
        ig.strict_normals, ig.hide_emitters = int(strict_normals), int(hide_emitters)
        ig.sample_count = int(sample_count)

    def set_sampler(self, kind="independent", sample_count=None, dimension=4, scramble=None):
        """Sample generator: `independent`, `ldsampler` (sample count rounded up to a power of two,
        src/samplers/ldsampler.cpp:75-86) `halton` (src/samplers/halton.cpp:121-130)
        or `sobol` (src/samplers/sobol.cpp:85-107)."""
        kinds = {"independent": abi.SAMPLER_INDEPENDENT, "ldsampler": abi.SAMPLER_LDSAMPLER, "halton": abi.SAMPLER_HALTON,
                 "sobol": abi.SAMPLER_SOBOL}
        self.sampler.type = kinds[kind]
        if scramble is None:
            scramble = 0 if kind == "sobol" else -1          # sobol.cpp:91, halton.cpp:126
        self.sampler.dimension, self.sampler.scramble = int(dimension), int(scramble)
        if sample_count is not None:
            self.integrator.sample_count = int(sample_count)
        if kind == "ldsampler":
            n = max(1, int(self.integrator.sample_count))
            self.integrator.sample_count = 1 << (n - 1).bit_length()

    def set_sensor(self, to_world, fov, width, height, fov_axis="x", near_clip=1e-2, far_clip=1e4,
                   rfilter="gaussian", stddev=0.5, box_radius=0.5, crop=None):
        """`perspective` sensor + `hdrfilm` + reconstruction filter
        (src/librender/sensor.cpp:225-309, src/librender/film.cpp:28-54, src/rfilters/{gaussian,box}.cpp)."""
        s = self.sensor
        tw = np.asarray(to_world, dtype=np.float32).reshape(16)
        s.to_world = (C.c_float * 16)(*[float(x) for x in tw])
        aspect = width / float(height)
        if fov_axis == "smaller":
            fov_axis = "y" if aspect > 1 else "x"
        elif fov_axis == "larger":
            fov_axis = "x" if aspect > 1 else "y"
        if fov_axis == "x":
            xfov = fov
        elif fov_axis == "y":
            xfov = math.degrees(2 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
        else:
            raise ValueError("fovAxis must be x, y, smaller or larger")
        s.fov_x_deg = xfov
        s.near_clip, s.far_clip = near_clip, far_clip
        s.film_width, s.film_height = int(width), int(height)
        if crop is None:
            crop = (0, 0, width, height)
        s.crop_x, s.crop_y, s.crop_width, s.crop_height = [int(c) for c in crop]
        if rfilter == "gaussian":
            s.rfilter, s.rfilter_stddev, s.rfilter_radius = abi.RFILTER_GAUSSIAN, stddev, 4 * stddev
        elif rfilter == "box":
            s.rfilter, s.rfilter_stddev = abi.RFILTER_BOX, 0.0
            s.rfilter_radius = float(np.float32(box_radius) + np.float32(1e-5))   # box.cpp:39
        else:
            raise ValueError("unsupported rfilter " + rfilter)

    def add_material(self, m):
        self.materials.append(m)
        return len(self.materials) - 1

    def add_twosided(self, front, back=None):
        """`twosided` adapter (src/bsdfs/twosided.cpp): front/back are Material records."""
        m = abi.Material.from_buffer_copy(front)
        m.flags |= abi.MAT_TWOSIDED
        if back is not None:
            self.materials.append(back)
            m.back = len(self.materials) - 1
        self.materials.append(m)
        return len(self.materials) - 1

    def add_mesh(self, positions, indices, material, normals=None, texcoords=None, area_radiance=None,
                 sampling_weight=1.0):
        """A triangle-mesh shape; `area_radiance` attaches an `area` emitter (src/emitters/area.cpp)."""
        positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        indices = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1, 3)
        nv, nt = positions.shape[0], indices.shape[0]
        if nt and int(indices.max()) >= nv:
            raise ValueError("triangle index out of range")
        sh = dict(first_vertex=self._nv, n_vertices=nv, first_tri=self._nt, n_tris=nt, material=material,
                  emitter=-1, has_normals=normals is not None, has_texcoords=texcoords is not None)
        self._pos.append(positions)
        self._nrm.append(np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3) if normals is not None
                         else np.zeros((nv, 3), np.float32))
        self._uv.append(np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2) if texcoords is not None
                        else np.zeros((nv, 2), np.float32))
        self._idx.append(indices + np.uint32(self._nv))
        self._nv += nv
        self._nt += nt
        self.shapes.append(sh)
        if area_radiance is not None:
            e = abi.Emitter()
            e.type, e.shape = abi.EMITTER_AREA, len(self.shapes) - 1
            e.radiance = _f3(area_radiance)
            e.sampling_weight = sampling_weight
            self.emitters.append(e)
            sh["emitter"] = len(self.emitters) - 1
        return len(self.shapes) - 1

    def add_point_light(self, position, intensity, sampling_weight=1.0):
        e = abi.Emitter()
        e.type, e.shape = abi.EMITTER_POINT, -1
        e.radiance, e.position, e.sampling_weight = _f3(intensity), _f3(position), sampling_weight
        self.emitters.append(e)

    def add_constant_env(self, radiance, sampling_weight=1.0):
        e = abi.Emitter()
        e.type, e.shape = abi.EMITTER_CONSTANT, -1
        e.radiance, e.sampling_weight = _f3(radiance), sampling_weight
        self.emitters.append(e)

    def add_envmap(self, image, to_world=None, scale=1.0, sampling_weight=1.0):
        """`envmap` emitter (src/emitters/envmap.cpp:105-189): `image` is a (height, width, 3) linear RGB array in
        latitude-longitude layout, row 0 at the +Y pole."""
        img = np.ascontiguousarray(image, dtype=np.float32)
        assert img.ndim == 3 and img.shape[2] == 3
        tw = np.eye(4, dtype=np.float64) if to_world is None else np.asarray(to_world, np.float64).reshape(4, 4)
        em = abi.Envmap()
        em.height, em.width = img.shape[0], img.shape[1]
        em.rgb = abi.ptr(img, C.c_float)
        em.to_world = (C.c_float * 16)(*[float(x) for x in tw.astype(np.float32).reshape(16)])
        em.to_local = (C.c_float * 16)(*[float(x) for x in np.linalg.inv(tw).astype(np.float32).reshape(16)])
        em.scale = float(scale)
        self.envmap, self._envmap_image = em, img
        e = abi.Emitter()
        e.type, e.shape = abi.EMITTER_ENVMAP, -1
        e.radiance, e.sampling_weight = _f3([0, 0, 0]), sampling_weight
        self.emitters.append(e)

    # -- convenience geometry --------------------------------------------------
    # -- textures (src/textures; the returned handle goes into Material.tex_reflectance / tex_transmittance) --------
    def _add_texture(self, t, uscale, vscale, uoffset, voffset, scale, keep=None):
        t.uv_scale = (C.c_float * 2)(float(uscale), float(vscale))
        t.uv_offset = (C.c_float * 2)(float(uoffset), float(voffset))
        t.scale = _f3(scale)
        t.max_anisotropy = t.max_anisotropy or 20.0
        if not hasattr(self, "textures"):
            self.textures, self._texture_keep = [], []
        self.textures.append(t)
        self._texture_keep.append(keep)
        return len(self.textures)

    def add_checkerboard(self, color0=0.4, color1=0.2, uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0, scale=1.0):
        """`checkerboard` texture (src/textures/checkerboard.cpp:44-69); `scale` folds an enclosing `scale` plugin."""
        t = abi.Texture()
        t.type, t.color0, t.color1 = abi.TEX_CHECKERBOARD, _f3(color0), _f3(color1)
        return self._add_texture(t, uscale, vscale, uoffset, voffset, scale)

    def add_gridtexture(self, color0=0.2, color1=0.4, line_width=0.01, uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0,
                        scale=1.0):
        """`gridtexture` (src/textures/gridtexture.cpp:45-79)."""
        t = abi.Texture()
        t.type, t.color0, t.color1, t.line_width = abi.TEX_GRID, _f3(color0), _f3(color1), float(line_width)
        return self._add_texture(t, uscale, vscale, uoffset, voffset, scale)

    def add_bitmap(self, image, filter_type="ewa", wrap_mode="repeat", wrap_mode_u=None, wrap_mode_v=None,
                   max_anisotropy=20.0, uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0, scale=1.0):
        """`bitmap` texture (src/textures/bitmap.cpp:179-303): `image` is (height, width) luminance or
        (height, width, 3) linear RGB, row 0 = v 0."""
        img = np.ascontiguousarray(np.asarray(image, dtype=np.float32))
        if img.ndim == 2:
            img = img[:, :, None]
        if img.ndim != 3 or img.shape[2] not in (1, 3):
            raise ValueError("bitmap: expected (h, w) or (h, w, 3)")
        filt = {"nearest": abi.TEXFILTER_NEAREST, "bilinear": abi.TEXFILTER_BILINEAR, "trilinear": abi.TEXFILTER_TRILINEAR,
                "ewa": abi.TEXFILTER_EWA}
        wrap = {"repeat": abi.WRAP_REPEAT, "clamp": abi.WRAP_CLAMP, "mirror": abi.WRAP_MIRROR, "zero": abi.WRAP_ZERO,
                "black": abi.WRAP_ZERO, "one": abi.WRAP_ONE, "white": abi.WRAP_ONE}
        t = abi.Texture()
        t.type, t.filter = abi.TEX_BITMAP, filt[filter_type]
        t.wrap_u, t.wrap_v = wrap[wrap_mode_u or wrap_mode], wrap[wrap_mode_v or wrap_mode]
        t.max_anisotropy = float(max_anisotropy)
        t.height, t.width, t.channels = img.shape
        t.texels = abi.ptr(img, C.c_float)
        return self._add_texture(t, uscale, vscale, uoffset, voffset, scale, keep=img)

    def add_quad(self, p0, p1, p2, p3, material, **kw):
        """Two triangles (p0,p1,p2), (p0,p2,p3)."""
        return self.add_mesh(np.array([p0, p1, p2, p3], np.float32), np.array([[0, 1, 2], [0, 2, 3]], np.uint32),
                             material, **kw)

    # -- flatten ------------------------------------------------------------------
    @property
    def n_tris(self):
        return self._nt

    def arrays(self):
        pos = np.ascontiguousarray(np.concatenate(self._pos)) if self._pos else np.zeros((0, 3), np.float32)
        nrm = np.ascontiguousarray(np.concatenate(self._nrm)) if self._nrm else np.zeros((0, 3), np.float32)
        uv = np.ascontiguousarray(np.concatenate(self._uv)) if self._uv else np.zeros((0, 2), np.float32)
        idx = np.ascontiguousarray(np.concatenate(self._idx)) if self._idx else np.zeros((0, 3), np.uint32)
        return pos, nrm, uv, idx

    def desc(self):
        """Returns (MtsbSceneDesc, keepalive)."""
        pos, nrm, uv, idx = self.arrays()
        shapes = (abi.Shape * max(1, len(self.shapes)))()
        for i, s in enumerate(self.shapes):
            shapes[i].first_vertex, shapes[i].n_vertices = s["first_vertex"], s["n_vertices"]
            shapes[i].first_tri, shapes[i].n_tris = s["first_tri"], s["n_tris"]
            shapes[i].material, shapes[i].emitter = s["material"], s["emitter"]
            shapes[i].has_normals, shapes[i].has_texcoords = int(s["has_normals"]), int(s["has_texcoords"])
        mats = (abi.Material * max(1, len(self.materials)))(*self.materials)
        ems = (abi.Emitter * max(1, len(self.emitters)))(*self.emitters)
        d = abi.SceneDesc()
        d.abi_version = abi.ABI_VERSION
        d.n_shapes, d.n_materials, d.n_emitters = len(self.shapes), len(self.materials), len(self.emitters)
        d.n_vertices, d.n_tris = pos.shape[0], idx.shape[0]
        d.shapes, d.materials, d.emitters = shapes, mats, ems
        d.positions = abi.ptr(pos, C.c_float)
        d.normals = abi.ptr(nrm, C.c_float)
        d.texcoords = abi.ptr(uv, C.c_float)
        d.indices = abi.ptr(idx, C.c_uint32)
        d.sensor = self.sensor
        d.integrator = self.integrator
        d.sampler = self.sampler
        if getattr(self, "envmap", None) is not None:
            d.envmap = C.pointer(self.envmap)
        texs = None
        if getattr(self, "textures", None):
            texs = (abi.Texture * len(self.textures))(*self.textures)
            d.textures, d.n_textures = texs, len(self.textures)
        keep = (pos, nrm, uv, idx, shapes, mats, ems, texs)
        return d, keep
