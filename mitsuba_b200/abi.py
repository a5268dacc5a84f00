"""ctypes mirror of include/mtsb200.h (the C ABI of the B200 `path` hot path).

Only layout definitions live here -- no compute.  Both the product binding
(`mitsuba_b200.lib`) and the test-only oracle binding (`oracle/`) describe scenes
with these structures so that they consume byte-identical inputs.
"""
import ctypes as C

import numpy as np

ABI_VERSION = 6

OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_CANCELLED, ERR_IO, ERR_PARSE = range(7)
BSDF_DIFFUSE, BSDF_ROUGHCONDUCTOR, BSDF_ROUGHDIELECTRIC, BSDF_NULL, BSDF_DIELECTRIC, BSDF_CONDUCTOR, BSDF_PLASTIC = range(7)
DISTR_BECKMANN, DISTR_GGX = 0, 1
MAT_TWOSIDED, MAT_SAMPLE_VISIBLE, MAT_NONLINEAR = 1, 2, 4
EMITTER_AREA, EMITTER_POINT, EMITTER_CONSTANT, EMITTER_ENVMAP = 0, 1, 2, 3
RFILTER_GAUSSIAN, RFILTER_BOX = 0, 1
TRACE_CLOSEST, TRACE_ANY = 0, 1
NO_HIT = 0xFFFFFFFF
RENDER_KEEP_FILM = 1
SAMPLER_INDEPENDENT, SAMPLER_LDSAMPLER, SAMPLER_HALTON, SAMPLER_SOBOL = 0, 1, 2, 3
TEX_BITMAP, TEX_CHECKERBOARD, TEX_GRID = 0, 1, 2
TEXFILTER_NEAREST, TEXFILTER_BILINEAR, TEXFILTER_TRILINEAR, TEXFILTER_EWA = 0, 1, 2, 3
WRAP_CLAMP, WRAP_REPEAT, WRAP_MIRROR, WRAP_ZERO, WRAP_ONE = 0, 1, 2, 3, 4


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("flags", C.c_uint32), ("distribution", C.c_uint32), ("back", C.c_int32),
                ("alpha_u", C.c_float), ("alpha_v", C.c_float), ("eta", C.c_float), ("fdr_int", C.c_float),
                ("reflectance", C.c_float * 3), ("transmittance", C.c_float * 3),
                ("cond_eta", C.c_float * 3), ("cond_k", C.c_float * 3), ("spec_weight", C.c_float),
                ("tex_reflectance", C.c_uint32), ("tex_transmittance", C.c_uint32), ("tex_alpha", C.c_uint32)]


class Shape(C.Structure):
    _fields_ = [("first_vertex", C.c_uint32), ("n_vertices", C.c_uint32), ("first_tri", C.c_uint32),
                ("n_tris", C.c_uint32), ("material", C.c_int32), ("emitter", C.c_int32),
                ("has_normals", C.c_uint32), ("has_texcoords", C.c_uint32)]


class Emitter(C.Structure):
    _fields_ = [("type", C.c_uint32), ("shape", C.c_int32), ("radiance", C.c_float * 3),
                ("position", C.c_float * 3), ("sampling_weight", C.c_float), ("_pad", C.c_float * 3)]


class Sensor(C.Structure):
    _fields_ = [("to_world", C.c_float * 16), ("fov_x_deg", C.c_float), ("near_clip", C.c_float),
                ("far_clip", C.c_float), ("film_width", C.c_int32), ("film_height", C.c_int32),
                ("crop_x", C.c_int32), ("crop_y", C.c_int32), ("crop_width", C.c_int32), ("crop_height", C.c_int32),
                ("rfilter", C.c_uint32), ("rfilter_radius", C.c_float), ("rfilter_stddev", C.c_float)]


class Integrator(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("strict_normals", C.c_int32),
                ("hide_emitters", C.c_int32), ("sample_count", C.c_int32), ("_pad", C.c_uint32)]


class Sampler(C.Structure):
    _fields_ = [("type", C.c_uint32), ("dimension", C.c_int32), ("scramble", C.c_int32), ("_pad", C.c_uint32)]


class Envmap(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb", C.POINTER(C.c_float)), ("to_world", C.c_float * 16),
                ("to_local", C.c_float * 16), ("scale", C.c_float), ("_pad", C.c_uint32 * 3)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("filter", C.c_uint32), ("wrap_u", C.c_uint32), ("wrap_v", C.c_uint32),
                ("uv_scale", C.c_float * 2), ("uv_offset", C.c_float * 2), ("color0", C.c_float * 3), ("color1", C.c_float * 3),
                ("line_width", C.c_float), ("scale", C.c_float * 3), ("max_anisotropy", C.c_float),
                ("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32), ("texels", C.POINTER(C.c_float))]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("n_shapes", C.c_uint32), ("n_materials", C.c_uint32),
                ("n_emitters", C.c_uint32), ("n_vertices", C.c_uint64), ("n_tris", C.c_uint64),
                ("shapes", C.POINTER(Shape)), ("materials", C.POINTER(Material)), ("emitters", C.POINTER(Emitter)),
                ("positions", C.POINTER(C.c_float)), ("normals", C.POINTER(C.c_float)),
                ("texcoords", C.POINTER(C.c_float)), ("indices", C.POINTER(C.c_uint32)),
                ("sensor", Sensor), ("integrator", Integrator), ("sampler", Sampler),
                ("envmap", C.POINTER(Envmap)), ("textures", C.POINTER(Texture)), ("n_textures", C.c_uint32), ("_pad", C.c_uint32)]


class RenderParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("spp_offset", C.c_int32), ("seed", C.c_uint64),
                ("tile_index", C.c_int32), ("tile_count", C.c_int32), ("spp_per_pass", C.c_int32),
                ("flags", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("shade_events", C.c_uint64), ("path_length_sum", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("bad_samples", C.c_uint64), ("ms_total", C.c_float), ("ms_trace_closest", C.c_float),
                ("ms_trace_any", C.c_float), ("ms_shade", C.c_float), ("ms_raygen", C.c_float),
                ("ms_splat", C.c_float), ("ms_h2d", C.c_float), ("ms_d2h", C.c_float), ("n_passes", C.c_uint32),
                ("closest_launches", C.c_uint32), ("any_launches", C.c_uint32), ("n_devices", C.c_uint32),
                ("ms_reduce", C.c_float), ("_pad", C.c_uint32),
                ("closest_rays_bounce", C.c_uint64 * 16), ("shadow_rays_bounce", C.c_uint64 * 16)]


class SceneInfo(C.Structure):
    _fields_ = [("n_tris", C.c_uint64), ("n_bvh_nodes", C.c_uint64), ("bvh_bytes", C.c_uint64),
                ("tri_bytes", C.c_uint64), ("aabb_min", C.c_float * 3), ("aabb_max", C.c_float * 3),
                ("build_ms", C.c_float), ("n_emitters", C.c_uint32)]


# numpy views of the array-of-struct records that cross the ABI
RAY_DTYPE = np.dtype([("o", np.float32, 3), ("mint", np.float32), ("d", np.float32, 3), ("maxt", np.float32)])
HIT_DTYPE = np.dtype([("prim", np.uint32), ("u", np.float32), ("v", np.float32), ("t", np.float32)])
BSDF_QUERY_DTYPE = np.dtype([("wi", np.float32, 3), ("wo", np.float32, 3), ("u", np.float32, 3)])
BSDF_RESULT_DTYPE = np.dtype([("eval", np.float32, 3), ("pdf", np.float32), ("weight", np.float32, 3),
                              ("spdf", np.float32), ("swo", np.float32, 3), ("seta", np.float32)])
DIRECT_QUERY_DTYPE = np.dtype([("ref", np.float32, 3), ("ref_n", np.float32, 3), ("u", np.float32, 2)])
DIRECT_RESULT_DTYPE = np.dtype([("value", np.float32, 3), ("pdf", np.float32), ("d", np.float32, 3),
                                ("dist", np.float32), ("p", np.float32, 3), ("emitter", np.int32),
                                ("n", np.float32, 3), ("measure", np.int32)])
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 16
assert BSDF_QUERY_DTYPE.itemsize == 36 and BSDF_RESULT_DTYPE.itemsize == 48
assert DIRECT_QUERY_DTYPE.itemsize == 32 and DIRECT_RESULT_DTYPE.itemsize == 64
assert C.sizeof(Material) == 96 and C.sizeof(Shape) == 32 and C.sizeof(Emitter) == 48


def ptr(arr, ctype):
    """ctypes pointer to a C-contiguous numpy array (or NULL)."""
    if arr is None:
        return C.cast(None, C.POINTER(ctype))
    assert arr.flags["C_CONTIGUOUS"]
    return arr.ctypes.data_as(C.POINTER(ctype))
