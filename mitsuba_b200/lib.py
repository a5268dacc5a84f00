"""ctypes binding of mitsuba_b200/libmtsb200.so -- the C ABI declared in include/mtsb200.h.

There is no fallback of any kind: if the CUDA library has not been built, or no sm_100-class
GPU is present, the calls raise.  (The CPU oracle under oracle/ is test infrastructure and is
never imported from here.)
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# MTSB200_LIB: load another build of the same library (kernel experiments, tools/build_variant.sh)
LIB_PATH = os.environ.get("MTSB200_LIB") or os.path.join(_HERE, "libmtsb200.so")
CSRC = os.path.join(_HERE, "csrc")

# every symbol include/mtsb200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "mtsb_abi_version", "mtsb_last_error", "mtsb_init", "mtsb_device_count", "mtsb_scene_create", "mtsb_scene_info",
    "mtsb_scene_destroy", "mtsb_trace", "mtsb_trace_device", "mtsb_bsdf_query", "mtsb_emitter_sample",
    "mtsb_fill_intersection", "mtsb_sensor_rays", "mtsb_film_clear", "mtsb_film_put", "mtsb_film_size",
    "mtsb_film_download", "mtsb_film_device_ptr", "mtsb_film_develop", "mtsb_render", "mtsb_cancel",
    "mtsb_gen_rays_device", "mtsb_multi_create", "mtsb_multi_render", "mtsb_multi_cancel", "mtsb_multi_scene",
    "mtsb_multi_device_count", "mtsb_multi_destroy", "mtsb_trace_static", "mtsb_trace_static_max_rays",
]


class MtsbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("mtsb200 error %d: %s" % (code, msg))
        self.code = code


def build(verbose=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", CSRC, "-j8"], stdout=None if verbose else subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MtsbError(abi.ERR_NO_DEVICE, "libmtsb200.so has not been built (run `python -c 'import __graft_entry__ as g; "
                                               "g.build()'`); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.mtsb_last_error.restype = C.c_char_p
        L.mtsb_init.argtypes = [C.c_int]
        L.mtsb_scene_create.argtypes = [C.POINTER(abi.SceneDesc), C.POINTER(C.c_void_p)]
        L.mtsb_scene_info.argtypes = [C.c_void_p, C.POINTER(abi.SceneInfo)]
        L.mtsb_scene_destroy.argtypes = [C.c_void_p]
        L.mtsb_scene_destroy.restype = None
        L.mtsb_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.mtsb_trace_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                                        C.POINTER(C.c_float)]
        L.mtsb_bsdf_query.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.mtsb_emitter_sample.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.mtsb_fill_intersection.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.mtsb_sensor_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.mtsb_film_clear.argtypes = [C.c_void_p]
        L.mtsb_film_put.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.mtsb_film_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.mtsb_film_download.argtypes = [C.c_void_p, C.c_void_p]
        L.mtsb_film_device_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        L.mtsb_film_develop.argtypes = [C.c_void_p, C.c_void_p]
        L.mtsb_render.argtypes = [C.c_void_p, C.POINTER(abi.RenderParams), C.POINTER(abi.Stats)]
        L.mtsb_cancel.argtypes = [C.c_void_p]
        L.mtsb_gen_rays_device.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_uint64, C.c_void_p]
        L.mtsb_trace_static.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.mtsb_multi_create.argtypes = [C.POINTER(abi.SceneDesc), C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p)]
        L.mtsb_multi_render.argtypes = [C.c_void_p, C.POINTER(abi.RenderParams), C.POINTER(abi.Stats)]
        L.mtsb_multi_cancel.argtypes = [C.c_void_p]
        L.mtsb_multi_scene.argtypes = [C.c_void_p, C.c_int]
        L.mtsb_multi_scene.restype = C.c_void_p
        L.mtsb_multi_device_count.argtypes = [C.c_void_p]
        L.mtsb_multi_destroy.argtypes = [C.c_void_p]
        L.mtsb_multi_destroy.restype = None
        _lib = L
    return _lib


def _check(rc):
    if rc != abi.OK:
        raise MtsbError(rc, lib().mtsb_last_error().decode(errors="replace"))


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def init(device=-1):
    _check(lib().mtsb_init(device))


def device_count():
    return lib().mtsb_device_count()


class Scene:
    """A scene resident on one GPU (mtsb_scene_t).  `builder_or_desc` is a SceneBuilder or the
    (MtsbSceneDesc, keepalive) pair its desc() returns."""

    def __init__(self, builder_or_desc):
        L = lib()
        self._desc, self._keep = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        h = C.c_void_p()
        _check(L.mtsb_scene_create(C.byref(self._desc), C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().mtsb_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        i = abi.SceneInfo()
        _check(lib().mtsb_scene_info(self.h, C.byref(i)))
        return i

    # -- ray queries (ShapeKDTree::rayIntersect) --------------------------------------
    def trace(self, rays, mode=abi.TRACE_CLOSEST):
        rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=abi.HIT_DTYPE)
        _check(lib().mtsb_trace(self.h, _vp(rays), rays.shape[0], mode, _vp(hits)))
        return hits

    def trace_into(self, rays_ptr, n, hits_ptr, mode=abi.TRACE_CLOSEST):
        """mtsb_trace on caller-owned host buffers given by address (e.g. pinned memory): no allocation in the binding."""
        _check(lib().mtsb_trace(self.h, C.c_void_p(rays_ptr), int(n), mode, C.c_void_p(hits_ptr)))

    def trace_static(self, rays, mode=abi.TRACE_CLOSEST):
        """mtsb_trace_static: the same query with all scratch in a fixed device array (no allocation; n <= 262144)."""
        rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=abi.HIT_DTYPE)
        _check(lib().mtsb_trace_static(self.h, _vp(rays), rays.shape[0], mode, _vp(hits)))
        return hits

    def trace_device(self, d_rays_ptr, n, d_hits_ptr, mode=abi.TRACE_CLOSEST, stream=None, timed=True):
        ms = C.c_float(0.0)
        _check(lib().mtsb_trace_device(self.h, C.c_void_p(d_rays_ptr), n, mode, C.c_void_p(d_hits_ptr),
                                       C.c_void_p(stream) if stream else None, C.byref(ms) if timed else None))
        return ms.value

    def gen_rays_device(self, kind, n, seed, d_rays_ptr):
        _check(lib().mtsb_gen_rays_device(self.h, kind, n, seed, C.c_void_p(d_rays_ptr)))

    # -- record-level queries (BSDF / Emitter / Intersection / Sensor) -----------------
    def bsdf_query(self, material, q):
        q = np.ascontiguousarray(q, dtype=abi.BSDF_QUERY_DTYPE)
        out = np.zeros(q.shape[0], dtype=abi.BSDF_RESULT_DTYPE)
        _check(lib().mtsb_bsdf_query(self.h, material, _vp(q), q.shape[0], _vp(out)))
        return out

    def emitter_sample(self, q):
        q = np.ascontiguousarray(q, dtype=abi.DIRECT_QUERY_DTYPE)
        out = np.zeros(q.shape[0], dtype=abi.DIRECT_RESULT_DTYPE)
        _check(lib().mtsb_emitter_sample(self.h, _vp(q), q.shape[0], _vp(out)))
        return out

    def fill_intersection(self, rays, hits):
        rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
        hits = np.ascontiguousarray(hits, dtype=abi.HIT_DTYPE)
        out = np.zeros((rays.shape[0], 17), np.float32)
        _check(lib().mtsb_fill_intersection(self.h, _vp(rays), _vp(hits), rays.shape[0], _vp(out)))
        return out

    def sensor_rays(self, pos):
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(-1, 2)
        rays = np.zeros(pos.shape[0], dtype=abi.RAY_DTYPE)
        _check(lib().mtsb_sensor_rays(self.h, _vp(pos), pos.shape[0], _vp(rays)))
        return rays

    # -- film (ImageBlock / HDRFilm) ------------------------------------------------------
    def film_size(self):
        w, h, b = C.c_int(), C.c_int(), C.c_int()
        _check(lib().mtsb_film_size(self.h, C.byref(w), C.byref(h), C.byref(b)))
        return w.value, h.value, b.value

    def film_clear(self):
        _check(lib().mtsb_film_clear(self.h))

    def film_put(self, samples6):
        s = np.ascontiguousarray(samples6, dtype=np.float32).reshape(-1, 6)
        _check(lib().mtsb_film_put(self.h, _vp(s), s.shape[0]))

    def film_download(self):
        w, h, b = self.film_size()
        out = np.zeros((h + 2 * b, w + 2 * b, 5), np.float32)
        _check(lib().mtsb_film_download(self.h, _vp(out)))
        return out

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(lib().mtsb_film_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def film_develop(self):
        w, h, _ = self.film_size()
        out = np.zeros((h, w, 3), np.float32)
        _check(lib().mtsb_film_develop(self.h, _vp(out)))
        return out

    # -- the integrator (SamplingIntegrator::render) ----------------------------------------
    def render(self, spp=0, seed=1, spp_offset=0, tile_index=0, tile_count=0, spp_per_pass=0, keep_film=False):
        p = abi.RenderParams()
        p.spp, p.spp_offset, p.seed = int(spp), int(spp_offset), int(seed)
        p.tile_index, p.tile_count, p.spp_per_pass = int(tile_index), int(tile_count), int(spp_per_pass)
        p.flags = abi.RENDER_KEEP_FILM if keep_film else 0
        st = abi.Stats()
        _check(lib().mtsb_render(self.h, C.byref(p), C.byref(st)))
        return st

    def cancel(self):
        _check(lib().mtsb_cancel(self.h))


class _SceneView(Scene):
    """Scene methods on a handle that somebody else owns (a replica inside a MultiScene)."""

    def __init__(self, handle, owner):
        self.h, self._owner = handle, owner

    def close(self):
        self.h = None


class MultiScene:
    """One scene replicated over several GPUs of this node inside the library (mtsb_multi_t): one frame's tiles are
    interleaved over the devices and the films are merged by one NCCL reduce onto the first device."""

    def __init__(self, builder_or_desc, n_devices, device_ids=None):
        L = lib()
        self._desc, self._keep = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        ids = (C.c_int * n_devices)(*device_ids) if device_ids is not None else None
        h = C.c_void_p()
        _check(L.mtsb_multi_create(C.byref(self._desc), int(n_devices), ids, C.byref(h)))
        self.h = h

    def device_count(self):
        return lib().mtsb_multi_device_count(self.h)

    def scene(self, i=0):
        """A non-owning Scene view of the replica on the i-th device (i = 0 holds the merged film)."""
        hh = lib().mtsb_multi_scene(self.h, int(i))
        if not hh:
            raise MtsbError(abi.ERR_INVALID, "no such device")
        return _SceneView(C.c_void_p(hh), self)

    def render(self, spp=0, seed=1, spp_offset=0, spp_per_pass=0, keep_film=False):
        p = abi.RenderParams()
        p.spp, p.spp_offset, p.seed, p.spp_per_pass = int(spp), int(spp_offset), int(seed), int(spp_per_pass)
        p.flags = abi.RENDER_KEEP_FILM if keep_film else 0
        st = abi.Stats()
        _check(lib().mtsb_multi_render(self.h, C.byref(p), C.byref(st)))
        return st

    def film_develop(self):
        return self.scene(0).film_develop()

    def cancel(self):
        _check(lib().mtsb_multi_cancel(self.h))

    def close(self):
        if getattr(self, "h", None):
            lib().mtsb_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
