"""ORACLE -- TEST INFRASTRUCTURE ONLY.

ctypes bindings of oracle/liboracle.so (CPU restatement of the reference `path` hot path)
and, when it was built, oracle/_ref/libmtsref.so (kd-tree / TriAccel / microfacet / warp code
compiled from the reference tree itself).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / `--impl reference` legs may import this package; the product
package `mitsuba_b200` never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from mitsuba_b200 import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")
_REF = os.path.join(_HERE, "_ref", "libmtsref.so")


def build(ref=True):
    """Compile the oracle (and oracle/_ref when the reference tree is present)."""
    subprocess.check_call(["make", "-C", _HERE, "all"] + (["ref"] if ref else []), stdout=subprocess.DEVNULL)


def _host_cpu():
    """(model name, has AVX-512F, has AVX2+FMA) of the machine this process runs on."""
    model, flags = "unknown", set()
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and model == "unknown":
                model = line.split(":", 1)[1].strip()
            elif line.startswith("flags"):
                flags = set(line.split(":", 1)[1].split())
                break
    except OSError:
        pass
    return model, "avx512f" in flags and "avx512vl" in flags and "avx512bw" in flags and "avx512dq" in flags, "avx2" in flags and "fma" in flags


_fast_flags = None


def use_timed_build():
    """Switch this process to the TIMED builds (bench.py's CPU legs only; parity always uses the strict build):
    liboracle_fast.so compiled on this machine with the reference's release flags at -march=native, and the
    prebuilt _ref/libmtsref_fast_{v3,v4}.so this host can run.  Must be called before the first lib() / ref().
    Returns a description of the flags for the bench line."""
    global _LIB, _REF, _fast_flags
    if _fast_flags is not None:
        return _fast_flags
    if _lib is not None or _ref is not None:
        raise RuntimeError("oracle.use_timed_build() must run before the strict libraries are loaded")
    model, v4, v3 = _host_cpu()
    stamp = os.path.join(_HERE, "liboracle_fast.host")
    fast = os.path.join(_HERE, "liboracle_fast.so")
    same_host = os.path.exists(stamp) and os.path.exists(fast) and open(stamp).read() == model + ("/v4" if v4 else "/v3" if v3 else "/v2")
    subprocess.check_call(["make", "-C", _HERE] + ([] if same_host else ["-B"]) + ["fast"], stdout=subprocess.DEVNULL)
    open(stamp, "w").write(model + ("/v4" if v4 else "/v3" if v3 else "/v2"))
    _LIB = fast
    out = {"oracle": open(os.path.join(_HERE, "liboracle_fast.flags")).read().strip(), "host_cpu": model}
    for lvl, ok in (("v4", v4), ("v3", v3)):
        cand = os.path.join(_HERE, "_ref", "libmtsref_fast_%s.so" % lvl)
        if ok and os.path.exists(cand):
            _REF = cand
            out["reference_kdtree"] = open(cand[:-3] + ".flags").read().strip()
            break
    else:
        if os.path.exists(_REF):
            out["reference_kdtree"] = "-O2 -ffp-contract=off (strict build: no release build for this host)"
    _fast_flags = out
    return out


class OrcStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("shade_events", C.c_uint64), ("path_length_sum", C.c_uint64), ("bad_samples", C.c_uint64),
                ("lum_sum", C.c_double), ("lum_sq_sum", C.c_double), ("seconds", C.c_double), ("threads", C.c_int)]


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build(ref=False)
        L = C.CDLL(_LIB)
        L.orc_last_error.restype = C.c_char_p
        L.orc_scene_create.argtypes = [C.POINTER(abi.SceneDesc), C.POINTER(C.c_void_p)]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_scene_set_external_accel.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int]
        L.orc_bsdf_query.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_emitter_sample.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_fill_intersection.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_sensor_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_film_clear.argtypes = [C.c_void_p]
        L.orc_film_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_film_put.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_film_download.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_film_develop.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_render.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_int, C.POINTER(OrcStats)]
        L.orc_render2.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_int, C.POINTER(OrcStats), C.c_int, C.c_int]
        L.orc_li.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p]
        L.orc_texture_create.restype = C.c_void_p
        L.orc_texture_create.argtypes = [C.POINTER(abi.Texture)]
        L.orc_texture_destroy.argtypes = [C.c_void_p]
        L.orc_texture_levels.argtypes = [C.c_void_p]
        L.orc_texture_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_texture_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_texture_eval.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_compute_partials.argtypes = [C.c_size_t, C.c_void_p, C.c_void_p]
        L.orc_sensor_differentials.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.orc_scene_bounds.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_triaccel.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_erf.argtypes = [C.c_float]
        L.orc_erf.restype = C.c_float
        L.orc_erfinv.argtypes = [C.c_float]
        L.orc_erfinv.restype = C.c_float
        L.orc_fresnel_dielectric_ext.argtypes = [C.c_float, C.c_float, C.POINTER(C.c_float)]
        L.orc_fresnel_dielectric_ext.restype = C.c_float
        L.orc_fresnel_conductor_exact.argtypes = [C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_microfacet.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_warp.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.orc_float_to_half.argtypes = [C.c_float]
        L.orc_float_to_half.restype = C.c_uint16
        L.orc_half_to_float.argtypes = [C.c_uint16]
        L.orc_half_to_float.restype = C.c_float
        L.orc_square_to_tent.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_envmap_query.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_halton_sequence.argtypes = [C.c_int] * 9 + [C.c_void_p]
        L.orc_sobol_sequence.argtypes = [C.c_int] * 9 + [C.c_void_p]
        L.orc_sobol_matrix32.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_sobol_matrix32.restype = C.c_uint32
        L.orc_sobol_vdc.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_sobol_sample.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32]
        L.orc_sobol_sample.restype = C.c_float
        L.orc_radical_inverse.argtypes = [C.c_int, C.c_uint64]
        L.orc_radical_inverse.restype = C.c_float
        L.orc_radical_inverse_incremental.argtypes = [C.c_int, C.c_float]
        L.orc_radical_inverse_incremental.restype = C.c_float
        L.orc_radical_inverse_fast.argtypes = [C.c_int, C.c_uint64]
        L.orc_radical_inverse_fast.restype = C.c_float
        L.orc_scrambled_radical_inverse_fast.argtypes = [C.c_int, C.c_uint64, C.c_void_p]
        L.orc_scrambled_radical_inverse_fast.restype = C.c_float
        L.orc_permutation.argtypes = [C.c_int, C.c_int, C.c_void_p]
        L.orc_sample02.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_sample_tea.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        L.orc_sample_tea.restype = C.c_uint64
        L.orc_ldsampler_sets.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def ref_available():
    return os.path.exists(_REF)


def ref():
    """The library compiled from the reference tree (None when it was never built)."""
    global _ref
    if _ref is None:
        if not ref_available():
            return None
        R = C.CDLL(_REF)
        R.ref_kd_build.argtypes = [C.POINTER(abi.SceneDesc), C.c_int]
        R.ref_kd_build.restype = C.c_void_p
        R.ref_kd_destroy.argtypes = [C.c_void_p]
        R.ref_kd_bounds.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        R.ref_kd_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        R.ref_all_hits.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_size_t]
        R.ref_all_hits.restype = C.c_size_t
        R.ref_triaccel_load.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        R.ref_triaccel_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        R.ref_clipped_aabb.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        for f in ("ref_erf", "ref_erfinv"):
            getattr(R, f).argtypes = [C.c_float]
            getattr(R, f).restype = C.c_float
        R.ref_hypot2.argtypes = [C.c_float, C.c_float]
        R.ref_hypot2.restype = C.c_float
        R.ref_microfacet.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        R.ref_microfacet_eval.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p]
        R.ref_warp.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        R.ref_set_log_level.argtypes = [C.c_int]
        R.ref_set_core_count.argtypes = [C.c_int]
        _ref = R
    return _ref


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


class RefKDTree:
    """The reference's own SAH kd-tree + TriAccel, built over a scene description."""

    def __init__(self, builder_or_desc, verbose=False):
        R = ref()
        if R is None:
            raise RuntimeError("oracle/_ref/libmtsref.so was not built (reference tree absent)")
        self._desc, self._keep = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        self.h = R.ref_kd_build(C.byref(self._desc), int(verbose))

    def trace(self, rays, mode=abi.TRACE_CLOSEST):
        rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=abi.HIT_DTYPE)
        ref().ref_kd_trace(self.h, _vp(rays), rays.shape[0], mode, _vp(hits))
        return hits

    def bounds(self):
        t, e = np.zeros(6, np.float32), np.zeros(6, np.float32)
        ref().ref_kd_bounds(self.h, _vp(t), _vp(e))
        return t, e

    def all_hits(self, ray, mint, maxt, cap=64):
        ray = np.ascontiguousarray(ray, dtype=abi.RAY_DTYPE).reshape(1)
        out = np.zeros(cap, dtype=abi.HIT_DTYPE)
        n = ref().ref_all_hits(self.h, _vp(ray), mint, maxt, _vp(out), cap)
        return out[:min(n, cap)], n

    def __del__(self):
        try:
            if self.h:
                ref().ref_kd_destroy(self.h)
                self.h = None
        except Exception:
            pass


class Oracle:
    """CPU restatement of the hot path over one scene (same entry points as the product ABI)."""

    def __init__(self, builder_or_desc, use_ref_kdtree=False):
        L = lib()
        self._desc, self._keep = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        h = C.c_void_p()
        rc = L.orc_scene_create(C.byref(self._desc), C.byref(h))
        if rc != 0:
            raise RuntimeError("oracle: " + L.orc_last_error().decode())
        self.h = h
        self.kd = None
        if use_ref_kdtree:
            self.kd = RefKDTree((self._desc, self._keep))
            fn = C.cast(ref().ref_kd_search, C.c_void_p)
            L.orc_scene_set_external_accel(self.h, fn, self.kd.h)

    def __del__(self):
        try:
            if self.h:
                lib().orc_scene_set_external_accel(self.h, None, None)
                lib().orc_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def trace(self, rays, mode=abi.TRACE_CLOSEST, threads=0):
        rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=abi.HIT_DTYPE)
        lib().orc_trace(self.h, _vp(rays), rays.shape[0], mode, _vp(hits), threads)
        return hits

    def bsdf_query(self, material, q):
        q = np.ascontiguousarray(q, dtype=abi.BSDF_QUERY_DTYPE)
        out = np.zeros(q.shape[0], dtype=abi.BSDF_RESULT_DTYPE)
        rc = lib().orc_bsdf_query(self.h, material, _vp(q), q.shape[0], _vp(out))
        if rc:
            raise RuntimeError(lib().orc_last_error().decode())
        return out

    def emitter_sample(self, q):
        q = np.ascontiguousarray(q, dtype=abi.DIRECT_QUERY_DTYPE)
        out = np.zeros(q.shape[0], dtype=abi.DIRECT_RESULT_DTYPE)
        lib().orc_emitter_sample(self.h, _vp(q), q.shape[0], _vp(out))
        return out

    def envmap_query(self, kind, q):
        """kind 0: evalEnvironment(direction); 1: internalSampleDirection(u) -> (d, value, pdf); 2: internalPdfDirection(d)."""
        q = np.ascontiguousarray(q, dtype=np.float32).reshape(-1, 3)
        out = np.zeros((q.shape[0], 7), np.float32)
        if lib().orc_envmap_query(self.h, kind, _vp(q), q.shape[0], _vp(out)):
            raise RuntimeError(lib().orc_last_error().decode())
        return out

    def fill_intersection(self, rays, hits):
        rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
        hits = np.ascontiguousarray(hits, dtype=abi.HIT_DTYPE)
        out = np.zeros((rays.shape[0], 17), np.float32)
        lib().orc_fill_intersection(self.h, _vp(rays), _vp(hits), rays.shape[0], _vp(out))
        return out

    def sensor_rays(self, pos):
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(-1, 2)
        rays = np.zeros(pos.shape[0], dtype=abi.RAY_DTYPE)
        lib().orc_sensor_rays(self.h, _vp(pos), pos.shape[0], _vp(rays))
        return rays

    def film_size(self):
        w, h, b = C.c_int(), C.c_int(), C.c_int()
        lib().orc_film_size(self.h, C.byref(w), C.byref(h), C.byref(b))
        return w.value, h.value, b.value

    def film_clear(self):
        lib().orc_film_clear(self.h)

    def film_put(self, samples6):
        s = np.ascontiguousarray(samples6, dtype=np.float32).reshape(-1, 6)
        lib().orc_film_put(self.h, _vp(s), s.shape[0])

    def film_download(self):
        w, h, b = self.film_size()
        out = np.zeros((h + 2 * b, w + 2 * b, 5), np.float32)
        lib().orc_film_download(self.h, _vp(out))
        return out

    def film_develop(self):
        w, h, _ = self.film_size()
        out = np.zeros((h, w, 3), np.float32)
        lib().orc_film_develop(self.h, _vp(out))
        return out

    def render(self, spp=0, seed=1, threads=0, max_blocks=0, clear=True, counter_rng=False, spp_offset=0):
        """counter_rng=True draws every sample from the counter-based stream keyed by
        (seed, pixel, spp_offset + j) -- the numbering the CUDA renderer uses."""
        if clear:
            self.film_clear()
        st = OrcStats()
        lib().orc_render2(self.h, spp, seed, threads, max_blocks, C.byref(st), int(counter_rng), spp_offset)
        return st

    def li(self, sample_pos, seed=1):
        p = np.ascontiguousarray(sample_pos, dtype=np.float32).reshape(-1, 2)
        out = np.zeros((p.shape[0], 4), np.float32)
        lib().orc_li(self.h, _vp(p), p.shape[0], seed, _vp(out))
        return out

    def bounds(self):
        t, e, s = np.zeros(6, np.float32), np.zeros(6, np.float32), np.zeros(4, np.float32)
        lib().orc_scene_bounds(self.h, _vp(t), _vp(e), _vp(s))
        return t, e, s

    def triaccel(self, prim):
        out = np.zeros(12, np.uint32)
        lib().orc_triaccel(self.h, prim, _vp(out))
        return out
