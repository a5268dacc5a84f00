"""ctypes binding of mitsuba_b200/libmtsb200host.so (include/mtsb200_host.h): the C++ scene loader
and plugin mirror in mitsuba_b200/host/.  Loading and flattening run on the CPU (they are host
logic, like the reference's SceneHandler); rendering goes through libmtsb200.so and needs the GPU.
"""
import ctypes as C
import os

import numpy as np

from . import abi
from .lib import MtsbError

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmtsb200host.so")
CLI_PATH = os.path.join(_HERE, "mtsb200")

SYMBOLS = [
    "mtsbh_last_error", "mtsbh_load_file", "mtsbh_load_string", "mtsbh_scene_free", "mtsbh_scene_desc", "mtsbh_scene_seed",
    "mtsbh_warning_count", "mtsbh_warning", "mtsbh_render", "mtsbh_render_multi", "mtsbh_write_pfm",
]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MtsbError(abi.ERR_IO, "libmtsb200host.so has not been built (make -C mitsuba_b200/host)")
        L = C.CDLL(LIB_PATH)
        L.mtsbh_last_error.restype = C.c_char_p
        L.mtsbh_load_file.argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.POINTER(C.c_void_p)]
        L.mtsbh_load_string.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p), C.POINTER(C.c_void_p)]
        L.mtsbh_scene_free.argtypes = [C.c_void_p]
        L.mtsbh_scene_free.restype = None
        L.mtsbh_scene_desc.argtypes = [C.c_void_p]
        L.mtsbh_scene_desc.restype = C.POINTER(abi.SceneDesc)
        L.mtsbh_scene_seed.argtypes = [C.c_void_p]
        L.mtsbh_warning_count.argtypes = [C.c_void_p]
        L.mtsbh_warning.argtypes = [C.c_void_p, C.c_int]
        L.mtsbh_warning.restype = C.c_char_p
        L.mtsbh_render.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(abi.Stats)]
        L.mtsbh_render_multi.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(abi.Stats)]
        L.mtsbh_write_pfm.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
        _lib = L
    return _lib


def _defines(params):
    if not params:
        return None
    items = [("%s=%s" % (k, v)).encode() for k, v in params.items()]
    arr = (C.c_char_p * (len(items) + 1))(*items, None)
    return arr


class HostScene:
    """A loaded scene (mtsbh_scene_t): the configured plugin graph flattened to MtsbSceneDesc."""

    def __init__(self, handle):
        self.h = handle
        L = lib()
        self.desc_ptr = L.mtsbh_scene_desc(self.h)
        self.seed = L.mtsbh_scene_seed(self.h)
        self.warnings = [L.mtsbh_warning(self.h, i).decode() for i in range(L.mtsbh_warning_count(self.h))]

    @classmethod
    def load_file(cls, path, params=None):
        h = C.c_void_p()
        rc = lib().mtsbh_load_file(os.fspath(path).encode(), _defines(params), C.byref(h))
        if rc:
            raise MtsbError(rc, lib().mtsbh_last_error().decode(errors="replace"))
        return cls(h)

    @classmethod
    def load_string(cls, xml, name="<string>", params=None):
        h = C.c_void_p()
        rc = lib().mtsbh_load_string(xml.encode(), name.encode(), _defines(params), C.byref(h))
        if rc:
            raise MtsbError(rc, lib().mtsbh_last_error().decode(errors="replace"))
        return cls(h)

    def close(self):
        if getattr(self, "h", None):
            lib().mtsbh_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def desc(self):
        return self.desc_ptr.contents

    def desc_pair(self):
        """(MtsbSceneDesc, keepalive) as mitsuba_b200.lib.Scene takes it."""
        return self.desc, self

    # -- numpy views of the flattened arrays (for tests / tools) -----------------------------
    def arrays(self):
        d = self.desc
        nv, nt = int(d.n_vertices), int(d.n_tris)
        pos = np.ctypeslib.as_array(d.positions, shape=(nv, 3)).copy() if nv else np.zeros((0, 3), np.float32)
        idx = np.ctypeslib.as_array(d.indices, shape=(nt, 3)).copy() if nt else np.zeros((0, 3), np.uint32)
        nrm = np.ctypeslib.as_array(d.normals, shape=(nv, 3)).copy() if (nv and d.normals) else None
        uv = np.ctypeslib.as_array(d.texcoords, shape=(nv, 2)).copy() if (nv and d.texcoords) else None
        return pos, nrm, uv, idx

    def shapes(self):
        d = self.desc
        return [d.shapes[i] for i in range(d.n_shapes)]

    def materials(self):
        d = self.desc
        return [d.materials[i] for i in range(d.n_materials)]

    def emitters(self):
        d = self.desc
        return [d.emitters[i] for i in range(d.n_emitters)]

    def render(self, spp=0, gpus=1):
        """RenderJob: device scene + mtsb_render + develop.  Returns (rgb[h,w,3], Stats).
        gpus != 1: the frame's blocks are distributed over that many GPUs of the node (0 = all)."""
        s = self.desc.sensor
        rgb = np.zeros((s.crop_height, s.crop_width, 3), np.float32)
        st = abi.Stats()
        if gpus == 1:
            rc = lib().mtsbh_render(self.h, int(spp), rgb.ctypes.data_as(C.c_void_p), C.byref(st))
        else:
            rc = lib().mtsbh_render_multi(self.h, int(spp), int(gpus), rgb.ctypes.data_as(C.c_void_p), C.byref(st))
        if rc:
            raise MtsbError(rc, lib().mtsbh_last_error().decode(errors="replace"))
        return rgb, st


def write_pfm(path, rgb):
    rgb = np.ascontiguousarray(rgb, np.float32)
    rc = lib().mtsbh_write_pfm(os.fspath(path).encode(), rgb.ctypes.data_as(C.c_void_p), rgb.shape[1], rgb.shape[0])
    if rc:
        raise MtsbError(rc, lib().mtsbh_last_error().decode(errors="replace"))
