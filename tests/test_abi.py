"""CPU: the C-ABI library loads, exports every symbol include/mtsb200.h declares, refuses to
compute without a GPU, and the product package never touches the oracle."""
import ctypes as C
import os
import re

import pytest

from mitsuba_b200 import abi, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "mtsb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mtsb_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _declared() == sorted(lib.SYMBOLS)


def test_library_exports_every_symbol():
    assert os.path.exists(lib.LIB_PATH), "build the library first (__graft_entry__.build())"
    L = C.CDLL(lib.LIB_PATH)
    for s in lib.SYMBOLS:
        assert hasattr(L, s), s
    assert L.mtsb_abi_version() == abi.ABI_VERSION


def test_struct_sizes_match_header():
    assert C.sizeof(abi.Material) == 96 and C.sizeof(abi.Shape) == 32 and C.sizeof(abi.Emitter) == 48
    assert C.sizeof(abi.Texture) == 96 and abi.Material.tex_reflectance.offset == 84
    assert abi.RAY_DTYPE.itemsize == 32 and abi.HIT_DTYPE.itemsize == 16


def test_no_cpu_fallback():
    """Without a device the product must fail loudly (MTSB_ERR_NO_DEVICE), never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from mitsuba_b200 import scenes
    with pytest.raises(lib.MtsbError) as e:
        lib.Scene(scenes.cbox(width=8, height=8, spp=1))
    assert e.value.code == abi.ERR_NO_DEVICE


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "mitsuba_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", txt, flags=re.M), f
                assert "liboracle" not in txt and "libmtsref" not in txt and "orc_" not in txt, f


def test_host_library_header_and_exports():
    """include/mtsb200_host.h <-> libmtsb200host.so (the C++ scene loader / plugin mirror)."""
    from mitsuba_b200 import hostlib
    src = open(os.path.join(ROOT, "include", "mtsb200_host.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    declared = sorted(set(re.findall(r"\b(mtsbh_[a-z_0-9]+)\s*\(", src)))
    assert declared == sorted(hostlib.SYMBOLS)
    assert os.path.exists(hostlib.LIB_PATH), "build the host library first (__graft_entry__.build())"
    L = C.CDLL(hostlib.LIB_PATH)
    for s in hostlib.SYMBOLS:
        assert hasattr(L, s), s


def test_host_render_refuses_without_gpu():
    """The render job goes through libmtsb200.so: without a device it must fail, not fall back."""
    import numpy as np
    from mitsuba_b200 import hostlib
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present")
    except ImportError:
        pass
    hs = hostlib.HostScene.load_file(os.path.join(ROOT, "scenes", "cbox.xml"), {"res": 16, "spp": 1})
    with pytest.raises(lib.MtsbError) as ei:
        hs.render()
    assert ei.value.code == abi.ERR_NO_DEVICE


def test_ctypes_mirror_matches_the_compiled_header(tmp_path):
    """Every record that crosses the C ABI has the same size (and the statistics record the same field offsets) in the
    ctypes mirror (mitsuba_b200/abi.py) as in include/mtsb200.h compiled by the C compiler."""
    import subprocess
    names = ["MtsbMaterial", "MtsbTexture", "MtsbShape", "MtsbEmitter", "MtsbEnvmap", "MtsbSensor", "MtsbIntegrator", "MtsbSampler",
             "MtsbSceneDesc", "MtsbRay", "MtsbHit", "MtsbBsdfQuery", "MtsbBsdfResult", "MtsbDirectQuery", "MtsbDirectResult",
             "MtsbRenderParams", "MtsbStats", "MtsbSceneInfo"]
    src = '#include <stdio.h>\n#include <stddef.h>\n#include "mtsb200.h"\nint main(void) {\n'
    src += "".join('  printf("%s %%zu\\n", sizeof(%s));\n' % (n, n) for n in names)
    for f in ("n_passes", "closest_launches", "ms_reduce", "closest_rays_bounce", "shadow_rays_bounce"):
        src += '  printf("MtsbStats.%s %%zu\\n", offsetof(MtsbStats, %s));\n' % (f, f)
    src += '  printf("MtsbMaterial.tex_alpha %zu\\n", offsetof(MtsbMaterial, tex_alpha));\n  return 0;\n}\n'
    (tmp_path / "sz.c").write_text(src)
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(tmp_path / "sz"), str(tmp_path / "sz.c")])
    got = dict(l.split() for l in subprocess.check_output([str(tmp_path / "sz")], text=True).splitlines())
    mirror = {"MtsbMaterial": abi.Material, "MtsbTexture": abi.Texture, "MtsbShape": abi.Shape, "MtsbEmitter": abi.Emitter,
              "MtsbEnvmap": abi.Envmap, "MtsbSensor": abi.Sensor, "MtsbIntegrator": abi.Integrator, "MtsbSampler": abi.Sampler,
              "MtsbSceneDesc": abi.SceneDesc, "MtsbRenderParams": abi.RenderParams, "MtsbStats": abi.Stats, "MtsbSceneInfo": abi.SceneInfo}
    for n, t in mirror.items():
        assert C.sizeof(t) == int(got[n]), (n, C.sizeof(t), got[n])
    for n, dt in (("MtsbRay", abi.RAY_DTYPE), ("MtsbHit", abi.HIT_DTYPE), ("MtsbBsdfQuery", abi.BSDF_QUERY_DTYPE), ("MtsbBsdfResult", abi.BSDF_RESULT_DTYPE),
                  ("MtsbDirectQuery", abi.DIRECT_QUERY_DTYPE), ("MtsbDirectResult", abi.DIRECT_RESULT_DTYPE)):
        assert dt.itemsize == int(got[n]), (n, dt.itemsize, got[n])
    for f in ("n_passes", "closest_launches", "ms_reduce", "closest_rays_bounce", "shadow_rays_bounce"):
        assert getattr(abi.Stats, f).offset == int(got["MtsbStats." + f]), f
    assert abi.Material.tex_alpha.offset == int(got["MtsbMaterial.tex_alpha"])
