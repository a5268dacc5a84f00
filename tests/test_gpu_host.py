"""GPU: the C++ host side end to end -- scene XML -> plugin graph -> MtsbSceneDesc -> mtsb_render.
The image rendered through the loader must be the image rendered from the Python SceneBuilder
description of the same scene (same seed, same counter-based streams: the same samples), and
must agree with the oracle under the same random stream."""
import os
import subprocess

import numpy as np
import pytest

from mitsuba_b200 import hostlib, scenes

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CBOX = os.path.join(ROOT, "scenes", "cbox.xml")


def test_loader_render_equals_builder_render(gpu_lib):
    hs = hostlib.HostScene.load_file(CBOX, {"res": 96, "spp": 8})
    rgb, st = hs.render()                                     # RenderJob: sampleCount spp, sampler seed 1004
    assert st.samples == 96 * 96 * 8 and st.bad_samples == 0
    sc = gpu_lib.Scene(scenes.cbox(width=96, height=96, spp=8))
    st2 = sc.render(spp=8, seed=hs.seed)
    ref = sc.film_develop()
    assert (st.closest_rays, st.shadow_rays, st.shade_events) == (st2.closest_rays, st2.shadow_rays, st2.shade_events)
    # identical samples; the film sums them in queue order (atomic slot allocation), so only fp32 reassociation differs
    assert np.allclose(rgb, ref, rtol=2e-4, atol=1e-7)
    sc.close()


def test_loader_scene_matches_oracle_same_stream(gpu_lib, oracle_mod):
    hs = hostlib.HostScene.load_file(CBOX, {"res": 48, "spp": 4, "depth": 5})
    sc = gpu_lib.Scene(hs.desc_pair())
    st = sc.render(spp=4, seed=3)
    fg = sc.film_download().astype(np.float64)
    b = scenes.cbox(width=48, height=48, spp=4, max_depth=5)
    orc = oracle_mod.Oracle(b)
    ost = orc.render(spp=4, seed=3, counter_rng=True)
    fo = orc.film_download().astype(np.float64) if hasattr(orc, "film_download") else None
    assert st.samples == ost.samples
    assert abs(int(st.closest_rays) - int(ost.closest_rays)) <= 2e-4 * ost.closest_rays
    if fo is not None:
        denom = np.maximum(np.abs(fo), 1e-3)
        bad = np.abs(fg - fo) / denom > 1e-3
        assert bad.any(axis=-1).mean() < 2e-3
    sc.close()


def test_command_line_tool_writes_a_pfm(tmp_path):
    out = tmp_path / "cbox.pfm"
    r = subprocess.run([hostlib.CLI_PATH, "-D", "res=64", "-Dspp=4", "-o", str(out), CBOX], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "32 triangles" in r.stdout
    data = out.read_bytes()
    assert data.startswith(b"PF\n64 64\n-1\n")
    img = np.frombuffer(data[len(b"PF\n64 64\n-1\n"):], np.float32).reshape(64, 64, 3)[::-1]
    hs = hostlib.HostScene.load_file(CBOX, {"res": 64, "spp": 4})
    rgb, _ = hs.render()
    assert np.allclose(img, rgb, rtol=2e-4, atol=1e-7)
    # errors come back on stderr with a non-zero status, like the reference binary
    r = subprocess.run([hostlib.CLI_PATH, "-Dspp=0", CBOX], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "sampleCount" in r.stderr


def test_textured_xml_scene_matches_oracle_same_stream(gpu_lib, oracle_mod, tmp_path):
    """Texture plugins through the loader (bitmap EWA / nearest, checkerboard, gridtexture, scale; PFM input): the GPU
    render of the flattened scene equals the oracle's under the same counter-based streams."""
    from tests.test_host_textures import write_pfm
    write_pfm(tmp_path / "t.pfm", scenes.pattern_image(64, 64))
    xml = """<scene version="0.5.0"><integrator type="path"><integer name="maxDepth" value="5"/></integrator>
<sensor type="perspective"><transform name="toWorld"><lookat origin="0 1.2 5" target="0 0.4 0" up="0 1 0"/></transform><float name="fov" value="42"/>
  <sampler type="independent"><integer name="sampleCount" value="4"/></sampler>
  <film type="hdrfilm"><integer name="width" value="64"/><integer name="height" value="48"/></film></sensor>
<shape type="rectangle"><transform name="toWorld"><rotate x="1" angle="-90"/><scale value="3"/></transform>
  <bsdf type="diffuse"><texture type="bitmap" name="reflectance"><string name="filename" value="%(f)s"/><float name="uvscale" value="8"/></texture></bsdf></shape>
<shape type="rectangle"><transform name="toWorld"><scale value="3"/><translate z="-3" y="3"/></transform>
  <bsdf type="diffuse"><texture type="checkerboard" name="reflectance"><float name="uvscale" value="6"/></texture></bsdf></shape>
<shape type="cube"><transform name="toWorld"><scale value=".6"/><rotate y="1" angle="30"/><translate x="-1" y=".6"/></transform>
  <bsdf type="plastic"><texture type="scale" name="diffuseReflectance"><rgb name="scale" value=".9 .5 .4"/>
    <texture type="bitmap"><string name="filename" value="%(f)s"/><string name="filterType" value="nearest"/><string name="wrapMode" value="mirror"/></texture></texture></bsdf></shape>
<shape type="cube"><transform name="toWorld"><scale value=".5"/><rotate y="1" angle="-20"/><translate x="1.1" y=".5"/></transform>
  <bsdf type="twosided"><bsdf type="diffuse"><texture type="gridtexture" name="reflectance"><float name="uvscale" value="3"/><float name="lineWidth" value=".08"/></texture></bsdf></bsdf></shape>
<shape type="rectangle"><transform name="toWorld"><rotate x="1" angle="90"/><translate y="2.9"/></transform><emitter type="area"><rgb name="radiance" value="12 11 9"/></emitter></shape>
</scene>""" % dict(f=tmp_path / "t.pfm")
    hs = hostlib.HostScene.load_string(xml, "textured.xml")
    assert hs.desc.n_textures == 4
    sc = gpu_lib.Scene(hs.desc_pair())
    st = sc.render(spp=4, seed=3)
    fg = sc.film_download().astype(np.float64)
    orc = oracle_mod.Oracle(hs.desc_pair())
    ost = orc.render(spp=4, seed=3, counter_rng=True)
    fo = orc.film_download().astype(np.float64)
    assert st.samples == ost.samples and st.bad_samples == 0
    assert abs(int(st.closest_rays) - int(ost.closest_rays)) <= 2e-4 * ost.closest_rays
    assert fo[..., :3].max() > 0.05
    denom = np.maximum(np.abs(fo), 1e-3)
    bad = np.abs(fg - fo) / denom > 1e-3
    assert bad.any(axis=-1).mean() < 2e-3
    sc.close()
