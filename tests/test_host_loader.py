"""Host side (mitsuba_b200/host): scene XML loading, plugin defaults and errors, mesh readers and
the flattening to MtsbSceneDesc -- the behaviour of the reference's SceneHandler / plugin
constructors (src/librender/scenehandler.cpp and the plugin sources cited in the host files).
Pure CPU: nothing here needs a GPU."""
import math
import os
import struct

import numpy as np
import pytest

from mitsuba_b200 import abi, hostlib, scenedesc, scenes
from mitsuba_b200.lib import MtsbError

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

HEAD = '<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"/>'


def load(body, head=HEAD, name="mem.xml", params=None):
    return hostlib.HostScene.load_string(head + body + "</scene>", name, params)


def tri_positions(hs):
    pos, _, _, idx = hs.arrays()
    return pos[idx]


# ---------------------------------------------------------------------------------------------
def test_cbox_xml_equals_builder():
    """scenes/cbox.xml through the C++ loader == scenes.cbox() through SceneBuilder."""
    hs = hostlib.HostScene.load_file(os.path.join(ROOT, "scenes", "cbox.xml"))
    assert hs.warnings == []
    b = scenes.cbox()
    bp, _, _, bi = b.arrays()
    assert np.array_equal(tri_positions(hs), bp[bi])
    d, (bd, keep) = hs.desc, b.desc()
    assert (d.n_shapes, d.n_emitters, d.n_tris) == (bd.n_shapes, bd.n_emitters, bd.n_tris)
    assert bytes(d.sensor) == bytes(bd.sensor)
    assert bytes(d.integrator) == bytes(bd.integrator)
    for i in range(d.n_shapes):
        s, t = d.shapes[i], bd.shapes[i]
        assert (s.first_tri, s.n_tris, s.emitter, s.has_normals) == (t.first_tri, t.n_tris, t.emitter, t.has_normals)
        assert bytes(d.materials[s.material]) == bytes(bd.materials[t.material])
    assert bytes(d.emitters[0]) == bytes(bd.emitters[0])
    assert d.normals is None or not d.normals


def test_defaults_of_a_minimal_scene():
    hs = load('<shape type="rectangle"/>')
    d = hs.desc
    s, ig = d.sensor, d.integrator
    assert (s.film_width, s.film_height) == (768, 576)                     # film.cpp:30-33
    assert (s.crop_x, s.crop_y, s.crop_width, s.crop_height) == (0, 0, 768, 576)
    assert s.rfilter == abi.RFILTER_GAUSSIAN and s.rfilter_stddev == 0.5 and s.rfilter_radius == 2.0
    assert s.near_clip == np.float32(1e-2) and s.far_clip == np.float32(1e4)
    # focalLength 50mm -> diagonal fov -> x fov (sensor.cpp:266-282, 305-309)
    aspect = 768 / 576
    dfov = 2 * math.atan(math.sqrt(36 * 36 + 24 * 24) / (2 * 50.0))
    width = 2 * math.tan(0.5 * dfov) / math.sqrt(1 + 1 / aspect ** 2)
    assert s.fov_x_deg == pytest.approx(math.degrees(2 * math.atan(width * 0.5)), rel=1e-6)
    assert list(s.to_world) == [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]
    assert (ig.max_depth, ig.rr_depth, ig.strict_normals, ig.hide_emitters, ig.sample_count) == (-1, 5, 0, 0, 4)
    assert hs.seed == 1004
    # shape.cpp:49-62: no BSDF -> diffuse 0.5
    m = d.materials[d.shapes[0].material]
    assert m.type == abi.BSDF_DIFFUSE and list(m.reflectance) == [0.5, 0.5, 0.5]


def test_emitter_shape_gets_black_bsdf_and_links():
    hs = load('<shape type="rectangle"><emitter type="area"><rgb name="radiance" value="1,2,3"/>'
              '<float name="samplingWeight" value="2.5"/></emitter></shape>'
              '<emitter type="point"><point name="position" x="1" y="2" z="3"/><rgb name="intensity" value="5"/></emitter>'
              '<emitter type="constant"/>')
    d = hs.desc
    assert d.n_emitters == 3
    e0, e1, e2 = d.emitters[0], d.emitters[1], d.emitters[2]
    assert (e0.type, e0.shape, list(e0.radiance), e0.sampling_weight) == (abi.EMITTER_AREA, 0, [1, 2, 3], 2.5)
    assert (e1.type, e1.shape, list(e1.position), list(e1.radiance)) == (abi.EMITTER_POINT, -1, [1, 2, 3], [5, 5, 5])
    assert (e2.type, list(e2.radiance)) == (abi.EMITTER_CONSTANT, [1, 1, 1])          # constant.cpp:49, D65 == 1 in RGB
    assert d.shapes[0].emitter == 0
    assert list(d.materials[d.shapes[0].material].reflectance) == [0, 0, 0]


def test_fov_axes():
    def fov(body, w=400, h=200):
        hs = load("", head='<scene version="0.5.0"><integrator type="path"/><sensor type="perspective">%s'
                            '<film type="hdrfilm"><integer name="width" value="%d"/><integer name="height" value="%d"/></film>'
                            '</sensor>' % (body, w, h))
        return hs.desc.sensor.fov_x_deg
    assert fov('<float name="fov" value="40"/>') == 40
    yf = math.degrees(2 * math.atan(math.tan(math.radians(20)) * 2))
    assert fov('<float name="fov" value="40"/><string name="fovAxis" value="y"/>') == pytest.approx(yf, rel=1e-6)
    assert fov('<float name="fov" value="40"/><string name="fovAxis" value="smaller"/>') == pytest.approx(yf, rel=1e-6)
    assert fov('<float name="fov" value="40"/><string name="fovAxis" value="larger"/>') == 40
    assert fov('<float name="fov" value="40"/><string name="fovAxis" value="smaller"/>', 200, 400) == 40
    diag = 2 * math.tan(math.radians(20))
    xf = math.degrees(2 * math.atan(0.5 * diag / math.sqrt(1 + 1 / 4.0)))
    assert fov('<float name="fov" value="40"/><string name="fovAxis" value="diagonal"/>') == pytest.approx(xf, rel=1e-6)
    with pytest.raises(MtsbError, match="either a focal length"):
        fov('<float name="fov" value="40"/><string name="focalLength" value="35mm"/>')
    with pytest.raises(MtsbError, match="fovAxis"):
        fov('<float name="fov" value="40"/><string name="fovAxis" value="z"/>')
    with pytest.raises(MtsbError, match=r"interval \(0, 180\)"):
        fov('<float name="fov" value="180"/>')


def test_transform_composition_order():
    """Later tags are applied after earlier ones: M = T_last * ... * T_first (scenehandler.cpp:348-441)."""
    hs = load('<shape type="rectangle"><transform name="toWorld"><scale x="2" y="3"/><rotate z="1" angle="90"/>'
              '<translate x="10" y="20" z="30"/></transform></shape>')
    pos, nrm, uv, idx = hs.arrays()
    # (-1,-1,0) -> scale (-2,-3,0) -> rot90z (3,-2,0) -> translate (13,18,30)
    assert np.allclose(pos[0], [13, 18, 30], atol=1e-5)
    assert np.allclose(pos[2], [7, 22, 30], atol=1e-5)
    assert np.allclose(nrm, [[0, 0, 1]] * 4, atol=1e-6)
    assert idx.tolist() == [[0, 1, 2], [2, 3, 0]]
    hs = load('<shape type="rectangle"><transform name="toWorld"><matrix value="1 0 0 5  0 1 0 6  0 0 1 7  0 0 0 1"/>'
              '</transform></shape>')
    assert np.allclose(hs.arrays()[0][0], [4, 5, 7])
    hs = load('<shape type="rectangle"><transform name="toWorld"><scale value="4"/></transform><boolean name="flipNormals" value="true"/></shape>')
    pos, nrm, _, idx = hs.arrays()
    assert np.allclose(nrm, [[0, 0, -1]] * 4) and np.allclose(pos[2], [4, 4, 0])
    # winding follows the flipped normal
    p = pos[idx]
    assert np.all(np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])[:, 2] < 0)


def test_lookat_matches_builder_and_default_up():
    hs = load("", head='<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><transform name="toWorld">'
                       '<lookat origin="1,2,3" target="4,6,-2" up="0,1,0"/></transform></sensor>')
    ref = scenedesc.look_at([1, 2, 3], [4, 6, -2], [0, 1, 0]).reshape(16)
    assert np.allclose(list(hs.desc.sensor.to_world), ref, atol=1e-6)
    hs = load("", head='<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><transform name="toWorld">'
                       '<lookat origin="0,0,0" target="0,0,1"/></transform></sensor>')
    m = np.array(list(hs.desc.sensor.to_world)).reshape(4, 4)
    assert np.allclose(m[:3, 2], [0, 0, 1]) and np.allclose(m[:3, :3] @ m[:3, :3].T, np.eye(3), atol=1e-6)
    with pytest.raises(MtsbError, match="Scale factors"):
        load("", head='<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><transform name="toWorld">'
                      '<scale value="2"/></transform></sensor>')


def test_parameters_and_defaults():
    body = '<default name="r" value="0.25"/><shape type="rectangle"><bsdf type="diffuse"><rgb name="reflectance" value="$r"/></bsdf></shape>'
    assert list(load(body).materials()[0].reflectance) == [0.25] * 3
    assert list(load(body, params={"r": "0.75"}).materials()[0].reflectance) == [0.75] * 3     # -Dr=0.75 wins over <default>
    with pytest.raises(MtsbError, match="undefined parameter"):
        load('<shape type="rectangle"><bsdf type="diffuse"><rgb name="reflectance" value="$nope"/></bsdf></shape>')
    # longer names are substituted first (reverse map order, scenehandler.cpp:209-216)
    body = ('<default name="a" value="0.1"/><default name="ab" value="0.2"/>'
            '<shape type="rectangle"><bsdf type="diffuse"><rgb name="reflectance" value="$ab"/></bsdf></shape>')
    assert list(load(body).materials()[0].reflectance) == pytest.approx([0.2] * 3)


def test_colour_values():
    def refl(tag):
        return list(load('<shape type="rectangle"><bsdf type="diffuse">%s</bsdf></shape>' % tag).materials()[0].reflectance)
    assert refl('<rgb name="reflectance" value="#ff8000"/>') == pytest.approx([1.0, 128 / 255, 0.0])
    assert refl('<rgb name="reflectance" value="#804020"/>') == pytest.approx([128 / 255, 64 / 255, 32 / 255])
    assert refl('<rgb name="reflectance" value="0.3"/>') == pytest.approx([0.3] * 3)
    assert refl('<spectrum name="reflectance" value="0.4"/>') == pytest.approx([0.4] * 3)
    assert refl('<float name="reflectance" value="0.6"/>') == pytest.approx([0.6] * 3)
    s = 0.5
    lin = ((s + 0.055) / 1.055) ** 2.4
    assert refl('<srgb name="reflectance" value="0.5, 0.02, 1"/>') == pytest.approx([lin, 0.02 / 12.92, 1.0], rel=1e-6)
    assert refl('<rgb name="diffuseReflectance" value="0.1,0.2,0.3"/>') == pytest.approx([0.1, 0.2, 0.3])
    # ensureEnergyConservation (bsdf.cpp:96-113)
    r = refl('<rgb name="reflectance" value="2, 1, 0.5"/>')
    assert r == pytest.approx([0.99, 0.495, 0.2475])
    with pytest.raises(MtsbError, match="Invalid rgb value"):
        refl('<rgb name="reflectance" value="1,2"/>')
    with pytest.raises(MtsbError, match="wrong type"):
        refl('<string name="reflectance" value="red"/>')


def test_bsdf_records_match_python_builder():
    hs = load('<shape type="rectangle"><bsdf type="roughconductor"><string name="material" value="Au"/><float name="alpha" value="0.2"/>'
              '<string name="distribution" value="ggx"/></bsdf></shape>'
              '<shape type="rectangle"><bsdf type="roughdielectric"><float name="intIOR" value="1.5"/><float name="extIOR" value="1.0"/>'
              '<float name="alphaU" value="0.05"/><float name="alphaV" value="0.3"/><boolean name="sampleVisible" value="false"/></bsdf></shape>'
              '<shape type="rectangle"><bsdf type="roughdielectric"/></shape>'
              '<shape type="rectangle"><bsdf type="roughconductor"><string name="material" value="none"/>'
              '<rgb name="eta" value="1,2,3"/><rgb name="k" value="4,5,6"/><float name="extEta" value="2"/><float name="alpha" value="0"/></bsdf></shape>')
    m = hs.materials()
    eta, k = scenes.METALS["Au"]
    assert bytes(m[0]) == bytes(scenedesc.roughconductor(eta, k, alpha=0.2, distribution="ggx"))
    assert bytes(m[1]) == bytes(scenedesc.roughdielectric(1.5, 1.0, distribution="beckmann", sample_visible=False, alpha_u=0.05, alpha_v=0.3))
    assert bytes(m[2]) == bytes(scenedesc.roughdielectric())                              # bk7 / air defaults
    assert list(m[3].cond_eta) == [0.5, 1.0, 1.5] and list(m[3].cond_k) == [2.0, 2.5, 3.0]
    assert m[3].alpha_u == np.float32(1e-4)                                               # microfacet.h:135-136
    assert any("clamped" in w for w in hs.warnings)
    with pytest.raises(MtsbError, match="either 'alpha' or 'alphaU'"):
        load('<shape type="rectangle"><bsdf type="roughconductor"><float name="alpha" value="0.1"/><float name="alphaU" value="0.1"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="both 'alphaU' and 'alphaV'"):
        load('<shape type="rectangle"><bsdf type="roughconductor"><float name="alphaU" value="0.1"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="invalid distribution"):
        load('<shape type="rectangle"><bsdf type="roughconductor"><string name="distribution" value="blinn"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="must be positive and differ"):
        load('<shape type="rectangle"><bsdf type="roughdielectric"><float name="intIOR" value="1.2"/><float name="extIOR" value="1.2"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="Unable to find an IOR"):
        load('<shape type="rectangle"><bsdf type="roughdielectric"><string name="intIOR" value="unobtainium"/></bsdf></shape>')


def test_smooth_dielectric_and_conductor_records():
    """dielectric.cpp:141-168 / conductor.cpp:146-174 property surface."""
    hs = load('<shape type="rectangle"><bsdf type="dielectric"/></shape>'
              '<shape type="rectangle"><bsdf type="dielectric"><string name="intIOR" value="water"/><float name="extIOR" value="1.0"/>'
              '<rgb name="specularReflectance" value="2,1,1"/><rgb name="specularTransmittance" value="0.5,0.25,0.125"/></bsdf></shape>'
              '<shape type="rectangle"><bsdf type="conductor"><string name="material" value="Au"/></bsdf></shape>'
              '<shape type="rectangle"><bsdf type="conductor"><string name="material" value="none"/></bsdf></shape>'
              '<shape type="rectangle"><bsdf type="twosided"><bsdf type="conductor"><rgb name="eta" value="1,2,3"/><rgb name="k" value="4,5,6"/>'
              '<float name="extEta" value="2"/></bsdf></bsdf></shape>')
    m = hs.materials()
    eta, k = scenes.METALS["Au"]
    assert bytes(m[0]) == bytes(scenedesc.dielectric())                                    # bk7 / air defaults
    assert bytes(m[1]) == bytes(scenedesc.dielectric(1.3330, 1.0, [2, 1, 1], [0.5, 0.25, 0.125]))
    assert list(m[1].reflectance) == pytest.approx([0.99, 0.495, 0.495])                  # ensureEnergyConservation
    assert any("specularReflectance" in w and "conserve energy" in w for w in hs.warnings)
    assert bytes(m[2]) == bytes(scenedesc.conductor(eta, k))
    assert bytes(m[3]) == bytes(scenedesc.conductor(0.0, 1.0))
    assert bytes(m[4]) == bytes(scenedesc.conductor([1, 2, 3], [4, 5, 6], ext_eta=2.0, twosided=True))
    with pytest.raises(MtsbError, match="without a transmission component"):
        load('<shape type="rectangle"><bsdf type="twosided"><bsdf type="dielectric"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="must be positive"):
        load('<shape type="rectangle"><bsdf type="dielectric"><float name="intIOR" value="-1"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="specify 'eta' and 'k'"):
        load('<shape type="rectangle"><bsdf type="conductor"><string name="material" value="unobtainium"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="not found"):
        load('<shape type="rectangle"><bsdf type="phong"/></shape>')


def test_plastic_records():
    """plastic.cpp:141-165, 185-206: IOR defaults, energy clamp, fdrInt, specular sampling weight, nonlinear."""
    hs = load('<shape type="rectangle"><bsdf type="plastic"/></shape>'
              '<shape type="rectangle"><bsdf type="twosided"><bsdf type="plastic"><float name="intIOR" value="1.9"/><string name="extIOR" value="water"/>'
              '<rgb name="diffuseReflectance" value="0.1,0.2,0.9"/><rgb name="specularReflectance" value="0.5"/>'
              '<boolean name="nonlinear" value="true"/></bsdf></bsdf></shape>')
    m = hs.materials()
    for got, want in ((m[0], scenedesc.plastic()),
                      (m[1], scenedesc.plastic(1.9, 1.3330, [0.1, 0.2, 0.9], 0.5, nonlinear=True, twosided=True))):
        assert got.fdr_int == pytest.approx(want.fdr_int, rel=1e-6) and 0.3 < got.fdr_int < 0.7
        got.fdr_int = want.fdr_int                      # two double-precision sums in different order: equal to rounding
        assert bytes(got) == bytes(want)
    assert m[1].flags == abi.MAT_TWOSIDED | abi.MAT_NONLINEAR and m[0].eta == np.float32(np.float32(1.49) / np.float32(1.000277))


def test_twosided_and_shared_references():
    hs = load('<bsdf type="twosided" id="two"><bsdf type="diffuse"><rgb name="reflectance" value="0.1"/></bsdf>'
              '<bsdf type="roughconductor"><string name="material" value="Al"/></bsdf></bsdf>'
              '<bsdf type="twosided" id="one"><bsdf type="diffuse"/></bsdf>'
              '<shape type="rectangle"><ref id="two"/></shape><shape type="cube"><ref id="two"/></shape><shape type="rectangle"><ref id="one"/></shape>')
    d = hs.desc
    assert d.n_materials == 3                       # back record + front record, shared by two shapes; one single-BSDF adapter
    sh = hs.shapes()
    assert sh[0].material == sh[1].material
    front = d.materials[sh[0].material]
    assert front.type == abi.BSDF_DIFFUSE and front.flags & abi.MAT_TWOSIDED and front.back >= 0
    assert d.materials[front.back].type == abi.BSDF_ROUGHCONDUCTOR
    single = d.materials[sh[2].material]
    assert single.flags & abi.MAT_TWOSIDED and single.back == -1
    with pytest.raises(MtsbError, match="without a transmission"):
        load('<shape type="rectangle"><bsdf type="twosided"><bsdf type="roughdielectric"/></bsdf></shape>')
    with pytest.raises(MtsbError, match="nested one-sided material is required"):
        load('<shape type="rectangle"><bsdf type="twosided"/></shape>')
    with pytest.raises(MtsbError, match="No more than two"):
        load('<shape type="rectangle"><bsdf type="twosided"><bsdf type="diffuse"/><bsdf type="diffuse"/><bsdf type="diffuse"/></bsdf></shape>')


ERRORS = [
    ('<scene><integrator type="path"/></scene>', "missing version information"),
    ('<scene version="0.4.0"><integrator type="path"/></scene>', "older version"),
    ('<scene version="1.0.0"><integrator type="path"/></scene>', "future version"),
    (HEAD + "<bogus/></scene>", 'Unhandled tag "bogus"'),
    (HEAD + '<shape type="teapot"/></scene>', 'Plugin "teapot" not found'),
    (HEAD + '<shape type="sphere"/></scene>', "analytic primitive"),
    (HEAD + '<shape type="rectangle" id="a"/><shape type="rectangle" id="a"/></scene>', "Duplicate ID 'a'"),
    (HEAD + '<shape type="rectangle"><ref id="nope"/></shape></scene>', "Referenced object 'nope' not found"),
    (HEAD + '<shape type="rectangle"><boolean name="flipNormals" value="yes"/></shape></scene>', "Unsupported boolean constant"),
    (HEAD + '<shape type="rectangle"><integer name="x" value="1.5"/></shape></scene>', "Invalid integer value"),
    (HEAD + '<shape type="rectangle"><float name="x" value="abc"/></shape></scene>', "Invalid floating point value"),
    (HEAD + '<shape type="rectangle"><emitter type="area"/><emitter type="area"/></shape></scene>', "multiple emitters"),
    (HEAD + '<emitter type="area"/></scene>', "must be child of a shape"),
    (HEAD + '<shape type="rectangle"><emitter type="area"><transform name="toWorld"><translate x="1"/></transform></emitter></shape></scene>',
     "'toWorld' transformation -- this is not allowed"),
    (HEAD + '<emitter type="constant"/><emitter type="constant"/></scene>', "one environment emitter"),
    ('<scene version="0.5.0"><sensor type="perspective"/></scene>', "no <integrator>"),
    ('<scene version="0.5.0"><integrator type="bdpt"/><sensor type="perspective"/></scene>', 'Plugin "bdpt" not found'),
    ('<scene version="0.5.0"><integrator type="path"><integer name="rrDepth" value="0"/></integrator><sensor type="perspective"/></scene>',
     "'rrDepth' must be set to a value greater than zero"),
    ('<scene version="0.5.0"><integrator type="path"><integer name="maxDepth" value="0"/></integrator><sensor type="perspective"/></scene>',
     "'maxDepth' must be set to -1"),
    ('<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><film type="hdrfilm"><integer name="cropWidth" value="9999"/>'
     '</film></sensor></scene>', "Invalid crop window"),
    ('<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><float name="nearClip" value="0"/></sensor></scene>',
     "'nearClip' parameter must be greater than zero"),
    ('<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><film type="hdrfilm"><rfilter type="lanczos"/></film>'
     '</sensor></scene>', 'Plugin "lanczos" not found'),
    ('<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"><sampler type="stratified"/></sensor></scene>', 'Plugin "stratified" not found'),
    (HEAD + '<shape type="rectangle"><float name="a" value="1"/><float name="a" value="2"/></shape></scene>', "specified multiple times"),
    (HEAD + '<shape type="obj"><string name="filename" value="does_not_exist.obj"/></shape></scene>', "not found"),
    (HEAD + '<shape type="rectangle"><medium type="homogeneous"/></shape></scene>', "participating media"),
    (HEAD + "<shape type='rectangle'></scene>", "mismatched closing tag"),
    ("<scene version='0.5.0'", "unterminated tag"),
]


@pytest.mark.parametrize("xml,msg", ERRORS, ids=[e[1][:28] for e in ERRORS])
def test_errors_are_reported_like_the_reference(xml, msg):
    with pytest.raises(MtsbError) as ei:
        hostlib.HostScene.load_string(xml, "mem.xml")
    assert msg in str(ei.value)
    assert ei.value.code == abi.ERR_PARSE
    assert 'mem.xml' in str(ei.value)            # file + line, as XMLLog prints them


def test_missing_file_is_an_io_error():
    with pytest.raises(MtsbError) as ei:
        hostlib.HostScene.load_file("/nonexistent/scene.xml")
    assert ei.value.code == abi.ERR_IO


def test_unqueried_properties_are_warned_about():
    hs = load('<shape type="rectangle"><float name="radius" value="3"/></shape>')
    assert hs.warnings == ['Unqueried attribute "radius" in element "shape"']


def test_sampler_plugins():
    """independent (independent.cpp:55-60), ldsampler (ldsampler.cpp:75-86: the sample count is rounded up to a
    power of two with a warning) and halton (halton.cpp:121-130)."""
    base = '<scene version="0.5.0"><integrator type="path"/><sensor type="perspective">%s</sensor></scene>'
    hs = hostlib.HostScene.load_string(base % '<sampler type="ldsampler"><integer name="sampleCount" value="20"/></sampler>', "mem.xml")
    d = hs.desc
    assert (d.sampler.type, d.sampler.dimension, d.integrator.sample_count) == (abi.SAMPLER_LDSAMPLER, 4, 32)
    assert hs.warnings == ["Sample count should be a power of two -- rounding to 32"]
    hs = hostlib.HostScene.load_string(base % '<sampler type="ldsampler"><integer name="sampleCount" value="64"/><integer name="dimension" value="7"/></sampler>', "mem.xml")
    assert (hs.desc.sampler.dimension, hs.desc.integrator.sample_count, hs.warnings) == (7, 64, [])
    hs = hostlib.HostScene.load_string(base % '<sampler type="halton"><integer name="sampleCount" value="12"/></sampler>', "mem.xml")
    assert (hs.desc.sampler.type, hs.desc.sampler.scramble, hs.desc.integrator.sample_count) == (abi.SAMPLER_HALTON, -1, 12)
    hs = hostlib.HostScene.load_string(base % '<sampler type="halton"><integer name="scramble" value="0"/></sampler>', "mem.xml")
    assert (hs.desc.sampler.scramble, hs.desc.integrator.sample_count) == (0, 4)
    hs = hostlib.HostScene.load_string(base % '<sampler type="sobol"><integer name="sampleCount" value="9"/><integer name="scramble" value="3"/></sampler>', "mem.xml")
    assert (hs.desc.sampler.type, hs.desc.sampler.scramble, hs.desc.integrator.sample_count) == (abi.SAMPLER_SOBOL, 3, 9)
    hs = hostlib.HostScene.load_string(base % '<sampler type="independent"/>', "mem.xml")
    assert hs.desc.sampler.type == abi.SAMPLER_INDEPENDENT


def _write_pfm(path, img, big_endian=False, scale=1.0):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n%s\n" % (w, h, (b"%g" % scale) if big_endian else (b"-%g" % scale)))
        f.write(img[::-1].astype(">f4" if big_endian else "<f4").tobytes())


def _write_rgbe(path, img, rle=True):
    """Radiance RGBE writer for the test (flat or run-length encoded scanlines)."""
    h, w, _ = img.shape
    m = img.max(-1)
    e = np.where(m > 1e-32, np.floor(np.log2(np.maximum(m, 1e-38))) + 1, 0).astype(np.int32)
    mant = np.where((m > 1e-32)[..., None], img / np.exp2(e - 8.0)[..., None], 0).astype(np.int32).clip(0, 255)
    rgbe = np.concatenate([mant, np.where(m > 1e-32, e + 128, 0)[..., None]], -1).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w))
        for y in range(h):
            if not rle:
                f.write(rgbe[y].tobytes())
                continue
            f.write(bytes([2, 2, w >> 8, w & 255]))
            for c in range(4):
                row, i = rgbe[y, :, c], 0
                while i < w:
                    run = 1
                    while i + run < w and run < 127 and row[i + run] == row[i]:
                        run += 1
                    if run >= 4:
                        f.write(bytes([128 + run, int(row[i])]))
                        i += run
                    else:
                        j = i
                        while j < w and j - i < 128 and not (j + 3 < w and row[j] == row[j + 1] == row[j + 2] == row[j + 3]):
                            j += 1
                        j = max(j, i + 1)
                        f.write(bytes([j - i]) + row[i:j].tobytes())
                        i = j
    # what a reader must return (RGBE_ToFloat, bitmap.cpp:3522-3530)
    return np.where(rgbe[..., 3:4] > 0, rgbe[..., :3].astype(np.float32) * np.exp2(rgbe[..., 3:4].astype(np.float32) - 136.0), 0).astype(np.float32)


def _envmap_rgb(hs):
    em = hs.desc.envmap.contents
    return np.ctypeslib.as_array(em.rgb, shape=(em.height, em.width, 3)).copy(), em


def test_envmap_plugin_reads_pfm_and_rgbe(tmp_path):
    """envmap.cpp:105-189 through the loader: filename relative to the scene file, `scale`, `toWorld`; PFM (either byte
    order, bottom row first: bitmap.cpp:3764-3814) and Radiance RGBE, flat and run-length encoded (:3590-3690)."""
    from mitsuba_b200 import scenes
    img = scenes.synthetic_sky(24, 12)
    _write_pfm(tmp_path / "le.pfm", img)
    _write_pfm(tmp_path / "be.pfm", img, big_endian=True, scale=2.0)
    want_rle = _write_rgbe(tmp_path / "sky.hdr", img)
    want_flat = _write_rgbe(tmp_path / "flat.rgbe", img, rle=False)
    body = ('<emitter type="envmap"><string name="filename" value="%s"/><float name="scale" value="2.5"/>'
            '<transform name="toWorld"><rotate y="1" angle="90"/></transform></emitter>')
    def load_scene(fn):
        p = tmp_path / ("scene_%s.xml" % fn)
        p.write_text(HEAD + body % fn + "</scene>")
        return hostlib.HostScene.load_file(str(p))
    hs = load_scene("le.pfm")
    rgb, em = _envmap_rgb(hs)
    assert np.array_equal(rgb, img) and (em.width, em.height) == (24, 12) and em.scale == 2.5
    tw = np.array(em.to_world[:]).reshape(4, 4)
    tl = np.array(em.to_local[:]).reshape(4, 4)
    assert np.allclose(tw[:3, :3], [[0, 0, 1], [0, 1, 0], [-1, 0, 0]], atol=1e-6) and np.allclose(tw @ tl, np.eye(4), atol=1e-6)
    assert [e.type for e in hs.emitters()] == [abi.EMITTER_ENVMAP]
    assert np.array_equal(_envmap_rgb(load_scene("be.pfm"))[0], img * np.float32(2.0))
    assert np.array_equal(_envmap_rgb(load_scene("sky.hdr"))[0], want_rle)
    assert np.array_equal(_envmap_rgb(load_scene("flat.rgbe"))[0], want_flat)
    assert np.abs(want_rle - img).max() <= img.max() / 128
    for fn, msg in (("missing.pfm", "could not be found"), ("scene_le.pfm.xml", "unsupported image format")):
        with pytest.raises(MtsbError) as ei:
            load_scene(fn)
        assert msg in str(ei.value)
    with pytest.raises(MtsbError) as ei:
        hostlib.HostScene.load_string(HEAD + '<emitter type="constant"/>' + body % str(tmp_path / "le.pfm") + "</scene>", "mem.xml")
    assert "one environment emitter" in str(ei.value)
