"""Host side: texture plugins of the scene loader -- `bitmap`, `checkerboard`, `gridtexture`,
`scale` as children of the BSDFs on the path (src/textures/*.cpp, src/librender/texture.cpp:81-96,
BSDF::ensureEnergyConservation bsdf.cpp:96-113), and the PNG reader that feeds `bitmap`
(Bitmap::readPNG bitmap.cpp:2450-2580 + FormatConverter's linearisation fmtconv.cpp:151-157, 1092-1101).
CPU: the records the C++ loader flattens must equal the records the Python SceneBuilder writes for the same scene."""
import ctypes as C
import struct
import zlib

import numpy as np
import pytest

from mitsuba_b200 import abi, hostlib, scenedesc, scenes
from mitsuba_b200.lib import MtsbError

HEAD = '<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"/>'


def load(body, name="mem.xml"):
    return hostlib.HostScene.load_string(HEAD + body + "</scene>", name)


def write_pfm(path, img):
    img = np.asarray(img, np.float32)
    h, w = img.shape[:2]
    with open(path, "wb") as f:
        f.write(b"%s\n%d %d\n-1.0\n" % (b"PF" if img.ndim == 3 else b"Pf", w, h))
        f.write(np.ascontiguousarray(img[::-1]).astype("<f4").tobytes())


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def write_png(path, rows, width, height, depth, ctype, chunks=(), palette=None):
    """`rows`: height byte strings (packed samples, big endian for 16 bit); scanline filter y % 5 is applied."""
    channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    bpp = max(1, channels * depth // 8)
    out, prev = bytearray(), bytes(len(rows[0]))
    for y, row in enumerate(rows):
        f = y % 5
        enc = bytearray()
        for i, x in enumerate(row):
            a = row[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            pred = (0, a, b, (a + b) >> 1, _paeth(a, b, c))[f]
            enc.append((x - pred) & 255)
        out += bytes([f]) + enc
        prev = row

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    comp = zlib.compress(bytes(out))
    with open(path, "wb") as fh:
        fh.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", width, height, depth, ctype, 0, 0, 0)))
        for tag, data in chunks:
            fh.write(chunk(tag, data))
        if palette is not None:
            fh.write(chunk(b"PLTE", bytes(palette)))
        half = len(comp) // 2                    # two IDAT chunks: the stream continues across chunk boundaries
        fh.write(chunk(b"IDAT", comp[:half]) + chunk(b"IDAT", comp[half:]) + chunk(b"IEND", b""))


def srgb_to_linear(v, maxv):
    v = np.float32(v) * (np.float32(1.0) / np.float32(maxv))
    lo = v * np.float32(1.0 / 12.92)
    hi = np.power(((v + np.float32(0.055)) * np.float32(1.0 / 1.055)).astype(np.float32), np.float32(2.4)).astype(np.float32)
    return np.where(v <= np.float32(0.04045), lo, hi).astype(np.float32)


def close1(a, b):
    """Equal up to the last bit of powf (numpy's float32 power and libm's differ by <= 1 ulp)."""
    return a.shape == b.shape and np.allclose(a, b, rtol=2.5e-7, atol=0)


def texels_of(t):
    n = t.width * t.height * t.channels
    return np.ctypeslib.as_array(t.texels, shape=(n,)).reshape(t.height, t.width, t.channels).copy()


def same_record(a, b):
    """Two MtsbTexture records are equal apart from the texel pointer, and so are the texels."""
    ba, bb = bytearray(bytes(a)), bytearray(bytes(b))
    off = abi.Texture.texels.offset
    ba[off:off + 8] = bb[off:off + 8] = b"\0" * 8
    if ba != bb:
        return False
    return a.type != abi.TEX_BITMAP or np.array_equal(texels_of(a), texels_of(b))


TEXTURED = """
<texture type="bitmap" id="wood"><string name="filename" value="%(rgb)s"/><float name="uscale" value="3"/><float name="vscale" value="2"/>
    <float name="maxAnisotropy" value="8"/></texture>
<shape type="rectangle"><bsdf type="diffuse"><texture type="checkerboard" name="reflectance">
    <rgb name="color0" value=".7 .7 .65"/><rgb name="color1" value=".15 .2 .3"/><float name="uvscale" value="4"/></texture></bsdf></shape>
<shape type="rectangle"><bsdf type="diffuse"><ref name="diffuseReflectance" id="wood"/></bsdf></shape>
<shape type="rectangle"><bsdf type="diffuse"><texture type="gridtexture" name="reflectance"><rgb name="color0" value=".14 .45 .091"/>
    <rgb name="color1" value=".9 .9 .2"/><float name="lineWidth" value=".04"/><float name="uscale" value="5"/><float name="vscale" value="5"/>
    <float name="uoffset" value=".3"/></texture></bsdf></shape>
<shape type="rectangle"><bsdf type="diffuse"><texture type="scale" name="reflectance"><rgb name="scale" value=".9 .3 .25"/>
    <texture type="bitmap"><string name="filename" value="%(rgb)s"/><string name="filterType" value="nearest"/>
    <string name="wrapMode" value="mirror"/><float name="uvscale" value="1.5"/></texture></texture></bsdf></shape>
<shape type="cube"><bsdf type="plastic"><texture type="bitmap" name="diffuseReflectance"><string name="filename" value="%(grey)s"/>
    <string name="filterType" value="trilinear"/><string name="wrapModeU" value="clamp"/><string name="wrapModeV" value="black"/></texture></bsdf></shape>
<shape type="cube"><bsdf type="twosided">
    <bsdf type="roughconductor"><string name="material" value="none"/><rgb name="eta" value=".2 .92 1.1"/><rgb name="k" value="3.9 2.45 2.14"/>
        <float name="alpha" value=".15"/><texture type="scale" name="specularReflectance"><float name="scale" value=".95"/>
        <texture type="bitmap"><string name="filename" value="%(rgb)s"/><string name="filterType" value="bilinear"/></texture></texture></bsdf>
    <bsdf type="diffuse"><texture type="scale" name="reflectance"><rgb name="scale" value=".5 1 .5"/><texture type="checkerboard">
        <spectrum name="color0" value=".8"/><spectrum name="color1" value=".1"/><float name="uvscale" value="2"/></texture></texture></bsdf>
</bsdf></shape>
<shape type="rectangle"><transform name="toWorld"><translate y="3"/></transform><emitter type="area"><rgb name="radiance" value="18 15 8"/></emitter></shape>
"""


@pytest.fixture()
def textured(tmp_path):
    img = scenes.pattern_image(32, 24)
    grey = scenes.pattern_image(16, 32, seed=5)[:, :, 1]
    write_pfm(tmp_path / "rgb.pfm", img)
    write_pfm(tmp_path / "grey.pfm", grey)
    hs = load(TEXTURED % dict(rgb=tmp_path / "rgb.pfm", grey=tmp_path / "grey.pfm"))
    return hs, img, grey


def test_texture_records_equal_the_python_builder(textured):
    hs, img, grey = textured
    assert hs.warnings == []
    d = hs.desc
    b = scenedesc.SceneBuilder()
    want = [b.add_checkerboard([.7, .7, .65], [.15, .2, .3], uscale=4, vscale=4),
            b.add_bitmap(img, "ewa", "repeat", uscale=3, vscale=2, max_anisotropy=8.0),
            b.add_gridtexture([.14, .45, .091], [.9, .9, .2], .04, uscale=5, vscale=5, uoffset=.3),
            b.add_bitmap(img, "nearest", "mirror", uscale=1.5, vscale=1.5, scale=[.9, .3, .25]),
            b.add_bitmap(grey, "trilinear", wrap_mode_u="clamp", wrap_mode_v="zero", max_anisotropy=1.0),
            b.add_bitmap(img, "bilinear", "repeat", scale=.95, max_anisotropy=1.0),
            b.add_checkerboard(.8, .1, uscale=2, vscale=2, scale=[.5, 1, .5])]
    # non-EWA filters reset maxAnisotropy to 1 (bitmap.cpp:234-235)
    b.textures[3].max_anisotropy = 1.0
    assert d.n_textures == len(want)
    for i in range(d.n_textures):
        assert same_record(d.textures[i], b.textures[i]), i
    # every textured material points at its record: shapes in order, the twosided adapter's front side first
    mats = [d.materials[d.shapes[i].material] for i in range(d.n_shapes)]
    assert [m.tex_reflectance for m in mats[:4]] == [1, 2, 3, 4]
    plastic = mats[4]
    assert plastic.type == abi.BSDF_PLASTIC and plastic.tex_reflectance == 5 and plastic.tex_transmittance == 0
    front = mats[5]
    assert front.flags & abi.MAT_TWOSIDED and front.tex_reflectance == 6 and front.back >= 0
    assert d.materials[front.back].tex_reflectance == 7
    # the cube and rectangle plugins carry texture coordinates
    assert all(d.shapes[i].has_texcoords for i in range(d.n_shapes))


def test_plastic_sampling_weight_uses_the_texture_average(textured):
    hs, img, grey = textured
    d = hs.desc
    m = [d.materials[d.shapes[i].material] for i in range(d.n_shapes)][4]
    # plastic.cpp:199-204 with barray.h:103-126's float running sum
    acc = np.float32(0)
    for v in grey.ravel():
        acc = np.float32(acc + v)
    d_avg = np.float32(acc / np.float32(grey.size))
    assert m.spec_weight == pytest.approx(1.0 / (float(d_avg) + 1.0), rel=2e-6)


def test_energy_conservation_wraps_the_texture_in_a_scale():
    hs = load('<shape type="rectangle"><bsdf type="diffuse"><texture type="checkerboard" name="reflectance">'
              '<rgb name="color0" value="2 .5 .5"/></texture></bsdf></shape>')
    assert any("conserve energy" in w for w in hs.warnings)
    t = hs.desc.textures[0]
    assert list(t.scale) == [np.float32(0.99) * (np.float32(1.0) / np.float32(2.0))] * 3
    # roughconductor is left alone in this fork (roughconductor.cpp:243)
    hs = load('<shape type="rectangle"><bsdf type="roughconductor"><texture type="checkerboard" name="specularReflectance">'
              '<rgb name="color0" value="2 .5 .5"/></texture></bsdf></shape>')
    assert hs.warnings == [] and list(hs.desc.textures[0].scale) == [1, 1, 1]


def test_scale_around_a_constant_stays_a_constant():
    hs = load('<shape type="rectangle"><bsdf type="diffuse"><texture type="scale" name="reflectance"><float name="scale" value=".5"/>'
              '<rgb name="value" value=".8 .4 .2"/></texture></bsdf></shape>')
    d = hs.desc
    assert d.n_textures == 0
    m = d.materials[d.shapes[0].material]
    assert m.tex_reflectance == 0 and np.allclose(list(m.reflectance), [.4, .2, .1], rtol=1e-7)


@pytest.mark.parametrize("body,msg", [
    ('<texture type="bitmap" name="reflectance"><string name="filename" value="/nonexistent.pfm"/></texture>', "could not be found"),
    ('<texture type="checkerboard" name="reflectance"><string name="coordinates" value="object"/></texture>', "Only UV coordinates"),
    ('<texture type="scale" name="reflectance"/>', "needs a nested texture"),
    ('<texture type="wrinkle" name="reflectance"/>', "not found"),
    ('<texture type="checkerboard" name="alpha"/>', "does not accept children"),
])
def test_errors(body, msg):
    with pytest.raises(MtsbError) as e:
        load('<shape type="rectangle"><bsdf type="diffuse">%s</bsdf></shape>' % body)
    assert msg in str(e.value)


def test_bitmap_property_errors(tmp_path):
    write_pfm(tmp_path / "a.pfm", scenes.pattern_image(8, 8))
    for prop, val, msg in (("filterType", "cubic", "Unknown filter type"), ("wrapMode", "tile", "Unknown wrap mode")):
        with pytest.raises(MtsbError) as e:
            load('<shape type="rectangle"><bsdf type="diffuse"><texture type="bitmap" name="reflectance"><string name="filename" value="%s"/>'
                 '<string name="%s" value="%s"/></texture></bsdf></shape>' % (tmp_path / "a.pfm", prop, val))
        assert msg in str(e.value)
    with pytest.raises(MtsbError) as e:
        load('<shape type="rectangle"><bsdf type="roughconductor"><texture type="checkerboard" name="alphaV"/></bsdf></shape>')
    assert "not part of the B200 path" in str(e.value)


def _bitmap_texels(path, extra=""):
    hs = load('<shape type="rectangle"><bsdf type="diffuse"><texture type="bitmap" name="reflectance"><string name="filename" value="%s"/>%s'
              '</texture></bsdf></shape>' % (path, extra))
    return texels_of(hs.desc.textures[0]), hs


def test_png_reader(tmp_path):
    rng = np.random.default_rng(7)
    w, h = 13, 11
    # 8-bit RGB, sRGB by default; every scanline filter type occurs
    rgb = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    write_png(tmp_path / "rgb8.png", [bytes(r.tobytes()) for r in rgb], w, h, 8, 2)
    tx, _ = _bitmap_texels(tmp_path / "rgb8.png")
    assert tx.shape == (h, w, 3) and close1(tx, srgb_to_linear(rgb, 255))
    # RGBA: alpha dropped (bitmap.cpp:274-277)
    rgba = np.concatenate([rgb, rng.integers(0, 256, (h, w, 1), dtype=np.uint8)], -1)
    write_png(tmp_path / "rgba8.png", [bytes(r.tobytes()) for r in rgba], w, h, 8, 6)
    assert close1(_bitmap_texels(tmp_path / "rgba8.png")[0], srgb_to_linear(rgb, 255))
    # 16-bit grey with a gAMA chunk (1/2.2): luminance texture, pow(v, 2.2)
    g16 = rng.integers(0, 65536, (h, w), dtype=np.uint16)
    write_png(tmp_path / "g16.png", [bytes(r.astype(">u2").tobytes()) for r in g16], w, h, 16, 0, chunks=[(b"gAMA", struct.pack(">I", 45455))])
    tx, _ = _bitmap_texels(tmp_path / "g16.png")
    gamma = np.float32(1.0) / np.float32(45455 / 100000.0)
    want = np.power(np.float32(g16) * (np.float32(1.0) / np.float32(65535)), gamma).astype(np.float32)
    assert tx.shape == (h, w, 1) and np.allclose(tx[:, :, 0], want, rtol=1e-6, atol=0)
    # palette image, 4 bits per index; sRGB chunk
    pal = rng.integers(0, 256, (16, 3), dtype=np.uint8)
    idx = rng.integers(0, 16, (h, w), dtype=np.uint8)
    rows = []
    for r in idx:
        r = np.concatenate([r, np.zeros((-w) % 2, np.uint8)])
        rows.append(bytes(((r[0::2] << 4) | r[1::2]).astype(np.uint8).tobytes()))
    write_png(tmp_path / "pal4.png", rows, w, h, 4, 3, chunks=[(b"sRGB", b"\0")], palette=pal.ravel().tolist())
    assert close1(_bitmap_texels(tmp_path / "pal4.png")[0], srgb_to_linear(pal[idx], 255))
    # 1-bit grey expands to 0 / 255 (255 * float(1/255) is not exactly 1: fmtconv.cpp:1143-1144 computes it that way)
    bits = rng.integers(0, 2, (h, w), dtype=np.uint8)
    rows = [bytes(np.packbits(r).tobytes()) for r in bits]
    write_png(tmp_path / "g1.png", rows, w, h, 1, 0)
    assert close1(_bitmap_texels(tmp_path / "g1.png")[0][:, :, 0], srgb_to_linear(bits * np.uint8(255), 255))
    # `gamma` overrides what the file declares (bitmap.cpp:218, 251-252 -> Bitmap::setGamma; fmtconv.cpp:1091-1100):
    # 1 = the stored values ARE linear (data maps), 2.2 = pow(v, 2.2), -1 = sRGB
    tx, hs = _bitmap_texels(tmp_path / "rgb8.png", '<float name="gamma" value="1"/>')
    assert not hs.warnings and np.allclose(tx, np.float32(rgb) * (np.float32(1.0) / np.float32(255)), rtol=1e-6, atol=0)
    tx, _ = _bitmap_texels(tmp_path / "rgb8.png", '<float name="gamma" value="2.2"/>')
    assert np.allclose(tx, np.power(np.float32(rgb) * (np.float32(1.0) / np.float32(255)), np.float32(2.2)), rtol=2e-6, atol=0)
    tx, _ = _bitmap_texels(tmp_path / "g16.png", '<float name="gamma" value="-1"/>')
    assert close1(tx[:, :, 0], srgb_to_linear(g16, 65535))
    # a single channel of an RGB image
    tx, _ = _bitmap_texels(tmp_path / "rgb8.png", '<string name="channel" value="g"/>')
    assert tx.shape == (h, w, 1) and close1(tx[:, :, 0], srgb_to_linear(rgb[:, :, 1], 255))


def test_negative_texels_are_clamped(tmp_path):
    img = scenes.pattern_image(8, 8) - np.float32(0.3)
    write_pfm(tmp_path / "neg.pfm", img)
    tx, hs = _bitmap_texels(tmp_path / "neg.pfm")
    assert any("negative pixel values" in w for w in hs.warnings)
    assert np.array_equal(tx, np.maximum(img, 0))


def test_oracle_renders_the_loaded_scene_like_the_builder_scene(tmp_path):
    """The flattened description is complete: the oracle shades the XML scene exactly as it shades the same scene
    written through the SceneBuilder (same counter-based streams -> same film, bit for bit)."""
    import oracle
    img = scenes.pattern_image(32, 32)
    write_pfm(tmp_path / "t.pfm", img)
    xml = ('<scene version="0.5.0"><integrator type="path"><integer name="maxDepth" value="4"/></integrator>'
           '<sensor type="perspective"><transform name="toWorld"><lookat origin="0 0 4" target="0 0 0" up="0 1 0"/></transform>'
           '<float name="fov" value="40"/><sampler type="independent"><integer name="sampleCount" value="2"/></sampler>'
           '<film type="hdrfilm"><integer name="width" value="24"/><integer name="height" value="16"/></film></sensor>'
           '<shape type="rectangle"><bsdf type="diffuse"><texture type="bitmap" name="reflectance"><string name="filename" value="%s"/>'
           '<float name="uvscale" value="6"/></texture></bsdf></shape>'
           '<emitter type="constant"><rgb name="radiance" value="1 1 1"/></emitter></scene>' % (tmp_path / "t.pfm"))
    hs = hostlib.HostScene.load_string(xml, "t.xml")
    o1 = oracle.Oracle(hs.desc_pair())
    o1.render(spp=2, seed=5, counter_rng=True)
    f1 = o1.film_download()
    # same scene through the builder, geometry taken from the loader
    pos, nrm, uv, idx = hs.arrays()
    b = scenedesc.SceneBuilder()
    d = hs.desc
    b.set_integrator(max_depth=4, rr_depth=d.integrator.rr_depth, sample_count=2)
    m = scenedesc.diffuse(0.5)
    m.tex_reflectance = b.add_bitmap(img, "ewa", "repeat", uscale=6, vscale=6)
    b.add_mesh(pos, idx, b.add_material(m), normals=nrm if nrm is not None and len(nrm) else None, texcoords=uv)
    b.add_constant_env([1, 1, 1])
    desc2, keep2 = b.desc()
    desc2.sensor = d.sensor
    o2 = oracle.Oracle((desc2, keep2))
    o2.render(spp=2, seed=5, counter_rng=True)
    f2 = o2.film_download()
    assert np.isfinite(f1).all() and f1[..., :3].max() > 0.05
    assert np.array_equal(f1, f2)


def test_textured_roughness_goes_to_the_alpha_slot():
    """<texture name="alpha"> on a rough BSDF (roughconductor.cpp:241-247): the third texture slot of the record; a constant
    under a `scale` stays a constant roughness; separately textured alphaU / alphaV are refused."""
    hs = load('<shape type="rectangle"><bsdf type="roughconductor"><string name="distribution" value="ggx"/>'
              '<texture type="checkerboard" name="alpha"><spectrum name="color0" value="0.05"/><spectrum name="color1" value="0.4"/></texture>'
              '</bsdf></shape>')
    m = hs.materials()[0]
    assert m.tex_alpha == 1 and m.tex_reflectance == 0 and hs.desc.n_textures == 1
    t = hs.desc.textures[0]
    assert t.type == abi.TEX_CHECKERBOARD and abs(t.color0[0] - 0.05) < 1e-7 and abs(t.color1[0] - 0.4) < 1e-7
    with pytest.raises(MtsbError):
        load('<shape type="rectangle"><bsdf type="roughdielectric"><texture type="checkerboard" name="alphaU"/></bsdf></shape>')
