"""Writes scenes/cbox.xml + scenes/meshes/cbox_*.obj from the same Cornell measurements that
mitsuba_b200.scenes.cbox() uses (the reference tree ships no cbox scene).  Run from the repo root:
    python scenes/make_cbox.py
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from mitsuba_b200 import scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = ["floor", "ceiling", "back", "greenwall", "redwall", "smallbox", "largebox", "luminaire"]
BSDF = ["white", "white", "white", "green", "red", "white", "white", None]


def main():
    b = scenes.cbox()
    pos, _, _, idx = b.arrays()
    for name, sh in zip(NAMES, b.shapes):
        v0, nv, t0, nt = sh["first_vertex"], sh["n_vertices"], sh["first_tri"], sh["n_tris"]
        with open(os.path.join(HERE, "meshes", "cbox_%s.obj" % name), "w") as f:
            f.write("# Cornell box: %s\n" % name)
            for p in pos[v0:v0 + nv]:
                f.write("v %s %s %s\n" % tuple(repr(float(x)) for x in p))
            # quads are written as 4-gons: the loader fans them into (0,1,2),(0,2,3) like SceneBuilder.add_quad
            tris = idx[t0:t0 + nt] - v0 + 1
            for k in range(0, nt, 2):
                a, b_, c = tris[k]
                a2, c2, d = tris[k + 1]
                assert a == a2 and c == c2
                f.write("f %d %d %d %d\n" % (a, b_, c, d))
    with open(os.path.join(HERE, "cbox.xml"), "w") as f:
        f.write(XML)


XML = """<?xml version="1.0" encoding="utf-8"?>
<!-- Cornell box (BASELINE.json configs[0]): path integrator, 512x512, 64 spp, maxDepth 8.
     Usage: mtsb200 scenes/cbox.xml -o cbox.pfm   (override with -Dspp=16 -Dres=256 -Ddepth=5) -->
<scene version="0.5.0">
	<default name="spp" value="64"/>
	<default name="res" value="512"/>
	<default name="depth" value="8"/>

	<integrator type="path">
		<integer name="maxDepth" value="$depth"/>
		<integer name="rrDepth" value="5"/>
	</integrator>

	<sensor type="perspective">
		<float name="fov" value="39.3077"/>
		<string name="fovAxis" value="x"/>
		<float name="nearClip" value="10"/>
		<float name="farClip" value="2800"/>
		<transform name="toWorld">
			<lookat origin="278, 273, -800" target="278, 273, -799" up="0, 1, 0"/>
		</transform>
		<sampler type="independent">
			<integer name="sampleCount" value="$spp"/>
		</sampler>
		<film type="hdrfilm">
			<integer name="width" value="$res"/>
			<integer name="height" value="$res"/>
			<rfilter type="gaussian"/>
		</film>
	</sensor>

	<bsdf type="diffuse" id="white">
		<rgb name="reflectance" value="0.725, 0.71, 0.68"/>
	</bsdf>
	<bsdf type="diffuse" id="red">
		<rgb name="reflectance" value="0.63, 0.065, 0.05"/>
	</bsdf>
	<bsdf type="diffuse" id="green">
		<rgb name="reflectance" value="0.14, 0.45, 0.091"/>
	</bsdf>
%s
</scene>
"""

SHAPE = """
	<shape type="obj">
		<string name="filename" value="meshes/cbox_%s.obj"/>
		<boolean name="faceNormals" value="true"/>
%s	</shape>"""

_body = ""
for _n, _m in zip(NAMES, BSDF):
    if _m:
        _body += SHAPE % (_n, '\t\t<ref id="%s"/>\n' % _m)
    else:
        _body += SHAPE % (_n, '\t\t<emitter type="area">\n\t\t\t<rgb name="radiance" value="18.4, 15.6, 8.0"/>\n\t\t</emitter>\n')
XML = XML % _body

if __name__ == "__main__":
    main()
