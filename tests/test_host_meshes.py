"""Mesh readers and shape plugins of the host side: OBJ / PLY parsing order (which fixes the global
primitive ids), vertex merging, normal generation (trimesh.cpp:608-681), cube / rectangle data,
<include>.  Pure CPU."""
import os
import struct

import numpy as np
import pytest

from mitsuba_b200 import hostlib
from mitsuba_b200.lib import MtsbError

HEAD = '<scene version="0.5.0"><integrator type="path"/><sensor type="perspective"/>'


def load_dir(tmp_path, body):
    p = tmp_path / "scene.xml"
    p.write_text(HEAD + body + "</scene>")
    return hostlib.HostScene.load_file(str(p))


def test_obj_faces_fans_negative_indices_and_merging(tmp_path):
    (tmp_path / "m.obj").write_text(
        "# comment\n"
        "v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv 0.5 2 0\n"
        "f 1 2 3 4 5\n"          # pentagon -> fan (1,2,3) (1,3,4) (1,4,5)
        "f -5 -4 -3\n"           # negative = relative to the end: (1,2,3) again
        "f 3 \\\n 4 5\n")        # line continuation
    hs = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="m.obj"/><boolean name="faceNormals" value="true"/></shape>')
    pos, nrm, uv, idx = hs.arrays()
    assert pos.shape == (5, 3)                               # merged: 5 unique vertices in order of first use
    assert idx.tolist() == [[0, 1, 2], [0, 2, 3], [0, 3, 4], [0, 1, 2], [2, 3, 4]]
    assert nrm is None and uv is None
    assert hs.shapes()[0].has_normals == 0


def test_obj_normals_texcoords_and_groups(tmp_path):
    (tmp_path / "g.obj").write_text(
        "v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\n"
        "vn 0 0 2\nvn 0 0 -1\n"
        "vt 0.25 0.75\nvt 1 0\n"
        "g first\n"
        "f 1/1/1 2/2/1 3/1/1\n"
        "f 1//2 2//2 3//2\n"        # same positions, other normal -> new vertices
        "g second\n"
        "f 1/2 2/2 4/2\n")
    hs = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="g.obj"/>'
                            '<transform name="toWorld"><scale x="2"/><translate y="5"/></transform></shape>')
    d = hs.desc
    assert d.n_shapes == 2                                    # one mesh per group (obj.cpp:250-275)
    s0, s1 = hs.shapes()
    assert (s0.n_tris, s0.n_vertices, s0.has_normals, s0.has_texcoords) == (2, 6, 1, 1)
    assert (s1.first_tri, s1.n_tris, s1.n_vertices, s1.has_texcoords) == (2, 1, 3, 1)
    assert s0.material == s1.material                          # both use the shape's BSDF
    pos, nrm, uv, idx = hs.arrays()
    assert np.allclose(pos[1], [2, 5, 0])                      # toWorld applied
    # normals: inverse transpose then normalised (scale x2 leaves +z)
    assert np.allclose(nrm[0], [0, 0, 1]) and np.allclose(nrm[3], [0, 0, -1])
    assert np.allclose(uv[0], [0.25, 0.25])                    # flipTexCoords default true: v -> 1-v (obj.cpp:210, 301-302)
    assert np.all(idx[2] >= s1.first_vertex)                   # global indices
    # second mesh has no normals in the file: smooth normals are generated for it
    assert s1.has_normals == 1
    hs = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="g.obj"/><boolean name="collapse" value="true"/>'
                            '<boolean name="flipTexCoords" value="false"/></shape>')
    assert hs.desc.n_shapes == 1 and hs.shapes()[0].n_tris == 3
    assert np.allclose(hs.arrays()[2][0], [0.25, 0.75])
    hs = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="g.obj"/><integer name="shapeIndex" value="1"/></shape>')
    assert hs.desc.n_shapes == 1 and hs.shapes()[0].n_tris == 1
    with pytest.raises(MtsbError, match="multiple objects"):
        load_dir(tmp_path, '<shape type="obj"><string name="filename" value="g.obj"/><emitter type="area"/></shape>')


def test_obj_out_of_range_index(tmp_path):
    (tmp_path / "bad.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 7\n")
    with pytest.raises(MtsbError, match="Out of bounds: tried to access vertex 7"):
        load_dir(tmp_path, '<shape type="obj"><string name="filename" value="bad.obj"/></shape>')


def _angle_weighted_normals(pos, idx):
    n = np.zeros_like(pos, dtype=np.float64)
    for t in idx:
        p = pos[t].astype(np.float64)
        fn = np.cross(p[1] - p[0], p[2] - p[0])
        if np.linalg.norm(fn) == 0:
            continue
        fn /= np.linalg.norm(fn)
        for i in range(3):
            a, b = p[(i + 1) % 3] - p[i], p[(i + 2) % 3] - p[i]
            a, b = a / np.linalg.norm(a), b / np.linalg.norm(b)
            n[t[i]] += fn * np.arccos(np.clip(a @ b, -1, 1))
    return n / np.linalg.norm(n, axis=1, keepdims=True)


def test_generated_normals_are_angle_weighted(tmp_path):
    rng = np.random.default_rng(5)
    pos = rng.normal(size=(12, 3))
    idx = np.array([[0, 1, 2], [2, 1, 3], [3, 4, 5], [5, 1, 0], [6, 7, 8], [8, 9, 10], [10, 11, 6], [0, 6, 3]])
    with open(tmp_path / "n.obj", "w") as f:
        for p in pos:
            f.write("v %r %r %r\n" % tuple(float(x) for x in p))
        for t in idx:
            f.write("f %d %d %d\n" % tuple(int(x) + 1 for x in t))
    hs = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="n.obj"/></shape>')
    hpos, nrm, _, hidx = hs.arrays()
    ref = _angle_weighted_normals(hpos, hidx)
    assert np.allclose(nrm, ref, atol=2e-5)
    hs2 = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="n.obj"/><boolean name="flipNormals" value="true"/></shape>')
    assert np.allclose(hs2.arrays()[1], -ref, atol=2e-5)
    # faceNormals + flipNormals swaps the winding instead (trimesh.cpp:616-622)
    hs3 = load_dir(tmp_path, '<shape type="obj"><string name="filename" value="n.obj"/><boolean name="faceNormals" value="true"/>'
                             '<boolean name="flipNormals" value="true"/></shape>')
    assert hs3.arrays()[1] is None
    assert np.array_equal(hs3.arrays()[3], hidx[:, [1, 0, 2]])


def _write_ply(path, fmt):
    verts = [(0, 0, 0, 0, 0, 1, 0, 0), (1, 0, 0, 0, 0, 1, 1, 0), (1, 1, 0, 0, 0, 1, 1, 1), (0, 1, 0, 0, 0, 1, 0, 1), (2, 2, 2, 0, 1, 0, .5, .5)]
    faces = [(0, 1, 2, 3), (0, 2, 4)]
    hdr = ("ply\nformat %s 1.0\ncomment test\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\n"
           "property float nx\nproperty float ny\nproperty float nz\nproperty float u\nproperty float v\n"
           "element face 2\nproperty list uchar int vertex_indices\nend_header\n") % fmt
    with open(path, "wb") as f:
        f.write(hdr.encode())
        if fmt == "ascii":
            for v in verts:
                f.write((" ".join(repr(float(x)) for x in v) + "\n").encode())
            for fc in faces:
                f.write(("%d %s\n" % (len(fc), " ".join(str(i) for i in fc))).encode())
        else:
            e = "<" if fmt == "binary_little_endian" else ">"
            for v in verts:
                f.write(struct.pack(e + "8f", *v))
            for fc in faces:
                f.write(struct.pack(e + "B%di" % len(fc), len(fc), *fc))


@pytest.mark.parametrize("fmt", ["ascii", "binary_little_endian", "binary_big_endian"])
def test_ply_formats(tmp_path, fmt):
    _write_ply(tmp_path / "m.ply", fmt)
    hs = load_dir(tmp_path, '<shape type="ply"><string name="filename" value="m.ply"/></shape>')
    pos, nrm, uv, idx = hs.arrays()
    assert pos.shape == (5, 3) and np.allclose(pos[4], [2, 2, 2])
    assert idx.tolist() == [[0, 1, 2], [0, 2, 3], [0, 2, 4]]           # quad -> (0,1,2),(0,2,3)
    assert np.allclose(nrm[4], [0, 1, 0]) and np.allclose(uv[4], [0.5, 0.5])


def test_ply_errors(tmp_path):
    (tmp_path / "x.ply").write_text("plx\n")
    with pytest.raises(MtsbError, match="not a PLY file"):
        load_dir(tmp_path, '<shape type="ply"><string name="filename" value="x.ply"/></shape>')
    (tmp_path / "y.ply").write_text("ply\nformat ascii 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\n"
                                    "element face 1\nproperty list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n0 1 0\n3 0 1 5\n")
    with pytest.raises(MtsbError, match="out of range"):
        load_dir(tmp_path, '<shape type="ply"><string name="filename" value="y.ply"/></shape>')


def test_cube_layout():
    hs = hostlib.HostScene.load_string(HEAD + '<shape type="cube"><transform name="toWorld"><scale x="1" y="2" z="3"/></transform></shape></scene>')
    pos, nrm, uv, idx = hs.arrays()
    assert pos.shape == (24, 3) and idx.shape == (12, 3)
    assert pos.min(0).tolist() == [-1, -2, -3] and pos.max(0).tolist() == [1, 2, 3]
    # face order -y +y +x +z -x -z, two triangles (0,1,2),(3,0,2) per face (cube.cpp:24-30)
    assert idx[:2].tolist() == [[0, 1, 2], [3, 0, 2]] and idx[10:].tolist() == [[20, 21, 22], [23, 20, 22]]
    face_n = [[0, -1, 0], [0, 1, 0], [1, 0, 0], [0, 0, 1], [-1, 0, 0], [0, 0, -1]]
    assert np.allclose(nrm.reshape(6, 4, 3), np.array(face_n, np.float32)[:, None, :])
    # winding agrees with the outward normal
    p = pos[idx]
    g = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
    assert np.all(np.einsum("ij,ij->i", g, nrm[idx[:, 0]]) > 0)
    assert np.allclose(uv[:4], [[0, 1], [1, 1], [1, 0], [0, 0]])


def test_include_merges_the_child_scene(tmp_path):
    (tmp_path / "geo.xml").write_text('<scene version="0.5.0"><bsdf type="diffuse" id="grey"><rgb name="reflectance" value="0.3"/></bsdf>'
                                      '<shape type="cube"><ref id="grey"/></shape>'
                                      '<emitter type="point"><point name="position" x="0" y="5" z="0"/></emitter></scene>')
    hs = load_dir(tmp_path, '<shape type="rectangle"/><include filename="geo.xml"/><shape type="rectangle"><ref id="grey"/></shape>')
    d = hs.desc
    assert d.n_shapes == 3 and d.n_emitters == 1 and d.n_tris == 2 + 12 + 2
    sh = hs.shapes()
    # ids defined in the included file are visible afterwards (shared NamedObjectMap, scenehandler.cpp:675)
    assert list(d.materials[sh[2].material].reflectance) == pytest.approx([0.3] * 3)
    assert sh[1].material == sh[2].material


def _serialized_blob(version, pos, idx, nrm=None, uv=None, col=None, double=False, name="m"):
    """One mesh of a .serialized container (TriMesh::serialize, src/librender/trimesh.cpp:838-905)."""
    import zlib
    flags = (0x2000 if double else 0x1000) | (1 if nrm is not None else 0) | (2 if uv is not None else 0) | (8 if col is not None else 0)
    ft = "<f8" if double else "<f4"
    body = struct.pack("<I", flags)
    if version == 4:
        body += name.encode() + b"\0"
    body += struct.pack("<QQ", pos.shape[0], idx.shape[0]) + pos.astype(ft).tobytes()
    for a in (nrm, uv, col):
        if a is not None:
            body += a.astype(ft).tobytes()
    body += idx.astype("<u4").tobytes()
    return struct.pack("<HH", 0x041C, version) + zlib.compress(body)


def _write_serialized(path, version, blobs):
    offsets, data = [], b""
    for b in blobs:
        offsets.append(len(data))
        data += b
    data += b"".join(struct.pack("<Q" if version == 4 else "<I", o) for o in offsets) + struct.pack("<I", len(blobs))
    path.write_bytes(data)


@pytest.mark.parametrize("version,double", [(3, False), (4, False), (4, True)])
def test_serialized_container(tmp_path, version, double):
    """`serialized` shape plugin (serialized.cpp:148-210, trimesh.cpp:175-298): both container versions, float and double
    payloads, the shape dictionary at the end of the file, toWorld incl. the handedness flip, optional arrays."""
    rng = np.random.default_rng(version * 2 + double)
    meshes = []
    for k in range(3):
        nv, nt = 5 + 3 * k, 4 + k
        pos = rng.normal(size=(nv, 3)).astype(np.float32)
        idx = np.stack([rng.permutation(nv)[:3] for _ in range(nt)]).astype(np.uint32)
        nrm = rng.normal(size=(nv, 3)).astype(np.float32)
        nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
        uv = rng.random((nv, 2)).astype(np.float32)
        meshes.append((pos, idx, nrm if k != 1 else None, uv if k != 2 else None, rng.random((nv, 3)).astype(np.float32) if k == 0 else None))
    _write_serialized(tmp_path / "m.serialized", version,
                      [_serialized_blob(version, p, i, n, u, c, double, "mesh%d" % k) for k, (p, i, n, u, c) in enumerate(meshes)])
    for k, (p, i, n, u, c) in enumerate(meshes):
        hs = load_dir(tmp_path, '<shape type="serialized"><string name="filename" value="m.serialized"/><integer name="shapeIndex" value="%d"/></shape>' % k)
        pos, nrm, uv, idx = hs.arrays()
        assert np.array_equal(pos, p) and np.array_equal(idx, i)
        sh = hs.shapes()[0]
        assert sh.has_texcoords == (u is not None)
        if u is not None:
            assert np.array_equal(uv, u)
        if n is not None:
            assert np.allclose(nrm, n, atol=1e-6)
        else:                              # no normals in the file: smooth normals are generated (trimesh.cpp:446-530)
            assert sh.has_normals and np.allclose(np.linalg.norm(nrm, axis=1), 1, atol=1e-5)
    # mirrored toWorld: positions transformed, winding swapped (serialized.cpp:197-202)
    hs = load_dir(tmp_path, '<shape type="serialized"><string name="filename" value="m.serialized"/>'
                            '<transform name="toWorld"><scale x="-2"/><translate y="1"/></transform></shape>')
    pos, nrm, uv, idx = hs.arrays()
    p, i = meshes[0][0], meshes[0][1]
    assert np.allclose(pos, p * [-2, 1, 1] + [0, 1, 0], atol=1e-6)
    assert np.array_equal(idx, i[:, [1, 0, 2]])
    for body, msg in (('<integer name="shapeIndex" value="7"/>', "shape index is out of range"), ('<integer name="shapeIndex" value="-1"/>', "nonnegative")):
        with pytest.raises(MtsbError) as ei:
            load_dir(tmp_path, '<shape type="serialized"><string name="filename" value="m.serialized"/>%s</shape>' % body)
        assert msg in str(ei.value)
    (tmp_path / "bad.serialized").write_bytes(struct.pack("<HH", 0x1C04, 3) + b"xx")
    with pytest.raises(MtsbError) as ei:
        load_dir(tmp_path, '<shape type="serialized"><string name="filename" value="bad.serialized"/></shape>')
    assert "old version of Mitsuba" in str(ei.value)
