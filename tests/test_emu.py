"""CPU: the kernel bodies of the BVH builder and the traversal (mitsuba_b200/csrc/bvh.h,
traverse.h), compiled with g++ by tests/emu, against the reference kd-tree's hit records
(tests/golden) and against the oracle.  This checks the LOGIC the GPU runs; the GPU parity
tests proper are tests/test_gpu_*.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from mitsuba_b200 import abi, scenes
from tests import util

EMU_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu")


@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-C", EMU_DIR], stdout=subprocess.DEVNULL)
    E = C.CDLL(os.path.join(EMU_DIR, "libemu.so"))
    E.emu_build.restype = C.c_void_p
    E.emu_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    E.emu_trace.restype = C.c_uint64
    E.emu_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    E.emu_info.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    E.emu_destroy.argtypes = [C.c_void_p]
    return E


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def build(E, b):
    pos, _, _, idx = b.arrays()
    return E.emu_build(vp(pos), vp(idx), idx.shape[0]), (pos, idx)


def trace(E, h, rays, mode):
    rays = np.ascontiguousarray(rays, dtype=abi.RAY_DTYPE)
    hits = np.zeros(rays.shape[0], dtype=abi.HIT_DTYPE)
    steps = E.emu_trace(h, vp(rays), rays.shape[0], mode, vp(hits))
    return hits, steps


@pytest.mark.parametrize("name", ["cbox", "mesh"])
def test_bvh_traversal_vs_reference_kdtree(emu, name):
    g = util.load_golden("hits_%s.npz" % name)
    h, keep = build(emu, util.golden_scenes()[name])
    nn, t6, e6 = C.c_uint32(), np.zeros(6, np.float32), np.zeros(6, np.float32)
    emu.emu_info(h, C.byref(nn), vp(t6), vp(e6))
    assert np.array_equal(t6, g["tight"]) and np.array_equal(e6, g["enlarged"])
    for rs in ("chord", "camera", "secondary"):
        hits, _ = trace(emu, h, g["rays_" + rs], abi.TRACE_CLOSEST)
        st = util.compare_hits(hits, g["closest_" + rs], label=name + "/" + rs)
        assert st["bit_exact"] >= 0.995 * st["n"], st
        occl, _ = trace(emu, h, g["rays_" + rs], abi.TRACE_ANY)
        assert ((occl["prim"] != abi.NO_HIT) != g["any_" + rs]).sum() <= max(3, 2e-3 * len(hits))
    emu.emu_destroy(h)


@pytest.mark.parametrize("n_tris", [0, 1, 2, 3, 4, 5, 9, 33])
def test_tiny_and_ragged_scenes(emu, oracle_mod, n_tris):
    """empty scene, fewer triangles than one leaf, and counts around the leaf / node width"""
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    rng = np.random.default_rng(n_tris)
    b = SceneBuilder()
    m = b.add_material(diffuse(0.5))
    if n_tris:
        tri = rng.normal(size=(n_tris, 3, 3)).astype(np.float32)
        b.add_mesh(tri.reshape(-1, 3), np.arange(n_tris * 3, dtype=np.uint32).reshape(-1, 3), m)
    h, keep = build(emu, b)
    rays = scenes.chord_rays([0, 0, 0], 4.0, 4000, seed=3)
    hits, _ = trace(emu, h, rays, abi.TRACE_CLOSEST)
    ref = oracle_mod.Oracle(b).trace(rays, abi.TRACE_CLOSEST)
    util.compare_hits(hits, ref, label="n=%d" % n_tris)
    if n_tris == 0:
        assert (hits["prim"] == abi.NO_HIT).all()
    emu.emu_destroy(h)


def test_duplicate_and_degenerate_triangles(emu, oracle_mod):
    """identical Morton keys (duplicated triangles) and zero-area triangles (k = 3 never hit)"""
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    rng = np.random.default_rng(5)
    base = rng.normal(size=(20, 3, 3)).astype(np.float32)
    tri = np.concatenate([base, base, base[:5]])
    tri[3, 2] = tri[3, 1]
    b = SceneBuilder()
    b.add_mesh(tri.reshape(-1, 3), np.arange(tri.shape[0] * 3, dtype=np.uint32).reshape(-1, 3), b.add_material(diffuse(0.5)))
    h, keep = build(emu, b)
    rays = scenes.chord_rays([0, 0, 0], 4.0, 4000, seed=9)
    hits, _ = trace(emu, h, rays, abi.TRACE_CLOSEST)
    ref = oracle_mod.Oracle(b).trace(rays, abi.TRACE_CLOSEST)
    # duplicates tie at equal t by construction: only t/u/v and hit/miss can be compared
    th, rh = hits["prim"] != abi.NO_HIT, ref["prim"] != abi.NO_HIT
    assert (th == rh).all()
    assert np.array_equal(hits["t"][th], ref["t"][rh])
    assert (hits["prim"][th] != 3).all()            # the zero-area copy is never reported (triaccel.h:84-87)
    emu.emu_destroy(h)


def test_philox_known_answers(emu):
    """Random123 kat_vectors for philox4x32-10 -- the counter-based generator that replaces
    IndependentSampler's SFMT stream (rng.h)."""
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
        ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
        ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0], [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
    ]
    for ctr, key, exp in kat:
        out = np.zeros(4, np.uint32)
        emu.emu_philox(vp(np.array(ctr, np.uint32)), vp(np.array(key, np.uint32)), vp(out))
        assert out.tolist() == exp


def test_sample_floats_are_in_unit_interval_and_uniform(emu):
    emu.emu_sample_block.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]
    vals = np.zeros((4096, 4), np.float32)
    for i in range(4096):
        emu.emu_sample_block(1004, (i << 24) | 3, 1, 0, vp(vals[i]))
    assert vals.min() >= 0.0 and vals.max() < 1.0
    assert abs(vals.mean() - 0.5) < 0.01 and abs(np.corrcoef(vals[:, 0], vals[:, 1])[0, 1]) < 0.05


def test_full_size_1m_scene_vs_reference_kdtree_records(emu, oracle_mod):
    """BASELINE configs[1] geometry at full size (1 048 580 triangles): the emulated BVH build + traversal against the
    committed records of the reference kd-tree (tests/golden/hits_big_1m.npz) on the 65 536-ray subsamples of the
    coherent and the incoherent set.  (The CUDA path runs the same comparison, and all 4 M rays, in
    tests/test_gpu_trace_big.py.)"""
    import hashlib
    from tests.golden.make_golden_big import sensor_rays_oracle
    b = scenes.scene_1m()
    g = util.load_golden("hits_big_1m.npz")
    assert util.scene_digest(b).encode() == g["digest"].tobytes(), "the scene generator does not reproduce the golden scene on this host"
    h, keep = build(emu, b)
    nn, t6, e6 = C.c_uint32(), np.zeros(6, np.float32), np.zeros(6, np.float32)
    emu.emu_info(h, C.byref(nn), vp(t6), vp(e6))
    assert np.array_equal(t6, g["tight"])
    sets = util.big_ray_sets(b, sensor_rays_oracle(b), t6[:3], t6[3:])
    for sname, rays in sets.items():
        assert hashlib.sha256(rays.tobytes()).hexdigest().encode() == g["sha_" + sname].tobytes(), sname
        sub = rays[util.golden_subset(len(rays))]
        hits, _ = trace(emu, h, sub, abi.TRACE_CLOSEST)
        rep = util.hit_parity_report(hits, g["closest_" + sname])
        assert rep["hard_mismatches"] == 0 and rep["excluded_fraction"] <= 2e-3 and rep["bit_exact_fraction"] >= 0.995, (sname, rep)
        occl, _ = trace(emu, h, sub, abi.TRACE_ANY)
        g_any = np.unpackbits(g["any_" + sname])[:len(sub)].astype(bool)
        assert ((occl["prim"] != abi.NO_HIT) != g_any).sum() <= max(3, 2e-3 * len(sub))
    emu.emu_destroy(h)
