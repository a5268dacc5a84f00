"""GPU: first-hit parity of the CUDA traversal through the C ABI (mtsb_trace / mtsb_trace_device)
against the reference kd-tree's records (tests/golden) and against the oracle."""
import numpy as np
import pytest

from mitsuba_b200 import abi, scenes
from tests import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["cbox", "mesh"])
def test_hits_vs_reference_kdtree_golden(gpu_lib, name):
    g = util.load_golden("hits_%s.npz" % name)
    sc = gpu_lib.Scene(util.golden_scenes()[name])
    info = sc.info()
    assert np.array_equal(np.array(info.aabb_min[:] + info.aabb_max[:], np.float32), g["tight"])
    for rs in ("chord", "camera", "secondary"):
        st = util.compare_hits(sc.trace(g["rays_" + rs], abi.TRACE_CLOSEST), g["closest_" + rs], label=name + "/" + rs)
        assert st["bit_exact"] >= 0.995 * st["n"], st
        occl = sc.trace(g["rays_" + rs], abi.TRACE_ANY)["prim"] != abi.NO_HIT
        assert (occl != g["any_" + rs]).sum() <= max(3, 2e-3 * len(occl))
    sc.close()


def test_hits_vs_oracle_65k_tris(gpu_lib, oracle_mod):
    b = scenes.scene_1m(width=320, height=180, spp=1, nu=64, nv=32, n_obj=16)
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    tight, _, _ = orc.bounds()
    c = (tight[:3] + tight[3:]) / 2
    rays = {"chord": scenes.chord_rays(c, float(np.linalg.norm(tight[3:] - c)), 300000)}
    ys, xs = np.mgrid[0:180, 0:320]
    rays["camera"] = orc.sensor_rays(np.stack([xs.ravel() + 0.5, ys.ravel() + 0.5], -1).astype(np.float32))
    hits = orc.trace(rays["camera"])
    ok = hits["prim"] != abi.NO_HIT
    sec = np.zeros(int(ok.sum()), dtype=abi.RAY_DTYPE)
    sec["o"] = rays["camera"]["o"][ok] + rays["camera"]["d"][ok] * hits["t"][ok][:, None]
    d = np.random.default_rng(3).normal(size=(sec.shape[0], 3))
    sec["d"] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    sec["mint"], sec["maxt"] = np.float32(1e-4), np.inf
    rays["secondary"] = sec
    for k, r in rays.items():
        st = util.compare_hits(sc.trace(r, abi.TRACE_CLOSEST), orc.trace(r, abi.TRACE_CLOSEST), label=k)
        assert st["bit_exact"] >= 0.995 * st["n"], st
        a = sc.trace(r, abi.TRACE_ANY)["prim"] != abi.NO_HIT
        o = orc.trace(r, abi.TRACE_ANY)["prim"] != abi.NO_HIT
        assert (a != o).sum() <= max(3, 2e-3 * len(r))
    sc.close()


@pytest.mark.parametrize("n_tris", [0, 1, 2, 3, 4, 5, 9, 33])
def test_tiny_and_ragged_scenes(gpu_lib, oracle_mod, n_tris):
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    rng = np.random.default_rng(n_tris)
    b = SceneBuilder()
    m = b.add_material(diffuse(0.5))
    if n_tris:
        tri = rng.normal(size=(n_tris, 3, 3)).astype(np.float32)
        b.add_mesh(tri.reshape(-1, 3), np.arange(n_tris * 3, dtype=np.uint32).reshape(-1, 3), m)
    sc = gpu_lib.Scene(b)
    rays = scenes.chord_rays([0, 0, 0], 4.0, 4001, seed=3)     # ragged count: not a multiple of the warp size
    util.compare_hits(sc.trace(rays), oracle_mod.Oracle(b).trace(rays), label="n=%d" % n_tris)
    assert sc.trace(rays[:0]).shape == (0,)
    sc.close()


def test_device_resident_rays_and_properties_1m(gpu_lib):
    """BASELINE configs[1] geometry at full size (1M triangles): size-independent properties.
    (a) any-hit <=> closest-hit exists; (b) re-tracing with maxt just below the reported t finds
    nothing closer (the reported hit is the first one); (c) re-tracing with maxt == t returns the
    same record (inclusive interval, idempotence)."""
    import torch
    b = scenes.scene_1m(width=64, height=36, spp=1)
    sc = gpu_lib.Scene(b)
    assert sc.info().n_tris == b.n_tris >= 1_000_000
    n = 1 << 20
    for kind in (0, 1):
        rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
        hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
        sc.gen_rays_device(kind, n, 1004, rays.data_ptr())
        torch.cuda.synchronize()
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(rays.data_ptr(), n, hits.data_ptr(), abi.TRACE_CLOSEST)
        prim = hits.view(torch.int32)[:, 0].clone()
        t = hits[:, 3].clone()
        hit = prim != -1
        assert 0.05 < hit.float().mean().item() < 1.0
        any_hits = torch.empty_like(hits)
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(rays.data_ptr(), n, any_hits.data_ptr(), abi.TRACE_ANY)
        assert torch.equal(any_hits.view(torch.int32)[:, 0] != -1, hit)
        r2 = rays.clone()
        r2[:, 7] = torch.where(hit, t * (1 - 1e-4), r2[:, 7])
        h2 = torch.empty_like(hits)
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(r2.data_ptr(), n, h2.data_ptr(), abi.TRACE_CLOSEST)
        closer = (h2.view(torch.int32)[:, 0] != -1) & hit
        # anything still found must be a different primitive strictly in front (never the case for a true first hit)
        assert closer.sum().item() == 0
        r3 = rays.clone()
        r3[:, 7] = torch.where(hit, t, r3[:, 7])
        h3 = torch.empty_like(hits)
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(r3.data_ptr(), n, h3.data_ptr(), abi.TRACE_CLOSEST)
        same = h3.view(torch.int32)[:, 0] == prim
        assert (hit & ~same).sum().item() <= 4       # equal-t ties only
        assert torch.equal(h3[:, 3][hit & same], t[hit & same])
    sc.close()


def _brute_force_first_hit(pos, idx, rays, chunk=1 << 20):
    """Independent check at full size: Moeller-Trumbore in fp64 over ALL triangles (torch, on the GPU).
    Returns (t_min, prim) per ray."""
    import torch
    o = rays[:, 0:3].double(); d = rays[:, 4:7].double()
    mint = rays[:, 3].double(); maxt = rays[:, 7].double()
    best_t = torch.full((rays.shape[0],), float("inf"), dtype=torch.float64, device="cuda")
    best_p = torch.full((rays.shape[0],), -1, dtype=torch.int64, device="cuda")
    for c0 in range(0, idx.shape[0], chunk):
        tri = pos[idx[c0:c0 + chunk].long()].double()              # (m, 3, 3)
        e1, e2 = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
        for r0 in range(0, rays.shape[0], 64):
            oo, dd = o[r0:r0 + 64, None, :], d[r0:r0 + 64, None, :]
            pv = torch.cross(dd.expand(-1, tri.shape[0], -1), e2[None].expand(dd.shape[0], -1, -1), dim=-1)
            det = (e1[None] * pv).sum(-1)
            inv = 1.0 / det
            tv = oo - tri[None, :, 0]
            u = (tv * pv).sum(-1) * inv
            qv = torch.cross(tv, e1[None].expand(tv.shape[0], -1, -1), dim=-1)
            v = (dd * qv).sum(-1) * inv
            t = (e2[None] * qv).sum(-1) * inv
            ok = (det.abs() > 1e-30) & (u >= 0) & (v >= 0) & (u + v <= 1) & (t >= mint[r0:r0 + 64, None]) & (t <= maxt[r0:r0 + 64, None])
            t = torch.where(ok, t, torch.full_like(t, float("inf")))
            tm, pm = t.min(dim=1)
            upd = tm < best_t[r0:r0 + 64]
            best_t[r0:r0 + 64] = torch.where(upd, tm, best_t[r0:r0 + 64])
            best_p[r0:r0 + 64] = torch.where(upd, pm + c0, best_p[r0:r0 + 64])
    return best_t, best_p


def test_full_size_10m_properties_and_brute_force(gpu_lib):
    """BASELINE configs[2]/[3] geometry at full size (10.5M triangles, cfg 4's two ray sets).
    Size-independent properties as in the 1M test, plus an algorithm-independent check: a sample of the
    rays is intersected with ALL triangles by brute force in fp64 and must report the same first hit."""
    import torch
    free, _ = torch.cuda.mem_get_info()
    if free < 24 << 30:
        pytest.skip("needs ~24 GB of free device memory")
    b = scenes.scene_10m(width=64, height=36, spp=1)
    sc = gpu_lib.Scene(b)
    assert sc.info().n_tris == b.n_tris >= 10_000_000
    pos_h, _, _, idx_h = b.arrays()
    pos = torch.from_numpy(pos_h).cuda()
    idx = torch.from_numpy(idx_h.astype(np.int64)).cuda()
    n = 1 << 22
    for kind in (0, 1):
        rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
        hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
        sc.gen_rays_device(kind, n, 1004, rays.data_ptr())
        torch.cuda.synchronize()
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(rays.data_ptr(), n, hits.data_ptr(), abi.TRACE_CLOSEST)
        prim = hits.view(torch.int32)[:, 0].clone()
        t = hits[:, 3].clone()
        hit = prim != -1
        assert 0.05 < hit.float().mean().item() <= 1.0
        any_hits = torch.empty_like(hits)
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(rays.data_ptr(), n, any_hits.data_ptr(), abi.TRACE_ANY)
        assert torch.equal(any_hits.view(torch.int32)[:, 0] != -1, hit)
        r2 = rays.clone()
        r2[:, 7] = torch.where(hit, t * (1 - 1e-4), r2[:, 7])
        h2 = torch.empty_like(hits)
        torch.cuda.synchronize()        # torch's stream -> the scene's own stream (see mtsb200.h)
        sc.trace_device(r2.data_ptr(), n, h2.data_ptr(), abi.TRACE_CLOSEST)
        assert ((h2.view(torch.int32)[:, 0] != -1) & hit).sum().item() == 0
        # brute force on a strided sample of the rays
        sel = torch.arange(0, n, n // 192, device="cuda")[:192]
        bt, bp = _brute_force_first_hit(pos, idx, rays[sel])
        gp, gt = prim[sel].long(), t[sel].double()
        miss_b, miss_g = torch.isinf(bt), gp == -1
        # a disagreement about hit/miss is only acceptable for grazing hits the two arithmetics may see differently: none expected
        assert (miss_b != miss_g).sum().item() <= 1
        both = ~miss_b & ~miss_g
        assert (torch.abs(bt[both] - gt[both]) <= 1e-4 * gt[both].abs() + 1e-5).all()
        agree = (bp[both] == gp[both]).float().mean().item()
        assert agree > 0.97, agree            # different primitive only at shared edges (equal t within rounding)
    sc.close()


def test_many_triangles_sharing_one_centroid(gpu_lib, oracle_mod):
    """100 000 triangles with the SAME centroid: identical Morton keys (the radix tree falls back to the index
    tie-break), maximal overlap, long leaf chains.  Hits must equal the oracle's and the traversal stack must not
    overflow (an overflow is reported as an error, never as a silently missing subtree)."""
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    rng = np.random.default_rng(21)
    n = 100000
    tri = rng.normal(size=(n, 3, 3)) * rng.uniform(0.05, 1.0, (n, 1, 1))
    tri -= tri.mean(axis=1, keepdims=True)                     # every centroid at the origin
    b = SceneBuilder()
    b.add_mesh(tri.reshape(-1, 3).astype(np.float32), np.arange(n * 3, dtype=np.uint32).reshape(-1, 3), b.add_material(diffuse(0.5)))
    sc = gpu_lib.Scene(b)
    rays = scenes.chord_rays([0, 0, 0], 3.0, 3000, seed=4)
    got = sc.trace(rays)
    ref = oracle_mod.Oracle(b).trace(rays)
    st = util.compare_hits(got, ref, label="one centroid", max_degenerate_frac=5e-3)
    assert st["hit_frac"] > 0.5
    sc.close()


def test_invalid_indices_are_rejected(gpu_lib):
    """A description whose triangle indices leave the shape's vertex range is refused (MTSB_ERR_INVALID), not traced."""
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    b = SceneBuilder()
    m = b.add_material(diffuse(0.5))
    b.add_quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), m)
    b.add_quad((-1, 1, -1), (-1, 1, 1), (1, 1, 1), (1, 1, -1), m)
    b._idx[1] = b._idx[1].copy()
    b._idx[1][0, 0] = 0                                        # second shape points into the first shape's vertices
    with pytest.raises(gpu_lib.MtsbError) as e:
        gpu_lib.Scene(b)
    assert e.value.code == abi.ERR_INVALID
