"""Shared helpers of the parity tests."""
import ctypes as C
import hashlib
import os

import numpy as np

from mitsuba_b200 import abi, scenes

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_scenes():
    return {
        "cbox": scenes.cbox(width=64, height=64, spp=4),
        "mesh": scenes.scene_1m(width=96, height=54, spp=4, nu=48, nv=24, n_obj=4),
    }


def scene_digest(b):
    pos, nrm, uv, idx = b.arrays()
    h = hashlib.sha256()
    h.update(pos.tobytes())
    h.update(idx.tobytes())
    return h.hexdigest()


def load_golden(name):
    g = np.load(os.path.join(GOLDEN, name))
    return {k: g[k] for k in g.files}


def compare_hits(test, ref, tol=1e-5, max_degenerate_frac=2e-3, label=""):
    """First-hit parity.  A ray is "degenerate" when the two searches accept
    different primitives at (nearly) the same distance, when the hit lies on an edge
    (min barycentric < tol), or when only one of them accepts a grazing hit; everything else
    must agree: primitive index bit-exact, t/u/v within `tol` relative."""
    assert test.shape == ref.shape
    n = test.shape[0]
    th, rh = test["prim"] != abi.NO_HIT, ref["prim"] != abi.NO_HIT
    both = th & rh
    same_prim = both & (test["prim"] == ref["prim"])
    tt, rt = test["t"], ref["t"]
    scale = np.maximum(np.where(both, np.abs(rt), 1.0), 1.0)
    with np.errstate(invalid="ignore"):
        t_close = both & (np.abs(np.where(both, tt - rt, 0.0)) <= tol * scale)
    # exact agreement on the bulk
    ok = (~th & ~rh) | (same_prim & t_close &
                        (np.abs(test["u"] - ref["u"]) <= tol) & (np.abs(test["v"] - ref["v"]) <= tol))
    bad = ~ok
    # classify the rest
    def edge(h):
        return np.minimum(np.minimum(h["u"], h["v"]), 1.0 - h["u"] - h["v"]) < 10 * tol
    tie = both & ~same_prim & t_close                              # (i) two prims at the same distance
    on_edge = (th & edge(test)) | (rh & edge(ref))                 # (ii) edge / vertex hits
    degenerate = bad & (tie | on_edge)
    hard = bad & ~degenerate
    stats = dict(n=n, hit_frac=float(rh.mean()), bit_exact=int((same_prim & (tt == rt) & (test["u"] == ref["u"]) &
                 (test["v"] == ref["v"])).sum() + (~th & ~rh).sum()), degenerate=int(degenerate.sum()), hard=int(hard.sum()))
    assert hard.sum() == 0, "%s: %d non-degenerate mismatches, e.g. test=%s ref=%s" % (
        label, hard.sum(), test[hard][:3], ref[hard][:3])
    assert degenerate.sum() <= max(3, max_degenerate_frac * n), "%s: too many degenerate rays: %s" % (label, stats)
    return stats


def rel_close(a, b, rtol, atol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b) <= atol + rtol * np.maximum(np.abs(a), np.abs(b))


def frac_close(a, b, rtol, atol):
    return float(rel_close(a, b, rtol, atol).mean())


# ---- full-size ray sets of BASELINE configs[3] (tests/test_gpu_trace_big.py, tests/golden/make_golden_big.py) ----
def exact_chord_rays(center, radius, n, seed):
    """Chords between points on the bounding sphere (the ray set of src/tests/test_kd.cpp:112-117) built from
    correctly rounded operations only (+, *, /, sqrt on PCG64 doubles): bit-identical on every host, unlike the
    sin/cos construction of scenes.chord_rays."""
    rng = np.random.default_rng(seed)

    def sphere(k):
        out = np.zeros((0, 3))
        while len(out) < k:
            p = rng.random((k, 3)) * 2.0 - 1.0
            r2 = p[:, 0] * p[:, 0] + p[:, 1] * p[:, 1] + p[:, 2] * p[:, 2]
            ok = (r2 <= 1.0) & (r2 > 1e-4)
            out = np.concatenate([out, p[ok] / np.sqrt(r2[ok])[:, None]])
        return out[:k]
    c = np.asarray(center, np.float64)
    p1, p2 = c + radius * sphere(n), c + radius * sphere(n)
    d = p2 - p1
    ln = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])
    rays = np.zeros(n, dtype=abi.RAY_DTYPE)
    rays["o"] = p1.astype(np.float32)
    rays["d"] = (d / ln[:, None]).astype(np.float32)
    rays["mint"] = 0.0
    rays["maxt"] = ln.astype(np.float32)
    return rays


def camera_grid_positions(width, height, nx, ny):
    """nx x ny sample positions on a regular grid over the film (pixel centres of a coarser image), row-major."""
    xs = ((np.arange(nx, dtype=np.float64) + 0.5) * (width / nx)).astype(np.float32)
    ys = ((np.arange(ny, dtype=np.float64) + 0.5) * (height / ny)).astype(np.float32)
    X, Y = np.meshgrid(xs, ys)
    return np.stack([X.ravel(), Y.ravel()], -1).astype(np.float32)


BIG_SCENES = {"1m": ("scene_1m", 1048580), "10m": ("scene_10m", 10485770)}
BIG_COHERENT = (2048, 2048)          # 4 M camera rays
BIG_INCOHERENT = 1 << 22             # 4 M chords
BIG_GOLDEN_N = 65536                 # rays of each set whose reference records are committed


def big_ray_sets(builder, sensor_rays, aabb_min, aabb_max):
    """The two ray sets of the cfg-3 microbenchmark at parity-test size: 4 M coherent camera rays and 4 M incoherent
    bounding-sphere chords.  `sensor_rays(pos)` is Scene.sensor_rays of the product or of the oracle (bit-identical)."""
    s = builder.sensor
    cam = sensor_rays(camera_grid_positions(s.crop_width, s.crop_height, *BIG_COHERENT))
    lo, hi = np.asarray(aabb_min, np.float64), np.asarray(aabb_max, np.float64)
    c = (lo + hi) * 0.5
    e = hi - c
    radius = float(np.sqrt(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]))
    return {"coherent": cam, "incoherent": exact_chord_rays(c, radius, BIG_INCOHERENT, seed=1004)}


def golden_subset(n):
    """Indices of the rays whose reference records are committed: an even stride through the set."""
    return (np.arange(BIG_GOLDEN_N, dtype=np.int64) * (n // BIG_GOLDEN_N))


def kd_trace_parallel(kd, rays, mode, threads=0):
    """RefKDTree.trace over a thread pool (the traversal is read-only and the call releases the GIL)."""
    from concurrent.futures import ThreadPoolExecutor
    threads = threads or (os.cpu_count() or 8)
    chunks = np.array_split(np.arange(len(rays)), threads * 4)
    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(lambda ix: kd.trace(rays[ix[0]:ix[-1] + 1], mode) if len(ix) else np.zeros(0, abi.HIT_DTYPE), chunks))
    return np.concatenate(parts)


def hit_parity_report(test, ref, tol=1e-5):
    """Row T of SURVEY.md 8a, with the numbers the row asks to be reported: rays, bit-exact records, rays excluded as
    degenerate (equal-t ties between two primitives, hits on an edge or vertex, grazing hits only one search accepts)
    and hard mismatches (must be zero)."""
    n = test.shape[0]
    th, rh = test["prim"] != abi.NO_HIT, ref["prim"] != abi.NO_HIT
    both = th & rh
    same_prim = both & (test["prim"] == ref["prim"])
    tt, rt = test["t"], ref["t"]
    scale = np.maximum(np.where(both, np.abs(rt), 1.0), 1.0)
    with np.errstate(invalid="ignore"):
        t_close = both & (np.abs(np.where(both, tt - rt, 0.0)) <= tol * scale)
    ok = (~th & ~rh) | (same_prim & t_close & (np.abs(test["u"] - ref["u"]) <= tol) & (np.abs(test["v"] - ref["v"]) <= tol))
    bit = (~th & ~rh) | (same_prim & (tt == rt) & (test["u"] == ref["u"]) & (test["v"] == ref["v"]))

    def edge(h):
        return np.minimum(np.minimum(h["u"], h["v"]), 1.0 - h["u"] - h["v"]) < 10 * tol
    tie = both & ~same_prim & t_close
    on_edge = (th & edge(test)) | (rh & edge(ref))
    degenerate = ~ok & (tie | on_edge)
    hard = ~ok & ~degenerate
    return dict(rays=int(n), hit_fraction=float(rh.mean()), bit_exact=int(bit.sum()), bit_exact_fraction=float(bit.mean()),
                within_tolerance=int(ok.sum()), excluded_degenerate=int(degenerate.sum()), excluded_fraction=float(degenerate.mean()),
                excluded_ties=int((~ok & tie).sum()), excluded_edge_hits=int((~ok & on_edge & ~tie).sum()), hard_mismatches=int(hard.sum()))
