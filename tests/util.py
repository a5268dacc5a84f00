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
