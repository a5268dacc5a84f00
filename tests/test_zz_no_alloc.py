"""GPU: mtsb_trace_static -- the allocation-free variant of mtsb_trace (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Collected last on purpose (pytest -x).

Checks, through the C ABI: the hit records equal mtsb_trace's bit for bit (same kernel, same arithmetic; the parity tests
of mtsb_trace use exactly this bar for t, u, v on equal primitives), closest and any-hit, on TWO calls in a row over
different ray sets (the static scratch keeps the previous call's contents), at the maximum size and at a ragged one; and
sizes beyond the maximum are rejected with a clear error while the scene stays usable."""
import numpy as np
import pytest

from mitsuba_b200 import abi, scenes

pytestmark = pytest.mark.gpu


def _same(a, b):
    """Bit-identical records; an exact equal-t tie between two triangles (a ray through a shared edge) may be broken
    either way by the order in which a warp tests its queued triangles, so at most 3 such rays may name the other triangle."""
    if not np.array_equal(a["t"], b["t"]):
        return False
    differ = (a["prim"] != b["prim"]) | (a["u"] != b["u"]) | (a["v"] != b["v"])
    return int(differ.sum()) <= 3


def test_trace_static_matches_trace_on_two_calls_in_a_row(gpu_lib):
    L = gpu_lib.lib()
    max_rays = L.mtsb_trace_static_max_rays()
    assert max_rays == 262144
    b = scenes.scene_1m(width=96, height=54, spp=1, nu=48, nv=24, n_obj=4)
    sc = gpu_lib.Scene(b)
    info = sc.info()
    lo, hi = np.array(info.aabb_min[:]), np.array(info.aabb_max[:])
    c, r = (lo + hi) / 2, float(np.linalg.norm(hi - (lo + hi) / 2))
    first = scenes.chord_rays(c, r, max_rays, seed=11)                 # call 1: the maximum size
    second = scenes.chord_rays(c, r, 100003, seed=12)                  # call 2: fewer rays, ragged count, other rays
    for mode in (abi.TRACE_CLOSEST, abi.TRACE_ANY):
        for rays in (first, second):
            want = sc.trace(rays, mode)
            got = sc.trace_static(rays, mode)
            if mode == abi.TRACE_CLOSEST:
                assert _same(got, want)
                assert (got["prim"] != abi.NO_HIT).mean() > 0.1
            else:
                assert np.array_equal(got["prim"] != abi.NO_HIT, want["prim"] != abi.NO_HIT)
    # beyond the maximum: rejected, nothing traced, the scene still works
    too_many = scenes.chord_rays(c, r, max_rays + 1, seed=13)
    with pytest.raises(gpu_lib.MtsbError) as e:
        sc.trace_static(too_many)
    assert e.value.code == abi.ERR_INVALID and "MTSB_STATIC_TRACE_MAX_RAYS" in str(e.value)
    assert sc.trace_static(first[:0]).shape == (0,)
    assert _same(sc.trace_static(second), sc.trace(second))
    sc.close()
