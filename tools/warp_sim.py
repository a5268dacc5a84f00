"""Warp-scheduling model of the traversal kernel (design tool, CPU only).

The g++ build of the kernel bodies (tests/emu) records, for every ray, how many triangle tests
follow each node visit.  This script replays those per-ray sequences through a 32-lane warp with
the kernel's refill rule and counts warp-instructions for several lock-step policies, so that a
policy can be chosen before spending GPU time.  Costs per warp-level phase are the SASS counts of
the current kernel (node visit ~340, triangle test ~75, loop bookkeeping ~35).

    python tools/warp_sim.py [--res 320x180]
"""
import argparse
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mitsuba_b200 import abi, scenes  # noqa: E402

EMU_DIR = os.path.join(ROOT, "tests", "emu")
C_NODE, C_TRI, C_LOOP = 340, 75, 35


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def load_emu():
    subprocess.check_call(["make", "-C", EMU_DIR], stdout=subprocess.DEVNULL)
    E = C.CDLL(os.path.join(EMU_DIR, "libemu.so"))
    E.emu_build.restype = C.c_void_p
    E.emu_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    E.emu_trace.restype = C.c_uint64
    E.emu_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    E.emu_trace_profile.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p]
    return E


def camera_rays(b, w, h):
    s = b.sensor
    m = np.array(list(s.to_world), np.float64).reshape(4, 4)
    tx = np.tan(np.radians(s.fov_x_deg) / 2)
    ty = tx * s.film_height / s.film_width
    ys, xs = np.mgrid[0:h, 0:w]
    # 8x8 tile order, like the renderer's pixel order inside a pass
    x = (1 - 2 * (xs.ravel() + 0.5) / w) * tx
    y = (1 - 2 * (ys.ravel() + 0.5) / h) * ty
    d = np.stack([x, y, np.ones_like(x)], -1)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d @ m[:3, :3].T
    rays = np.zeros(w * h, abi.RAY_DTYPE)
    rays["o"] = m[:3, 3]
    rays["d"] = d
    rays["mint"] = 1e-4
    rays["maxt"] = np.inf
    return rays


def bounce_rays(rng, rays, hits, pos, idx, light_c):
    ok = hits["prim"] != abi.NO_HIT
    r, h = rays[ok], hits[ok]
    p = r["o"] + r["d"] * h["t"][:, None]
    tri = pos[idx[h["prim"]]]
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    n *= -np.sign(np.einsum("ij,ij->i", n, r["d"]))[:, None]
    # cosine-weighted direction about n
    u = rng.random((len(p), 2))
    ph, rr = 2 * np.pi * u[:, 0], np.sqrt(u[:, 1])
    a = np.where(np.abs(n[:, :1]) > 0.9, [[0, 1, 0]], [[1, 0, 0]])
    t1 = np.cross(n, a)
    t1 /= np.linalg.norm(t1, axis=1, keepdims=True)
    t2 = np.cross(n, t1)
    d = t1 * (rr * np.cos(ph))[:, None] + t2 * (rr * np.sin(ph))[:, None] + n * np.sqrt(1 - u[:, 1])[:, None]
    out = np.zeros(len(p), abi.RAY_DTYPE)
    out["o"], out["d"], out["mint"], out["maxt"] = p, d, 1e-4, np.inf
    sh = np.zeros(len(p), abi.RAY_DTYPE)
    to = light_c + rng.uniform(-2.5, 2.5, (len(p), 3)) * [1, 0, 1] - p
    ln = np.linalg.norm(to, axis=1)
    sh["o"], sh["d"], sh["mint"], sh["maxt"] = p, to / ln[:, None], 1e-4, ln * (1 - 1e-4)
    return out, sh


def profile(E, hnd, rays, mode, max_len=255):
    rays = np.ascontiguousarray(rays)
    seq = np.zeros((len(rays), max_len), np.uint8)
    lens = np.zeros(len(rays), np.uint32)
    E.emu_trace_profile(hnd, vp(rays), len(rays), mode, vp(seq), max_len, vp(lens))
    return seq, lens


def simulate(seq, lens, policy, refill=26, k_tri=1, theta=0):
    """Returns (warp-instructions per ray, mean lanes active in node phases, in tri phases)."""
    n = len(lens)
    nxt = 0
    # lane state
    ray = np.full(32, -1)
    k = np.zeros(32, np.int64)          # node visits done
    r = np.zeros(32, np.int64)          # pending triangle tests
    cost = 0
    node_lanes = node_phases = tri_lanes = tri_phases = 0
    while True:
        idle = ray < 0
        if idle.any() and nxt < n:
            for ln in np.nonzero(idle)[0]:
                while nxt < n and lens[nxt] == 0:
                    nxt += 1
                if nxt >= n:
                    break
                ray[ln], k[ln], r[ln] = nxt, 0, 0
                nxt += 1
        if not (ray >= 0).any():
            break
        while True:
            act = ray >= 0
            want = act & (r == 0)
            cost += C_LOOP
            run_node = want.any()
            if policy == "vote" and want.sum() < theta and (act & (r > 0)).any():
                run_node = False
            if run_node:
                cost += C_NODE
                node_phases += 1
                node_lanes += want.sum()
                for ln in np.nonzero(want)[0]:
                    r[ln] = seq[ray[ln], k[ln]]
                    k[ln] += 1
            steps = 1 if policy in ("one", "vote") else (k_tri if policy == "k" else 10 ** 6)
            for _ in range(steps):
                tri = act & (r > 0)
                if not tri.any():
                    break
                cost += C_TRI
                tri_phases += 1
                tri_lanes += tri.sum()
                r[tri] -= 1
            for ln in np.nonzero(act)[0]:
                if r[ln] == 0 and k[ln] >= lens[ray[ln]]:
                    ray[ln] = -1
            live = (ray >= 0).sum()
            if live == 0 or (nxt < n and live < refill):
                break
    return cost / max(1, (lens > 0).sum()), node_lanes / max(1, node_phases), tri_lanes / max(1, tri_phases)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", default="256x144")
    ap.add_argument("--max-rays", type=int, default=20000)
    a = ap.parse_args()
    w, h = [int(x) for x in a.res.split("x")]
    E = load_emu()
    b = scenes.scene_1m()
    pos, _, _, idx = b.arrays()
    hnd = E.emu_build(vp(pos), vp(idx), idx.shape[0])
    rng = np.random.default_rng(3)
    prim = camera_rays(b, w, h)
    hits = np.zeros(len(prim), abi.HIT_DTYPE)
    E.emu_trace(hnd, vp(prim), len(prim), 0, vp(hits))
    sec, shadow = bounce_rays(rng, prim, hits, pos, idx, np.array([0, 7.0, 0]))
    sets = {"primary": (prim, 0), "bounce1": (sec, 0), "shadow": (shadow, 1)}
    for name, (rays, mode) in sets.items():
        lo = max(0, len(rays) // 2 - a.max_rays // 2)
        rays = rays[lo:lo + a.max_rays]          # a contiguous block from the middle of the image
        seq, lens = profile(E, hnd, rays, mode)
        tot_tri = sum(int(seq[i, :lens[i]].sum()) for i in range(len(lens)))
        print("%s: %d rays, %.1f node visits/ray, %.1f tri tests/ray" % (name, len(rays), lens.mean(), tot_tri / len(rays)))
        for label, kw in [("one tri / iteration (current)", dict(policy="one")),
                          ("2 tris / iteration", dict(policy="k", k_tri=2)),
                          ("3 tris / iteration", dict(policy="k", k_tri=3)),
                          ("4 tris / iteration", dict(policy="k", k_tri=4)),
                          ("all tris (while-while)", dict(policy="all")),
                          ("vote theta=12", dict(policy="vote", theta=12)),
                          ("vote theta=20", dict(policy="vote", theta=20)),
                          ("one, refill 20", dict(policy="one", refill=20)),
                          ("one, refill 30", dict(policy="one", refill=30))]:
            c, nl, tl = simulate(seq, lens, **kw)
            print("   %-32s %8.0f warp-instr/ray   node lanes %.1f  tri lanes %.1f" % (label, c, nl, tl))


if __name__ == "__main__":
    main()


def simulate_defer(seq, lens, depth=2, refill=26):
    """Policy 'defer': a lane keeps up to `depth` pending triangle batches and goes on visiting nodes
    while the oldest batch is tested one triangle per iteration."""
    n = len(lens)
    nxt = 0
    ray = [-1] * 32
    k = [0] * 32
    q = [[] for _ in range(32)]
    cost = node_lanes = node_phases = tri_lanes = tri_phases = 0
    while True:
        for ln in range(32):
            if ray[ln] < 0:
                while nxt < n and lens[nxt] == 0:
                    nxt += 1
                if nxt < n:
                    ray[ln], k[ln], q[ln] = nxt, 0, []
                    nxt += 1
        if all(r < 0 for r in ray):
            break
        while True:
            cost += C_LOOP
            want = [ln for ln in range(32) if ray[ln] >= 0 and k[ln] < lens[ray[ln]] and len(q[ln]) < depth]
            if want:
                cost += C_NODE
                node_phases += 1
                node_lanes += len(want)
                for ln in want:
                    c = int(seq[ray[ln], k[ln]])
                    k[ln] += 1
                    if c:
                        q[ln].append(c)
            tri = [ln for ln in range(32) if ray[ln] >= 0 and q[ln]]
            if tri:
                cost += C_TRI
                tri_phases += 1
                tri_lanes += len(tri)
                for ln in tri:
                    q[ln][0] -= 1
                    if q[ln][0] == 0:
                        q[ln].pop(0)
            for ln in range(32):
                if ray[ln] >= 0 and not q[ln] and k[ln] >= lens[ray[ln]]:
                    ray[ln] = -1
            live = sum(1 for r in ray if r >= 0)
            if live == 0 or (nxt < n and live < refill):
                break
    return cost / max(1, (lens > 0).sum()), node_lanes / max(1, node_phases), tri_lanes / max(1, tri_phases)
