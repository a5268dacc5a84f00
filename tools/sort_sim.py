"""Does reordering the bounce rays pay?  Warp-divergence model (CPU only, see tools/warp_sim.py).
A 32x32 tile x S samples of cfg 1 / cfg 2 is traced with the emulated kernel; the bounce rays are replayed through the
lock-step warp model in (a) shading order, (b) stably binned by direction octant, (c) binned by a finer direction key.
    python tools/sort_sim.py [--scene 10m] [--spp 8] [--tile 16]"""
import argparse, sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import warp_sim as W
from mitsuba_b200 import abi, scenes

def morton_tile_rays(b, x0, y0, tile, spp, rng):
    s = b.sensor
    m = np.array(list(s.to_world), np.float64).reshape(4, 4)
    w, h = s.film_width, s.film_height
    tx = np.tan(np.radians(s.fov_x_deg) / 2); ty = tx * h / w
    def c1(x):
        x = x & 0x55555555; x = (x ^ (x >> 1)) & 0x33333333; x = (x ^ (x >> 2)) & 0x0f0f0f0f; x = (x ^ (x >> 4)) & 0x00ff00ff; return x
    pix = np.arange(tile * tile)
    px = x0 + c1(pix); py = y0 + c1(pix >> 1)
    px = np.repeat(px, spp) + rng.random(tile * tile * spp); py = np.repeat(py, spp) + rng.random(tile * tile * spp)
    x = (1 - 2 * px / w) * tx; y = (1 - 2 * py / h) * ty
    d = np.stack([x, y, np.ones_like(x)], -1); d /= np.linalg.norm(d, axis=1, keepdims=True); d = d @ m[:3, :3].T
    rays = np.zeros(len(x), abi.RAY_DTYPE); rays["o"] = m[:3, 3]; rays["d"] = d; rays["mint"] = 1e-4; rays["maxt"] = np.inf
    return rays

def dir_key(d, fine):
    octant = (d[:, 0] >= 0).astype(int) | ((d[:, 1] >= 0).astype(int) << 1) | ((d[:, 2] >= 0).astype(int) << 2)
    if not fine: return octant
    a = np.abs(d); ax = np.argmax(a, 1); n = len(d); r = np.arange(n)
    u = d[r, (ax + 1) % 3] / a[r, ax]; v = d[r, (ax + 2) % 3] / a[r, ax]
    face = ax * 2 + (d[r, ax] < 0)
    return (face * 4 + np.clip(((u + 1) * 2).astype(int), 0, 3)) * 4 + np.clip(((v + 1) * 2).astype(int), 0, 3)

def run(E, hnd, rays, mode, label):
    seq, lens = W.profile(E, hnd, rays, mode)
    c, nl, tl = W.simulate_defer(seq, lens)
    print("   %-40s %7.0f warp-instr/ray  node lanes %.1f  tri lanes %.1f  (%.1f visits, %d rays)" % (label, c, nl, tl, lens.mean(), len(rays)))
    return c

def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--scene", default="10m"); ap.add_argument("--spp", type=int, default=8); ap.add_argument("--tile", type=int, default=16)
    a = ap.parse_args()
    W.C_NODE, W.C_TRI, W.C_LOOP = 280, 85, 35
    E = W.load_emu()
    b = scenes.scene_10m() if a.scene == "10m" else scenes.scene_1m()
    pos, _, _, idx = b.arrays()
    hnd = E.emu_build(W.vp(pos), W.vp(idx), idx.shape[0])
    rng = np.random.default_rng(5)
    for (x0, y0) in [(960, 540), (400, 700)]:
        prim = morton_tile_rays(b, x0, y0, a.tile, a.spp, rng)
        hits = np.zeros(len(prim), abi.HIT_DTYPE); E.emu_trace(hnd, W.vp(prim), len(prim), 0, W.vp(hits))
        sec, shadow = W.bounce_rays(rng, prim, hits, pos, idx, np.array([0, 8.0, 0]))
        print("tile at (%d,%d): %d primary, %d bounce rays" % (x0, y0, len(prim), len(sec)))
        run(E, hnd, prim, 0, "primary")
        for name, rays, mode in (("bounce", sec, 0), ("shadow", shadow, 1)):
            run(E, hnd, rays, mode, name + " shading order")
            for fine in (False, True):
                k = dir_key(rays["d"], fine)
                order = np.argsort(k, kind="stable")
                run(E, hnd, rays[order], mode, name + (" binned by 96 direction cells" if fine else " binned by octant"))
            # second bounce of the unsorted set
        h2 = np.zeros(len(sec), abi.HIT_DTYPE); E.emu_trace(hnd, W.vp(sec), len(sec), 0, W.vp(h2))
        sec2, _ = W.bounce_rays(rng, sec, h2, pos, idx, np.array([0, 8.0, 0]))
        run(E, hnd, sec2, 0, "bounce 2 shading order")
        k = dir_key(sec2["d"], False); run(E, hnd, sec2[np.argsort(k, kind="stable")], 0, "bounce 2 binned by octant")
        k = dir_key(sec2["d"], True); run(E, hnd, sec2[np.argsort(k, kind="stable")], 0, "bounce 2 binned by 96 cells")
main()
