"""BASELINE.json configs[3]: primary + shadow ray intersection microbenchmark -- 64 M coherent (camera) and 64 M
incoherent (bounding-sphere chord) rays against the 10 M-triangle BVH, closest-hit and any-hit, rays and hit records
resident in HBM (32 B in, 16 B out per ray).  Times come from the library's own CUDA events
around the k_trace launch (mtsb_trace_device's `ms`).  Hit records of the same ray sets are checked against the
reference kd-tree / a fp64 brute force in tests/test_gpu_trace.py; here only the any-hit <=> closest-hit property is
re-checked on the full sets.  An `e2e` figure traces a host-resident ray set through mtsb_trace (H2D + trace + D2H).

usage: python tools/bench_rays.py [--rays 67108864] [--reps 3] [--small]      -> one JSON line per ray set"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=1 << 26)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--e2e-rays", type=int, default=1 << 24)
    ap.add_argument("--small", action="store_true", help="1 M-triangle scene (testing)")
    a = ap.parse_args()
    import torch
    from mitsuba_b200 import abi, lib, scenes
    lib.init(0)
    b = (scenes.scene_1m if a.small else scenes.scene_10m)(width=64, height=36, spp=1)
    sc = lib.Scene(b)
    info = sc.info()
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    n = a.rays
    rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    hits_any = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    flush = torch.empty(160 << 20, dtype=torch.uint8, device="cuda")
    for kind, name in ((0, "coherent"), (1, "incoherent")):
        sc.gen_rays_device(kind, n, 1004, rays.data_ptr())
        torch.cuda.synchronize()
        out = {"metric": "mrays_per_s", "unit": "Mrays/s", "config": {"workload": "cfg3_rays_" + name, "rays": n,
               "triangles": int(info.n_tris), "bvh_nodes": int(info.n_bvh_nodes), "reps": a.reps,
               "l2": "flushed between launches (160 MB write); 3 GB of rays + hits per launch >> L2"}, "n_gpus": 1}
        for mode, key, dst in ((abi.TRACE_CLOSEST, "closest", hits), (abi.TRACE_ANY, "any", hits_any)):
            sc.trace_device(rays.data_ptr(), n, dst.data_ptr(), mode)          # warm-up
            ms = []
            for _ in range(a.reps):
                flush.zero_()
                torch.cuda.synchronize()
                ms.append(sc.trace_device(rays.data_ptr(), n, dst.data_ptr(), mode))
            t = float(np.mean(ms))
            out[key] = {"ms": t, "mrays_per_s": n / t / 1e3, "ms_all": [round(x, 3) for x in ms],
                        # streaming floor of this launch: 32 B ray in + 16 B hit out per ray
                        "io_gbs": n * 48 / t / 1e6, "io_frac_of_hbm_peak": n * 48 / t / 1e6 / peak}
        torch.cuda.synchronize()
        prim = hits.view(torch.int32)[:, 0]
        out["hit_fraction"] = float((prim != -1).float().mean())
        out["any_equals_closest"] = bool(torch.equal(hits_any.view(torch.int32)[:, 0] != -1, prim != -1))
        # host-resident rays through the C ABI: H2D + trace + D2H
        m = min(a.e2e_rays, n)
        h_rays = rays[:m].cpu().numpy().view(abi.RAY_DTYPE).reshape(-1)
        sc.trace(h_rays[: 1 << 16])
        h_hits = np.empty(m, dtype=abi.HIT_DTYPE)
        h_hits["prim"] = 0                                      # pages touched: the caller's buffer exists before the call
        dts = []
        for _ in range(3):
            t0 = time.time()
            sc.trace_into(h_rays.ctypes.data, m, h_hits.ctypes.data)
            dts.append(time.time() - t0)
        dt = min(dts)
        out["e2e"] = {"rays": m, "mrays_per_s": m / dt / 1e6, "h2d_bytes": int(h_rays.nbytes), "d2h_bytes": int(h_hits.nbytes),
                      "includes": "mtsb_trace with PAGEABLE host buffers: chunked H2D of the rays || k_trace<closest> || D2H of the hit records (pinned staging inside)"}
        # same records as the device-resident launch: t bit-identical; an exact equal-t tie between two triangles may name either
        d_hits = hits[:m].cpu().numpy()
        assert np.array_equal(h_hits["t"], d_hits[:, 3])
        out["e2e"]["prim_differs_at_equal_t"] = int((h_hits["prim"] != np.ascontiguousarray(d_hits).view(np.uint32)[:, 0]).sum())
        assert out["e2e"]["prim_differs_at_equal_t"] <= 1e-5 * m
        # the same with buffers the caller pinned (torch pinned tensors): used in place, one cudaMemcpyAsync per chunk and direction
        p_rays = rays[:m].cpu().pin_memory()
        p_hits = torch.empty((m, 4), dtype=torch.float32).pin_memory()
        dts = []
        for _ in range(3):
            t0 = time.time()
            sc.trace_into(p_rays.data_ptr(), m, p_hits.data_ptr())
            dts.append(time.time() - t0)
        out["e2e_pinned"] = {"rays": m, "mrays_per_s": m / min(dts) / 1e6, "includes": "mtsb_trace with caller-pinned host buffers"}
        assert np.array_equal(p_hits.numpy()[:, 3], d_hits[:, 3])
        print(json.dumps(out), flush=True)
    sc.close()


if __name__ == "__main__":
    main()
