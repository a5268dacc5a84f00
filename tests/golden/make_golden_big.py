"""Generates tests/golden/hits_big_{1m,10m}.npz: first-hit records of the REFERENCE kd-tree (oracle/_ref, compiled from
the reference tree: SAHKDTree3D + TriAccel, src/librender/skdtree.cpp:112-148, 213-232) on the full-size scenes of
BASELINE.json configs[1] / [2], for an even subsample (65 536 rays) of the two 4 M-ray sets of configs[3]:
coherent camera rays and incoherent bounding-sphere chords (ray-set template src/tests/test_kd.cpp:86-131).

    make -C oracle all ref && python tests/golden/make_golden_big.py        (needs /root/reference; ~3 min, ~12 GB)

tests/test_gpu_trace_big.py compares the CUDA traversal with these records on every box, and with the kd-tree itself
(all 4 M rays of each set) where oracle/_ref is present."""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from mitsuba_b200 import abi, scenes  # noqa: E402
from tests import util  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sensor_rays_oracle(builder):
    """The oracle's sensor over a scene that keeps only the camera (no geometry: nothing to build)."""
    from mitsuba_b200.scenedesc import SceneBuilder, diffuse
    b = SceneBuilder()
    b.sensor, b.integrator, b.sampler = builder.sensor, builder.integrator, builder.sampler
    b.add_material(diffuse(0.5))
    b.add_quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), 0)
    return oracle.Oracle(b).sensor_rays


def main():
    for name, (fn, n_tris) in util.BIG_SCENES.items():
        t0 = time.time()
        b = getattr(scenes, fn)()
        kd = oracle.RefKDTree(b)
        tight, _ = kd.bounds()
        print(name, "kd-tree built in %.1f s" % (time.time() - t0), flush=True)
        sets = util.big_ray_sets(b, sensor_rays_oracle(b), tight[:3], tight[3:])
        out = {"digest": np.frombuffer(util.scene_digest(b).encode(), dtype=np.uint8), "tight": tight, "n_tris": np.int64(n_tris)}
        for sname, rays in sets.items():
            ix = util.golden_subset(len(rays))
            sub = np.ascontiguousarray(rays[ix])
            out["sha_" + sname] = np.frombuffer(hashlib.sha256(rays.tobytes()).hexdigest().encode(), dtype=np.uint8)
            closest = util.kd_trace_parallel(kd, sub, abi.TRACE_CLOSEST)
            occl = util.kd_trace_parallel(kd, sub, abi.TRACE_ANY)["prim"] != abi.NO_HIT
            out["closest_" + sname] = closest
            out["any_" + sname] = np.packbits(occl)
            print("  %s: %d of %d rays, hit fraction %.3f, occluded %.3f" % (sname, len(sub), len(rays), (closest["prim"] != abi.NO_HIT).mean(), occl.mean()), flush=True)
        np.savez_compressed(os.path.join(HERE, "hits_big_%s.npz" % name), **out)
        del kd


if __name__ == "__main__":
    main()
