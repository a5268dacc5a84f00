"""Wall-clock split of one end-to-end step (scene create / render / develop / destroy)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_b200 import lib, scenes
lib.init(0)
b = scenes.scene_1m(spp=int(sys.argv[1]) if len(sys.argv) > 1 else 64)
pair = b.desc()
for it in range(3):
    t0 = time.time(); sc = lib.Scene(pair); t1 = time.time()
    st = sc.render(seed=it + 1); t2 = time.time()
    img = sc.film_develop(); t3 = time.time()
    sc.close(); t4 = time.time()
    print("create %.1f ms (bvh %.1f)  render %.1f ms (device %.1f)  develop %.1f ms  destroy %.1f ms" % (
        (t1 - t0) * 1e3, sc.info().build_ms if False else 0, (t2 - t1) * 1e3, st.ms_total, (t3 - t2) * 1e3, (t4 - t3) * 1e3))
