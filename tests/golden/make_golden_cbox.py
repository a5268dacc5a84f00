"""Generates tests/golden/cbox_512_4096spp.npz: BASELINE.json configs[0] (Cornell box, `path`, 512x512, maxDepth 8,
independent sampler) rendered by the CPU oracle at 4096 spp -- the converged image the equal-spp comparison of
tests/test_gpu_render_full.py measures relMSE against.  ~5 min on 8 cores.

    python tests/golden/make_golden_cbox.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from mitsuba_b200 import scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    b = scenes.cbox(width=512, height=512, spp=64, max_depth=8)
    orc = oracle.Oracle(b)
    st = orc.render(spp=4096, seed=20261105)
    img = orc.film_develop()
    n = st.samples
    np.savez_compressed(os.path.join(HERE, "cbox_512_4096spp.npz"), image=img.astype(np.float16), samples=np.int64(n),
                        lum_mean=np.float64(st.lum_sum / n), lum_var=np.float64(st.lum_sq_sum / n - (st.lum_sum / n) ** 2),
                        rays_per_sample=np.float64(st.closest_rays / n), path_length=np.float64(st.path_length_sum / n))
    print("samples", n, "seconds", st.seconds, "mean luminance", st.lum_sum / n)


if __name__ == "__main__":
    main()
