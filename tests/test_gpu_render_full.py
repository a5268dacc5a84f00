"""GPU: BASELINE.json configs[0] at its FULL size -- Cornell box, `path`, 512 x 512, 64 spp, maxDepth 8, independent
sampler -- against the converged 4096-spp image of the CPU oracle (tests/golden/cbox_512_4096spp.npz, made by
tests/golden/make_golden_cbox.py).

Bar: relMSE of the CUDA render against the converged image no worse than 1.25 x the relMSE of the CPU
oracle's own render at the same sample count (both are Monte Carlo noise of the same estimator, so the ratio sits at
1 up to the sampling error of the error estimate itself), and the mean luminance inside the 3-sigma interval of the
per-sample estimator.  The GPU draws counter-based Philox numbers, the oracle its sequential generator: this is the
independent-stream comparison; the same-stream (bit-level) comparison is tests/test_gpu_render.py."""
import numpy as np
import pytest

from mitsuba_b200 import scenes
from tests import util

pytestmark = pytest.mark.gpu
LUM = np.array([0.212671, 0.715160, 0.072169])


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def test_cbox_512_64spp_against_converged_oracle_image(gpu_lib, oracle_mod):
    g = util.load_golden("cbox_512_4096spp.npz")
    ref = g["image"].astype(np.float64)
    b = scenes.cbox(width=512, height=512, spp=64, max_depth=8)
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    st = sc.render(spp=64, seed=1004)
    gpu = sc.film_develop().astype(np.float64)
    cst = orc.render(spp=64, seed=77)
    cpu = orc.film_develop().astype(np.float64)
    assert st.samples == cst.samples == 512 * 512 * 64 and st.bad_samples == 0 and np.isfinite(gpu).all()
    e_gpu, e_cpu = rel_mse(gpu, ref), rel_mse(cpu, ref)
    assert e_gpu <= 1.25 * e_cpu, (e_gpu, e_cpu)
    # mean luminance: 3 sigma of the per-sample estimator (variance from the converged run)
    sigma = np.sqrt(float(g["lum_var"]) * (1.0 / st.samples + 1.0 / float(g["samples"])))
    lum_gpu, lum_ref = float((gpu @ LUM).mean()), float((ref @ LUM).mean())
    assert abs(lum_gpu - lum_ref) <= 3 * sigma + 5e-4 * lum_ref, (lum_gpu, lum_ref, sigma)     # 5e-4: the half-precision storage of the golden image
    # the same amount of work per sample as the converged run (reference statistics counters)
    assert abs(st.closest_rays / st.samples - float(g["rays_per_sample"])) <= 0.01 * float(g["rays_per_sample"])
    assert abs(st.path_length_sum / st.samples - float(g["path_length"])) <= 0.01 * float(g["path_length"])
    sc.close()
