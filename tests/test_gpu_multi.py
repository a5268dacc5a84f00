"""GPU: one frame on several GPUs inside the library (mtsb_multi_*), the cancel path, and the per-bounce statistics.

Reference behaviour: the scheduler hands the blocks of ONE frame to its workers and merges them into one film
(src/librender/renderproc.cpp:149-156, imageproc.cpp:43-78); Integrator::cancel stops a running job
(src/librender/integrator.cpp:90-93).  The multi-GPU tests need two devices and skip on a one-GPU box."""
import os
import threading
import time

import numpy as np
import pytest

from mitsuba_b200 import abi, scenes
from mitsuba_b200.scenedesc import roughconductor

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bounce_statistics_add_up(gpu_lib):
    b = scenes.cbox(width=96, height=96, spp=16)
    sc = gpu_lib.Scene(b)
    st = sc.render(spp=16, seed=3)
    cb, sb = list(st.closest_rays_bounce), list(st.shadow_rays_bounce)
    assert sum(cb) == st.closest_rays and sum(sb) == st.shadow_rays
    assert cb[0] == st.samples                      # one camera ray per sample
    assert all(cb[i] >= cb[i + 1] for i in range(15))        # a ray spawns at most one ray
    assert cb[8] == 0                                # maxDepth 8: bounces 0..7 only
    assert 1 <= st.closest_launches <= 8 * st.n_passes and st.any_launches <= st.closest_launches
    assert st.n_devices == 1
    sc.close()


def test_unbounded_depth_terminates_by_polling(gpu_lib, oracle_mod):
    """maxDepth = -1: the host keeps enqueueing bounces until the poll ring shows an empty queue."""
    b = scenes.cbox(width=48, height=48, spp=8, max_depth=-1)
    sc, orc = gpu_lib.Scene(b), oracle_mod.Oracle(b)
    st = sc.render(spp=8, seed=5)
    ost = orc.render(spp=8, seed=5, counter_rng=True)
    assert st.closest_rays == ost.closest_rays and st.path_length_sum == ost.path_length_sum
    assert sum(st.closest_rays_bounce) == st.closest_rays and st.closest_rays_bounce[15] > 0     # paths longer than 15 exist
    sc.close()


def test_cancel_from_second_thread(gpu_lib):
    """mtsb_cancel during a render -> MTSB_ERR_CANCELLED; the next render of the same scene is clean."""
    b = scenes.scene_1m(width=1920, height=1080, spp=64, n_obj=4, nu=64, nv=32)
    sc = gpu_lib.Scene(b)
    ref = sc.render(spp=2, seed=1, spp_per_pass=1)
    img_ref = sc.film_develop()
    result = {}

    def work():
        try:
            result["st"] = sc.render(spp=64, seed=2, spp_per_pass=1)       # 64 passes: plenty of cancellation points
        except gpu_lib.MtsbError as e:
            result["err"] = e
    t = threading.Thread(target=work)
    t.start()
    time.sleep(0.05)
    sc.cancel()
    t.join(timeout=120)
    assert not t.is_alive()
    assert "err" in result and result["err"].code == abi.ERR_CANCELLED, result
    again = sc.render(spp=2, seed=1, spp_per_pass=1)
    assert again.samples == ref.samples and again.closest_rays == ref.closest_rays
    np.testing.assert_allclose(sc.film_develop(), img_ref, rtol=1e-5, atol=1e-6)      # film atomics: fp32 addition order
    sc.close()


def _two_gpus(gpu_lib):
    if gpu_lib.device_count() < 2:
        pytest.skip("needs two GPUs (run under gpurun --gpus 2)")


def test_multi_render_equals_single_gpu(gpu_lib):
    _two_gpus(gpu_lib)
    b = scenes.cbox(width=160, height=128, spp=8, short_material=roughconductor(*scenes.METALS["Cu"], alpha=0.1))
    single = gpu_lib.Scene(b)
    st1 = single.render(spp=8, seed=4)
    f1 = single.film_download().astype(np.float64)
    for n in (2, min(4, gpu_lib.device_count())):
        multi = gpu_lib.MultiScene(b, n)
        assert multi.device_count() == n
        stn = multi.render(spp=8, seed=4)
        fn = multi.scene(0).film_download().astype(np.float64)
        # same samples, same random numbers; only the order of the fp32 additions into the film differs
        assert stn.samples == st1.samples and stn.closest_rays == st1.closest_rays and stn.shadow_rays == st1.shadow_rays
        assert stn.path_length_sum == st1.path_length_sum and stn.n_devices == n and stn.ms_reduce > 0
        np.testing.assert_allclose(fn, f1, rtol=2e-5, atol=1e-5)
        np.testing.assert_allclose(multi.film_develop(), single.film_develop(), rtol=1e-4, atol=1e-5)
        # the partition belongs to the library
        with pytest.raises(gpu_lib.MtsbError):
            p = abi.RenderParams(); p.spp = 1; p.tile_index = 1; p.tile_count = 2
            import ctypes as C
            gpu_lib._check(gpu_lib.lib().mtsb_multi_render(multi.h, C.byref(p), None))
        multi.close()
    single.close()


def test_host_render_job_on_two_gpus(gpu_lib):
    """scenes/cbox.xml through the C++ loader and the render job on 2 GPUs (mtsbh_render_multi) == on one."""
    _two_gpus(gpu_lib)
    from mitsuba_b200 import hostlib
    sc = hostlib.HostScene.load_file(os.path.join(ROOT, "scenes", "cbox.xml"), {"res": 128, "spp": 4})
    img1, st1 = sc.render(spp=4)
    img2, st2 = sc.render(spp=4, gpus=2)
    assert st2.n_devices == 2 and st2.samples == st1.samples and st2.closest_rays == st1.closest_rays
    np.testing.assert_allclose(img2, img1, rtol=1e-4, atol=1e-5)
