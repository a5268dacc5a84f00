"""GPU: the closed-form furnace tests (tests/furnace.py) on the CUDA path (mtsb_render): radiance L through a lossless white
scene under a uniform sky, albedo x L on a convex diffuse body -- checks of MIS, roulette, cosine / pdf factors and the film
normalisation that do not involve the oracle at all."""
import pytest

from tests import furnace

pytestmark = pytest.mark.gpu


def _render(gpu_lib):
    def render(b, spp, seed):
        sc = gpu_lib.Scene(b)
        st = sc.render(spp=spp, seed=seed)
        assert st.bad_samples == 0
        img = sc.film_develop()
        sc.close()
        return img
    return render


def test_white_furnace(gpu_lib):
    furnace.check_furnace(_render(gpu_lib), spp=256)


def test_convex_diffuse_body(gpu_lib):
    furnace.check_convex(_render(gpu_lib), spp=256)
