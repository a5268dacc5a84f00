"""CPU: the closed-form furnace tests (tests/furnace.py) on the ORACLE's Li -- a pin of the restatement of
MIPathTracer::Li (path.cpp:121-296) that does not come from the restatement's author reading the same code twice."""
from tests import furnace


def _render(oracle_mod):
    def render(b, spp, seed):
        orc = oracle_mod.Oracle(b)
        orc.render(spp=spp, seed=seed)
        return orc.film_develop()
    return render


def test_white_furnace_oracle(oracle_mod):
    furnace.check_furnace(_render(oracle_mod), spp=48)


def test_convex_diffuse_body_oracle(oracle_mod):
    furnace.check_convex(_render(oracle_mod), spp=48)
