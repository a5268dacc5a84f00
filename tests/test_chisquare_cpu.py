"""CPU: the chi-square machinery of tests/chisq.py exercised on the oracle's BSDF records (a reduced run: 6 incident
directions, 60 k samples each) -- checks that the port of ChiSquare accepts correct sampling routines and REJECTS a
broken one, so that the GPU run of the same suite (tests/test_gpu_chisquare.py) means something."""
import numpy as np

from mitsuba_b200 import abi, scenes
from mitsuba_b200.scenedesc import SceneBuilder, diffuse, look_at, roughconductor, roughdielectric
from tests import chisq


def _scene():
    b = SceneBuilder()
    eta, k = scenes.METALS["Au"]
    mats = {
        "diffuse": (b.add_material(diffuse(0.5)), False),
        "roughconductor_beckmann": (b.add_material(roughconductor(eta, k, alpha=0.3, distribution="beckmann")), False),
        "roughdielectric_ggx": (b.add_material(roughdielectric(1.5, 1.0, alpha=0.3, distribution="ggx")), True),
    }
    b.add_quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), 0)
    b.set_sensor(look_at([0, 1, -4], [0, 0.5, 0], [0, 1, 0]), 40.0, 16, 16)
    return b, mats


def test_chi_square_accepts_the_oracle_bsdfs(oracle_mod):
    b, mats = _scene()
    orc = oracle_mod.Oracle(b)
    for name, (mid, back) in mats.items():
        res = chisq.run_bsdf_suite(orc.bsdf_query, mid, back, seed=11, wi_samples=6, n_samples=60000)
        for r in res:
            assert r["accept"], (name, r)
            assert abs(r["integral"] - r["sampled_frac"]) < 0.02, (name, r)       # the pdf integrates to the sampling success rate
            assert r["weight_mismatch"] < 2e-3 and r["pdf_mismatch"] < 2e-3, (name, r)


def test_chi_square_rejects_a_wrong_pdf(oracle_mod):
    """Sampling a Beckmann lobe while reporting the pdf of a narrower one must be rejected."""
    b, mats = _scene()
    eta, k = scenes.METALS["Au"]
    wrong = b.add_material(roughconductor(eta, k, alpha=0.22, distribution="beckmann"))
    orc = oracle_mod.Oracle(b)
    right = mats["roughconductor_beckmann"][0]

    def mixed(material, q):
        out = orc.bsdf_query(right, q)                 # sample() of alpha = 0.3 ...
        other = orc.bsdf_query(wrong, q)               # ... eval()/pdf() of alpha = 0.22
        out["eval"], out["pdf"] = other["eval"], other["pdf"]
        return out
    res = chisq.run_bsdf_suite(mixed, right, False, seed=3, wi_samples=3, n_samples=60000)
    assert not any(r["accept"] for r in res), res
