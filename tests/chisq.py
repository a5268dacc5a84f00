"""Chi-square test of BSDF sampling routines -- the procedure of the reference's own statistical tests
(src/tests/test_chisquare.cpp:94-243 BSDFAdapter, :391-506 test01_BSDF; src/libcore/chisquare.cpp:94-263 ChiSquare),
restated over the record-query interface of the C ABI (mtsb_bsdf_query) so that it runs on the CUDA path:

 * `sample()` draws N directions for a fixed wi; they are binned in a thetaBins x phiBins table of spherical
   coordinates (chisquare.cpp:113-126);
 * the reference table is N x the integral of pdf(wo) sin(theta) over every bin (chisquare.cpp:148-176; a fixed
   composite Gauss-Legendre rule per bin replaces the adaptive cubature) -- pdf() reports 0 where eval() is zero, as the
   adapter does (test_chisquare.cpp:216-222);
 * cells are pooled until their expected count reaches 5, the statistic is compared with the chi-square
   distribution at the Sidak-corrected significance level (chisquare.cpp:190-263);
 * for every sample the three ways of getting its weight must agree: sample(pdf) == sample() == eval/pdf and
   sampledPDF == pdf(wo), absolute / relative error 1e-2 (test_chisquare.cpp:158-198, ERROR_REQ single precision).
`query(material, records)` is Scene.bsdf_query of the product binding or of the oracle (same record types)."""
import numpy as np
from scipy import stats

from mitsuba_b200 import abi

SIGNIFICANCE_LEVEL = 0.0025          # test_chisquare.cpp:30
ERROR_REQ = 1e-2                     # test_chisquare.cpp:35 (single precision build)
MIN_EXP_FREQUENCY = 5                # chisquare.h CHISQR_MIN_EXP_FREQUENCY


def to_spherical(d):
    """toSphericalCoordinates (util.cpp): theta = acos(z), phi = atan2(y, x) in [0, 2 pi)."""
    theta = np.arccos(np.clip(d[:, 2], -1.0, 1.0))
    phi = np.arctan2(d[:, 1], d[:, 0])
    phi = np.where(phi < 0, phi + 2 * np.pi, phi)
    return theta, phi


def square_to_cosine_hemisphere(u):
    """warp::squareToCosineHemisphere via the concentric disk map (warp.cpp)."""
    r1, r2 = 2 * u[:, 0] - 1, 2 * u[:, 1] - 1
    with np.errstate(divide="ignore", invalid="ignore"):
        a = np.abs(r1) > np.abs(r2)
        r = np.where(a, r1, r2)
        phi = np.where(a, (np.pi / 4) * (r2 / r1), (np.pi / 2) - (r1 / r2) * (np.pi / 4))
    phi = np.where((r1 == 0) & (r2 == 0), 0.0, phi)
    x, y = r * np.cos(phi), r * np.sin(phi)
    z = np.sqrt(np.maximum(0.0, 1 - x * x - y * y))
    return np.stack([x, y, z], -1)


def square_to_uniform_sphere(u):
    z = 1 - 2 * u[:, 1]
    r = np.sqrt(np.maximum(0.0, 1 - z * z))
    return np.stack([r * np.cos(2 * np.pi * u[:, 0]), r * np.sin(2 * np.pi * u[:, 0]), z], -1)


def _records(wi, wo=None, u=None, n=None):
    n = n if n is not None else (len(wo) if wo is not None else len(u))
    q = np.zeros(n, dtype=abi.BSDF_QUERY_DTYPE)
    q["wi"] = np.asarray(wi, np.float32)
    if wo is not None:
        q["wo"] = np.asarray(wo, np.float32)
    if u is not None:
        q["u"] = np.asarray(u, np.float32)
    return q


def chi_square_bsdf(query, material, wi, rng, n_samples=200000, theta_bins=10, phi_bins=20, n_tests=20, gl_order=8, gl_panels=6):
    """One contingency-table test for one incident direction.  Returns a dict with the p-value, the corrected
    significance level, the verdict and the fraction of samples whose weights / pdfs disagree."""
    wi = np.asarray(wi, np.float64)
    # ---- sampled table (+ the consistency checks of BSDFAdapter::generateSample) ----
    u = rng.random((n_samples, 3))
    s = query(material, _records(wi, u=u))
    ok = (s["spdf"] > 0) & np.any(s["weight"] != 0, axis=1)
    wo = s["swo"][ok].astype(np.float64)
    e = query(material, _records(wi, wo=wo))
    ev, pdf = e["eval"].astype(np.float64), e["pdf"].astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        manual = np.where(pdf[:, None] > 0, ev / pdf[:, None], 0.0)
    sampled = s["weight"][ok].astype(np.float64)

    def mismatch(a, b):
        mn, err = np.minimum(a, b), np.abs(a - b)
        return ((mn < ERROR_REQ) & (err > ERROR_REQ)) | ((mn > ERROR_REQ) & (err / np.maximum(mn, 1e-30) > ERROR_REQ))
    w_bad = mismatch(sampled, manual).any(axis=1)
    p_bad = mismatch(s["spdf"][ok].astype(np.float64), pdf)
    theta, phi = to_spherical(wo)
    tb = np.clip(np.floor(theta * theta_bins / np.pi).astype(int), 0, theta_bins - 1)
    pb = np.clip(np.floor(phi * phi_bins / (2 * np.pi)).astype(int), 0, phi_bins - 1)
    table = np.bincount(tb * phi_bins + pb, minlength=theta_bins * phi_bins).astype(np.float64)
    # ---- reference table: N * integral of pdf(wo) sin(theta) over each bin ----
    # composite Gauss-Legendre rule: gl_panels panels of order gl_order per bin and axis (48 x 48 points per bin; a
    # GGX lobe at grazing incidence needs that many for the 1e-3 accuracy the test's power asks for)
    x0, w0 = np.polynomial.legendre.leggauss(gl_order)
    x = np.concatenate([(x0 + 1 + 2 * p) / gl_panels - 1 for p in range(gl_panels)])
    wq = np.tile(w0 / gl_panels, gl_panels)
    gl_order = gl_order * gl_panels
    dth, dph = np.pi / theta_bins, 2 * np.pi / phi_bins
    th = ((np.arange(theta_bins)[:, None] + 0.5 * (x[None, :] + 1)) * dth).ravel()            # (tb * g)
    ph = ((np.arange(phi_bins)[:, None] + 0.5 * (x[None, :] + 1)) * dph).ravel()              # (pb * g)
    T, P = np.meshgrid(th, ph, indexing="ij")
    d = np.stack([np.sin(T) * np.cos(P), np.sin(T) * np.sin(P), np.cos(T)], -1).reshape(-1, 3)
    r = query(material, _records(wi, wo=d))
    f = np.where(np.any(r["eval"] != 0, axis=1), r["pdf"].astype(np.float64), 0.0) * np.sin(T).ravel()
    W = (np.tile(wq, theta_bins)[:, None] * np.tile(wq, phi_bins)[None, :]).ravel() * (0.25 * dth * dph)
    cell = (np.repeat(np.arange(theta_bins), gl_order)[:, None] * phi_bins + np.repeat(np.arange(phi_bins), gl_order)[None, :]).ravel()
    ref = np.bincount(cell, weights=f * W, minlength=theta_bins * phi_bins) * n_samples
    # ---- ChiSquare::runTest ----
    order = np.argsort(ref, kind="stable")
    chsq, df, pooled_c, pooled_r, pooled_n = 0.0, 0, 0.0, 0.0, 0
    zero_cell_hits = 0.0
    for i in order:
        if ref[i] == 0:
            zero_cell_hits += table[i]
        elif ref[i] < MIN_EXP_FREQUENCY or (0 < pooled_r < MIN_EXP_FREQUENCY):
            pooled_c += table[i]; pooled_r += ref[i]; pooled_n += 1
        else:
            chsq += (table[i] - ref[i]) ** 2 / ref[i]; df += 1
    if pooled_n:
        chsq += (pooled_c - pooled_r) ** 2 / pooled_r; df += 1
    df -= 1
    alpha = 1 - (1 - SIGNIFICANCE_LEVEL) ** (1.0 / n_tests)
    pval = float(stats.chi2.sf(chsq, df)) if df > 0 else 1.0
    return dict(pval=pval, alpha=alpha, df=df, chsq=chsq, accept=bool(pval >= alpha and zero_cell_hits <= n_samples * 1e-6),
                integral=float(ref.sum() / n_samples), sampled_frac=float(ok.mean()), weight_mismatch=float(w_bad.mean()) if ok.any() else 0.0,
                pdf_mismatch=float(p_bad.mean()) if ok.any() else 0.0, zero_cell_hits=float(zero_cell_hits))


def run_bsdf_suite(query, material, back_side, seed, wi_samples=20, n_samples=200000):
    """test01_BSDF for one material: wi_samples incident directions (uniform sphere when the BSDF has a back side,
    cosine hemisphere otherwise: test_chisquare.cpp:418-421)."""
    rng = np.random.default_rng(seed)
    u = rng.random((wi_samples, 2))
    wis = square_to_uniform_sphere(u) if back_side else square_to_cosine_hemisphere(u)
    return [dict(wi=wi.tolist(), **chi_square_bsdf(query, material, wi, rng, n_samples=n_samples, n_tests=wi_samples)) for wi in wis]
