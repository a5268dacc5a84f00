"""Closed-form radiance tests for the `path` integrator (independent of the oracle restatement of MIPathTracer::Li).

 * White furnace: surfaces that absorb nothing (diffuse reflectance 1) inside a uniform environment of radiance L and no
   other light: the radiance arriving at the camera is L along EVERY path (energy is neither created nor lost), whatever the
   geometry, with unbounded path length (Russian roulette is unbiased: path.cpp:278-288).  Any error in the MIS weights between
   BSDF sampling and emitter sampling (path.cpp:174-266), in the cosine / pdf factors, in roulette or in the film normalisation
   shows up as a deviation from L.
 * Convex diffuse body of albedo rho under the same environment: no interreflection, so the outgoing radiance is rho * L at
   every visible point, and L where the camera sees the sky.
`render(builder, spp, seed)` returns (image[h, w, 3], per-sample luminance mean, variance or None)."""
import numpy as np

from mitsuba_b200 import scenes
from mitsuba_b200.scenedesc import SceneBuilder, diffuse, look_at

LUM = np.array([0.212671, 0.715160, 0.072169])


def convex_polyhedron(n_lat=10, n_lon=16, radius=1.0, center=(0.0, 0.0, 0.0)):
    """A UV sphere with flat facets: convex, so no point of it sees another."""
    th = np.linspace(0, np.pi, n_lat + 1)
    ph = np.linspace(0, 2 * np.pi, n_lon, endpoint=False)
    P = [[0, radius, 0]]
    for t in th[1:-1]:
        for p in ph:
            P.append([radius * np.sin(t) * np.cos(p), radius * np.cos(t), radius * np.sin(t) * np.sin(p)])
    P.append([0, -radius, 0])
    P = np.asarray(P, np.float32) + np.asarray(center, np.float32)
    T = []
    ring = lambda r, k: 1 + r * n_lon + (k % n_lon)
    for k in range(n_lon):
        T.append([0, ring(0, k + 1), ring(0, k)])
        T.append([len(P) - 1, ring(n_lat - 2, k), ring(n_lat - 2, k + 1)])
    for r in range(n_lat - 2):
        for k in range(n_lon):
            a, b, c, d = ring(r, k), ring(r, k + 1), ring(r + 1, k + 1), ring(r + 1, k)
            T += [[a, b, c], [a, c, d]]
    return P, np.asarray(T, np.uint32)


def furnace_scene(width=96, height=54, spp=64, radiance=(0.8, 1.0, 1.2)):
    """The 1 M-triangle family's geometry (displaced spheres and tori, a ground plane) FLAT-shaded -- with interpolated vertex
    normals a sampled direction can enter the geometric surface and is absorbed by the back side, a real (and intended) loss --
    and without the one-sided light quad; every surface diffuse white, uniform environment, unbounded depth."""
    src = scenes.scene_1m(width=width, height=height, spp=spp, max_depth=-1, nu=32, nv=16, n_obj=9)
    b = SceneBuilder()
    b.set_integrator(max_depth=-1, rr_depth=5, sample_count=spp)
    b.sensor = src.sensor
    white = b.add_material(diffuse(1.0))
    for k, sh in enumerate(src.shapes):
        if sh["emitter"] >= 0:
            continue
        idx = src._idx[k] - np.uint32(sh["first_vertex"])
        p = src._pos[k]
        fn = np.cross(p[idx[:, 1]] - p[idx[:, 0]], p[idx[:, 2]] - p[idx[:, 0]])
        b.add_mesh(p, idx[np.linalg.norm(fn, axis=1) > 0], white)          # the collapsed pole triangles carry no area
    b.add_constant_env(list(radiance))
    return b


def convex_scene(width=96, height=72, spp=64, albedo=(0.6, 0.3, 0.9), radiance=2.0):
    b = SceneBuilder()
    b.set_integrator(max_depth=-1, rr_depth=5, sample_count=spp)
    P, T = convex_polyhedron()
    b.add_mesh(P, T, b.add_material(diffuse(list(albedo))))
    b.add_constant_env(radiance)
    b.set_sensor(look_at([0, 0.4, -4], [0, 0, 0], [0, 1, 0]), 40.0, width, height)
    return b, P, T


def check_furnace(render, spp=64, tol_sigma=4.0):
    L = np.array([0.8, 1.0, 1.2])
    b = furnace_scene(spp=spp, radiance=L)
    img = render(b, spp, 3).astype(np.float64)
    assert np.isfinite(img).all()
    ratio = img / L                                   # expected: 1 in every pixel and channel
    # per pixel: a mean of `spp` samples of a bounded estimator; the whole image: ~5000 of those
    assert abs(ratio.mean() - 1.0) < 2e-3, ratio.mean()
    assert np.percentile(np.abs(ratio - 1.0), 99) < 0.25, np.percentile(np.abs(ratio - 1.0), 99)
    # no region is systematically off: 6 x 6 pixel block means stay within a few percent
    h, w = ratio.shape[:2]
    blocks = ratio[: h // 6 * 6, : w // 6 * 6].reshape(h // 6, 6, w // 6, 6, 3).mean(axis=(1, 3, 4))
    tol = 0.65 / np.sqrt(spp)                       # ~4 sigma of a 36-pixel block mean (per-sample sigma ~1); 0.094 at 48 spp, 0.041 at 256
    assert np.abs(blocks - 1.0).max() < tol, (np.abs(blocks - 1.0).max(), tol)
    return float(ratio.mean())


def check_convex(render, spp=64):
    albedo, Lenv = np.array([0.6, 0.3, 0.9]), 2.0
    b, P, T = convex_scene(spp=spp, albedo=albedo, radiance=Lenv)
    img = render(b, spp, 5).astype(np.float64)
    h, w = img.shape[:2]
    # classify pixels by what the pixel CENTRE ray sees, well inside / well outside the silhouette (radius 1 around the origin)
    cam = np.array([0, 0.4, -4.0])
    fwd = -cam / np.linalg.norm(cam)
    right = np.cross([0, 1, 0], fwd); right /= np.linalg.norm(right)
    up = np.cross(fwd, right)
    tx = np.tan(np.radians(40.0) / 2); ty = tx * h / w
    ys, xs = np.mgrid[0:h, 0:w]
    d = fwd[None, None] + (1 - 2 * (xs[..., None] + 0.5) / w) * tx * right + (1 - 2 * (ys[..., None] + 0.5) / h) * ty * up
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    dist = np.linalg.norm(np.cross(np.broadcast_to(-cam, d.shape), d), axis=-1)      # distance of the ray from the origin
    inside, outside = dist < 0.80, dist > 1.15
    assert inside.sum() > 200 and outside.sum() > 500
    sky = img[outside]
    assert np.abs(sky / Lenv - 1.0).max() < 1e-4, np.abs(sky / Lenv - 1.0).max()          # deterministic: the environment itself
    body = img[inside] / (albedo * Lenv)
    assert abs(body.mean() - 1.0) < 3e-3, body.mean()
    assert np.percentile(np.abs(body - 1.0), 99) < 0.2
    return float(body.mean())
