// shading.h -- intersection record, emitter sampling and film splatting (per-thread bodies).
//
//   fill_intersection      ShapeKDTree::fillIntersectionRecord<true>   include/mitsuba/render/skdtree.h:343-430
//   pmf_sample_reuse       DiscreteDistribution::sampleReuse           include/mitsuba/core/pmf.h:124-136,183-188
//   emitter_sample_direct  AreaLight/PointEmitter/ConstantBackground   src/emitters/{area.cpp:158-173, point.cpp:131-147, constant.cpp:173-215}
//                          Shape::sampleDirect src/librender/shape.cpp:102-115, TriMesh::samplePosition src/librender/trimesh.cpp:412-423,
//                          Triangle::sample src/libcore/triangle.cpp:24-62
//   emitter_pdf_direct     area.cpp:175-182 / shape.cpp:117-126, point.cpp:149-151, constant.cpp:217-232
//   sample_emitter_direct  Scene::sampleEmitterDirect src/librender/scene.cpp:828-855 (visibility is a separate shadow-ray pass)
//   film_weights           ImageBlock::put include/mitsuba/render/imageblock.h:143-206
#pragma once
#include "bsdf.h"
#include "scene_view.h"

namespace mtsb {

struct Frame3 {
    V3 s, t, n;
    MTSB_HD V3 toLocal(V3 v) const { return V3(dot(v, s), dot(v, t), dot(v, n)); }
    MTSB_HD V3 toWorld(V3 v) const { return s * v.x + t * v.y + n * v.z; }
};
// util.cpp:592-601
MTSB_HD void coordinateSystem(V3 a, V3 &b, V3 &c) {
    if (fabsf(a.x) > fabsf(a.y)) {
        float invLen = 1.0f / sqrtf(a.x * a.x + a.z * a.z);
        c = V3(a.z * invLen, 0.0f, -a.x * invLen);
    } else {
        float invLen = 1.0f / sqrtf(a.y * a.y + a.z * a.z);
        c = V3(0.0f, a.z * invLen, -a.y * invLen);
    }
    b = cross(c, a);
}

struct Its {
    V3 p, geoN; Frame3 sh; P2 uv; V3 wi; uint32_t shape;
    V3 dpdu, dpdv;          // position partials (skdtree.h:373-380): uv tangents where the mesh has texture coordinates
};

// Per-triangle shading record, 128 B = one L2 line (built once at scene creation by k_pack_tris, render.cu): everything
// fillIntersectionRecord reads for a hit sits behind ONE index, instead of the chain prim -> tri_shape -> shape ->
// indices -> three vertices -> three normals of the flat arrays.
//   r0 = p0.xyz, shape   r1 = p1.xyz, flags (1: vertex normals, 2: texture coordinates)   r2 = p2.xyz, -
//   r3 = n0.xyz, uv0.x   r4 = n1.xyz, uv0.y   r5 = n2.xyz, uv1.x   r6 = uv1.y, uv2.x, uv2.y, -   (r3.. only where flagged)
#define MTSB_SHADE_TRI_U4 8
MTSB_HD void fill_intersection(const SceneView &sc, V3 ray_d, uint32_t prim, float hu, float hv, Its &its) {
    const U4 *rec = sc.shade_tris + (size_t) MTSB_SHADE_TRI_U4 * prim;
    const U4 r0 = ld16(rec), r1 = ld16(rec + 1), r2 = ld16(rec + 2);
    const uint32_t s = r0.w, flags = r1.w;
    const V3 b(1 - hu - hv, hu, hv);
    const V3 p0(u2f(r0.x), u2f(r0.y), u2f(r0.z)), p1(u2f(r1.x), u2f(r1.y), u2f(r1.z)), p2(u2f(r2.x), u2f(r2.y), u2f(r2.z));
    its.p = p0 * b.x + p1 * b.y + p2 * b.z;
    V3 side1 = p1 - p0, side2 = p2 - p0;
    V3 faceNormal = cross(side1, side2);
    float len = length(faceNormal);
    if (!isZero(faceNormal)) faceNormal = faceNormal / len;
    U4 r3, r4, r5;
    r3.x = r3.y = r3.z = r3.w = 0; r4 = r3; r5 = r3;
    if (flags & 3u) { r3 = ld16(rec + 3); r4 = ld16(rec + 4); r5 = ld16(rec + 5); }
    V3 shN;
    if (flags & 1u) {
        const V3 n0(u2f(r3.x), u2f(r3.y), u2f(r3.z)), n1(u2f(r4.x), u2f(r4.y), u2f(r4.z)), n2(u2f(r5.x), u2f(r5.y), u2f(r5.z));
        shN = normalize(n0 * b.x + n1 * b.y + n2 * b.z);
        if (dot(faceNormal, shN) < 0) faceNormal = -faceNormal;
    } else shN = faceNormal;
    its.geoN = faceNormal;
    if (flags & 2u) {
        const U4 r6 = ld16(rec + 6);
        const float t0[2] = {u2f(r3.w), u2f(r4.w)}, t1[2] = {u2f(r5.w), u2f(r6.x)}, t2[2] = {u2f(r6.y), u2f(r6.z)};
        its.uv.x = t0[0] * b.x + t1[0] * b.y + t2[0] * b.z; its.uv.y = t0[1] * b.x + t1[1] * b.y + t2[1] * b.z;
        // TriMesh::computeUVTangents (trimesh.cpp:683-735): this fork's TriMesh::configure always builds the tangents of a
        // mesh that has texture coordinates (trimesh.cpp:383-385), and fillIntersectionRecord then takes dpdu / dpdv from them
        const float d1x = t1[0] - t0[0], d1y = t1[1] - t0[1], d2x = t2[0] - t0[0], d2y = t2[1] - t0[1];
        const float determinant = d1x * d2y - d1y * d2x;
        if (len == 0) { its.dpdu = V3(0.0f); its.dpdv = V3(0.0f); }
        else if (determinant == 0) coordinateSystem(cross(side1, side2) / len, its.dpdu, its.dpdv);
        else {
            const float invDet = 1.0f / determinant;
            its.dpdu = (side1 * d2y - side2 * d1y) * invDet;
            its.dpdv = (side1 * -d2x + side2 * d1x) * invDet;
        }
    } else { its.uv.x = b.y; its.uv.y = b.z; its.dpdu = side1; its.dpdv = side2; }
    its.shape = s;
    // computeShadingFrame(n, dpdu): Gram-Schmidt on dpdu (util.cpp:603-608)
    its.sh.n = shN;
    its.sh.s = normalize(its.dpdu - shN * dot(shN, its.dpdu));
    its.sh.t = cross(shN, its.sh.s);
    its.wi = its.sh.toLocal(-ray_d);
}

// DiscreteDistribution::sample on a cdf with cdf[0] = 0 and n+1 entries
MTSB_HD uint32_t pmf_sample(const float *cdf, uint32_t n, float v) {
    // std::lower_bound(cdf, cdf + n + 1, v)
    uint32_t lo = 0, cnt = n + 1;
    while (cnt > 0) {
        uint32_t step = cnt / 2, mid = lo + step;
        if (cdf[mid] < v) { lo = mid + 1; cnt -= step + 1; } else cnt = step;
    }
    int64_t idx = (int64_t) lo - 1;
    if (idx < 0) idx = 0;
    if (idx > (int64_t) n - 1) idx = (int64_t) n - 1;
    uint32_t index = (uint32_t) idx;
    while (index < n - 1 && cdf[index + 1] - cdf[index] == 0) ++index;
    return index;
}
MTSB_HD uint32_t pmf_sample_reuse(const float *cdf, uint32_t n, float &v, float &pdf) {
    uint32_t index = pmf_sample(cdf, n, v);
    pdf = cdf[index + 1] - cdf[index];
    v = (v - cdf[index]) / (cdf[index + 1] - cdf[index]);
    return index;
}

struct DRec {
    V3 ref, refN, p, n, d; float dist, pdf; int measure; int emitter;
};
enum { MTSB_MEASURE_SOLID_ANGLE = 0, MTSB_MEASURE_DISCRETE = 1 };

MTSB_HD bool solve_quadratic(float a, float b, float c, float &x0, float &x1) {       // util.cpp:447-485
    if (a == 0) { if (b != 0) { x0 = x1 = -c / b; return true; } return false; }
    float discrim = b * b - 4.0f * a * c;
    if (discrim < 0) return false;
    float temp, sqrtDiscrim = sqrtf(discrim);
    if (b < 0) temp = -0.5f * (b - sqrtDiscrim); else temp = -0.5f * (b + sqrtDiscrim);
    x0 = temp / a; x1 = c / temp;
    if (x0 > x1) { float t = x0; x0 = x1; x1 = t; }
    return true;
}
MTSB_HD bool bsphere_intersect(const float *c3, float r, V3 o, V3 d, float &nearT, float &farT) {   // bsphere.h:88-95
    V3 oc = o - ld3(c3);
    float A = lengthSquared(d), B = 2 * dot(oc, d), C = lengthSquared(oc) - r * r;
    return solve_quadratic(A, B, C, nearT, farT);
}

MTSB_HD V3 area_eval(const MtsbEmitter &e, V3 shN, V3 d) {       // area.cpp:104-109
    if (dot(shN, d) <= 0) return V3(0.0f);
    return ld3(e.radiance);
}

// Scene::evalEnvironment (scene.h:914-917): radiance arriving along a ray of direction d that left the scene
MTSB_HD V3 environment_eval(const SceneView &sc, V3 d) {
    const MtsbEmitter &e = sc.emitters[sc.env_emitter];
    if (e.type == MTSB_EMITTER_ENVMAP) { const EnvmapView &ev = *sc.envmap; return envmap_eval(ev, d); }
    return ld3(e.radiance);                                                  // constant.cpp:242-244
}

// The same for a camera ray, which carries differentials: EnvironmentMap::evalEnvironment filters the lookup through
// MIPMap::eval (envmap.cpp:393-407); rxd / ryd are the directions of the ray's x / y offset rays.
MTSB_HD V3 environment_eval_filtered(const SceneView &sc, V3 d, V3 rxd, V3 ryd) {
    const MtsbEmitter &e = sc.emitters[sc.env_emitter];
    if (e.type == MTSB_EMITTER_ENVMAP) { const EnvmapView &ev = *sc.envmap; return envmap_eval_filtered(ev, d, rxd, ryd); }
    return ld3(e.radiance);
}

MTSB_HD V3 emitter_sample_direct(const SceneView &sc, int ei, DRec &dRec, P2 u) {
    const MtsbEmitter &e = sc.emitters[ei];
    if (e.type == MTSB_EMITTER_AREA) {
        const MtsbShape &sh = sc.shapes[e.shape];
        float sy = u.y, tri_pdf;
        const uint32_t index = pmf_sample_reuse(sc.area_cdf + sc.area_cdf_offset[e.shape], sh.n_tris, sy, tri_pdf);
        const uint32_t *ix = sc.indices + 3 * ((size_t) sh.first_tri + index);
        const uint32_t i0 = ix[0], i1 = ix[1], i2 = ix[2];
        const V3 p0 = ld3(sc.positions + 3 * (size_t) i0), p1 = ld3(sc.positions + 3 * (size_t) i1), p2 = ld3(sc.positions + 3 * (size_t) i2);
        P2 uu; uu.x = u.x; uu.y = sy;
        const P2 bary = squareToUniformTriangle(uu);
        const V3 sideA = p1 - p0, sideB = p2 - p0;
        dRec.p = p0 + (sideA * bary.x) + (sideB * bary.y);
        if (sh.has_normals) {
            const V3 n0 = ld3(sc.normals + 3 * (size_t) i0), n1 = ld3(sc.normals + 3 * (size_t) i1), n2 = ld3(sc.normals + 3 * (size_t) i2);
            dRec.n = normalize(n0 * (1.0f - bary.x - bary.y) + n1 * bary.x + n2 * bary.y);
        } else dRec.n = normalize(cross(sideA, sideB));
        dRec.pdf = sc.inv_area[e.shape];
        dRec.d = dRec.p - dRec.ref;
        const float distSquared = lengthSquared(dRec.d);
        dRec.dist = sqrtf(distSquared);
        dRec.d = dRec.d / dRec.dist;
        const float dp = absDot(dRec.d, dRec.n);
        dRec.pdf *= dp != 0 ? (distSquared / dp) : 0.0f;
        dRec.measure = MTSB_MEASURE_SOLID_ANGLE;
        if (dot(dRec.d, dRec.refN) >= 0 && dot(dRec.d, dRec.n) < 0 && dRec.pdf != 0)
            return ld3(e.radiance) / dRec.pdf;
        dRec.pdf = 0.0f;
        return V3(0.0f);
    } else if (e.type == MTSB_EMITTER_POINT) {
        dRec.p = ld3(e.position);
        dRec.pdf = 1.0f; dRec.measure = MTSB_MEASURE_DISCRETE;
        dRec.d = dRec.p - dRec.ref;
        dRec.dist = length(dRec.d);
        const float invDist = 1.0f / dRec.dist;
        dRec.d *= invDist;
        dRec.n = V3(0.0f);
        return ld3(e.radiance) * (invDist * invDist);
    } else if (e.type == MTSB_EMITTER_ENVMAP) {
        // EnvironmentMap::sampleDirect (envmap.cpp:516-543)
        const EnvmapView &ev = *sc.envmap;
        V3 dl, value; float pdf;
        envmap_sample_direction(ev, u.x, u.y, dl, value, pdf);
        const V3 d = envmap_xform(ev.to_world, dl);
        float nearT, farT;
        dRec.pdf = 0.0f;
        if (isZero(value) || pdf == 0 || !bsphere_intersect(sc.bsphere_c, sc.bsphere_r, dRec.ref, d, nearT, farT) || nearT >= 0 || farT <= 0)
            return V3(0.0f);
        dRec.pdf = pdf;
        dRec.p = dRec.ref + d * farT;
        dRec.n = normalize(ld3(sc.bsphere_c) - dRec.p);
        dRec.dist = farT; dRec.d = d; dRec.measure = MTSB_MEASURE_SOLID_ANGLE;
        return value / pdf;
    } else {
        V3 d; float pdf;
        if (!isZero(dRec.refN)) {
            d = squareToCosineHemisphere(u);
            pdf = squareToCosineHemispherePdf(d);
            Frame3 f; f.n = dRec.refN; coordinateSystem(f.n, f.s, f.t);
            d = f.toWorld(d);
        } else {
            d = squareToUniformSphere(u);
            pdf = MTSB_INV_FOURPI;
        }
        float nearT, farT;
        dRec.pdf = 0.0f;
        if (!bsphere_intersect(sc.bsphere_c, sc.bsphere_r, dRec.ref, d, nearT, farT)) return V3(0.0f);
        if (!(nearT < 0 && farT > 0)) return V3(0.0f);
        dRec.p = dRec.ref + d * farT;
        dRec.n = normalize(ld3(sc.bsphere_c) - dRec.p);
        dRec.measure = MTSB_MEASURE_SOLID_ANGLE;
        dRec.d = d; dRec.dist = farT; dRec.pdf = pdf;
        if (!isZero(dRec.refN) && dot(dRec.d, dRec.refN) <= 0) return V3(0.0f);
        return ld3(e.radiance) / pdf;
    }
}

MTSB_HD float emitter_pdf_direct(const SceneView &sc, int ei, const DRec &dRec) {
    const MtsbEmitter &e = sc.emitters[ei];
    if (e.type == MTSB_EMITTER_AREA) {
        if (dot(dRec.d, dRec.refN) >= 0 && dot(dRec.d, dRec.n) < 0) {
            const float pdfPos = sc.inv_area[e.shape];
            if (dRec.measure == MTSB_MEASURE_SOLID_ANGLE) return pdfPos * (dRec.dist * dRec.dist) / absDot(dRec.d, dRec.n);
        }
        return 0.0f;
    } else if (e.type == MTSB_EMITTER_POINT) {
        return dRec.measure == MTSB_MEASURE_DISCRETE ? 1.0f : 0.0f;
    } else if (e.type == MTSB_EMITTER_ENVMAP) {
        // EnvironmentMap::pdfDirect (envmap.cpp:545-556)
        const EnvmapView &ev = *sc.envmap;
        const float pdfSA = envmap_pdf_direction(ev, envmap_xform(ev.to_local, dRec.d));
        return dRec.measure == MTSB_MEASURE_SOLID_ANGLE ? pdfSA : 0.0f;
    }
    float pdfSA;
    if (!isZero(dRec.refN)) pdfSA = MTSB_INV_PI * fmaxf(0.0f, dot(dRec.d, dRec.refN));
    else pdfSA = MTSB_INV_FOURPI;
    return dRec.measure == MTSB_MEASURE_SOLID_ANGLE ? pdfSA : 0.0f;
}

MTSB_HD V3 sample_emitter_direct(const SceneView &sc, DRec &dRec, P2 u) {
    dRec.emitter = -1; dRec.pdf = 0;
    if (sc.n_emitters == 0) return V3(0.0f);
    float emPdf, sx = u.x;
    const uint32_t index = pmf_sample_reuse(sc.emitter_cdf, sc.n_emitters, sx, emPdf);
    P2 uu; uu.x = sx; uu.y = u.y;
    V3 value = emitter_sample_direct(sc, (int) index, dRec, uu);
    if (dRec.pdf != 0) {
        dRec.emitter = (int) index;
        dRec.pdf *= emPdf;
        value = value / emPdf;
        return value;
    }
    return V3(0.0f);
}
// Scene::pdfEmitterDirect scene.cpp:975-978 * pdfEmitterDiscrete scene.h:852-854
MTSB_HD float pdf_emitter_direct(const SceneView &sc, const DRec &dRec) {
    return emitter_pdf_direct(sc, dRec.emitter, dRec) * sc.emitters[dRec.emitter].sampling_weight * sc.emitter_norm;
}

// path.cpp:385-389
MTSB_HD float mi_weight(float pdfA, float pdfB) { pdfA *= pdfA; pdfB *= pdfB; return pdfA / (pdfA + pdfB); }

// ---- film -----------------------------------------------------------------------------------
// Pixel footprint and separable weights of one sample inside a bordered buffer whose pixel
// (0,0) is film pixel (ox - border, oy - border) and whose size is sx x sy.
struct Footprint { int minx, miny, nx, ny; float wx[16], wy[16]; };
MTSB_HD float filter_discretized(const SceneView &sc, float x) {
    int i = (int) fabsf(x * sc.filter_scale);
    return sc.filter_table[i < 31 ? i : 31];
}
MTSB_HD void film_footprint(const SceneView &sc, float pos_x, float pos_y, int ox, int oy, int sx, int sy, Footprint &f) {
    const float px = pos_x - 0.5f - (float) (ox - sc.film_border), py = pos_y - 0.5f - (float) (oy - sc.film_border);
    int minx = (int) ceilf(px - sc.filter_radius), miny = (int) ceilf(py - sc.filter_radius);
    int maxx = (int) floorf(px + sc.filter_radius), maxy = (int) floorf(py + sc.filter_radius);
    if (minx < 0) minx = 0;
    if (miny < 0) miny = 0;
    if (maxx > sx - 1) maxx = sx - 1;
    if (maxy > sy - 1) maxy = sy - 1;
    f.minx = minx; f.miny = miny;
    f.nx = maxx - minx + 1; f.ny = maxy - miny + 1;
    if (f.nx > 16) f.nx = 16;
    if (f.ny > 16) f.ny = 16;
    for (int i = 0; i < f.nx; ++i) f.wx[i] = filter_discretized(sc, (float) (minx + i) - px);
    for (int i = 0; i < f.ny; ++i) f.wy[i] = filter_discretized(sc, (float) (miny + i) - py);
}

}  // namespace mtsb
