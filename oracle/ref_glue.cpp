// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// Thin C wrapper around code that is COMPILED FROM /root/reference (nothing is copied):
//   - GenericKDTree / SAHKDTree3D builder + Havran traversal  include/mitsuba/render/{gkdtree,sahkdtree3}.h
//   - TriAccel                                               include/mitsuba/render/triaccel.h
//   - Triangle::{getAABB,getClippedAABB}                     src/libcore/triangle.cpp
//   - MicrofacetDistribution                                 src/bsdfs/microfacet.h
//   - warp::*, math::{erf,erfinv,hypot2}                     src/libcore/{warp,math}.cpp
//   - AABB::rayIntersect                                     include/mitsuba/core/aabb.h:308-338
//   - fresnelDielectricExt / fresnelConductorExact / fresnelDiffuseReflectance / reflect / refract
//                                                            src/libcore/util.cpp:651-860 (+ quad.cpp)
//   - InterpolatedSpectrum + Spectrum::fromContinuousSpectrum (SPD file -> RGB)
//                                                            src/libcore/spectrum.cpp:172-191, 575-600
// The CRTP `RefTree` below supplies what ShapeKDTree supplies (skdtree.h:205-337): primitive
// count, boxes, clipped boxes and the two TriAccel `intersect` callbacks.
// Built by oracle/Makefile into oracle/_ref/libmtsref.so when /root/reference is present.
#include <mitsuba/mitsuba.h>
#include <mitsuba/core/aabb.h>
#include <mitsuba/core/triangle.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/core/timer.h>
#include <mitsuba/render/triaccel.h>
#include <mitsuba/render/sahkdtree3.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/core/util.h>
#include <mitsuba/core/random.h>
#include <boost/filesystem.hpp>
#include <mutex>
#include "src/bsdfs/microfacet.h"
#include <mitsuba/core/qmc.h>
#include <mitsuba/core/half.h>
#include "src/samplers/faure.h"
#include "src/samplers/sobolseq.h"
#include "../include/mtsb200.h"
#include <vector>

using namespace mitsuba;

class RefTree : public SAHKDTree3D<RefTree> {
    friend class GenericKDTree<AABB, SurfaceAreaHeuristic3, RefTree>;
    friend class SAHKDTree3D<RefTree>;
public:
    struct Cache { SizeType shapeIndex, primIndex; Float u, v; };   // == ShapeKDTree::IntersectionCache

    RefTree(const float *pos, size_t nv, const uint32_t *idx, size_t nt, const uint32_t *triShape, const uint32_t *shapeFirst)
        : m_triAccel(NULL) {
        m_positions.resize(nv);
        for (size_t i = 0; i < nv; ++i) m_positions[i] = Point(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        m_tris.resize(nt);
        for (size_t i = 0; i < nt; ++i) { m_tris[i].idx[0] = idx[3 * i]; m_tris[i].idx[1] = idx[3 * i + 1]; m_tris[i].idx[2] = idx[3 * i + 2]; }
        m_triShape.assign(triShape, triShape + nt);
        m_shapeFirst = shapeFirst;
    }
    ~RefTree() { if (m_triAccel) freeAligned(m_triAccel); }

    inline SizeType getPrimitiveCount() const { return (SizeType) m_tris.size(); }
    inline AABB getAABB(IndexType i) const { return m_tris[i].getAABB(&m_positions[0]); }
    inline AABB getClippedAABB(IndexType i, const AABB &box) const { return m_tris[i].getClippedAABB(&m_positions[0], box); }

    // skdtree.cpp:68-110
    void build(int logLevel) {
        setLogLevel((ELogLevel) logLevel);
        buildInternal();
        SizeType n = getPrimitiveCount();
        m_triAccel = static_cast<TriAccel *>(allocAligned(std::max((size_t) 1, (size_t) n) * sizeof(TriAccel)));
        for (IndexType i = 0; i < n; ++i) {
            const Triangle &t = m_tris[i];
            m_triAccel[i].load(m_positions[t.idx[0]], m_positions[t.idx[1]], m_positions[t.idx[2]]);
            m_triAccel[i].shapeIndex = m_triShape[i];
            m_triAccel[i].primIndex = i - m_shapeFirst[m_triShape[i]];
        }
    }

    // skdtree.h:248-304 (TriAccel branch)
    FINLINE bool intersect(const Ray &ray, IndexType idx, Float mint, Float maxt, Float &t, void *temp) const {
        Cache *cache = static_cast<Cache *>(temp);
        const TriAccel &ta = m_triAccel[idx];
        Float tempU, tempV, tempT;
        if (ta.rayIntersect(ray, mint, maxt, tempU, tempV, tempT)) {
            t = tempT;
            cache->shapeIndex = ta.shapeIndex; cache->primIndex = ta.primIndex;
            cache->u = tempU; cache->v = tempV;
            return true;
        }
        return false;
    }
    // skdtree.h:310-337
    FINLINE bool intersect(const Ray &ray, IndexType idx, Float mint, Float maxt) const {
        Float tempU, tempV, tempT;
        return m_triAccel[idx].rayIntersect(ray, mint, maxt, tempU, tempV, tempT);
    }

    bool closest(const Ray &ray, Float mint, Float maxt, Float &t, Cache &c) const {
        return rayIntersectHavran<false>(ray, mint, maxt, t, &c);
    }
    bool any(const Ray &ray, Float mint, Float maxt) const {
        Float t = std::numeric_limits<Float>::infinity();
        return rayIntersectHavran<true>(ray, mint, maxt, t, NULL);
    }
    const AABB &box() const { return m_aabb; }
    const AABB &tightBox() const { return m_tightAABB; }
    const TriAccel *triAccel() const { return m_triAccel; }
    const uint32_t *shapeFirst() const { return m_shapeFirst; }

    MTS_DECLARE_CLASS()
private:
    std::vector<Point> m_positions;
    std::vector<Triangle> m_tris;
    std::vector<uint32_t> m_triShape;
    const uint32_t *m_shapeFirst;
    TriAccel *m_triAccel;
};
MTS_IMPLEMENT_CLASS(RefTree, false, KDTreeBase)

struct RefHandle {
    RefTree *tree;
    std::vector<uint32_t> shapeFirst;
};

extern "C" {

void ref_static_init();   // ref_runtime.cpp

// Build the reference SAH kd-tree over the triangles of a scene description.
void *ref_kd_build(const MtsbSceneDesc *sd, int verbose) {
    ref_static_init();
    RefHandle *h = new RefHandle();
    std::vector<uint32_t> triShape(sd->n_tris);
    h->shapeFirst.resize(sd->n_shapes);
    for (uint32_t s = 0; s < sd->n_shapes; ++s) {
        h->shapeFirst[s] = sd->shapes[s].first_tri;
        for (uint32_t j = 0; j < sd->shapes[s].n_tris; ++j) triShape[sd->shapes[s].first_tri + j] = s;
    }
    h->tree = new RefTree(sd->positions, sd->n_vertices, sd->indices, sd->n_tris, triShape.data(), h->shapeFirst.data());
    h->tree->incRef();
    h->tree->build(verbose ? EInfo : ETrace);
    return h;
}
void ref_kd_destroy(void *p) { RefHandle *h = static_cast<RefHandle *>(p); h->tree->decRef(); delete h; }

void ref_kd_bounds(void *p, float *tight6, float *enlarged6) {
    RefTree *t = static_cast<RefHandle *>(p)->tree;
    for (int i = 0; i < 3; ++i) {
        tight6[i] = t->tightBox().min[i]; tight6[3 + i] = t->tightBox().max[i];
        enlarged6[i] = t->box().min[i]; enlarged6[3 + i] = t->box().max[i];
    }
}

// Candidate search on a pre-clipped interval: the signature of orc::ExternalTraceFn, so the
// restated oracle can run its path tracer on top of the reference's own kd-tree.
int ref_kd_search(void *p, const MtsbRay *r, float mint, float maxt, int any, MtsbHit *hit) {
    RefHandle *h = static_cast<RefHandle *>(p);
    Ray ray(Point(r->o[0], r->o[1], r->o[2]), Vector(r->d[0], r->d[1], r->d[2]), r->mint, r->maxt, 0.0f);
    if (any) return h->tree->any(ray, mint, maxt) ? 1 : 0;
    RefTree::Cache c; Float t = std::numeric_limits<Float>::infinity();
    if (!h->tree->closest(ray, mint, maxt, t, c)) return 0;
    hit->prim = h->shapeFirst[c.shapeIndex] + c.primIndex; hit->u = c.u; hit->v = c.v; hit->t = t;
    return 1;
}

// ShapeKDTree::rayIntersect(ray, its) / rayIntersect(ray) interval logic (skdtree.cpp:112-148, 213-232)
// written against the reference's own AABB::rayIntersect and kd-tree.
void ref_kd_trace(void *p, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits) {
    RefHandle *h = static_cast<RefHandle *>(p);
    for (size_t i = 0; i < n; ++i) {
        const MtsbRay &r = rays[i];
        MtsbHit &hit = hits[i];
        hit.prim = MTSB_NO_HIT; hit.u = hit.v = 0; hit.t = std::numeric_limits<Float>::infinity();
        Ray ray(Point(r.o[0], r.o[1], r.o[2]), Vector(r.d[0], r.d[1], r.d[2]), r.mint, r.maxt, 0.0f);
        Float mint, maxt;
        if (!h->tree->box().rayIntersect(ray, mint, maxt)) continue;
        Float rayMinT = ray.mint;
        if (rayMinT == Epsilon) {
            Float m = std::max(std::max(std::abs(ray.o.x), std::abs(ray.o.y)), std::abs(ray.o.z));
            if (mode == MTSB_TRACE_CLOSEST) m = std::max(m, Epsilon);
            rayMinT *= m;
        }
        if (rayMinT > mint) mint = rayMinT;
        if (ray.maxt < maxt) maxt = ray.maxt;
        if (!(maxt > mint)) continue;
        if (mode == MTSB_TRACE_ANY) {
            if (h->tree->any(ray, mint, maxt)) { hit.prim = 0; hit.t = 0; }
        } else {
            RefTree::Cache c; Float t = std::numeric_limits<Float>::infinity();
            if (h->tree->closest(ray, mint, maxt, t, c) && t != std::numeric_limits<Float>::infinity()) {
                hit.prim = h->shapeFirst[c.shapeIndex] + c.primIndex; hit.u = c.u; hit.v = c.v; hit.t = t;
            }
        }
    }
}

// Every triangle TriAccel accepts on the clipped interval, by brute force over all primitives
// (used to classify "degenerate" rays: SURVEY.md 8a row T).  Returns the number of accepted
// primitives; the first `cap` are written.
size_t ref_all_hits(void *p, const MtsbRay *r, float mint, float maxt, MtsbHit *out, size_t cap) {
    RefHandle *h = static_cast<RefHandle *>(p);
    Ray ray(Point(r->o[0], r->o[1], r->o[2]), Vector(r->d[0], r->d[1], r->d[2]), r->mint, r->maxt, 0.0f);
    size_t n = 0, cnt = h->tree->getPrimitiveCount();
    for (size_t i = 0; i < cnt; ++i) {
        Float u, v, t;
        if (h->tree->triAccel()[i].rayIntersect(ray, mint, maxt, u, v, t)) {
            if (n < cap) { out[n].prim = (uint32_t) i; out[n].u = u; out[n].v = v; out[n].t = t; }
            ++n;
        }
    }
    return n;
}

void ref_triaccel_load(const float *A, const float *B, const float *C, uint32_t *out12) {
    TriAccel t; memset(&t, 0, sizeof(t));
    t.load(Point(A[0], A[1], A[2]), Point(B[0], B[1], B[2]), Point(C[0], C[1], C[2]));
    memcpy(out12, &t, 48);
}
int ref_triaccel_intersect(const uint32_t *rec12, const MtsbRay *r, float mint, float maxt, float *uvt) {
    TriAccel t; memcpy(&t, rec12, 48);
    Ray ray(Point(r->o[0], r->o[1], r->o[2]), Vector(r->d[0], r->d[1], r->d[2]), r->mint, r->maxt, 0.0f);
    return t.rayIntersect(ray, mint, maxt, uvt[0], uvt[1], uvt[2]) ? 1 : 0;
}

// Triangle::getClippedAABB (src/libcore/triangle.cpp:119-146) -- the known-answer vectors of
// src/tests/test_kd.cpp:34-84 are checked against this in tests/.
int ref_clipped_aabb(const float *v9, const float *box6, float *out6) {
    Point pts[3] = { Point(v9[0], v9[1], v9[2]), Point(v9[3], v9[4], v9[5]), Point(v9[6], v9[7], v9[8]) };
    Triangle t; t.idx[0] = 0; t.idx[1] = 1; t.idx[2] = 2;
    AABB clip(Point(box6[0], box6[1], box6[2]), Point(box6[3], box6[4], box6[5]));
    AABB r = t.getClippedAABB(pts, clip);
    for (int i = 0; i < 3; ++i) { out6[i] = r.min[i]; out6[3 + i] = r.max[i]; }
    return r.isValid() ? 1 : 0;
}

float ref_erf(float x) { return math::erf(x); }
float ref_erfinv(float x) { return math::erfinv(x); }
float ref_hypot2(float a, float b) { return math::hypot2(a, b); }

// MicrofacetDistribution::{sample,eval,smithG1} (src/bsdfs/microfacet.h)
void ref_microfacet(int type, float au, float av, int visible, const float *wi3, const float *u2, float *out) {
    MicrofacetDistribution d((MicrofacetDistribution::EType) type, au, av, visible != 0);
    Float pdf; Vector wi(wi3[0], wi3[1], wi3[2]);
    Normal m = d.sample(wi, Point2(u2[0], u2[1]), pdf);
    out[0] = m.x; out[1] = m.y; out[2] = m.z; out[3] = pdf; out[4] = d.eval(m); out[5] = d.smithG1(wi, m);
}
// D(m), pdf(wi,m), G(wi,wo,m) for an arbitrary m
void ref_microfacet_eval(int type, float au, float av, int visible, const float *wi3, const float *wo3, const float *m3, float *out3) {
    MicrofacetDistribution d((MicrofacetDistribution::EType) type, au, av, visible != 0);
    Vector wi(wi3[0], wi3[1], wi3[2]), wo(wo3[0], wo3[1], wo3[2]), m(m3[0], m3[1], m3[2]);
    out3[0] = d.eval(m); out3[1] = d.pdf(wi, m); out3[2] = d.G(wi, wo, m);
}
void ref_warp(int kind, const float *u2, float *out3) {
    Point2 u(u2[0], u2[1]);
    if (kind == 0) { Vector v = warp::squareToCosineHemisphere(u); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
    else if (kind == 1) { Vector v = warp::squareToUniformSphere(u); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
    else if (kind == 2) { Point2 p = warp::squareToUniformTriangle(u); out3[0] = p.x; out3[1] = p.y; out3[2] = 0; }
    else { Point2 p = warp::squareToUniformDiskConcentric(u); out3[0] = p.x; out3[1] = p.y; out3[2] = 0; }
}

// ---- Fresnel terms (util.cpp) -------------------------------------------------------------
float ref_fresnel_dielectric_ext(float cosThetaI, float eta, float *cosThetaT) {
    Float t; Float F = fresnelDielectricExt(cosThetaI, t, eta); *cosThetaT = t; return F;
}
void ref_fresnel_conductor_exact(float cosThetaI, const float *eta3, const float *k3, float *out3) {
    Spectrum eta, k; eta.fromLinearRGB(eta3[0], eta3[1], eta3[2]); k.fromLinearRGB(k3[0], k3[1], k3[2]);
    Spectrum r = fresnelConductorExact(cosThetaI, eta, k);
    r.toLinearRGB(out3[0], out3[1], out3[2]);
}
float ref_fresnel_diffuse_reflectance(float eta, int fast) { return fresnelDiffuseReflectance(eta, fast != 0); }
void ref_refract(const float *wi3, const float *n3, float eta, float *out3) {
    Vector r = refract(Vector(wi3[0], wi3[1], wi3[2]), Normal(n3[0], n3[1], n3[2]), eta);
    out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}

// ---- SPD file -> RGB, the conversion behind `material=` names (roughconductor.cpp:174-185) ----
int ref_spd_to_rgb(const char *path, float *out3) {
    static std::once_flag once;
    std::call_once(once, []() { Spectrum::staticInitialization(); });
    try {
        InterpolatedSpectrum spd((fs::path(path)));
        Spectrum s; s.fromContinuousSpectrum(spd);
        s.toLinearRGB(out3[0], out3[1], out3[2]);
        return 0;
    } catch (const std::exception &e) { return 1; }
}

// ---- quasi-Monte Carlo building blocks (src/libcore/qmc.cpp, include/mitsuba/core/qmc.h, src/samplers/faure.cpp) ----
float ref_radical_inverse(int base, uint64_t i) { return radicalInverse(base, i); }
float ref_radical_inverse_incremental(int base, float x) { return radicalInverseIncremental(base, x); }
float ref_radical_inverse_fast(int baseIndex, uint64_t index) { return radicalInverseFast((uint16_t) baseIndex, index); }
float ref_scrambled_radical_inverse_fast(int baseIndex, uint64_t index, uint16_t *perm) { return scrambledRadicalInverseFast((uint16_t) baseIndex, index, perm); }
int ref_prime(int i) { return primeTable[i]; }
// PermutationStorage(scramble)->getPermutation(baseIndex); inverse != 0: getInversePermutation (bases 2 and 3 only)
int ref_permutation(int scramble, int baseIndex, int inverse, uint16_t *out) {
    static std::mutex m; static ref<PermutationStorage> cache;
    std::lock_guard<std::mutex> g(m);
    if (cache == NULL || cache->getScramble() != scramble) cache = new PermutationStorage(scramble);
    const uint16_t *p = inverse ? cache->getInversePermutation(baseIndex) : cache->getPermutation(baseIndex);
    for (int j = 0; j < primeTable[baseIndex]; ++j) out[j] = p[j];
    return primeTable[baseIndex];
}
void ref_sample02(uint32_t n, uint32_t *scramble2, float *out2) { Point2f p = sample02Single(n, scramble2); out2[0] = p.x; out2[1] = p.y; }
float ref_radical_inverse2(uint32_t n, uint32_t scramble) { return radicalInverse2Single(n, scramble); }
float ref_sobol2(uint32_t n, uint32_t scramble) { return sobol2Single(n, scramble); }
uint64_t ref_sample_tea(uint32_t v0, uint32_t v1, int rounds) { return sampleTEA(v0, v1, rounds); }

// ---- Sobol' sequence (src/samplers/sobolseq.h, tables in sobolseq.cpp) ----
float ref_sobol_sample(uint64_t index, uint32_t dimension, uint64_t scramble) { return sobol::sample(index, dimension, scramble); }
uint64_t ref_sobol_look_up(uint32_t m, uint32_t frame, uint32_t px, uint32_t py, uint64_t scramble) { return sobol::look_up(m, frame, px, py, scramble); }
uint32_t ref_sobol_matrix32(uint32_t dimension, uint32_t column) { return sobol::Matrices::matrices32[dimension * sobol::Matrices::size + column]; }
uint64_t ref_sobol_vdc(uint32_t m, uint32_t column, int inverse) { return inverse ? sobol::Matrices::vdc_sobol_matrices_inv[m - 1][column] : sobol::Matrices::vdc_sobol_matrices[m - 1][column]; }

// ---- half precision (src/libcore/half.cpp, include/mitsuba/core/half.h): storage type of the environment map ----
uint16_t ref_float_to_half(float f) { half h(f); return h.bits(); }
float ref_half_to_float(uint16_t bits) { half h; h.setBits(bits); return (float) h; }
void ref_square_to_tent(const float *u2, float *out2) { Point2 p = warp::squareToTent(Point2(u2[0], u2[1])); out2[0] = p.x; out2[1] = p.y; }

}  // extern "C"
