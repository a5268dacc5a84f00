// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of the reference `path` hot path (winmad/mitsuba, Mitsuba 0.5 fork):
//   MonteCarloIntegrator::renderBlock   src/librender/integrator.cpp:196-244
//   MIPathTracer::Li                    src/integrators/path/path.cpp:121-296
//   ShapeKDTree::rayIntersect           src/librender/skdtree.cpp:112-148, 213-232 (interval logic)
//   TriAccel                            include/mitsuba/render/triaccel.h:64-168
//   fillIntersectionRecord              include/mitsuba/render/skdtree.h:343-430
//   BSDFs / MicrofacetDistribution      src/bsdfs/{diffuse,roughconductor,roughdielectric,twosided}.cpp, microfacet.h
//   emitters / direct sampling          src/emitters/{area,point,constant}.cpp, src/librender/scene.cpp:828-855
//   sensor, film                        src/sensors/perspective.cpp:286-317, include/mitsuba/render/imageblock.h:124-206
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  Pinning: the pieces that the reference itself compiles stand-alone
// (TriAccel, kd-tree build + Havran traversal, MicrofacetDistribution, warp, erf/erfinv,
// Triangle::getClippedAABB) are compiled from /root/reference by oracle/Makefile into
// oracle/_ref/libmtsref.so and compared with this restatement in tests/ (golden vectors in
// tests/golden/ were generated from that library).  The random number stream is NOT the
// reference's SFMT19937: image parity is statistical.
#include "orc_scene.h"
#include "orc_sampler.h"
#include <cstdio>
#include <cstdlib>
#include <string>
#include <chrono>
#include <memory>

using namespace orc;

namespace {

thread_local std::string g_err;

// Two interchangeable random sources.
//  * sequential: splitmix64 / xoshiro256** stream with Random::nextFloat's 23-bit mantissa construction
//    (src/libcore/random.cpp:630-640) -- the reference's model: one stream per worker (renderjob.cpp:59-70).
//  * counter: Philox4x32-10 (Salmon et al., SC'11; published constants) keyed by (seed, pixel<<24 | sample)
//    and numbered by (bounce, slot) -- the same numbering the CUDA wavefront uses, restated here
//    independently, so that a CPU render can be compared with the GPU film sample for sample.
struct Rng : orc::SeqRandom {
    orc::QmcSampler *qmc = nullptr;      // halton / ldsampler: draws go through the stateful sampler in call order
    uint64_t s[4];
    bool counter; uint32_t k0, k1, idLo, idHi;
    static uint64_t splitmix(uint64_t &x) {
        uint64_t z = (x += 0x9e3779b97f4a7c15ULL);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
        return z ^ (z >> 31);
    }
    explicit Rng(uint64_t seed) : counter(false), k0(0), k1(0), idLo(0), idHi(0) { for (int i = 0; i < 4; ++i) s[i] = splitmix(seed); }
    Rng(uint64_t seed, uint64_t sampleId) : counter(true), k0((uint32_t) seed), k1((uint32_t) (seed >> 32)),
        idLo((uint32_t) sampleId), idHi((uint32_t) (sampleId >> 32)) { s[0] = s[1] = s[2] = s[3] = 0; }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t nextULong() override {
        uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    static float toFloat(uint32_t v) { union { uint32_t u; float f; } x; x.u = (v >> 9) | 0x3f800000u; return x.f - 1.0f; }
    float nextFloat() override { return toFloat((uint32_t) (nextULong() & 0xFFFFFFFFull)); }
    P2 next2D() { float a = nextFloat(); float b = nextFloat(); return P2{a, b}; }
    void block(uint32_t blk, uint32_t sub, float out[4]) const {
        uint32_t c[4] = {idLo, idHi, blk, sub}, ka = k0, kb = k1;
        for (int r = 0; r < 10; ++r) {
            uint64_t p0 = (uint64_t) 0xD2511F53u * c[0], p1 = (uint64_t) 0xCD9E8D57u * c[2];
            uint32_t n0 = (uint32_t) (p1 >> 32) ^ c[1] ^ ka, n1 = (uint32_t) p1, n2 = (uint32_t) (p0 >> 32) ^ c[3] ^ kb, n3 = (uint32_t) p0;
            c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
            ka += 0x9E3779B9u; kb += 0xBB67AE85u;
        }
        for (int i = 0; i < 4; ++i) out[i] = toFloat(c[i]);
    }
    // the four places where the integrator draws numbers
    P2 q2() { float o[2]; qmc->next2D(o); return P2{o[0], o[1]}; }
    P2 pixel2D() { if (qmc) return q2(); if (!counter) return next2D(); float u[4]; block(0, 0, u); return P2{u[0], u[1]}; }                 // integrator.cpp:227
    P2 emitter2D(int depth) { if (qmc) return q2(); if (!counter) return next2D(); float u[4]; block((uint32_t) depth, 0, u); return P2{u[0], u[1]}; }   // path.cpp:178
    P2 bsdf2D(int depth) { if (qmc) return q2(); if (!counter) return next2D(); float u[4]; block((uint32_t) depth, 0, u); return P2{u[2], u[3]}; }      // path.cpp:211
    float extra1D(int depth) { if (qmc) return qmc->next1D(); if (!counter) return nextFloat(); float u[4]; block((uint32_t) depth, 1, u); return u[0]; }         // roughdielectric.cpp:554
    float rr1D(int depth) { if (qmc) return qmc->next1D(); if (!counter) return nextFloat(); float u[4]; block((uint32_t) depth, 1, u); return u[1]; }            // path.cpp:285
};

struct Counters { uint64_t closest = 0, shadow = 0, shade = 0, pathLen = 0, samples = 0; double lum = 0, lum2 = 0; };

inline float miWeight(float pdfA, float pdfB) { pdfA *= pdfA; pdfB *= pdfB; return pdfA / (pdfA + pdfB); }   // path.cpp:385-389

inline bool traceClosest(const Scene &sc, const MtsbRay &ray, Its &its, Counters &c) {
    MtsbHit hit;
    ++c.closest;
    sceneIntersect(sc.accel, ray, false, hit, sc.ext, sc.extUser);
    fillIntersection(sc, ray, hit, its);
    return its.valid;
}

inline MtsbRay makeRay(V3 o, V3 d, float mint, float maxt) {
    MtsbRay r; r.o[0] = o.x; r.o[1] = o.y; r.o[2] = o.z; r.d[0] = d.x; r.d[1] = d.y; r.d[2] = d.z; r.mint = mint; r.maxt = maxt; return r;
}

// MIPathTracer::Li, path.cpp:121-296.  rRec.type starts as ERadiance (all bits), medium-free.
V3 Li(const Scene &sc, const MtsbRay &primary, Rng &rng, Counters &cnt, float &alpha, const V3 *rxd = nullptr, const V3 *ryd = nullptr) {
    const MtsbIntegrator &ig = sc.sd.integrator;
    const int maxDepth = ig.max_depth, rrDepth = ig.rr_depth;
    const bool strictNormals = ig.strict_normals != 0, hideEmitters = ig.hide_emitters != 0;
    Its its;
    MtsbRay ray = primary;
    V3 L(0.0f);
    bool scattered = false;
    bool emittedRadiance = true;         // rRec.type & EEmittedRadiance
    int depth = 1;                       // rRec.depth (newQuery: integrator.h)
    traceClosest(sc, ray, its, cnt);     // rRec.rayIntersect(ray)
    alpha = its.valid ? 1.0f : 0.0f;     // records.inl:136-152 (no media)
    ray.mint = kEpsilon;
    V3 throughput(1.0f);
    float eta = 1.0f;

    while (depth <= maxDepth || maxDepth < 0) {
        if (!its.valid) {
            if (emittedRadiance && (!hideEmitters || scattered) && sc.envEmitter >= 0) {
                const V3 dir(ray.d[0], ray.d[1], ray.d[2]);
                // the camera ray still carries its differentials here (path.cpp:141-143; they are dropped at :223)
                L += throughput * ((depth == 1 && !scattered && rxd) ? evalEnvironmentFiltered(sc, dir, *rxd, *ryd) : evalEnvironment(sc, dir));
            }
            break;
        }
        const MtsbShape &shape = sc.shapes[its.shape];
        const int matId = shape.material;
        V3 rd(ray.d[0], ray.d[1], ray.d[2]);
        // its.getBSDF(ray) (records.inl:82-90): uv partials only for the camera ray, the only one that carries differentials
        if (depth == 1 && !scattered && rxd && matId >= 0 && sc.shapeNeedsPartials[its.shape]) {
            const UVPartials pt = computePartials(its.p, its.baseN, its.dpdu, its.dpdv, V3(primary.o[0], primary.o[1], primary.o[2]), *rxd, *ryd);
            its.hasUVPartials = true; its.dudx = pt.dudx; its.dvdx = pt.dvdx; its.dudy = pt.dudy; its.dvdy = pt.dvdy;
        }
        MtsbMaterial lm[2];
        const int mat = matId >= 0 ? resolveMaterials(sc, matId, its, lm) : -1;
        const MtsbMaterial *mats = lm;
        if (shape.emitter >= 0 && emittedRadiance && (!hideEmitters || scattered))
            L += throughput * areaEval(sc.emitters[shape.emitter], its.sh.n, -rd);     // its.Le(-ray.d)

        if ((depth >= maxDepth && maxDepth > 0) || (strictNormals && dot(rd, its.geoN) * cosTheta(its.wi) >= 0))
            break;
        ++cnt.shade;

        // ---- direct illumination (path.cpp:174-202) ----
        DRec dRec;
        dRec.ref = its.p; dRec.refN = V3(0.0f);
        const bool haveMat = mat >= 0;
        if (haveMat && !bsdf_has_back_or_transmission(mats[mat])) dRec.refN = its.sh.n;   // records.inl:175-179
        dRec.emitter = -1;
        if (haveMat && bsdf_is_smooth(mats, mats[mat])) {
            V3 value = sampleEmitterDirectNoVis(sc, dRec, rng.emitter2D(depth));
            if (dRec.pdf != 0) {
                // scene.cpp:840-846
                MtsbRay sray = makeRay(dRec.ref, dRec.d, kEpsilon, dRec.dist * (1 - kShadowEpsilon));
                MtsbHit sh; ++cnt.shadow;
                if (sceneIntersect(sc.accel, sray, true, sh, sc.ext, sc.extUser)) value = V3(0.0f);
            }
            if (!isZero(value)) {
                V3 wo = its.toLocal(dRec.d);
                const V3 bsdfVal = bsdf_eval(mats, mat, its.wi, wo);
                if (!isZero(bsdfVal) && (!strictNormals || dot(its.geoN, dRec.d) * cosTheta(wo) > 0)) {
                    const MtsbEmitter &em = sc.emitters[dRec.emitter];
                    bool onSurface = em.type != MTSB_EMITTER_POINT;
                    float bsdfPdf = (onSurface && dRec.measure == kSolidAngle) ? bsdf_pdf(mats, mat, its.wi, wo) : 0;
                    float weight = miWeight(dRec.pdf, bsdfPdf);
                    L += throughput * value * bsdfVal * weight;
                }
            }
        }

        // ---- BSDF sampling (path.cpp:208-221) ----
        if (!haveMat) break;
        BsdfSample bs;
        P2 u2 = rng.bsdf2D(depth);
        float u1 = 0.0f;
        {   // roughdielectric draws one more number inside sample() (roughdielectric.cpp:554)
            const MtsbMaterial *m = &mats[mat];
            if ((m->flags & MTSB_MAT_TWOSIDED) && cosTheta(its.wi) < 0 && m->back >= 0) m = &mats[m->back];
            if (m->type == MTSB_BSDF_ROUGHDIELECTRIC) u1 = rng.extra1D(depth);
        }
        bool ok = bsdf_sample(mats, mat, its.wi, u2, u1, bs);
        if (!ok || isZero(bs.weight)) break;
        scattered = true;
        const V3 wo = its.toWorld(bs.wo);
        float woDotGeoN = dot(its.geoN, wo);
        if (strictNormals && woDotGeoN * cosTheta(bs.wo) <= 0) break;

        bool hitEmitter = false;
        V3 value(0.0f);
        ray = makeRay(its.p, wo, kEpsilon, kInf);
        if (traceClosest(sc, ray, its, cnt)) {
            const MtsbShape &hs = sc.shapes[its.shape];
            if (hs.emitter >= 0) {
                value = areaEval(sc.emitters[hs.emitter], its.sh.n, -wo);
                // dRec.setQuery(ray, its) records.inl:185-193
                dRec.p = its.p; dRec.n = its.sh.n; dRec.measure = kSolidAngle; dRec.emitter = hs.emitter;
                dRec.d = wo; dRec.dist = its.t;
                hitEmitter = true;
            }
        } else {
            if (sc.envEmitter >= 0) {
                if (hideEmitters && !scattered) break;
                value = evalEnvironment(sc, wo);                                  // path.cpp:240
                // fillDirectSamplingRecord constant.cpp:246-260
                float nearT, farT;
                if (!bsphereIntersect(sc.bsphereC, sc.bsphereR, V3(ray.o[0], ray.o[1], ray.o[2]), wo, nearT, farT) || nearT > 0 || farT < 0) break;
                dRec.p = its.p;   // not used afterwards by `path`
                dRec.n = V3(0.0f); dRec.measure = kSolidAngle; dRec.emitter = sc.envEmitter; dRec.d = wo; dRec.dist = farT;
                hitEmitter = true;
            } else break;
        }

        throughput *= bs.weight;
        eta *= bs.eta;

        if (hitEmitter) {
            const float lumPdf = !bs.delta ? pdfEmitterDirect(sc, dRec) : 0;
            L += throughput * value * miWeight(bs.pdf, lumPdf);
        }
        if (!its.valid) break;
        emittedRadiance = false;          // rRec.type = ERadianceNoEmission

        const int d0 = depth++;
        if (d0 >= rrDepth) {
            float q = std::min(maxComp(throughput) * eta * eta, 0.95f);
            if (rng.rr1D(d0) >= q) break;
            throughput = throughput / q;
        }
    }
    cnt.pathLen += (uint64_t) depth;
    return L;
}

Scene *S(void *h) { return static_cast<Scene *>(h); }

}  // namespace

extern "C" {

const char *orc_last_error(void) { return g_err.c_str(); }

int orc_scene_create(const MtsbSceneDesc *d, void **out) {
    if (!d || !out || d->abi_version != MTSB_ABI_VERSION) { g_err = "bad scene description"; return MTSB_ERR_INVALID; }
    Scene *sc = new Scene();
    sc->shapes.assign(d->shapes, d->shapes + d->n_shapes);
    sc->materials.assign(d->materials, d->materials + d->n_materials);
    sc->emitters.assign(d->emitters, d->emitters + d->n_emitters);
    sc->positions.assign(d->positions, d->positions + 3 * d->n_vertices);
    if (d->normals) sc->normals.assign(d->normals, d->normals + 3 * d->n_vertices);
    else sc->normals.assign(3 * d->n_vertices, 0.0f);
    if (d->texcoords) sc->texcoords.assign(d->texcoords, d->texcoords + 2 * d->n_vertices);
    sc->indices.assign(d->indices, d->indices + 3 * d->n_tris);
    sc->sd = *d;
    sc->sd.shapes = sc->shapes.data(); sc->sd.materials = sc->materials.data(); sc->sd.emitters = sc->emitters.data();
    sc->sd.positions = sc->positions.data(); sc->sd.normals = sc->normals.data();
    sc->sd.texcoords = sc->texcoords.empty() ? nullptr : sc->texcoords.data(); sc->sd.indices = sc->indices.data();
    sc->textures.resize(d->n_textures);
    for (uint32_t i = 0; i < d->n_textures; ++i) sc->textures[i].build(d->textures[i]);
    sc->sd.textures = nullptr;
    {   // BSDF::usesRayDifferentials(): true when one of its textures is a bitmap (bitmap.cpp:547-549; twosided.cpp:98-99)
        auto bitmapTex = [&](const MtsbMaterial &m) {
            return (m.tex_reflectance && sc->textures[m.tex_reflectance - 1].rec.type == MTSB_TEX_BITMAP) ||
                   (m.tex_transmittance && sc->textures[m.tex_transmittance - 1].rec.type == MTSB_TEX_BITMAP) ||
                   (m.tex_alpha && sc->textures[m.tex_alpha - 1].rec.type == MTSB_TEX_BITMAP); };
        sc->shapeNeedsPartials.assign(d->n_shapes, false);
        for (uint32_t s = 0; s < d->n_shapes; ++s) {
            const int m = sc->shapes[s].material;
            if (m < 0) continue;
            const MtsbMaterial &mm = sc->materials[m];
            sc->shapeNeedsPartials[s] = bitmapTex(mm) || ((mm.flags & MTSB_MAT_TWOSIDED) && mm.back >= 0 && bitmapTex(sc->materials[mm.back]));
        }
    }
    sc->triShape.resize(d->n_tris);
    sc->areaPdf.resize(d->n_shapes); sc->invArea.assign(d->n_shapes, 0.0f);
    for (uint32_t s = 0; s < d->n_shapes; ++s) {
        const MtsbShape &sh = sc->shapes[s];
        for (uint32_t j = 0; j < sh.n_tris; ++j) sc->triShape[sh.first_tri + j] = s;
        if (sh.emitter >= 0) {
            // TriMesh::prepareSamplingTable trimesh.cpp:388-403; Triangle::surfaceArea triangle.cpp:64-70
            Pmf &pm = sc->areaPdf[s];
            for (uint32_t j = 0; j < sh.n_tris; ++j) {
                const uint32_t *ix = &sc->indices[3 * ((size_t) sh.first_tri + j)];
                V3 a = sc->pos(ix[1]) - sc->pos(ix[0]), b = sc->pos(ix[2]) - sc->pos(ix[0]);
                pm.append(0.5f * length(cross(a, b)));
            }
            float area = pm.normalize();
            sc->invArea[s] = 1.0f / area;
        }
    }
    sc->accel.build(sc->sd);
    for (size_t i = 0; i < sc->emitters.size(); ++i) {
        sc->emitterPdf.append(sc->emitters[i].sampling_weight);
        if (sc->emitters[i].type == MTSB_EMITTER_CONSTANT || sc->emitters[i].type == MTSB_EMITTER_ENVMAP) sc->envEmitter = (int) i;
        if (sc->emitters[i].type == MTSB_EMITTER_ENVMAP) {
            if (!d->envmap || !d->envmap->rgb) { g_err = "envmap emitter without image data"; delete sc; return MTSB_ERR_INVALID; }
            try { sc->envmap.build(*d->envmap); } catch (const std::exception &e) { g_err = e.what(); delete sc; return MTSB_ERR_INVALID; }
        }
    }
    if (!sc->emitters.empty()) sc->emitterPdf.normalize();
    // Scene::m_aabb: kd-tree box (+ sensor position, + emitter boxes) scene.cpp:388-414;
    // AABB::getBSphere: centre = box centre, radius = |centre - max|; constant.cpp:67-69 scales by 1.5
    {
        Box bx = sc->accel.enlarged;
        if (d->n_tris == 0) { bx.lo = V3(0.0f); bx.hi = V3(0.0f); }
        bx.grow(xformPoint(d->sensor.to_world, V3(0.0f)));
        for (size_t i = 0; i < sc->emitters.size(); ++i)
            if (sc->emitters[i].type == MTSB_EMITTER_POINT) bx.grow(rgb(sc->emitters[i].position));
        sc->bsphereC = (bx.hi + bx.lo) * 0.5f;
        sc->bsphereR = std::max(kEpsilon, length(sc->bsphereC - bx.hi) * 1.5f);
    }
    const MtsbSensor &se = d->sensor;
    double aspect = (double) se.film_width / (double) se.film_height;
    double tanHalf = std::tan(0.5 * (double) se.fov_x_deg * 3.14159265358979323846 / 180.0);
    sc->nearX = (float) ((double) se.near_clip * tanHalf);
    sc->nearY = (float) ((double) se.near_clip * tanHalf / aspect);
    filmInit(sc->film, se);
    *out = sc;
    return MTSB_OK;
}

void orc_scene_destroy(void *h) { delete S(h); }

// Plug an external candidate search (the reference kd-tree from oracle/_ref) under the
// ShapeKDTree::rayIntersect interval logic.
void orc_scene_set_external_accel(void *h, ExternalTraceFn fn, void *user) { S(h)->ext = fn; S(h)->extUser = user; }

int orc_trace(void *h, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits, int nthreads) {
    Scene *sc = S(h);
    if (nthreads <= 0) nthreads = (int) std::thread::hardware_concurrency();
    std::atomic<size_t> next(0);
    auto work = [&]() {
        for (;;) {
            size_t b = next.fetch_add(4096);
            if (b >= n) break;
            size_t e = std::min(n, b + 4096);
            for (size_t i = b; i < e; ++i) sceneIntersect(sc->accel, rays[i], mode == MTSB_TRACE_ANY, hits[i], sc->ext, sc->extUser);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; ++t) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
    return MTSB_OK;
}

int orc_bsdf_query(void *h, int material, const MtsbBsdfQuery *q, size_t n, MtsbBsdfResult *out) {
    Scene *sc = S(h);
    if (material < 0 || (size_t) material >= sc->materials.size()) { g_err = "bad material"; return MTSB_ERR_INVALID; }
    for (size_t i = 0; i < n; ++i) {
        V3 wi = rgb(q[i].wi), wo = rgb(q[i].wo);
        V3 e = bsdf_eval(sc->materials.data(), material, wi, wo);
        MtsbBsdfResult &r = out[i];
        r.eval[0] = e.x; r.eval[1] = e.y; r.eval[2] = e.z;
        r.pdf = bsdf_pdf(sc->materials.data(), material, wi, wo);
        BsdfSample bs; bs.wo = V3(0.0f); bs.weight = V3(0.0f); bs.pdf = 0; bs.eta = 1;
        bool ok = bsdf_sample(sc->materials.data(), material, wi, P2{q[i].u[0], q[i].u[1]}, q[i].u[2], bs);
        if (!ok) { bs.weight = V3(0.0f); bs.pdf = 0; bs.wo = V3(0.0f); bs.eta = 1; }
        r.weight[0] = bs.weight.x; r.weight[1] = bs.weight.y; r.weight[2] = bs.weight.z; r.spdf = bs.pdf;
        r.swo[0] = bs.wo.x; r.swo[1] = bs.wo.y; r.swo[2] = bs.wo.z; r.seta = bs.eta;
    }
    return MTSB_OK;
}

int orc_emitter_sample(void *h, const MtsbDirectQuery *q, size_t n, MtsbDirectResult *out) {
    Scene *sc = S(h);
    for (size_t i = 0; i < n; ++i) {
        DRec d; d.ref = rgb(q[i].ref); d.refN = rgb(q[i].ref_n);
        d.p = d.n = d.d = V3(0.0f); d.dist = 0; d.measure = 0;
        V3 v = sampleEmitterDirectNoVis(*sc, d, P2{q[i].u[0], q[i].u[1]});
        MtsbDirectResult &r = out[i];
        if (d.emitter < 0) { std::memset(&r, 0, sizeof(r)); r.emitter = -1; continue; }
        r.value[0] = v.x; r.value[1] = v.y; r.value[2] = v.z; r.pdf = d.pdf;
        r.d[0] = d.d.x; r.d[1] = d.d.y; r.d[2] = d.d.z; r.dist = d.dist;
        r.p[0] = d.p.x; r.p[1] = d.p.y; r.p[2] = d.p.z; r.emitter = d.emitter;
        r.n[0] = d.n.x; r.n[1] = d.n.y; r.n[2] = d.n.z; r.measure = d.measure;
    }
    return MTSB_OK;
}

int orc_fill_intersection(void *h, const MtsbRay *rays, const MtsbHit *hits, size_t n, float *out17) {
    Scene *sc = S(h);
    for (size_t i = 0; i < n; ++i) {
        Its its; float *o = out17 + 17 * i;
        fillIntersection(*sc, rays[i], hits[i], its);
        if (!its.valid) { for (int k = 0; k < 17; ++k) o[k] = 0; continue; }
        V3 v[5] = {its.p, its.geoN, its.sh.n, its.sh.s, its.sh.t};
        for (int k = 0; k < 5; ++k) { o[3 * k] = v[k].x; o[3 * k + 1] = v[k].y; o[3 * k + 2] = v[k].z; }
        o[15] = its.uv.x; o[16] = its.uv.y;
    }
    return MTSB_OK;
}

int orc_sensor_rays(void *h, const float *pos, size_t n, MtsbRay *rays) {
    for (size_t i = 0; i < n; ++i) sensorRay(*S(h), P2{pos[2 * i], pos[2 * i + 1]}, rays[i]);
    return MTSB_OK;
}

int orc_film_clear(void *h) { Film &f = S(h)->film; std::fill(f.data.begin(), f.data.end(), 0.0f); f.bad = 0; return MTSB_OK; }
int orc_film_size(void *h, int *w, int *hh, int *b) { Film &f = S(h)->film; *w = f.w; *hh = f.h; *b = f.border; return MTSB_OK; }
int orc_film_put(void *h, const float *s6, size_t n) {
    Film &f = S(h)->film;
    for (size_t i = 0; i < n; ++i) {
        const float *s = s6 + 6 * i;
        float v[5] = {s[2], s[3], s[4], s[5], 1.0f};
        if (!blockPut(f, f.data.data(), 0, 0, f.w, f.h, P2{s[0], s[1]}, v)) ++f.bad;
    }
    return MTSB_OK;
}
int orc_film_download(void *h, float *out) { Film &f = S(h)->film; std::memcpy(out, f.data.data(), f.data.size() * sizeof(float)); return MTSB_OK; }
// HDRFilm::develop -> Bitmap::convert from ESpectrumAlphaWeight: divide by the weight channel (bitmap.cpp:422-426)
int orc_film_develop(void *h, float *rgbOut) {
    Film &f = S(h)->film;
    for (int y = 0; y < f.h; ++y)
        for (int x = 0; x < f.w; ++x) {
            const float *p = &f.data[((size_t) (y + f.border) * f.stride() + x + f.border) * 5];
            float w = p[4], inv = w != 0 ? 1.0f / w : 0.0f;
            float *o = rgbOut + ((size_t) y * f.w + x) * 3;
            o[0] = p[0] * inv; o[1] = p[1] * inv; o[2] = p[2] * inv;
        }
    return MTSB_OK;
}

typedef struct OrcStats {
    uint64_t samples, closest_rays, shadow_rays, shade_events, path_length_sum, bad_samples;
    double lum_sum, lum_sq_sum;      // per-sample luminance moments (for the 3-sigma mean test)
    double seconds;                  // wall time of the render loop (kd/BVH build excluded)
    int threads;
} OrcStats;

// SamplingIntegrator::render + BlockRenderer: 32x32 blocks (scene.cpp:24), spiral order
// (imageproc.cpp:43-78), each block rendered into a private bordered buffer and merged
// under a mutex (renderproc.cpp:149-156).  Pixel order inside a block is row-major instead
// of Hilbert (order is irrelevant for an independent sampler).
int orc_render2(void *h, int spp, uint64_t seed, int nthreads, int max_blocks, OrcStats *stats, int counterRng, int sppOffset);
int orc_render(void *h, int spp, uint64_t seed, int nthreads, int max_blocks, OrcStats *stats) {
    return orc_render2(h, spp, seed, nthreads, max_blocks, stats, 0, 0);
}
// counterRng != 0: every sample uses the counter-based stream keyed by (seed, pixel << 24 | sppOffset + j)
int orc_render2(void *h, int spp, uint64_t seed, int nthreads, int max_blocks, OrcStats *stats, int counterRng, int sppOffset) {
    Scene *sc = S(h);
    Film &f = sc->film;
    if (spp <= 0) spp = sc->sd.integrator.sample_count;
    if (nthreads <= 0) nthreads = (int) std::thread::hardware_concurrency();
    const int BS = 32;
    int nbx = (f.w + BS - 1) / BS, nby = (f.h + BS - 1) / BS, total = nbx * nby;
    // spiral order
    std::vector<std::pair<int, int>> blocks;
    {
        int cx = nbx / 2, cy = nby / 2, dir = 0, numSteps = 1, stepsLeft = 1;   // imageproc.cpp:28-41
        blocks.emplace_back(cx, cy);
        while ((int) blocks.size() < total) {
            do {
                switch (dir) { case 0: ++cx; break; case 1: ++cy; break; case 2: --cx; break; case 3: --cy; break; }
                if (--stepsLeft == 0) { dir = (dir + 1) % 4; if (dir == 2 || dir == 0) ++numSteps; stepsLeft = numSteps; }
            } while (cx < 0 || cy < 0 || cx >= nbx || cy >= nby);
            blocks.emplace_back(cx, cy);
        }
    }
    if (max_blocks > 0 && max_blocks < total) {
        // bounded sample of the workload: every (total/max_blocks)-th block of the spiral
        std::vector<std::pair<int, int>> sub;
        for (int i = 0; i < max_blocks; ++i) sub.push_back(blocks[(size_t) i * total / max_blocks]);
        blocks.swap(sub);
    }
    // sample generator of the scene (src/samplers): halton / ldsampler objects are per worker, as in renderjob.cpp:59-70
    const MtsbSampler &smp = sc->sd.sampler;
    std::unique_ptr<orc::Permutations> perms;
    if (smp.type == MTSB_SAMPLER_HALTON && smp.scramble != 0) perms.reset(new orc::Permutations(smp.scramble));   // halton.cpp:139-149
    if (smp.type == MTSB_SAMPLER_LDSAMPLER && spp + sppOffset > (int) orc::LowDiscrepancySampler(std::max(1, sc->sd.integrator.sample_count), 0, nullptr).sampleCount) {
        g_err = "ldsampler: sample index exceeds sampleCount"; return MTSB_ERR_INVALID;
    }
    std::atomic<size_t> next(0);
    std::mutex filmMutex;
    std::vector<Counters> counters(nthreads);
    std::atomic<uint64_t> bad(0);
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](int tid) {
        Counters &cnt = counters[tid];
        std::vector<float> buf((size_t) (BS + 2 * f.border) * (BS + 2 * f.border) * 5);
        for (;;) {
            size_t bi = next.fetch_add(1);
            if (bi >= blocks.size()) break;
            int ox = blocks[bi].first * BS, oy = blocks[bi].second * BS;
            int bw = std::min(BS, f.w - ox), bh = std::min(BS, f.h - oy);
            std::fill(buf.begin(), buf.end(), 0.0f);
            Rng rng(seed * 0x100000001b3ULL + (uint64_t) (blocks[bi].second * nbx + blocks[bi].first) + 0x51ed27);
            std::unique_ptr<orc::QmcSampler> qmc;
            if (smp.type == MTSB_SAMPLER_HALTON) {
                orc::HaltonSampler *hs = new orc::HaltonSampler(perms.get());
                hs->setFilmResolution(sc->sd.sensor.crop_width, sc->sd.sensor.crop_height, true);       // integrator.cpp:38-42
                qmc.reset(hs);
            } else if (smp.type == MTSB_SAMPLER_SOBOL) {
                orc::SobolSampler *ss = new orc::SobolSampler((uint64_t) (uint32_t) smp.scramble);
                ss->setFilmResolution(sc->sd.sensor.crop_width, sc->sd.sensor.crop_height, true);
                qmc.reset(ss);
            } else if (smp.type == MTSB_SAMPLER_LDSAMPLER) {
                qmc.reset(new orc::LowDiscrepancySampler((size_t) std::max(1, sc->sd.integrator.sample_count), (size_t) std::max(0, smp.dimension), &rng));
            }
            rng.qmc = qmc.get();
            for (int y = 0; y < bh; ++y)
                for (int x = 0; x < bw; ++x) {
                    if (qmc) qmc->generate(ox + x + sc->sd.sensor.crop_x, oy + y + sc->sd.sensor.crop_y);       // integrator.cpp:222
                    for (int j = 0; j < spp; ++j) {
                        if (qmc) qmc->setSampleIndex((size_t) (sppOffset + j));                                // advance(), integrator.cpp:240
                        else if (counterRng) {
                            uint64_t pixel = (uint64_t) (oy + y + sc->sd.sensor.crop_y) * (uint64_t) sc->sd.sensor.film_width + (uint64_t) (ox + x + sc->sd.sensor.crop_x);
                            rng = Rng(seed, (pixel << 24) | (uint64_t) (sppOffset + j));
                        }
                        P2 u = rng.pixel2D();
                        P2 samplePos{(float) (ox + x) + u.x, (float) (oy + y) + u.y};
                        MtsbRay ray; sensorRay(*sc, samplePos, ray);
                        V3 rxd, ryd; sensorRayDifferentials(*sc, samplePos, rxd, ryd);
                        float alpha;
                        V3 L = Li(*sc, ray, rng, cnt, alpha, &rxd, &ryd);
                        ++cnt.samples;
                        float v[5] = {L.x, L.y, L.z, alpha, 1.0f};
                        if (!blockPut(f, buf.data(), ox, oy, bw, bh, samplePos, v)) bad++;
                        else { double l = luminance(L); cnt.lum += l; cnt.lum2 += l * l; }
                    }
                }
            // film->put(block): overlap-add incl. borders (imageblock.h:103-107)
            std::lock_guard<std::mutex> g(filmMutex);
            int bsx = bw + 2 * f.border;
            for (int y = 0; y < bh + 2 * f.border; ++y) {
                float *dst = &f.data[((size_t) (oy + y) * f.stride() + ox) * 5];
                const float *src = &buf[(size_t) y * bsx * 5];
                for (int k = 0; k < bsx * 5; ++k) dst[k] += src[k];
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto &t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    f.bad += bad.load();
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        for (auto &c : counters) {
            stats->samples += c.samples; stats->closest_rays += c.closest; stats->shadow_rays += c.shadow;
            stats->shade_events += c.shade; stats->path_length_sum += c.pathLen; stats->lum_sum += c.lum; stats->lum_sq_sum += c.lum2;
        }
        stats->bad_samples = f.bad;
        stats->seconds = std::chrono::duration<double>(t1 - t0).count();
        stats->threads = nthreads;
    }
    return MTSB_OK;
}

// Radiance of individual camera samples with a caller-supplied seed (unit tests).
int orc_li(void *h, const float *samplePos, size_t n, uint64_t seed, float *rgba) {
    Scene *sc = S(h);
    Counters cnt;
    for (size_t i = 0; i < n; ++i) {
        Rng rng(seed + i);
        MtsbRay ray; sensorRay(*sc, P2{samplePos[2 * i], samplePos[2 * i + 1]}, ray);
        V3 rxd, ryd; sensorRayDifferentials(*sc, P2{samplePos[2 * i], samplePos[2 * i + 1]}, rxd, ryd);
        float alpha; V3 L = Li(*sc, ray, rng, cnt, alpha, &rxd, &ryd);
        rgba[4 * i] = L.x; rgba[4 * i + 1] = L.y; rgba[4 * i + 2] = L.z; rgba[4 * i + 3] = alpha;
    }
    return MTSB_OK;
}

void orc_scene_bounds(void *h, float *tightLoHi6, float *enlargedLoHi6, float *bsphere4) {
    Scene *sc = S(h);
    for (int i = 0; i < 3; ++i) {
        tightLoHi6[i] = sc->accel.tight.lo[i]; tightLoHi6[3 + i] = sc->accel.tight.hi[i];
        enlargedLoHi6[i] = sc->accel.enlarged.lo[i]; enlargedLoHi6[3 + i] = sc->accel.enlarged.hi[i];
        bsphere4[i] = sc->bsphereC[i];
    }
    bsphere4[3] = sc->bsphereR;
}

// TriAccel record of one primitive (12 words) -- compared with the reference's own TriAccel::load
void orc_triaccel(void *h, uint32_t prim, uint32_t *out12) { std::memcpy(out12, &S(h)->accel.tris[prim], 48); }

// scalar helpers exposed for the known-answer tests
float orc_erf(float x) { return erf_as(x); }
float orc_erfinv(float x) { return erfinv(x); }
float orc_fresnel_dielectric_ext(float cosThetaI, float eta, float *cosThetaT) { return fresnelDielectricExt(cosThetaI, *cosThetaT, eta); }
void orc_fresnel_conductor_exact(float cosThetaI, const float *eta3, const float *k3, float *out3) {
    V3 r = fresnelConductorExact(cosThetaI, rgb(eta3), rgb(k3)); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
void orc_microfacet(int type, float au, float av, int visible, const float *wi3, const float *u2, float *out /*m3,pdf,D,G1*/) {
    Microfacet d(type, au, av, visible != 0);
    float pdf; V3 wi = rgb(wi3);
    V3 m = d.sample(wi, P2{u2[0], u2[1]}, pdf);
    out[0] = m.x; out[1] = m.y; out[2] = m.z; out[3] = pdf; out[4] = d.eval(m); out[5] = d.smithG1(wi, m);
}
void orc_warp(int kind, const float *u2, float *out3) {
    P2 u{u2[0], u2[1]};
    if (kind == 0) { V3 v = squareToCosineHemisphere(u); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
    else if (kind == 1) { V3 v = squareToUniformSphere(u); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
    else if (kind == 2) { P2 p = squareToUniformTriangle(u); out3[0] = p.x; out3[1] = p.y; out3[2] = 0; }
    else { P2 p = squareToUniformDiskConcentric(u); out3[0] = p.x; out3[1] = p.y; out3[2] = 0; }
}

// ---- sample generators (orc_sampler.h), exposed for the known-answer tests --------------------------
// Halton sampler driven like renderBlock drives it: generate(px,py); for each sample: setSampleIndex, then n1 x next2D
// followed by n2 x next1D.  res <= 0: setFilmResolution is not called (the state test_samplers.cpp:37-56 runs in).
int orc_halton_sequence(int resX, int resY, int scramble, int px, int py, int firstSample, int nSamples, int n2D, int n1D, float *out) {
    try {
        std::unique_ptr<orc::Permutations> perms;
        if (scramble != 0) perms.reset(new orc::Permutations(scramble));
        orc::HaltonSampler hs(perms.get());
        if (resX > 0) hs.setFilmResolution(resX, resY, true);
        hs.generate(px, py);
        for (int j = 0; j < nSamples; ++j) {
            hs.setSampleIndex((size_t) (firstSample + j));
            for (int k = 0; k < n2D; ++k) { hs.next2D(out); out += 2; }
            for (int k = 0; k < n1D; ++k) *out++ = hs.next1D();
        }
    } catch (const std::exception &e) { g_err = e.what(); return MTSB_ERR_INVALID; }
    return MTSB_OK;
}
// Sobol sampler driven the same way
int orc_sobol_sequence(int resX, int resY, int scramble, int px, int py, int firstSample, int nSamples, int n2D, int n1D, float *out) {
    try {
        orc::SobolSampler ss((uint64_t) (uint32_t) scramble);
        if (resX > 0) ss.setFilmResolution(resX, resY, true);
        ss.generate(px, py);
        for (int j = 0; j < nSamples; ++j) {
            ss.setSampleIndex((size_t) (firstSample + j));
            for (int k = 0; k < n2D; ++k) { ss.next2D(out); out += 2; }
            for (int k = 0; k < n1D; ++k) *out++ = ss.next1D();
        }
    } catch (const std::exception &e) { g_err = e.what(); return MTSB_ERR_INVALID; }
    return MTSB_OK;
}
uint32_t orc_sobol_matrix32(uint32_t dimension, uint32_t column) { return orc::SobolTables::get().matrices32[(size_t) dimension * 52 + column]; }
void orc_sobol_vdc(uint32_t m, uint64_t *vdc52, uint64_t *inv52) {
    std::vector<uint64_t> a, b; orc::SobolTables::get().pixelTables(m, a, b);
    std::copy(a.begin(), a.end(), vdc52); std::copy(b.begin(), b.end(), inv52);
}
float orc_sobol_sample(uint64_t index, uint32_t dimension, uint32_t scramble) { return orc::sobolSampleSingle(index, dimension, scramble); }
float orc_radical_inverse(int base, uint64_t i) { return orc::radicalInverse(base, i); }
float orc_radical_inverse_incremental(int base, float x) { return orc::radicalInverseIncremental(base, x); }
float orc_radical_inverse_fast(int baseIndex, uint64_t index) { return orc::radicalInverseFast(baseIndex, index); }
float orc_scrambled_radical_inverse_fast(int baseIndex, uint64_t index, const uint16_t *perm) { return orc::radicalInverseFast(baseIndex, index, perm); }
// permutation of prime `baseIndex` (faure.cpp); returns its length
int orc_permutation(int scramble, int baseIndex, uint16_t *out) {
    static std::mutex m; static std::unique_ptr<orc::Permutations> cache; static int cached = 0;
    std::lock_guard<std::mutex> g(m);
    if (!cache || cached != scramble) { cache.reset(new orc::Permutations(scramble)); cached = scramble; }
    const std::vector<uint16_t> &p = cache->perm[baseIndex];
    std::copy(p.begin(), p.end(), out);
    return (int) p.size();
}
void orc_sample02(uint32_t n, const uint32_t *scramble2, float *out2) { out2[0] = orc::radicalInverse2Single(n, scramble2[0]); out2[1] = orc::sobol2Single(n, scramble2[1]); }
uint64_t orc_sample_tea(uint32_t v0, uint32_t v1, int rounds) { return orc::sampleTEA(v0, v1, rounds); }
// LowDiscrepancySampler point sets of one pixel: out1[dim][count], out2[dim][count][2]
int orc_ldsampler_sets(int sampleCount, int dimension, uint64_t seed, float *out1, float *out2) {
    Rng rng(seed);
    orc::LowDiscrepancySampler ld((size_t) sampleCount, (size_t) dimension, &rng);
    ld.generate(0, 0);
    for (size_t d = 0; d < ld.maxDimension; ++d) {
        std::copy(ld.samples1D[d].begin(), ld.samples1D[d].end(), out1 + d * ld.sampleCount);
        std::copy(ld.samples2D[d].begin(), ld.samples2D[d].end(), out2 + d * 2 * ld.sampleCount);
    }
    return (int) ld.sampleCount;
}

// ---- environment map (orc_scene.h: Envmap), exposed for the unit tests --------------------------------------------
uint16_t orc_float_to_half(float f) { return floatToHalf(f); }
float orc_half_to_float(uint16_t h) { return halfToFloat(h); }
void orc_square_to_tent(const float *u2, float *out2) { out2[0] = Envmap::intervalToTent(u2[0]); out2[1] = Envmap::intervalToTent(u2[1]); }
// kind 0: evalEnvironment(d) -> rgb; kind 1: internalSampleDirection(u) -> d(3, local frame), value(3), pdf;
// kind 2: internalPdfDirection(d local) -> pdf.  `in` holds 3 floats per query, `out` 7 floats per query.
int orc_envmap_query(void *h, int kind, const float *in, size_t n, float *out) {
    Scene *sc = S(h);
    if (!sc->envmap.present()) { g_err = "scene has no envmap"; return MTSB_ERR_INVALID; }
    for (size_t i = 0; i < n; ++i) {
        const float *q = in + 3 * i; float *o = out + 7 * i;
        for (int k = 0; k < 7; ++k) o[k] = 0;
        if (kind == 0) { V3 v = sc->envmap.eval(V3(q[0], q[1], q[2])); o[0] = v.x; o[1] = v.y; o[2] = v.z; }
        else if (kind == 1) {
            V3 d, value; float pdf;
            sc->envmap.sampleDirection(P2{q[0], q[1]}, d, value, pdf);
            o[0] = d.x; o[1] = d.y; o[2] = d.z; o[3] = value.x; o[4] = value.y; o[5] = value.z; o[6] = pdf;
        } else o[6] = sc->envmap.pdfDirection(V3(q[0], q[1], q[2]));
    }
    return MTSB_OK;
}

// ---- textures (orc_texture.h) ----------------------------------------------------------------------------------
void *orc_texture_create(const MtsbTexture *t) { orc::Texture *x = new orc::Texture(); x->build(*t); return x; }
void orc_texture_destroy(void *h) { delete static_cast<orc::Texture *>(h); }
int orc_texture_levels(void *h) { return static_cast<orc::Texture *>(h)->mip.levels(); }
void orc_texture_level_size(void *h, int level, int *wh) { const orc::MipMap &m = static_cast<orc::Texture *>(h)->mip; wh[0] = m.w[level]; wh[1] = m.h[level]; }
// level texels as w*h*3 floats (the half-precision values widened)
void orc_texture_level(void *h, int level, float *out) { const orc::MipMap &m = static_cast<orc::Texture *>(h)->mip; std::copy(m.pyr[level].begin(), m.pyr[level].end(), out); }
// Texture2D::eval(its, filter): uv = hit uv (2 per query); partials = dudx, dvdx, dudy, dvdy per query or NULL (unfiltered)
void orc_texture_eval(void *h, size_t n, const float *uv, const float *partials, float *rgb) {
    const orc::Texture *t = static_cast<orc::Texture *>(h);
    for (size_t i = 0; i < n; ++i) {
        orc::UVPartials pt;
        if (partials) { pt.has = true; pt.dudx = partials[4 * i]; pt.dvdx = partials[4 * i + 1]; pt.dudy = partials[4 * i + 2]; pt.dvdy = partials[4 * i + 3]; }
        V3 v = t->eval(P2{uv[2 * i], uv[2 * i + 1]}, pt);
        rgb[3 * i] = v.x; rgb[3 * i + 1] = v.y; rgb[3 * i + 2] = v.z;
    }
}
// in: p(3) n(3) dpdu(3) dpdv(3) o(3) rxd(3) ryd(3) per query; out: dudx dvdx dudy dvdy
void orc_compute_partials(size_t n, const float *in, float *out) {
    for (size_t i = 0; i < n; ++i) {
        const float *q = in + 21 * i;
        auto v = [&](int k) { return V3(q[3 * k], q[3 * k + 1], q[3 * k + 2]); };
        const orc::UVPartials r = orc::computePartials(v(0), v(1), v(2), v(3), v(4), v(5), v(6));
        out[4 * i] = r.dudx; out[4 * i + 1] = r.dvdx; out[4 * i + 2] = r.dudy; out[4 * i + 3] = r.dvdy;
    }
}
// sensor ray differential directions (rxd, ryd) for sample positions of a scene
void orc_sensor_differentials(void *h, size_t n, const float *pos, float *out6) {
    for (size_t i = 0; i < n; ++i) {
        V3 rx, ry; sensorRayDifferentials(*S(h), P2{pos[2 * i], pos[2 * i + 1]}, rx, ry);
        float *o = out6 + 6 * i; o[0] = rx.x; o[1] = rx.y; o[2] = rx.z; o[3] = ry.x; o[4] = ry.y; o[5] = ry.z;
    }
}

}  // extern "C"
