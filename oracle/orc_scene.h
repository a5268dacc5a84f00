// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// Scene-level restatement: intersection record, emitters, direct sampling, sensor, film.
#pragma once
#include "orc_math.h"
#include "orc_bsdf.h"
#include "orc_accel.h"
#include <vector>
#include <mutex>
#include <stdexcept>
#include <cstring>
#include <algorithm>
#include <cmath>

namespace orc {

// Subset of Intersection (include/mitsuba/render/shape.h:30-190) that `path` reads.
struct Its {
    bool valid;
    float t;
    V3 p;
    V3 geoN;          // geoFrame.n
    Frame sh;         // shFrame
    P2 uv;
    V3 wi;            // local
    uint32_t shape, prim;   // prim = global primitive index
    V3 dpdu, dpdv;          // skdtree.h:373-380
    V3 baseN;               // baseFrame.n (the shading normal before any BSDF-side perturbation)
    bool hasUVPartials; float dudx, dvdx, dudy, dvdy;      // shape.h:175-184
    V3 toWorld(V3 v) const { return sh.toWorld(v); }
    V3 toLocal(V3 v) const { return sh.toLocal(v); }
};

// DirectSamplingRecord (include/mitsuba/render/common.h) fields used by `path`
struct DRec {
    V3 ref, refN;
    V3 p, n, d;
    float dist, pdf;
    int measure;      // 0 solid angle, 1 discrete, 2 area
    int emitter;      // index or -1
};
enum { kSolidAngle = 0, kDiscrete = 1 };

// DiscreteDistribution: include/mitsuba/core/pmf.h:101-188
struct Pmf {
    std::vector<float> cdf;   // cdf[0] = 0
    float sum, normalization;
    Pmf() : sum(0), normalization(0) { cdf.push_back(0.0f); }
    void append(float v) { cdf.push_back(cdf.back() + v); }
    float normalize() {
        sum = cdf.back();
        if (sum > 0) {
            normalization = 1.0f / sum;
            for (size_t i = 1; i < cdf.size(); ++i) cdf[i] *= normalization;
            cdf.back() = 1.0f;
        } else normalization = 0.0f;
        return sum;
    }
    float operator[](size_t i) const { return cdf[i + 1] - cdf[i]; }
    size_t sample(float v) const {                                   // pmf.h:124-136
        size_t entry = std::lower_bound(cdf.begin(), cdf.end(), v) - cdf.begin();
        size_t index = std::min(cdf.size() - 2, (size_t) std::max((ptrdiff_t) 0, (ptrdiff_t) entry - 1));
        while ((*this)[index] == 0 && index < cdf.size() - 1) ++index;
        return index;
    }
    size_t sampleReuse(float &v, float &pdf) const {                 // pmf.h:183-188
        size_t index = sample(v);
        pdf = (*this)[index];
        v = (v - cdf[index]) / (cdf[index + 1] - cdf[index]);
        return index;
    }
    size_t sampleReuse(float &v) const { float p; return sampleReuse(v, p); }
};

struct Film {
    int w, h, border;            // crop size, border = ceil(radius - 0.5)   (rfilter.cpp:50)
    float radius, scaleFactor;
    float table[33];             // MTS_FILTER_RESOLUTION = 31 -> 32 entries (+1 guard)
    std::vector<float> data;     // (h+2b) x (w+2b) x 5
    uint64_t bad;
    int stride() const { return w + 2 * border; }
};

// ---- EnvironmentMap (src/emitters/envmap.cpp) ---------------------------------------------------------------
// half (src/libcore/half.cpp:78-211, include/mitsuba/core/half.h): the map is stored as SpectrumHalf
inline uint16_t floatToHalf(float f) {
    union { float f; int i; } x; x.f = f;
    int i = x.i;
    int s = (i >> 16) & 0x00008000, e = ((i >> 23) & 0x000000ff) - (127 - 15), m = i & 0x007fffff;
    if (e <= 0) {
        if (e < -10) return (uint16_t) s;
        m = m | 0x00800000;
        int t = 14 - e, a = (1 << (t - 1)) - 1, b = (m >> t) & 1;
        m = (m + a + b) >> t;
        return (uint16_t) (s | m);
    } else if (e == 0xff - (127 - 15)) {
        if (m == 0) return (uint16_t) (s | 0x7c00);
        m >>= 13;
        return (uint16_t) (s | 0x7c00 | m | (m == 0));
    } else {
        m = m + 0x00000fff + ((m >> 13) & 1);
        if (m & 0x00800000) { m = 0; e += 1; }
        if (e > 30) return (uint16_t) (s | 0x7c00);
        return (uint16_t) (s | (e << 10) | (m >> 13));
    }
}
inline float halfToFloat(uint16_t h) {            // half.cpp:50-75 (the table the reference builds at start-up)
    int s = (h >> 15) & 1, e = (h >> 10) & 0x1f, m = h & 0x3ff;
    union { float f; uint32_t i; } x;
    if (e == 0) {
        if (m == 0) { x.i = (uint32_t) s << 31; return x.f; }
        while (!(m & 0x00000400)) { m <<= 1; e -= 1; }
        e += 1; m &= ~0x00000400;
    } else if (e == 31) { x.i = ((uint32_t) s << 31) | 0x7f800000u | ((uint32_t) m << 13); return x.f; }
    e = e + (127 - 15); m = m << 13;
    x.i = ((uint32_t) s << 31) | ((uint32_t) e << 23) | (uint32_t) m;
    return x.f;
}
}  // namespace orc
#include "orc_texture.h"
namespace orc {

struct Envmap {
    int w = 0, h = 0;
    std::vector<uint16_t> texels;             // 3 per texel
    std::vector<float> cdfRows, cdfCols, rowWeights;
    float normalization = 0, scale = 1; P2 pixelSize{0, 0};
    float toWorld[16], toLocal[16];
    MipMap mip;                               // the MIPMap the reference builds over the map (envmap.cpp:140-178): EWA, repeat / clamp, maxAnisotropy 10
    bool present() const { return w > 0; }
    // MIPMap::evalTexel (mipmap.h:503-563), ERepeat in u, EClamp in v
    V3 evalTexel(int x, int y) const {
        if (x < 0 || x >= w) { x = x % w; if (x < 0) x += w; }
        if (y < 0 || y >= h) y = std::min(std::max(y, 0), h - 1);
        const uint16_t *t = &texels[3 * ((size_t) y * w + x)];
        return V3(halfToFloat(t[0]), halfToFloat(t[1]), halfToFloat(t[2]));
    }
    // MIPMap::evalBilinear(0, uv) (mipmap.h:575-596)
    V3 evalBilinear(P2 uv) const {
        if (!std::isfinite(uv.x) || !std::isfinite(uv.y)) return V3(0.0f);
        float u = uv.x * w - 0.5f, v = uv.y * h - 0.5f;
        int xPos = (int) std::floor(u), yPos = (int) std::floor(v);
        float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
        return evalTexel(xPos, yPos) * dx2 * dy2 + evalTexel(xPos, yPos + 1) * dx2 * dy1
             + evalTexel(xPos + 1, yPos) * dx1 * dy2 + evalTexel(xPos + 1, yPos + 1) * dx1 * dy1;
    }
    // configure() (envmap.cpp:260-329)
    void build(const MtsbEnvmap &em) {
        w = em.width; h = em.height; scale = em.scale;
        std::memcpy(toWorld, em.to_world, sizeof(toWorld)); std::memcpy(toLocal, em.to_local, sizeof(toLocal));
        texels.resize((size_t) 3 * w * h);
        for (size_t i = 0; i < texels.size(); ++i) texels[i] = floatToHalf(em.rgb[i]);
        {   // the MIP pyramid for filtered lookups (envmap.cpp:140-178)
            MtsbTexture mt;
            std::memset(&mt, 0, sizeof(mt));
            mt.type = MTSB_TEX_BITMAP; mt.filter = MTSB_TEXFILTER_EWA; mt.wrap_u = MTSB_WRAP_REPEAT; mt.wrap_v = MTSB_WRAP_CLAMP;
            mt.uv_scale[0] = mt.uv_scale[1] = 1.0f; mt.scale[0] = mt.scale[1] = mt.scale[2] = 1.0f;
            mt.max_anisotropy = 10.0f; mt.width = w; mt.height = h; mt.channels = 3; mt.texels = em.rgb;
            mip.build(mt);
        }
        size_t nEntries = (size_t) (w + 1) * (size_t) h;
        cdfCols.assign(nEntries, 0); cdfRows.assign((size_t) h + 1, 0); rowWeights.assign((size_t) h, 0);
        size_t colPos = 0, rowPos = 0;
        float rowSum = 0.0f;
        cdfRows[rowPos++] = 0;
        for (int y = 0; y < h; ++y) {
            float colSum = 0;
            cdfCols[colPos++] = 0;
            for (int x = 0; x < w; ++x) {
                V3 value = evalTexel(x, y);
                colSum += luminance(value);
                cdfCols[colPos++] = (float) colSum;
            }
            float normalization = 1.0f / (float) colSum;
            for (int x = 1; x < w; ++x) cdfCols[colPos - x - 1] *= normalization;
            cdfCols[colPos - 1] = 1.0f;
            float weight = std::sin((y + 0.5f) * M_PI / h);
            rowWeights[y] = weight;
            rowSum += colSum * weight;
            cdfRows[rowPos++] = (float) rowSum;
        }
        float norm = 1.0f / (float) rowSum;
        for (int y = 1; y < h; ++y) cdfRows[rowPos - y - 1] *= norm;
        cdfRows[rowPos - 1] = 1.0f;
        if (rowSum == 0) throw std::runtime_error("The environment map is completely black -- this is not allowed.");
        if (!std::isfinite(rowSum)) throw std::runtime_error("The environment map contains an invalid floating point value (nan/inf) -- giving up.");
        normalization = 1.0f / (rowSum * (2 * M_PI / w) * (M_PI / h));
        pixelSize = P2{(float) (2 * M_PI / w), (float) (M_PI / h)};
    }
    // sampleReuse (envmap.cpp:655-661)
    static uint32_t sampleReuse(const float *cdf, uint32_t size, float &sample) {
        const float *entry = std::lower_bound(cdf, cdf + size + 1, (float) sample);
        uint32_t index = std::min((uint32_t) std::max((ptrdiff_t) 0, entry - cdf - 1), size - 1);
        sample = (sample - (float) cdf[index]) / (float) (cdf[index + 1] - cdf[index]);
        return index;
    }
    static float intervalToTent(float sample) {                   // warp.cpp:157-169
        float sign;
        if (sample < 0.5f) { sign = 1; sample *= 2; } else { sign = -1; sample = 2 * (sample - 0.5f); }
        return sign * (1 - std::sqrt(sample));
    }
    // internalSampleDirection (envmap.cpp:567-600)
    void sampleDirection(P2 sample, V3 &d, V3 &value, float &pdf) const {
        uint32_t row = sampleReuse(cdfRows.data(), (uint32_t) h, sample.y),
                 col = sampleReuse(cdfCols.data() + (size_t) row * (w + 1), (uint32_t) w, sample.x);
        P2 pos{(float) col + intervalToTent(sample.x), (float) row + intervalToTent(sample.y)};
        int xPos = (int) std::floor(pos.x), yPos = (int) std::floor(pos.y);
        float dx1 = pos.x - xPos, dx2 = 1.0f - dx1, dy1 = pos.y - yPos, dy2 = 1.0f - dy1;
        V3 value1 = evalTexel(xPos, yPos) * dx2 * dy2 + evalTexel(xPos + 1, yPos) * dx1 * dy2;
        V3 value2 = evalTexel(xPos, yPos + 1) * dx2 * dy1 + evalTexel(xPos + 1, yPos + 1) * dx1 * dy1;
        value = (value1 + value2) * scale;
        pdf = (luminance(value1) * rowWeights[std::min(std::max(yPos, 0), h - 1)] +
               luminance(value2) * rowWeights[std::min(std::max(yPos + 1, 0), h - 1)]) * normalization;
        float sinPhi = std::sin(pixelSize.x * (pos.x + 0.5f)), cosPhi = std::cos(pixelSize.x * (pos.x + 0.5f));
        float sinTheta = std::sin(pixelSize.y * (pos.y + 0.5f)), cosTheta = std::cos(pixelSize.y * (pos.y + 0.5f));
        d = V3(sinPhi * sinTheta, cosTheta, -cosPhi * sinTheta);
        pdf /= std::max(std::abs(sinTheta), kEpsilon);
    }
    // internalPdfDirection (envmap.cpp:603-633)
    float pdfDirection(V3 d) const {
        P2 uv{std::atan2(d.x, -d.z) * kInvTwoPi, std::acos(std::min(1.0f, std::max(-1.0f, d.y))) * kInvPi};
        if (!std::isfinite(uv.x) || !std::isfinite(uv.y)) return 0.0f;
        float u = uv.x * w - 0.5f, v = uv.y * h - 0.5f;
        int xPos = (int) std::floor(u), yPos = (int) std::floor(v);
        float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
        V3 value1 = evalTexel(xPos, yPos) * dx2 * dy2 + evalTexel(xPos + 1, yPos) * dx1 * dy2;
        V3 value2 = evalTexel(xPos, yPos + 1) * dx2 * dy1 + evalTexel(xPos + 1, yPos + 1) * dx1 * dy1;
        float sinTheta = std::sqrt(std::max(0.0f, 1 - d.y * d.y));
        return (luminance(value1) * rowWeights[std::min(std::max(yPos, 0), h - 1)] +
                luminance(value2) * rowWeights[std::min(std::max(yPos + 1, 0), h - 1)]) * normalization / std::max(std::abs(sinTheta), kEpsilon);
    }
    // evalEnvironment without differentials (envmap.cpp:380-392, 408-409); the EWA branch returns the same value
    // whenever the footprint is below one texel (mipmap.h:694-710)
    V3 eval(V3 dWorld) const {
        V3 v = xf(toLocal, dWorld);
        P2 uv{std::atan2(v.x, -v.z) * kInvTwoPi, std::acos(std::min(1.0f, std::max(-1.0f, v.y))) * kInvPi};
        return evalBilinear(uv) * scale;
    }
    // evalEnvironment for a ray with differentials (envmap.cpp:393-407): rxd / ryd = directions of the offset rays
    V3 evalFiltered(V3 dWorld, V3 rxd, V3 ryd) const {
        V3 v = xf(toLocal, dWorld);
        P2 uv{std::atan2(v.x, -v.z) * kInvTwoPi, std::acos(std::min(1.0f, std::max(-1.0f, v.y))) * kInvPi};
        V3 dvdx = xf(toLocal, rxd) - v, dvdy = xf(toLocal, ryd) - v;
        float t1 = kInvTwoPi / (v.x * v.x + v.z * v.z), t2 = -kInvPi / std::max(std::sqrt(std::max(0.0f, 1.0f - v.y * v.y)), kEpsilon);
        P2 dudx{t1 * (dvdx.z * v.x - dvdx.x * v.z), t2 * dvdx.y}, dudy{t1 * (dvdy.z * v.x - dvdy.x * v.z), t2 * dvdy.y};
        return mip.eval(uv, dudx, dudy) * scale;
    }
    static V3 xf(const float *m, V3 v) {
        return V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
    }
};

struct Scene {
    MtsbSceneDesc sd;            // pointers refer to the owned copies below
    std::vector<MtsbShape> shapes;
    std::vector<MtsbMaterial> materials;
    std::vector<MtsbEmitter> emitters;
    std::vector<float> positions, normals, texcoords;
    std::vector<uint32_t> indices;
    std::vector<uint32_t> triShape;           // global prim -> shape
    Accel accel;
    ExternalTraceFn ext; void *extUser;
    Pmf emitterPdf;                            // scene.cpp:375-381
    std::vector<Pmf> areaPdf;                  // per shape that is an emitter (trimesh.cpp:388-403)
    std::vector<float> invArea;                // per shape
    int envEmitter;
    Envmap envmap;                             // envmap.cpp (when the environment emitter is an image)
    std::vector<Texture> textures;             // src/textures (bitmap, checkerboard, gridtexture; scale folded in)
    std::vector<bool> shapeNeedsPartials;      // the shape's BSDF usesRayDifferentials() (a bitmap texture: bitmap.cpp:547-549)
    V3 bsphereC; float bsphereR;               // constant.cpp:67-71
    // sensor (closed form of perspective.cpp:127-150)
    float nearX, nearY;                        // near*tan(xfov/2), same / aspect
    Film film;
    Scene() : ext(nullptr), extUser(nullptr), envEmitter(-1) {}

    V3 pos(uint32_t i) const { return V3(positions[3 * i], positions[3 * i + 1], positions[3 * i + 2]); }
    V3 nrm(uint32_t i) const { return V3(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]); }
    P2 tex(uint32_t i) const { return P2{texcoords[2 * i], texcoords[2 * i + 1]}; }
};

inline V3 xformPoint(const float *m, V3 p) {     // affine part of a row-major 4x4
    return V3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3],
              m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
inline V3 xformVec(const float *m, V3 v) {
    return V3(m[0] * v.x + m[1] * v.y + m[2] * v.z,
              m[4] * v.x + m[5] * v.y + m[6] * v.z,
              m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

// ShapeKDTree::fillIntersectionRecord<true>: include/mitsuba/render/skdtree.h:343-430
inline void fillIntersection(const Scene &sc, const MtsbRay &ray, const MtsbHit &hit, Its &its) {
    its.valid = hit.prim != MTSB_NO_HIT;
    its.t = hit.t;
    if (!its.valid) return;
    uint32_t s = sc.triShape[hit.prim];
    const MtsbShape &sh = sc.shapes[s];
    const uint32_t *ix = &sc.indices[3 * (size_t) hit.prim];
    const V3 b(1 - hit.u - hit.v, hit.u, hit.v);
    const V3 p0 = sc.pos(ix[0]), p1 = sc.pos(ix[1]), p2 = sc.pos(ix[2]);
    its.p = p0 * b.x + p1 * b.y + p2 * b.z;
    V3 side1 = p1 - p0, side2 = p2 - p0;
    V3 faceNormal = cross(side1, side2);
    float len = length(faceNormal);
    if (!isZero(faceNormal)) faceNormal = faceNormal / len;
    V3 dpdu = side1, dpdv = side2;
    V3 shN;
    if (sh.has_normals) {
        V3 n0 = sc.nrm(ix[0]), n1 = sc.nrm(ix[1]), n2 = sc.nrm(ix[2]);
        shN = normalize(n0 * b.x + n1 * b.y + n2 * b.z);
        if (dot(faceNormal, shN) < 0) faceNormal = -faceNormal;
    } else shN = faceNormal;
    its.geoN = faceNormal;
    if (sh.has_texcoords && !sc.texcoords.empty()) {
        P2 t0 = sc.tex(ix[0]), t1 = sc.tex(ix[1]), t2 = sc.tex(ix[2]);
        its.uv = P2{t0.x * b.x + t1.x * b.y + t2.x * b.z, t0.y * b.x + t1.y * b.y + t2.y * b.z};
        // TriMesh::computeUVTangents (trimesh.cpp:683-735); this fork computes them for every mesh with texture
        // coordinates (TriMesh::configure, trimesh.cpp:383-385) and skdtree.h:373-376 then reads dpdu / dpdv from them
        V3 dP1 = p1 - p0, dP2 = p2 - p0;
        P2 dUV1{t1.x - t0.x, t1.y - t0.y}, dUV2{t2.x - t0.x, t2.y - t0.y};
        V3 n = cross(dP1, dP2);
        float nlen = length(n);
        if (nlen == 0) { dpdu = V3(0.0f); dpdv = V3(0.0f); }
        else {
            float determinant = dUV1.x * dUV2.y - dUV1.y * dUV2.x;
            if (determinant == 0) coordinateSystem(n / nlen, dpdu, dpdv);
            else {
                float invDet = 1.0f / determinant;
                dpdu = (dP1 * dUV2.y - dP2 * dUV1.y) * invDet;
                dpdv = (dP1 * -dUV2.x + dP2 * dUV1.x) * invDet;
            }
        }
    } else its.uv = P2{b.y, b.z};
    its.shape = s; its.prim = hit.prim;
    its.dpdu = dpdu; its.dpdv = dpdv; its.baseN = shN; its.hasUVPartials = false;
    its.sh = shadingFrame(shN, dpdu);
    its.wi = its.toLocal(-V3(ray.d[0], ray.d[1], ray.d[2]));
}

// ---- emitters ----------------------------------------------------------------
// AreaLight::eval  src/emitters/area.cpp:104-109
inline V3 areaEval(const MtsbEmitter &e, V3 shN, V3 d) {
    if (dot(shN, d) <= 0) return V3(0.0f);
    return rgb(e.radiance);
}

// BSphere::rayIntersect  include/mitsuba/core/bsphere.h:88-95
inline bool bsphereIntersect(V3 c, float r, V3 o, V3 d, float &nearT, float &farT) {
    V3 oc = o - c;
    float A = lengthSquared(d), B = 2 * dot(oc, d), C = lengthSquared(oc) - r * r;
    return solveQuadratic(A, B, C, nearT, farT);
}

// Emitter::sampleDirect for the three emitter types; returns value (radiance/pdf)
inline V3 emitterSampleDirect(const Scene &sc, int ei, DRec &dRec, P2 u) {
    const MtsbEmitter &e = sc.emitters[ei];
    if (e.type == MTSB_EMITTER_AREA) {
        // area.cpp:158-173 -> Shape::sampleDirect shape.cpp:102-115 -> TriMesh::samplePosition
        // trimesh.cpp:412-423 -> Triangle::sample triangle.cpp:24-62
        const MtsbShape &sh = sc.shapes[e.shape];
        float sy = u.y;
        size_t index = sc.areaPdf[e.shape].sampleReuse(sy);
        const uint32_t *ix = &sc.indices[3 * ((size_t) sh.first_tri + index)];
        V3 p0 = sc.pos(ix[0]), p1 = sc.pos(ix[1]), p2 = sc.pos(ix[2]);
        P2 bary = squareToUniformTriangle(P2{u.x, sy});
        V3 sideA = p1 - p0, sideB = p2 - p0;
        dRec.p = p0 + (sideA * bary.x) + (sideB * bary.y);
        if (sh.has_normals) {
            V3 n0 = sc.nrm(ix[0]), n1 = sc.nrm(ix[1]), n2 = sc.nrm(ix[2]);
            dRec.n = normalize(n0 * (1.0f - bary.x - bary.y) + n1 * bary.x + n2 * bary.y);
        } else dRec.n = normalize(cross(sideA, sideB));
        dRec.pdf = sc.invArea[e.shape];
        dRec.d = dRec.p - dRec.ref;
        float distSquared = lengthSquared(dRec.d);
        dRec.dist = std::sqrt(distSquared);
        dRec.d = dRec.d / dRec.dist;
        float dp = absDot(dRec.d, dRec.n);
        dRec.pdf *= dp != 0 ? (distSquared / dp) : 0.0f;
        dRec.measure = kSolidAngle;
        if (dot(dRec.d, dRec.refN) >= 0 && dot(dRec.d, dRec.n) < 0 && dRec.pdf != 0)
            return rgb(e.radiance) / dRec.pdf;
        dRec.pdf = 0.0f;
        return V3(0.0f);
    } else if (e.type == MTSB_EMITTER_POINT) {
        // point.cpp:131-147
        dRec.p = rgb(e.position);
        dRec.pdf = 1.0f; dRec.measure = kDiscrete;
        dRec.d = dRec.p - dRec.ref;
        dRec.dist = length(dRec.d);
        float invDist = 1.0f / dRec.dist;
        dRec.d *= invDist;
        dRec.n = V3(0.0f);
        return rgb(e.radiance) * (invDist * invDist);
    } else if (e.type == MTSB_EMITTER_ENVMAP) {
        // envmap.cpp:516-543
        V3 dl, value; float pdf;
        sc.envmap.sampleDirection(u, dl, value, pdf);
        V3 d = Envmap::xf(sc.envmap.toWorld, dl);
        float nearT, farT;
        if (isZero(value) || pdf == 0 || !bsphereIntersect(sc.bsphereC, sc.bsphereR, dRec.ref, d, nearT, farT) || nearT >= 0 || farT <= 0) {
            dRec.pdf = 0.0f;
            return V3(0.0f);
        }
        dRec.pdf = pdf;
        dRec.p = dRec.ref + d * farT;
        dRec.n = normalize(sc.bsphereC - dRec.p);
        dRec.dist = farT; dRec.d = d; dRec.measure = kSolidAngle;
        return value / pdf;
    } else {
        // constant.cpp:173-215
        V3 d; float pdf;
        if (!isZero(dRec.refN)) {
            d = squareToCosineHemisphere(u);
            pdf = squareToCosineHemispherePdf(d);
            d = frameFromNormal(dRec.refN).toWorld(d);
        } else {
            d = squareToUniformSphere(u);
            pdf = kInvFourPi;
        }
        float nearT, farT;
        dRec.pdf = 0.0f;
        if (!bsphereIntersect(sc.bsphereC, sc.bsphereR, dRec.ref, d, nearT, farT)) return V3(0.0f);
        if (!(nearT < 0 && farT > 0)) return V3(0.0f);
        dRec.p = dRec.ref + d * farT;
        dRec.n = normalize(sc.bsphereC - dRec.p);
        dRec.measure = kSolidAngle;
        dRec.d = d; dRec.dist = farT; dRec.pdf = pdf;
        if (!isZero(dRec.refN) && dot(dRec.d, dRec.refN) <= 0) return V3(0.0f);
        return rgb(e.radiance) / pdf;
    }
}

// Emitter::pdfDirect
inline float emitterPdfDirect(const Scene &sc, int ei, const DRec &dRec) {
    const MtsbEmitter &e = sc.emitters[ei];
    if (e.type == MTSB_EMITTER_AREA) {
        // area.cpp:175-182 -> Shape::pdfDirect shape.cpp:117-126
        if (dot(dRec.d, dRec.refN) >= 0 && dot(dRec.d, dRec.n) < 0) {
            float pdfPos = sc.invArea[e.shape];
            if (dRec.measure == kSolidAngle) return pdfPos * (dRec.dist * dRec.dist) / absDot(dRec.d, dRec.n);
            return 0.0f;
        }
        return 0.0f;
    } else if (e.type == MTSB_EMITTER_POINT) {
        return dRec.measure == kDiscrete ? 1.0f : 0.0f;                 // point.cpp:149-151
    } else if (e.type == MTSB_EMITTER_ENVMAP) {
        // envmap.cpp:545-556
        float pdfSA = sc.envmap.pdfDirection(Envmap::xf(sc.envmap.toLocal, dRec.d));
        return dRec.measure == kSolidAngle ? pdfSA : 0.0f;
    } else {
        // constant.cpp:217-232
        float pdfSA;
        if (!isZero(dRec.refN)) pdfSA = kInvPi * std::max(0.0f, dot(dRec.d, dRec.refN));
        else pdfSA = kInvFourPi;
        return dRec.measure == kSolidAngle ? pdfSA : 0.0f;
    }
}

// Scene::evalEnvironment for a ray that carries differentials (a camera ray): envmap.cpp:393-407 filters the lookup
inline V3 evalEnvironmentFiltered(const Scene &sc, V3 d, V3 rxd, V3 ryd) {
    const MtsbEmitter &e = sc.emitters[sc.envEmitter];
    return e.type == MTSB_EMITTER_ENVMAP ? sc.envmap.evalFiltered(d, rxd, ryd) : rgb(e.radiance);
}
// Scene::evalEnvironment (scene.h:914-917)
inline V3 evalEnvironment(const Scene &sc, V3 d) {
    const MtsbEmitter &e = sc.emitters[sc.envEmitter];
    return e.type == MTSB_EMITTER_ENVMAP ? sc.envmap.eval(d) : rgb(e.radiance);       // envmap.cpp:380-410 / constant.cpp:242-244
}

// Scene::sampleEmitterDirect scene.cpp:828-855 (visibility test done by the caller so that the
// unit tests can look at the unoccluded record)
inline V3 sampleEmitterDirectNoVis(const Scene &sc, DRec &dRec, P2 u) {
    dRec.emitter = -1; dRec.pdf = 0;
    if (sc.emitters.empty()) return V3(0.0f);
    float emPdf;
    float sx = u.x;
    size_t index = sc.emitterPdf.sampleReuse(sx, emPdf);
    V3 value = emitterSampleDirect(sc, (int) index, dRec, P2{sx, u.y});
    if (dRec.pdf != 0) {
        dRec.emitter = (int) index;
        dRec.pdf *= emPdf;
        value = value / emPdf;
        return value;
    }
    return V3(0.0f);
}
// Scene::pdfEmitterDirect scene.cpp:975-978, pdfEmitterDiscrete scene.h:852-854
inline float pdfEmitterDirect(const Scene &sc, const DRec &dRec) {
    return emitterPdfDirect(sc, dRec.emitter, dRec) * sc.emitters[dRec.emitter].sampling_weight * sc.emitterPdf.normalization;
}

// ---- sensor: PerspectiveCamera::sampleRayDifferential perspective.cpp:286-317 ----
inline void sensorRay(const Scene &sc, P2 samplePos, MtsbRay &ray) {
    const MtsbSensor &se = sc.sd.sensor;
    // near-plane point: closed form of m_sampleToCamera(Point(px/cropW, py/cropH, 0)) (:127-136,:296-298)
    float a = (samplePos.x + (float) se.crop_x) / (float) se.film_width;
    float b = (samplePos.y + (float) se.crop_y) / (float) se.film_height;
    V3 nearP((1.0f - 2.0f * a) * sc.nearX, (1.0f - 2.0f * b) * sc.nearY, se.near_clip);
    V3 d = normalize(nearP);
    float invZ = 1.0f / d.z;
    ray.mint = se.near_clip * invZ;
    ray.maxt = se.far_clip * invZ;
    V3 o = xformPoint(se.to_world, V3(0.0f));
    V3 wd = xformVec(se.to_world, d);
    ray.o[0] = o.x; ray.o[1] = o.y; ray.o[2] = o.z;
    ray.d[0] = wd.x; ray.d[1] = wd.y; ray.d[2] = wd.z;
}

// offset-ray directions of the camera ray (perspective.cpp:160-163, 310-314) after RayDifferential::scaleDifferential
// (ray.h:174-179) with diffScaleFactor = 1/sqrt(sampleCount) (integrator.cpp:200-201, 237)
inline void sensorRayDifferentials(const Scene &sc, P2 samplePos, V3 &rxd, V3 &ryd) {
    const MtsbSensor &se = sc.sd.sensor;
    float a = (samplePos.x + (float) se.crop_x) / (float) se.film_width;
    float b = (samplePos.y + (float) se.crop_y) / (float) se.film_height;
    V3 nearP((1.0f - 2.0f * a) * sc.nearX, (1.0f - 2.0f * b) * sc.nearY, se.near_clip);
    V3 dx(-2.0f * sc.nearX / (float) se.film_width, 0.0f, 0.0f), dy(0.0f, -2.0f * sc.nearY / (float) se.film_height, 0.0f);
    V3 d = xformVec(se.to_world, normalize(nearP));
    V3 rx = xformVec(se.to_world, normalize(nearP + dx)), ry = xformVec(se.to_world, normalize(nearP + dy));
    float amount = 1.0f / std::sqrt((float) std::max(1, sc.sd.integrator.sample_count));
    rxd = d + (rx - d) * amount;
    ryd = d + (ry - d) * amount;
}

// The material table a hit shades with: the shape's record (and the back side of a twosided adapter) with every textured
// spectrum replaced by the texture's value at the hit (e.g. diffuse.cpp:118: m_reflectance->eval(bRec.its)).
// Returns the index of the front record inside `lm`.
inline int resolveMaterials(const Scene &sc, int mat, const Its &its, MtsbMaterial lm[2]) {
    UVPartials pt;
    if (its.hasUVPartials) { pt.has = true; pt.dudx = its.dudx; pt.dvdx = its.dvdx; pt.dudy = its.dudy; pt.dvdy = its.dvdy; }
    lm[0] = sc.materials[mat];
    int n = 1;
    if ((lm[0].flags & MTSB_MAT_TWOSIDED) && lm[0].back >= 0) { lm[1] = sc.materials[lm[0].back]; lm[0].back = 1; n = 2; }
    for (int i = 0; i < n; ++i) {
        if (lm[i].tex_reflectance) { V3 v = sc.textures[lm[i].tex_reflectance - 1].eval(its.uv, pt); lm[i].reflectance[0] = v.x; lm[i].reflectance[1] = v.y; lm[i].reflectance[2] = v.z; }
        if (lm[i].tex_alpha) {         // roughconductor.cpp:379-384, roughdielectric.cpp:300-305; clamp of microfacet.h:135-136
            V3 v = sc.textures[lm[i].tex_alpha - 1].eval(its.uv, pt);
            float a = std::max((v.x + v.y + v.z) * (1.0f / 3.0f), 1e-4f);
            lm[i].alpha_u = a; lm[i].alpha_v = a;
        }
        if (lm[i].tex_transmittance) { V3 v = sc.textures[lm[i].tex_transmittance - 1].eval(its.uv, pt); lm[i].transmittance[0] = v.x; lm[i].transmittance[1] = v.y; lm[i].transmittance[2] = v.z; }
    }
    return 0;
}

// ---- film: ReconstructionFilter::configure rfilter.cpp:36-55; GaussianFilter::eval gaussian.cpp:50-55;
//      BoxFilter box.cpp:36-50; ImageBlock::put imageblock.h:143-206 --------------------------------
inline float rfilterEval(const MtsbSensor &se, float x) {
    if (se.rfilter == MTSB_RFILTER_GAUSSIAN) {
        float alpha = -1.0f / (2.0f * se.rfilter_stddev * se.rfilter_stddev);
        return std::max(0.0f, (float) std::exp((double) (alpha * x * x)) - (float) std::exp((double) (alpha * se.rfilter_radius * se.rfilter_radius)));   // math::fastexp (math.h:175-199)
    }
    return std::fabs(x) <= se.rfilter_radius ? 1.0f : 0.0f;
}
inline void filmInit(Film &f, const MtsbSensor &se) {
    const int RES = 31;
    f.w = se.crop_width; f.h = se.crop_height;
    f.radius = se.rfilter_radius;
    float sum = 0.0f;
    for (int i = 0; i < RES; ++i) { float v = rfilterEval(se, (f.radius * i) / RES); f.table[i] = v; sum += v; }
    f.table[RES] = 0.0f; f.table[RES + 1] = 0.0f;
    f.scaleFactor = RES / f.radius;
    f.border = (int) std::ceil(f.radius - 0.5f);
    sum *= 2 * f.radius / RES;
    float normalization = 1.0f / sum;
    for (int i = 0; i < RES; ++i) f.table[i] *= normalization;
    f.data.assign((size_t) (f.h + 2 * f.border) * (f.w + 2 * f.border) * 5, 0.0f);
    f.bad = 0;
}
inline float filmEvalDiscretized(const Film &f, float x) {
    return f.table[std::min((int) std::fabs(x * f.scaleFactor), 31)];
}
// put into a buffer covering [ox-border, ox+bw+border) x [oy-border, oy+bh+border) (block semantics)
inline bool blockPut(const Film &f, float *buf, int ox, int oy, int bw, int bh, P2 _pos, const float *value5) {
    for (int i = 0; i < 5; ++i) if (!std::isfinite(value5[i])) return false;
    const int sx = bw + 2 * f.border, sy = bh + 2 * f.border;
    const float px = _pos.x - 0.5f - (ox - f.border), py = _pos.y - 0.5f - (oy - f.border);
    int minx = std::max((int) std::ceil(px - f.radius), 0), miny = std::max((int) std::ceil(py - f.radius), 0);
    int maxx = std::min((int) std::floor(px + f.radius), sx - 1), maxy = std::min((int) std::floor(py + f.radius), sy - 1);
    float wx[16], wy[16];
    for (int x = minx, i = 0; x <= maxx; ++x) wx[i++] = filmEvalDiscretized(f, x - px);
    for (int y = miny, i = 0; y <= maxy; ++y) wy[i++] = filmEvalDiscretized(f, y - py);
    for (int y = miny, yr = 0; y <= maxy; ++y, ++yr) {
        float *dest = buf + ((size_t) y * sx + minx) * 5;
        for (int x = minx, xr = 0; x <= maxx; ++x, ++xr) {
            const float weight = wx[xr] * wy[yr];
            for (int k = 0; k < 5; ++k) *dest++ += weight * value5[k];
        }
    }
    return true;
}

}  // namespace orc
