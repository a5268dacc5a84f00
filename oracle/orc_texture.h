// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// Restatement of the reference's 2D textures:
//   Resampler<float>            include/mitsuba/core/rfilter.h:107-330 (resampling mode, resampleAndClamp)
//   LanczosSincFilter::eval     src/rfilters/lanczos.cpp:46-58
//   Bitmap::resample            src/libcore/bitmap.cpp:2230-2330 (X pass into a temporary, then Y pass)
//   TMIPMap                     include/mitsuba/render/mipmap.h (constructor :155-304, evalTexel :503-563, evalBox :565-568,
//                               evalBilinear :574-596, eval :628-716, evalEWA :760-840)
//   Checkerboard / GridTexture  src/textures/checkerboard.cpp:61-69, src/textures/gridtexture.cpp:66-79
//   Texture2D::eval             src/librender/texture.cpp:112-121
//   ScalingTexture::eval        src/textures/scale.cpp:69-71
//   Intersection::computePartials  src/librender/intersection.cpp:5-98
#pragma once
#include "orc_math.h"
#include "orc_bsdf.h"
#include "../include/mtsb200.h"
#include <vector>
#include <cmath>
#include <limits>
#include <algorithm>

namespace orc {

// floatToHalf / halfToFloat: orc_scene.h (half.cpp restatement), defined before this header is included

inline int texModulo(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }

// lookup with a boundary condition; `inside` = false means the constant `outside` applies
inline bool texBoundary(int &pos, int res, uint32_t bc, float &outside) {
    if (pos >= 0 && pos < res) return true;
    switch (bc) {
        case MTSB_WRAP_REPEAT: pos = texModulo(pos, res); return true;
        case MTSB_WRAP_CLAMP: pos = std::min(std::max(pos, 0), res - 1); return true;
        case MTSB_WRAP_MIRROR: pos = texModulo(pos, 2 * res); if (pos >= res) pos = 2 * res - pos - 1; return true;
        case MTSB_WRAP_ZERO: outside = 0.0f; return false;
        default: outside = 1.0f; return false;
    }
}

inline float lanczosEval(float x, float radius) {
    x = std::abs(x);
    if (x < kEpsilon) return 1.0f;
    else if (x > radius) return 0.0f;
    float x1 = (float) (M_PI * x);
    float x2 = x1 / radius;
    return (std::sin(x1) * std::sin(x2)) / (x1 * x2);
}

struct Resampler {
    uint32_t bc; int sourceRes, targetRes, taps;
    std::vector<int> start; std::vector<float> weights;
    Resampler(uint32_t bc_, int src, int trg, float radius = 2.0f) : bc(bc_), sourceRes(src), targetRes(trg) {
        float filterRadius = radius, scale = 1.0f, invScale = 1.0f;
        if (trg < src) { scale = (float) src / (float) trg; invScale = 1 / scale; filterRadius *= scale; }
        taps = (int) std::ceil(filterRadius * 2);
        start.resize(trg); weights.resize((size_t) taps * trg);
        for (int i = 0; i < trg; i++) {
            float center = (i + 0.5f) / trg * src;
            start[i] = (int) std::floor(center - filterRadius + 0.5f);
            float sum = 0;
            for (int j = 0; j < taps; j++) {
                float pos = start[i] + j + 0.5f - center;
                float weight = lanczosEval(pos * invScale, radius);
                weights[(size_t) i * taps + j] = weight;
                sum += weight;
            }
            float normalization = 1.0f / sum;
            for (int j = 0; j < taps; j++) weights[(size_t) i * taps + j] *= normalization;
        }
    }
    // one line of `channels`-interleaved samples; strides in pixels
    void resampleAndClamp(const float *source, size_t sourceStride, float *target, size_t targetStride, int channels, float mn, float mx) const {
        for (int i = 0; i < targetRes; ++i) {
            for (int ch = 0; ch < channels; ++ch) {
                float result = 0;
                for (int j = 0; j < taps; ++j) {
                    int pos = start[i] + j; float outside = 0.0f, v;
                    if (texBoundary(pos, sourceRes, bc, outside)) v = source[sourceStride * channels * (size_t) pos + ch]; else v = outside;
                    result += v * weights[(size_t) i * taps + j];
                }
                target[targetStride * channels * (size_t) i + ch] = std::min(mx, std::max(mn, result));
            }
        }
    }
};

// one downsampling step of the pyramid (mipmap.h:262: bitmap->resample(rfilter, bcu, bcv, size, 0, maxValue))
inline std::vector<float> resampleBitmap(const std::vector<float> &src, int sw, int sh, int tw, int th, int channels, uint32_t bcu, uint32_t bcv, float maxValue) {
    std::vector<float> cur = src; int cw = sw;
    if (sw != tw) {
        Resampler r(bcu, sw, tw);
        std::vector<float> tmp((size_t) tw * sh * channels);
        for (int y = 0; y < sh; ++y) r.resampleAndClamp(&src[(size_t) y * sw * channels], 1, &tmp[(size_t) y * tw * channels], 1, channels, 0.0f, maxValue);
        cur.swap(tmp); cw = tw;
    }
    if (sh != th) {
        Resampler r(bcv, sh, th);
        std::vector<float> out((size_t) tw * th * channels);
        for (int x = 0; x < cw; ++x) r.resampleAndClamp(&cur[(size_t) x * channels], cw, &out[(size_t) x * channels], tw, channels, 0.0f, maxValue);
        cur.swap(out);
    }
    return cur;
}

struct MipMap {
    uint32_t filter, bcu, bcv; float maxAnisotropy; int channels;
    std::vector<int> w, h;
    std::vector<std::vector<float>> pyr;      // per level: w*h*3 values after the round trip through half
    V3 minimum, maximum, average;
    float lut[64];
    int levels() const { return (int) pyr.size(); }

    void storeLevel(const std::vector<float> &img, int lw, int lh) {
        std::vector<float> out((size_t) lw * lh * 3);
        for (size_t i = 0; i < (size_t) lw * lh; ++i)
            for (int c = 0; c < 3; ++c) out[3 * i + c] = halfToFloat(floatToHalf(img[i * channels + (channels == 3 ? c : 0)]));
        w.push_back(lw); h.push_back(lh); pyr.push_back(out);
    }
    void build(const MtsbTexture &t) {
        filter = t.filter; bcu = t.wrap_u; bcv = t.wrap_v; channels = t.channels;
        maxAnisotropy = t.filter == MTSB_TEXFILTER_EWA ? t.max_anisotropy : 1.0f;                  // bitmap.cpp:232-235
        std::vector<float> bitmap(t.texels, t.texels + (size_t) t.width * t.height * channels);
        for (float &v : bitmap) v = std::max(v, 0.0f);                                            // clampNegative (mipmap.h:228-236)
        // Array2D::init(data, min, max, avg) (mipmap.h:222): statistics of the float values
        V3 mn(std::numeric_limits<float>::infinity()), mx(-std::numeric_limits<float>::infinity());
        double av[3] = {0, 0, 0};
        for (size_t i = 0; i < (size_t) t.width * t.height; ++i)
            for (int c = 0; c < 3; ++c) {
                float v = bitmap[i * channels + (channels == 3 ? c : 0)];
                float &lo = c == 0 ? mn.x : c == 1 ? mn.y : mn.z, &hi = c == 0 ? mx.x : c == 1 ? mx.y : mx.z;
                lo = std::min(lo, v); hi = std::max(hi, v); av[c] += v;
            }
        minimum = mn; maximum = mx;
        const double cnt = (double) t.width * t.height;
        average = V3((float) (av[0] / cnt), (float) (av[1] / cnt), (float) (av[2] / cnt));
        storeLevel(bitmap, t.width, t.height);
        if (filter != MTSB_TEXFILTER_NEAREST && filter != MTSB_TEXFILTER_BILINEAR) {
            int sx = t.width, sy = t.height;
            while (sx > 1 || sy > 1) {
                int nx = std::max(1, (sx + 1) / 2), ny = std::max(1, (sy + 1) / 2);
                bitmap = resampleBitmap(bitmap, sx, sy, nx, ny, channels, bcu, bcv, 1.0f);
                sx = nx; sy = ny;
                storeLevel(bitmap, sx, sy);
            }
        }
        for (int i = 0; i < 64; ++i) { float r2 = (float) i / 63.0f; lut[i] = (float) std::exp((double) (-2.0f * r2)) - (float) std::exp((double) -2.0f); }   // mipmap.h:296-302
    }
    V3 evalTexel(int level, int x, int y) const {
        float outside = 0.0f;
        if (!texBoundary(x, w[level], bcu, outside)) return V3(outside);
        if (!texBoundary(y, h[level], bcv, outside)) return V3(outside);
        const float *p = &pyr[level][3 * ((size_t) y * w[level] + x)];
        return V3(p[0], p[1], p[2]);
    }
    V3 evalBox(int level, P2 uv) const { return evalTexel(level, (int) std::floor(uv.x * w[level]), (int) std::floor(uv.y * h[level])); }
    V3 evalBilinear(int level, P2 uv) const {
        if (!std::isfinite(uv.x) || !std::isfinite(uv.y)) return V3(0.0f);
        if (level >= levels()) return evalBilinear(levels() - 1, uv);
        float u = uv.x * w[level] - 0.5f, v = uv.y * h[level] - 0.5f;
        int xPos = (int) std::floor(u), yPos = (int) std::floor(v);
        float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
        return evalTexel(level, xPos, yPos) * dx2 * dy2 + evalTexel(level, xPos, yPos + 1) * dx2 * dy1
             + evalTexel(level, xPos + 1, yPos) * dx1 * dy2 + evalTexel(level, xPos + 1, yPos + 1) * dx1 * dy1;
    }
    V3 evalEWA(int level, P2 uv, float A, float B, float C) const {
        if (!std::isfinite(A + B + C + uv.x + uv.y)) return V3(0.0f);
        if (level >= levels()) return evalBox(levels() - 1, uv);
        float u = uv.x * w[level] - 0.5f, v = uv.y * h[level] - 0.5f;
        float rx = (float) w[level] / (float) w[0], ry = (float) h[level] / (float) h[0];
        A /= rx * rx; B /= rx * ry; C /= ry * ry;
        float invDet = 1.0f / (-B * B + 4.0f * A * C), deltaU = 2.0f * std::sqrt(C * invDet), deltaV = 2.0f * std::sqrt(A * invDet);
        int u0 = (int) std::ceil(u - deltaU), u1 = (int) std::floor(u + deltaU), v0 = (int) std::ceil(v - deltaV), v1 = (int) std::floor(v + deltaV);
        float As = A * 64, Bs = B * 64, Cs = C * 64;
        V3 result(0.0f); float denominator = 0.0f;
        float ddq = 2 * As, uu0 = (float) u0 - u;
        for (int vt = v0; vt <= v1; ++vt) {
            const float vv = (float) vt - v;
            float q = As * uu0 * uu0 + (Bs * uu0 + Cs * vv) * vv;
            float dq = As * (2 * uu0 + 1) + Bs * vv;
            for (int ut = u0; ut <= u1; ++ut) {
                if (q < 64.0f) {
                    uint32_t qi = (uint32_t) q;
                    if (qi < 64) { const float weight = lut[(int) q]; result += evalTexel(level, ut, vt) * weight; denominator += weight; }
                }
                q += dq; dq += ddq;
            }
        }
        if (denominator == 0) return evalBilinear(level, uv);
        return result / denominator;
    }
    static float hypot2(float a, float b) {
        float r;
        if (std::abs(a) > std::abs(b)) { r = b / a; r = std::abs(a) * std::sqrt(1.0f + r * r); }
        else if (b != 0.0f) { r = a / b; r = std::abs(b) * std::sqrt(1.0f + r * r); }
        else r = 0.0f;
        return r;
    }
    static float log2f_(float v) { const float invLn2 = 1.0f / std::log(2.0f); return (float) std::log((double) v) * invLn2; }   // fastlog: math.h:175-199
    V3 eval(P2 uv, P2 d0, P2 d1) const {
        if (filter == MTSB_TEXFILTER_NEAREST) return evalBox(0, uv);
        else if (filter == MTSB_TEXFILTER_BILINEAR) return evalBilinear(0, uv);
        float du0 = d0.x * w[0], dv0 = d0.y * h[0], du1 = d1.x * w[0], dv1 = d1.y * h[0];
        float A = dv0 * dv0 + dv1 * dv1, B = -2.0f * (du0 * dv0 + du1 * dv1), C = du0 * du0 + du1 * du1, F = A * C - B * B * 0.25f;
        float root = hypot2(A - C, B), Aprime = 0.5f * (A + C - root), Cprime = 0.5f * (A + C + root),
              majorRadius = Aprime != 0 ? std::sqrt(F / Aprime) : 0, minorRadius = Cprime != 0 ? std::sqrt(F / Cprime) : 0;
        if (filter == MTSB_TEXFILTER_TRILINEAR || !(minorRadius > 0) || !(majorRadius > 0) || F < 0) {
            float level = log2f_(std::max(majorRadius, kEpsilon));
            int ilevel = (int) std::floor(level);
            if (ilevel < 0) return evalBilinear(0, uv);
            float a = level - ilevel;
            return evalBilinear(ilevel, uv) * (1.0f - a) + evalBilinear(ilevel + 1, uv) * a;
        }
        if (minorRadius * maxAnisotropy < majorRadius) {
            minorRadius = majorRadius / maxAnisotropy;
            float theta = 0.5f * std::atan(B / (A - C)), sinTheta, cosTheta;
            sincosf(theta, &sinTheta, &cosTheta);
            float a2 = majorRadius * majorRadius, b2 = minorRadius * minorRadius, sinTheta2 = sinTheta * sinTheta,
                  cosTheta2 = cosTheta * cosTheta, sin2Theta = 2 * sinTheta * cosTheta;
            A = a2 * cosTheta2 + b2 * sinTheta2; B = (a2 - b2) * sin2Theta; C = a2 * sinTheta2 + b2 * cosTheta2; F = a2 * b2;
        }
        float scale = 1.0f / F;
        A *= scale; B *= scale; C *= scale;
        float level = std::max(0.0f, log2f_(minorRadius));
        int ilevel = (int) level;
        float a = level - ilevel;
        if (majorRadius < 1 || !(A > 0 && C > 0)) return evalBilinear(ilevel, uv);
        return evalEWA(ilevel, uv, A, B, C) * (1.0f - a) + evalEWA(ilevel + 1, uv, A, B, C) * a;
    }
};

struct UVPartials { bool has; float dudx, dvdx, dudy, dvdy; UVPartials() : has(false), dudx(0), dvdx(0), dudy(0), dvdy(0) {} };

struct Texture {
    MtsbTexture rec;
    MipMap mip;
    void build(const MtsbTexture &t) { rec = t; if (t.type == MTSB_TEX_BITMAP) mip.build(t); rec.texels = nullptr; }
    V3 evalUV(P2 uv) const {            // <Texture>::eval(const Point2 &uv)
        if (rec.type == MTSB_TEX_CHECKERBOARD) {
            int x = 2 * texModulo((int) (uv.x * 2), 2) - 1, y = 2 * texModulo((int) (uv.y * 2), 2) - 1;
            return x * y == 1 ? rgb(rec.color0) : rgb(rec.color1);
        } else if (rec.type == MTSB_TEX_GRID) {
            float x = uv.x - (int) std::floor(uv.x), y = uv.y - (int) std::floor(uv.y);
            if (x > .5) x -= 1;
            if (y > .5) y -= 1;
            return (std::abs(x) < rec.line_width || std::abs(y) < rec.line_width) ? rgb(rec.color1) : rgb(rec.color0);
        }
        return rec.filter != MTSB_TEXFILTER_NEAREST ? mip.evalBilinear(0, uv) : mip.evalBox(0, uv);     // bitmap.cpp:431-453
    }
    V3 eval(P2 uvIts, const UVPartials &pt) const {      // Texture2D::eval(its, filter = true), then the scale plugin(s)
        P2 uv{uvIts.x * rec.uv_scale[0] + rec.uv_offset[0], uvIts.y * rec.uv_scale[1] + rec.uv_offset[1]};
        V3 value;
        if (pt.has && rec.type == MTSB_TEX_BITMAP)
            value = mip.eval(uv, P2{pt.dudx * rec.uv_scale[0], pt.dvdx * rec.uv_scale[1]}, P2{pt.dudy * rec.uv_scale[0], pt.dvdy * rec.uv_scale[1]});
        else value = evalUV(uv);
        return value * rgb(rec.scale);
    }
};

inline bool solveLinearSystem2x2(const float a[2][2], const float b[2], float x[2]) {     // util.cpp:527-539
    float det = a[0][0] * a[1][1] - a[0][1] * a[1][0];
    if (std::abs(det) <= 2.93873587705571876e-39f) return false;
    float inverse = 1.0f / det;
    x[0] = (a[1][1] * b[0] - a[0][1] * b[1]) * inverse;
    x[1] = (a[0][0] * b[1] - a[1][0] * b[0]) * inverse;
    return true;
}

// intersection.cpp:5-98; both offset rays start at the camera position `o` (perspective.cpp:310)
inline UVPartials computePartials(V3 p, V3 n, V3 dpdu, V3 dpdv, V3 o, V3 rxd, V3 ryd) {
    UVPartials r; r.has = true;
    if (isZero(dpdu) && isZero(dpdv)) return r;
    const float pp = dot(n, p), pox = dot(n, o), poy = dot(n, o), prx = dot(n, rxd), pry = dot(n, ryd);
    if (prx == 0 || pry == 0) return r;
    const float tx = (pp - pox) / prx, ty = (pp - poy) / pry;
    float absX = std::abs(n.x), absY = std::abs(n.y), absZ = std::abs(n.z);
    int axes[2];
    if (absX > absY && absX > absZ) { axes[0] = 1; axes[1] = 2; } else if (absY > absZ) { axes[0] = 0; axes[1] = 2; } else { axes[0] = 0; axes[1] = 1; }
    auto at = [](V3 v, int i) { return i == 0 ? v.x : i == 1 ? v.y : v.z; };
    float A[2][2], Bx[2], By[2], x[2];
    A[0][0] = at(dpdu, axes[0]); A[0][1] = at(dpdv, axes[0]); A[1][0] = at(dpdu, axes[1]); A[1][1] = at(dpdv, axes[1]);
    V3 px = o + rxd * tx, py = o + ryd * ty;
    Bx[0] = at(px, axes[0]) - at(p, axes[0]); Bx[1] = at(px, axes[1]) - at(p, axes[1]);
    By[0] = at(py, axes[0]) - at(p, axes[0]); By[1] = at(py, axes[1]) - at(p, axes[1]);
    if (solveLinearSystem2x2(A, Bx, x)) { r.dudx = x[0]; r.dvdx = x[1]; } else { r.dudx = 1; r.dvdx = 0; }
    if (solveLinearSystem2x2(A, By, x)) { r.dudy = x[0]; r.dvdy = x[1]; } else { r.dudy = 1; }
    return r;
}

}  // namespace orc
