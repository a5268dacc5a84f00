// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// CPU restatement of the reference's scalar math used on the `path` hot path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may call into oracle/.  Every function cites the reference lines it follows
// (paths relative to the reference tree).
#pragma once
#include <cmath>
#include <cstdint>
#include <algorithm>
#include <limits>

namespace orc {

// include/mitsuba/core/constants.h:28-31,62-68 (single precision build)
static const float kEpsilon = 1e-4f;
static const float kShadowEpsilon = 1e-3f;
static const float kPi = 3.14159265358979323846f;
static const float kInvPi = 0.31830988618379067154f;
static const float kInvTwoPi = 0.15915494309189533577f;
static const float kInvFourPi = 0.07957747154594766788f;
static const float kInf = std::numeric_limits<float>::infinity();

struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit V3(float a) : x(a), y(a), z(a) {}
    float operator[](int i) const { return (&x)[i]; }
    float &operator[](int i) { return (&x)[i]; }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, V3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
// TVector3::operator/(Float) multiplies by the reciprocal (include/mitsuba/core/vector.h)
inline V3 operator/(V3 a, float s) { float r = 1.0f / s; return V3(a.x * r, a.y * r, a.z * r); }
inline V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
inline V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
inline V3 &operator*=(V3 &a, float s) { a = a * s; return a; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float absDot(V3 a, V3 b) { return std::fabs(dot(a, b)); }
inline V3 cross(V3 a, V3 b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float lengthSquared(V3 a) { return dot(a, a); }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a / length(a); }
inline bool isZero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }
inline float maxComp(V3 a) { return std::max(a.x, std::max(a.y, a.z)); }
inline bool finite3(V3 a) { return std::isfinite(a.x) && std::isfinite(a.y) && std::isfinite(a.z); }
// Spectrum::getLuminance for linear RGB, include/mitsuba/core/spectrum.h:724-727
inline float luminance(V3 c) { return 0.212671f * c.x + 0.715160f * c.y + 0.072169f * c.z; }

inline float safe_sqrt(float v) { return std::sqrt(std::max(0.0f, v)); }   // core/math.h safe_sqrt
inline float signum(float v) { return std::copysign(1.0f, v); }            // core/math.h:269-278 (never zero)

// src/libcore/math.cpp:70-82
inline float hypot2(float a, float b) {
    float r;
    if (std::fabs(a) > std::fabs(b)) { r = b / a; r = std::fabs(a) * std::sqrt(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = std::fabs(b) * std::sqrt(1.0f + r * r); }
    else r = 0.0f;
    return r;
}

// src/libcore/math.cpp:25-52 (Giles' single-precision polynomial; fastlog == logf, core/math.h:201-215)
inline float erfinv(float x) {
    float w = -std::log((1.0f - x) * (1.0f + x));
    float p;
    if (w < 5.0f) {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = 3.43273939e-07f + p * w;
        p = -3.5233877e-06f + p * w;
        p = -4.39150654e-06f + p * w;
        p = 0.00021858087f + p * w;
        p = -0.00125372503f + p * w;
        p = -0.00417768164f + p * w;
        p = 0.246640727f + p * w;
        p = 1.50140941f + p * w;
    } else {
        w = std::sqrt(w) - 3.0f;
        p = -0.000200214257f;
        p = 0.000100950558f + p * w;
        p = 0.00134934322f + p * w;
        p = -0.00367342844f + p * w;
        p = 0.00573950773f + p * w;
        p = -0.0076224613f + p * w;
        p = 0.00943887047f + p * w;
        p = 1.00167406f + p * w;
        p = 2.83297682f + p * w;
    }
    return p * x;
}

// src/libcore/math.cpp:54-68 (Abramowitz-Stegun 7.1.26)
inline float erf_as(float x) {
    const float a1 = 0.254829592f, a2 = -0.284496736f, a3 = 1.421413741f,
                a4 = -1.453152027f, a5 = 1.061405429f, p = 0.3275911f;
    float sign = signum(x);
    x = std::fabs(x);
    float t = 1.0f / (1.0f + p * x);
    float y = 1.0f - (((((a5 * t + a4) * t) + a3) * t + a2) * t + a1) * t * std::exp(-x * x);
    return sign * y;
}

// ---- Frame: include/mitsuba/core/frame.h ---------------------------------
struct Frame {
    V3 s, t, n;
    V3 toLocal(V3 v) const { return V3(dot(v, s), dot(v, t), dot(v, n)); }
    V3 toWorld(V3 v) const { return s * v.x + t * v.y + n * v.z; }
};
// src/libcore/util.cpp:592-601
inline void coordinateSystem(V3 a, V3 &b, V3 &c) {
    if (std::fabs(a.x) > std::fabs(a.y)) {
        float invLen = 1.0f / std::sqrt(a.x * a.x + a.z * a.z);
        c = V3(a.z * invLen, 0.0f, -a.x * invLen);
    } else {
        float invLen = 1.0f / std::sqrt(a.y * a.y + a.z * a.z);
        c = V3(0.0f, a.z * invLen, -a.y * invLen);
    }
    b = cross(c, a);
}
inline Frame frameFromNormal(V3 n) { Frame f; f.n = n; coordinateSystem(n, f.s, f.t); return f; }
// src/libcore/util.cpp:603-608
inline Frame shadingFrame(V3 n, V3 dpdu) {
    Frame f; f.n = n;
    f.s = normalize(dpdu - n * dot(n, dpdu));
    f.t = cross(n, f.s);
    return f;
}
inline float cosTheta(V3 v) { return v.z; }
inline float cosTheta2(V3 v) { return v.z * v.z; }
inline float sinTheta2(V3 v) { return 1.0f - v.z * v.z; }
// frame.h tanTheta: temp = 1 - z^2; if (temp <= 0) return 0; return sqrt(temp)/z
inline float tanTheta(V3 v) { float t = 1.0f - v.z * v.z; if (t <= 0.0f) return 0.0f; return std::sqrt(t) / v.z; }

// ---- warp: src/libcore/warp.cpp ------------------------------------------
struct P2 { float x, y; };
inline void sincosf_(float a, float *s, float *c) { *s = std::sin(a); *c = std::cos(a); }
// warp.cpp:81-101
inline P2 squareToUniformDiskConcentric(P2 u) {
    float r1 = 2.0f * u.x - 1.0f, r2 = 2.0f * u.y - 1.0f;
    float phi, r;
    if (r1 == 0 && r2 == 0) { r = phi = 0; }
    else if (r1 * r1 > r2 * r2) { r = r1; phi = (kPi / 4.0f) * (r2 / r1); }
    else { r = r2; phi = (kPi / 2.0f) - (r1 / r2) * (kPi / 4.0f); }
    float s, c; sincosf_(phi, &s, &c);
    return P2{r * c, r * s};
}
// warp.cpp:43-52
inline V3 squareToCosineHemisphere(P2 u) {
    P2 p = squareToUniformDiskConcentric(u);
    float z = safe_sqrt(1.0f - p.x * p.x - p.y * p.y);
    if (z == 0) z = 1e-10f;
    return V3(p.x, p.y, z);
}
inline float squareToCosineHemispherePdf(V3 d) { return kInvPi * d.z; }   // warp.h:55
// warp.cpp:24-30
inline V3 squareToUniformSphere(P2 u) {
    float z = 1.0f - 2.0f * u.y;
    float r = safe_sqrt(1.0f - z * z);
    float s, c; sincosf_(2.0f * kPi * u.x, &s, &c);
    return V3(r * c, r * s, z);
}
// warp.cpp:76-79
inline P2 squareToUniformTriangle(P2 u) {
    float a = safe_sqrt(1.0f - u.x);
    return P2{1 - a, a * u.y};
}

// ---- Fresnel / reflect / refract: src/libcore/util.cpp -------------------
// util.cpp:651-681
inline float fresnelDielectricExt(float cosThetaI_, float &cosThetaT_, float eta) {
    if (eta == 1) { cosThetaT_ = -cosThetaI_; return 0.0f; }
    float scale = (cosThetaI_ > 0) ? 1 / eta : eta;
    float cosThetaTSqr = 1 - (1 - cosThetaI_ * cosThetaI_) * (scale * scale);
    if (cosThetaTSqr <= 0.0f) { cosThetaT_ = 0.0f; return 1.0f; }
    float cosThetaI = std::fabs(cosThetaI_);
    float cosThetaT = std::sqrt(cosThetaTSqr);
    float Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
    float Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
    cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
    return 0.5f * (Rs * Rs + Rp * Rp);
}
inline float fresnelDielectricExt(float cosThetaI, float eta) { float t; return fresnelDielectricExt(cosThetaI, t, eta); }

// util.cpp:715-737 (scalar), :739-761 (per channel) -- identical arithmetic per channel
inline float fresnelConductorExact(float cosThetaI, float eta, float k) {
    float cosThetaI2 = cosThetaI * cosThetaI, sinThetaI2 = 1 - cosThetaI2, sinThetaI4 = sinThetaI2 * sinThetaI2;
    float temp1 = eta * eta - k * k - sinThetaI2;
    float a2pb2 = safe_sqrt(temp1 * temp1 + 4 * k * k * eta * eta);
    float a = safe_sqrt(0.5f * (a2pb2 + temp1));
    float term1 = a2pb2 + cosThetaI2, term2 = 2 * a * cosThetaI;
    float Rs2 = (term1 - term2) / (term1 + term2);
    float term3 = a2pb2 * cosThetaI2 + sinThetaI4, term4 = term2 * sinThetaI2;
    float Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
    return 0.5f * (Rp2 + Rs2);
}
// Spectrum overload util.cpp:739-761: note `k*k*eta*eta*4` and `a*(2*cosThetaI)` ordering
inline V3 fresnelConductorExact(float cosThetaI, V3 eta, V3 k) {
    float cosThetaI2 = cosThetaI * cosThetaI, sinThetaI2 = 1 - cosThetaI2, sinThetaI4 = sinThetaI2 * sinThetaI2;
    V3 r;
    for (int c = 0; c < 3; ++c) {
        float e = eta[c], kk = k[c];
        float temp1 = e * e - kk * kk - sinThetaI2;
        float a2pb2 = safe_sqrt(temp1 * temp1 + kk * kk * e * e * 4);
        float a = safe_sqrt((a2pb2 + temp1) * 0.5f);
        float term1 = a2pb2 + cosThetaI2, term2 = a * (2 * cosThetaI);
        float Rs2 = (term1 - term2) / (term1 + term2);
        float term3 = a2pb2 * cosThetaI2 + sinThetaI4, term4 = term2 * sinThetaI2;
        float Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
        r[c] = 0.5f * (Rp2 + Rs2);
    }
    return r;
}
// util.cpp:763-765
inline V3 reflect(V3 wi, V3 n) { return 2 * dot(wi, n) * n - wi; }
// util.cpp:767-772
inline V3 refract(V3 wi, V3 n, float eta, float cosThetaT) {
    if (cosThetaT < 0) eta = 1 / eta;
    return n * (dot(wi, n) * eta + cosThetaT) - wi * eta;
}

// src/libcore/util.cpp:447-485
inline bool solveQuadratic(float a, float b, float c, float &x0, float &x1) {
    if (a == 0) { if (b != 0) { x0 = x1 = -c / b; return true; } return false; }
    float discrim = b * b - 4.0f * a * c;
    if (discrim < 0) return false;
    float temp, sqrtDiscrim = std::sqrt(discrim);
    if (b < 0) temp = -0.5f * (b - sqrtDiscrim); else temp = -0.5f * (b + sqrtDiscrim);
    x0 = temp / a; x1 = c / temp;
    if (x0 > x1) std::swap(x0, x1);
    return true;
}

}  // namespace orc
