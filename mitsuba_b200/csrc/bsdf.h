// bsdf.h -- device BSDF models of the `path` hot path.
//
//   MicrofacetDistribution (Beckmann / GGX)   reference src/bsdfs/microfacet.h
//   SmoothDiffuse                             reference src/bsdfs/diffuse.cpp:110-165
//   RoughConductor                            reference src/bsdfs/roughconductor.cpp:259-417
//   RoughDielectric                           reference src/bsdfs/roughdielectric.cpp:270-615
//   TwoSided adapter                          reference src/bsdfs/twosided.cpp:108-184
//   Fresnel / reflect / refract               reference src/libcore/util.cpp:651-772
//   warp::*                                   reference src/libcore/warp.cpp:24-101
//   math::erf / erfinv / hypot2               reference src/libcore/math.cpp:25-82
// The polynomial erf / erfinv and every numeric guard (1e-20, 1e-4, 0.99999, 1e-6) are the
// reference's own, so pdf <-> sample consistency holds at the same tolerances.
#pragma once
#include "hd.h"
#include "../../include/mtsb200.h"

namespace mtsb {

MTSB_HD float safe_sqrt(float v) { return sqrtf(fmaxf(0.0f, v)); }
MTSB_HD float signum(float v) { return copysignf(1.0f, v); }
MTSB_HD void sincos_(float a, float *s, float *c) {
#if MTSB_DEVICE_CODE
    sincosf(a, s, c);
#else
    *s = sinf(a); *c = cosf(a);
#endif
}
MTSB_HD float cosTheta(V3 v) { return v.z; }
MTSB_HD float cosTheta2(V3 v) { return v.z * v.z; }
MTSB_HD float sinTheta2(V3 v) { return 1.0f - v.z * v.z; }
MTSB_HD float tanTheta(V3 v) { float t = 1.0f - v.z * v.z; if (t <= 0.0f) return 0.0f; return sqrtf(t) / v.z; }

// math.cpp:70-82
MTSB_HD float hypot2(float a, float b) {
    float r;
    if (fabsf(a) > fabsf(b)) { r = b / a; r = fabsf(a) * sqrtf(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = fabsf(b) * sqrtf(1.0f + r * r); }
    else r = 0.0f;
    return r;
}
// math.cpp:25-52
MTSB_HD float erfinv_poly(float x) {
    float w = -logf((1.0f - x) * (1.0f + x));
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
        w = sqrtf(w) - 3.0f;
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
// math.cpp:54-68
MTSB_HD float erf_poly(float x) {
    const float a1 = 0.254829592f, a2 = -0.284496736f, a3 = 1.421413741f, a4 = -1.453152027f, a5 = 1.061405429f, p = 0.3275911f;
    float sign = signum(x);
    x = fabsf(x);
    float t = 1.0f / (1.0f + p * x);
    float y = 1.0f - (((((a5 * t + a4) * t) + a3) * t + a2) * t + a1) * t * expf(-x * x);
    return sign * y;
}

// ---- warp.cpp ---------------------------------------------------------------------
MTSB_HD P2 squareToUniformDiskConcentric(P2 u) {          // :81-101
    float r1 = 2.0f * u.x - 1.0f, r2 = 2.0f * u.y - 1.0f;
    float phi, r;
    if (r1 == 0 && r2 == 0) { r = phi = 0; }
    else if (r1 * r1 > r2 * r2) { r = r1; phi = (MTSB_PI / 4.0f) * (r2 / r1); }
    else { r = r2; phi = (MTSB_PI / 2.0f) - (r1 / r2) * (MTSB_PI / 4.0f); }
    float s, c; sincos_(phi, &s, &c);
    P2 o; o.x = r * c; o.y = r * s; return o;
}
MTSB_HD V3 squareToCosineHemisphere(P2 u) {                // :43-52
    P2 p = squareToUniformDiskConcentric(u);
    float z = safe_sqrt(1.0f - p.x * p.x - p.y * p.y);
    if (z == 0) z = 1e-10f;
    return V3(p.x, p.y, z);
}
MTSB_HD float squareToCosineHemispherePdf(V3 d) { return MTSB_INV_PI * d.z; }
MTSB_HD V3 squareToUniformSphere(P2 u) {                   // :24-30
    float z = 1.0f - 2.0f * u.y;
    float r = safe_sqrt(1.0f - z * z);
    float s, c; sincos_(2.0f * MTSB_PI * u.x, &s, &c);
    return V3(r * c, r * s, z);
}
MTSB_HD P2 squareToUniformTriangle(P2 u) {                 // :76-79
    float a = safe_sqrt(1.0f - u.x);
    P2 o; o.x = 1 - a; o.y = a * u.y; return o;
}

// ---- util.cpp Fresnel -----------------------------------------------------------------
MTSB_HD float fresnelDielectricExt(float cosThetaI_, float &cosThetaT_, float eta) {      // :651-681
    if (eta == 1) { cosThetaT_ = -cosThetaI_; return 0.0f; }
    float scale = (cosThetaI_ > 0) ? 1 / eta : eta;
    float cosThetaTSqr = 1 - (1 - cosThetaI_ * cosThetaI_) * (scale * scale);
    if (cosThetaTSqr <= 0.0f) { cosThetaT_ = 0.0f; return 1.0f; }
    float cosThetaI = fabsf(cosThetaI_);
    float cosThetaT = sqrtf(cosThetaTSqr);
    float Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
    float Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
    cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
    return 0.5f * (Rs * Rs + Rp * Rp);
}
MTSB_HD V3 fresnelConductorExact(float cosThetaI, V3 eta, V3 k) {                         // :739-761
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
        r.set(c, 0.5f * (Rp2 + Rs2));
    }
    return r;
}
MTSB_HD V3 reflect(V3 wi, V3 n) { return 2 * dot(wi, n) * n - wi; }                        // :763-765
MTSB_HD V3 refract(V3 wi, V3 n, float eta, float cosThetaT) {                              // :767-772
    if (cosThetaT < 0) eta = 1 / eta;
    return n * (dot(wi, n) * eta + cosThetaT) - wi * eta;
}

// ---- MicrofacetDistribution -------------------------------------------------------------
struct Microfacet {
    int type; float au, av; bool visible;
    MTSB_HD Microfacet(int t, float alphaU, float alphaV, bool sampleVisible)
        : type(t), au(fmaxf(alphaU, 1e-4f)), av(fmaxf(alphaV, 1e-4f)), visible(sampleVisible) {}      // microfacet.h:86-93
    MTSB_HD bool isotropic() const { return au == av; }
    MTSB_HD void scaleAlpha(float v) { au *= v; av *= v; }                                              // :178-183

    MTSB_HD float eval(V3 m) const {                                                                    // :191-234
        if (cosTheta(m) <= 0) return 0.0f;
        float ct2 = cosTheta2(m);
        float bexp = ((m.x * m.x) / (au * au) + (m.y * m.y) / (av * av)) / ct2;
        float result;
        if (type == MTSB_DISTR_BECKMANN) result = expf(-bexp) / (MTSB_PI * au * av * ct2 * ct2);
        else { float root = (1.0f + bexp) * ct2; result = 1.0f / (MTSB_PI * au * av * root * root); }
        if (result * cosTheta(m) < 1e-20f) result = 0;
        return result;
    }
    MTSB_HD float projectRoughness(V3 v) const {                                                        // :549-559
        float invSinTheta2 = 1 / sinTheta2(v);
        if (isotropic() || invSinTheta2 <= 0) return au;
        float cosPhi2 = v.x * v.x * invSinTheta2, sinPhi2 = v.y * v.y * invSinTheta2;
        return sqrtf(cosPhi2 * au * au + sinPhi2 * av * av);
    }
    MTSB_HD float smithG1(V3 v, V3 m) const {                                                           // :477-514
        if (dot(v, m) * cosTheta(v) <= 0) return 0.0f;
        float tanT = fabsf(tanTheta(v));
        if (tanT == 0.0f) return 1.0f;
        float alpha = projectRoughness(v);
        if (type == MTSB_DISTR_BECKMANN) {
            float a = 1.0f / (alpha * tanT);
            if (a >= 1.6f) return 1.0f;
            float aSqr = a * a;
            return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
        }
        float root = alpha * tanT;
        return 2.0f / (1.0f + hypot2(1.0f, root));
    }
    MTSB_HD float G(V3 wi, V3 wo, V3 m) const { return smithG1(wi, m) * smithG1(wo, m); }               // :527-530

    MTSB_HD V3 sampleAll(P2 u, float &pdf) const {                                                      // :287-395
        float cosThetaM = 0.0f, sinPhiM, cosPhiM, alphaSqr;
        if (isotropic()) {
            sincos_((2.0f * MTSB_PI) * u.y, &sinPhiM, &cosPhiM);
            alphaSqr = au * au;
        } else {
            float phiM = atanf(av / au * tanf(MTSB_PI + 2 * MTSB_PI * u.y)) + MTSB_PI * floorf(2 * u.y + 0.5f);
            sincos_(phiM, &sinPhiM, &cosPhiM);
            float cosSc = cosPhiM / au, sinSc = sinPhiM / av;
            alphaSqr = 1.0f / (cosSc * cosSc + sinSc * sinSc);
        }
        if (type == MTSB_DISTR_BECKMANN) {
            float tanThetaMSqr = alphaSqr * -logf(1.0f - u.x);
            cosThetaM = 1.0f / sqrtf(1.0f + tanThetaMSqr);
            pdf = (1.0f - u.x) / (MTSB_PI * au * av * cosThetaM * cosThetaM * cosThetaM);
        } else {
            float tanThetaMSqr = alphaSqr * u.x / (1.0f - u.x);
            cosThetaM = 1.0f / sqrtf(1.0f + tanThetaMSqr);
            float temp = 1 + tanThetaMSqr / alphaSqr;
            pdf = MTSB_INV_PI / (au * av * cosThetaM * cosThetaM * cosThetaM * temp * temp);
        }
        if (pdf < 1e-20f) pdf = 0;
        float sinThetaM = sqrtf(fmaxf(0.0f, 1 - cosThetaM * cosThetaM));
        return V3(sinThetaM * cosPhiM, sinThetaM * sinPhiM, cosThetaM);
    }
    MTSB_HD float pdfAll(V3 m) const { return eval(m) * cosTheta(m); }                                  // :404-407

    MTSB_HD void sampleVisible11(float thetaI, P2 u, float &sx, float &sy) const {                      // :581-705
        const float SQRT_PI_INV = 0.56418958354775628695f;
        if (type == MTSB_DISTR_BECKMANN) {
            if (thetaI < 1e-4f) {
                float s, c, r = sqrtf(-logf(1.0f - u.x));
                sincos_(2 * MTSB_PI * u.y, &s, &c);
                sx = r * c; sy = r * s; return;
            }
            float tanThetaI = tanf(thetaI), cotThetaI = 1 / tanThetaI;
            float a = -1, c = erf_poly(cotThetaI);
            float sample_x = fmaxf(u.x, 1e-6f);
            float fit = 1 + thetaI * (-0.876f + thetaI * (0.4265f - 0.0594f * thetaI));
            float b = c - (1 + c) * powf(1 - sample_x, fit);
            float normalization = 1 / (1 + c + SQRT_PI_INV * tanThetaI * expf(-cotThetaI * cotThetaI));
            int it = 0;
            while (++it < 10) {
                if (!(b >= a && b <= c)) b = 0.5f * (a + c);
                float invErf = erfinv_poly(b);
                float value = normalization * (1 + b + SQRT_PI_INV * tanThetaI * expf(-invErf * invErf)) - sample_x;
                float derivative = normalization * (1 - invErf * tanThetaI);
                if (fabsf(value) < 1e-5f) break;
                if (value > 0) c = b; else a = b;
                b -= value / derivative;
            }
            sx = erfinv_poly(b);
            sy = erfinv_poly(2.0f * fmaxf(u.y, 1e-6f) - 1.0f);
        } else {
            if (thetaI < 1e-4f) {
                float s, c, r = safe_sqrt(u.x / (1 - u.x));
                sincos_(2 * MTSB_PI * u.y, &s, &c);
                sx = r * c; sy = r * s; return;
            }
            float tanThetaI = tanf(thetaI);
            float a = 1 / tanThetaI;
            float G1 = 2.0f / (1.0f + safe_sqrt(1.0f + 1.0f / (a * a)));
            float A = 2.0f * u.x / G1 - 1.0f;
            if (fabsf(A) == 1) A -= signum(A) * MTSB_EPSILON;
            float tmp = 1.0f / (A * A - 1.0f);
            float B = tanThetaI;
            float D = safe_sqrt(B * B * tmp * tmp - (A * A - B * B) * tmp);
            float slope_x_1 = B * tmp - D, slope_x_2 = B * tmp + D;
            sx = (A < 0.0f || slope_x_2 > 1.0f / tanThetaI) ? slope_x_1 : slope_x_2;
            float S, y = u.y;
            if (y > 0.5f) { S = 1.0f; y = 2.0f * (y - 0.5f); }
            else { S = -1.0f; y = 2.0f * (0.5f - y); }
            float z = (y * (y * (y * (-0.365728915865723f) + 0.790235037209296f) - 0.424965825137544f) + 0.000152998850436920f) /
                      (y * (y * (y * (y * 0.169507819808272f - 0.397203533833404f) - 0.232500544458471f) + 1.0f) - 0.539825872510702f);
            sy = S * z * sqrtf(1.0f + sx * sx);
        }
    }
    MTSB_HD V3 sampleVisible(V3 _wi, P2 u) const {                                                      // :421-459
        V3 wi = normalize(V3(au * _wi.x, av * _wi.y, _wi.z));
        float theta = 0, phi = 0;
        if (wi.z < 0.99999f) { theta = acosf(wi.z); phi = atan2f(wi.y, wi.x); }
        float sinPhi, cosPhi; sincos_(phi, &sinPhi, &cosPhi);
        float sx, sy; sampleVisible11(theta, u, sx, sy);
        float rx = cosPhi * sx - sinPhi * sy, ry = sinPhi * sx + cosPhi * sy;
        rx *= au; ry *= av;
        float normalization = 1.0f / sqrtf(rx * rx + ry * ry + 1.0f);
        return V3(-rx * normalization, -ry * normalization, normalization);
    }
    MTSB_HD float pdfVisible(V3 wi, V3 m) const {                                                       // :462-466
        if (cosTheta(wi) == 0) return 0.0f;
        return smithG1(wi, m) * absDot(wi, m) * eval(m) / fabsf(cosTheta(wi));
    }
    MTSB_HD V3 sample(V3 wi, P2 u, float &pdf) const {                                                  // :240-249
        if (visible) { V3 m = sampleVisible(wi, u); pdf = pdfVisible(wi, m); return m; }
        return sampleAll(u, pdf);
    }
    MTSB_HD float pdf(V3 wi, V3 m) const { return visible ? pdfVisible(wi, m) : pdfAll(m); }            // :270-275
};

struct BsdfSample { V3 wo; V3 weight; float pdf; float eta; bool delta; };

MTSB_HD Microfacet distr_of(const MtsbMaterial &m) {
    return Microfacet((int) m.distribution, m.alpha_u, m.alpha_v, (m.flags & MTSB_MAT_SAMPLE_VISIBLE) != 0);
}

// ---- SmoothDiffuse ---------------------------------------------------------------------
MTSB_HD V3 diffuse_eval(const MtsbMaterial &m, V3 wi, V3 wo) {
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return V3(0.0f);
    return ld3(m.reflectance) * (MTSB_INV_PI * cosTheta(wo));
}
MTSB_HD float diffuse_pdf(V3 wi, V3 wo) {
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
    return squareToCosineHemispherePdf(wo);
}
MTSB_HD bool diffuse_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {
    if (cosTheta(wi) <= 0) return false;
    s.wo = squareToCosineHemisphere(u);
    s.eta = 1.0f; s.delta = false;
    s.pdf = squareToCosineHemispherePdf(s.wo);
    s.weight = ld3(m.reflectance);
    return true;
}

// ---- RoughConductor ----------------------------------------------------------------------
MTSB_HD V3 roughconductor_eval(const MtsbMaterial &m, V3 wi, V3 wo) {                      // :259-296
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return V3(0.0f);
    V3 H = normalize(wo + wi);
    Microfacet distr = distr_of(m);
    const float D = distr.eval(H);
    if (D == 0) return V3(0.0f);
    const V3 F = fresnelConductorExact(dot(wi, H), ld3(m.cond_eta), ld3(m.cond_k)) * ld3(m.reflectance);
    const float G = distr.G(wi, wo, H);
    float model = D * G / (4.0f * cosTheta(wi));
    return F * model;
}
MTSB_HD float roughconductor_pdf(const MtsbMaterial &m, V3 wi, V3 wo) {                    // :298-323
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
    V3 H = normalize(wo + wi);
    Microfacet distr = distr_of(m);
    if (distr.visible) return distr.eval(H) * distr.smithG1(wi, H) / (4.0f * cosTheta(wi));
    return distr.pdf(wi, H) / (4 * absDot(wo, H));
}
MTSB_HD bool roughconductor_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {    // :371-417
    if (cosTheta(wi) < 0) return false;          // `<`, not `<=` in the sample-with-pdf overload (:372)
    Microfacet distr = distr_of(m);
    float pdf;
    V3 mm = distr.sample(wi, u, pdf);
    if (pdf == 0) return false;
    s.wo = reflect(wi, mm);
    s.eta = 1.0f; s.delta = false;
    if (cosTheta(s.wo) <= 0) return false;
    V3 F = fresnelConductorExact(dot(wi, mm), ld3(m.cond_eta), ld3(m.cond_k)) * ld3(m.reflectance);
    float weight;
    if (distr.visible) weight = distr.smithG1(s.wo, mm);
    else weight = distr.eval(mm) * distr.G(wi, s.wo, mm) * dot(wi, mm) / (pdf * cosTheta(wi));
    pdf /= 4.0f * dot(s.wo, mm);
    s.pdf = pdf;
    s.weight = F * weight;
    return true;
}

// ---- RoughDielectric (radiance transport mode) -----------------------------------------------
MTSB_HD V3 roughdielectric_eval(const MtsbMaterial &m, V3 wi, V3 wo) {                     // :270-348
    if (cosTheta(wi) == 0) return V3(0.0f);
    const float m_eta = m.eta, m_invEta = 1 / m.eta;
    bool refl = cosTheta(wi) * cosTheta(wo) > 0;
    V3 H;
    if (refl) H = normalize(wo + wi);
    else { float eta = cosTheta(wi) > 0 ? m_eta : m_invEta; H = normalize(wi + wo * eta); }
    H *= signum(cosTheta(H));
    Microfacet distr = distr_of(m);
    const float D = distr.eval(H);
    if (D == 0) return V3(0.0f);
    float cosThetaT;
    const float F = fresnelDielectricExt(dot(wi, H), cosThetaT, m_eta);
    const float G = distr.G(wi, wo, H);
    if (refl) {
        float value = F * D * G / (4.0f * fabsf(cosTheta(wi)));
        return ld3(m.reflectance) * value;
    }
    float eta = cosTheta(wi) > 0.0f ? m_eta : m_invEta;
    float sqrtDenom = dot(wi, H) + eta * dot(wo, H);
    float value = ((1 - F) * D * G * eta * eta * dot(wi, H) * dot(wo, H)) / (cosTheta(wi) * sqrtDenom * sqrtDenom);
    float factor = cosTheta(wi) > 0 ? m_invEta : m_eta;
    return ld3(m.transmittance) * fabsf(value * factor * factor);
}
MTSB_HD float roughdielectric_pdf(const MtsbMaterial &m, V3 wi, V3 wo) {                   // :350-422
    const float m_eta = m.eta, m_invEta = 1 / m.eta;
    bool refl = cosTheta(wi) * cosTheta(wo) > 0;
    V3 H; float dwh_dwo;
    if (refl) {
        H = normalize(wo + wi);
        dwh_dwo = 1.0f / (4.0f * dot(wo, H));
    } else {
        float eta = cosTheta(wi) > 0 ? m_eta : m_invEta;
        H = normalize(wi + wo * eta);
        float sqrtDenom = dot(wi, H) + eta * dot(wo, H);
        dwh_dwo = (eta * eta * dot(wo, H)) / (sqrtDenom * sqrtDenom);
    }
    H *= signum(cosTheta(H));
    Microfacet sampleDistr = distr_of(m);
    if (!sampleDistr.visible) sampleDistr.scaleAlpha(1.2f - 0.2f * sqrtf(fabsf(cosTheta(wi))));
    float prob = sampleDistr.pdf(signum(cosTheta(wi)) * wi, H);
    float cosThetaT;
    float F = fresnelDielectricExt(dot(wi, H), cosThetaT, m_eta);
    prob *= refl ? F : (1 - F);
    return fabsf(prob * dwh_dwo);
}
// `u1` is the extra sampler->next1D() of roughdielectric.cpp:554
MTSB_HD bool roughdielectric_sample(const MtsbMaterial &m, V3 wi, P2 u, float u1, BsdfSample &s) {   // :513-615
    const float m_eta = m.eta, m_invEta = 1 / m.eta;
    Microfacet distr = distr_of(m);
    Microfacet sampleDistr = distr;
    if (!distr.visible) sampleDistr.scaleAlpha(1.2f - 0.2f * sqrtf(fabsf(cosTheta(wi))));
    float microfacetPDF;
    const V3 mm = sampleDistr.sample(signum(cosTheta(wi)) * wi, u, microfacetPDF);
    if (microfacetPDF == 0) return false;
    float pdf = microfacetPDF;
    float cosThetaT;
    float F = fresnelDielectricExt(dot(wi, mm), cosThetaT, m_eta);
    V3 weight(1.0f);
    bool sampleReflection = true;
    if (u1 > F) { sampleReflection = false; pdf *= 1 - F; } else pdf *= F;
    float dwh_dwo;
    s.delta = false;
    if (sampleReflection) {
        s.wo = reflect(wi, mm);
        s.eta = 1.0f;
        if (cosTheta(wi) * cosTheta(s.wo) <= 0) return false;
        weight *= ld3(m.reflectance);
        dwh_dwo = 1.0f / (4.0f * dot(s.wo, mm));
    } else {
        if (cosThetaT == 0) return false;
        s.wo = refract(wi, mm, m_eta, cosThetaT);
        s.eta = cosThetaT < 0 ? m_eta : m_invEta;
        if (cosTheta(wi) * cosTheta(s.wo) >= 0) return false;
        float factor = cosThetaT < 0 ? m_invEta : m_eta;
        weight *= ld3(m.transmittance) * (factor * factor);
        float sqrtDenom = dot(wi, mm) + s.eta * dot(s.wo, mm);
        dwh_dwo = (s.eta * s.eta * dot(s.wo, mm)) / (sqrtDenom * sqrtDenom);
    }
    if (distr.visible) weight *= distr.smithG1(s.wo, mm);
    else weight *= fabsf(distr.eval(mm) * distr.G(wi, s.wo, mm) * dot(wi, mm) / (microfacetPDF * cosTheta(wi)));
    pdf *= fabsf(dwh_dwo);
    s.pdf = pdf;
    s.weight = weight;
    return true;
}

// ---- SmoothDielectric (src/bsdfs/dielectric.cpp, mode == ERadiance) ----------------------------
// Both lobes are Dirac deltas: eval()/pdf() under the solid-angle measure of `path` are zero
// (:224-226, :256-259), so only sample() exists here.
MTSB_HD bool dielectric_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {           // :277-301
    const float m_eta = m.eta, m_invEta = 1 / m.eta;
    float cosThetaT;
    const float F = fresnelDielectricExt(cosTheta(wi), cosThetaT, m_eta);
    s.delta = true;
    if (u.x <= F) {
        s.wo = V3(-wi.x, -wi.y, wi.z);                                                        // :206-208
        s.eta = 1.0f;
        s.pdf = F;
        s.weight = ld3(m.reflectance);
    } else {
        const float scale = -(cosThetaT < 0 ? m_invEta : m_eta);                              // :211-214
        s.wo = V3(scale * wi.x, scale * wi.y, cosThetaT);
        s.eta = cosThetaT < 0 ? m_eta : m_invEta;
        s.pdf = 1 - F;
        const float factor = cosThetaT < 0 ? m_invEta : m_eta;
        s.weight = ld3(m.transmittance) * (factor * factor);
    }
    return true;
}
// ---- SmoothConductor (src/bsdfs/conductor.cpp:268-285) ----------------------------------------
MTSB_HD bool conductor_sample(const MtsbMaterial &m, V3 wi, BsdfSample &s) {
    if (cosTheta(wi) <= 0) return false;
    s.delta = true;
    s.wo = V3(-wi.x, -wi.y, wi.z);
    s.eta = 1.0f;
    s.pdf = 1.0f;
    s.weight = ld3(m.reflectance) * fresnelConductorExact(cosTheta(wi), ld3(m.cond_eta), ld3(m.cond_k));
    return true;
}

// ---- SmoothPlastic (src/bsdfs/plastic.cpp:245-440) -----------------------------------------------
// Dirac reflection lobe (specularReflectance in `transmittance`) over a diffuse base (diffuseReflectance in
// `reflectance`).  Under the solid-angle measure eval()/pdf() see only the diffuse component.
MTSB_HD float plastic_prob_specular(const MtsbMaterial &m, float Fi) {                          // :288-291
    return (Fi * m.spec_weight) / (Fi * m.spec_weight + (1 - Fi) * (1 - m.spec_weight));
}
MTSB_HD V3 plastic_diffuse_term(const MtsbMaterial &m) {                                        // :270-275
    V3 diff = ld3(m.reflectance);
    if (m.flags & MTSB_MAT_NONLINEAR) diff = diff / (V3(1.0f) - diff * m.fdr_int);
    else diff = diff / (1 - m.fdr_int);
    return diff;
}
MTSB_HD V3 plastic_eval(const MtsbMaterial &m, V3 wi, V3 wo) {                                  // :245-280
    if (cosTheta(wo) <= 0 || cosTheta(wi) <= 0) return V3(0.0f);
    float t;
    const float Fi = fresnelDielectricExt(cosTheta(wi), t, m.eta);
    const float Fo = fresnelDielectricExt(cosTheta(wo), t, m.eta);
    const float invEta2 = 1 / (m.eta * m.eta);
    return plastic_diffuse_term(m) * (squareToCosineHemispherePdf(wo) * invEta2 * (1 - Fi) * (1 - Fo));
}
MTSB_HD float plastic_pdf(const MtsbMaterial &m, V3 wi, V3 wo) {                                // :282-310
    if (cosTheta(wo) <= 0 || cosTheta(wi) <= 0) return 0.0f;
    float t;
    const float Fi = fresnelDielectricExt(cosTheta(wi), t, m.eta);
    return squareToCosineHemispherePdf(wo) * (1 - plastic_prob_specular(m, Fi));
}
MTSB_HD bool plastic_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {                // :373-412
    if (cosTheta(wi) <= 0) return false;
    float t;
    const float Fi = fresnelDielectricExt(cosTheta(wi), t, m.eta);
    s.eta = 1.0f;
    const float probSpecular = plastic_prob_specular(m, Fi);
    if (u.x < probSpecular) {
        s.delta = true;
        s.wo = V3(-wi.x, -wi.y, wi.z);
        s.pdf = probSpecular;
        s.weight = ld3(m.transmittance) * Fi / probSpecular;
    } else {
        s.delta = false;
        P2 v; v.x = (u.x - probSpecular) / (1 - probSpecular); v.y = u.y;
        s.wo = squareToCosineHemisphere(v);
        const float Fo = fresnelDielectricExt(cosTheta(s.wo), t, m.eta);
        const float invEta2 = 1 / (m.eta * m.eta);
        s.pdf = (1 - probSpecular) * squareToCosineHemispherePdf(s.wo);
        s.weight = plastic_diffuse_term(m) * (invEta2 * (1 - Fi) * (1 - Fo) / (1 - probSpecular));
    }
    return true;
}

// ---- dispatch + TwoSided ---------------------------------------------------------------------
// The per-material shading kernels call the *_one functions with a compile-time TYPE; the
// generic switch is used by the query entry point (mtsb_bsdf_query).
MTSB_HD V3 bsdf_eval_one(const MtsbMaterial &m, V3 wi, V3 wo) {
    switch (m.type) {
        case MTSB_BSDF_DIFFUSE: return diffuse_eval(m, wi, wo);
        case MTSB_BSDF_ROUGHCONDUCTOR: return roughconductor_eval(m, wi, wo);
        case MTSB_BSDF_ROUGHDIELECTRIC: return roughdielectric_eval(m, wi, wo);
        case MTSB_BSDF_PLASTIC: return plastic_eval(m, wi, wo);
        default: return V3(0.0f);
    }
}
MTSB_HD float bsdf_pdf_one(const MtsbMaterial &m, V3 wi, V3 wo) {
    switch (m.type) {
        case MTSB_BSDF_DIFFUSE: return diffuse_pdf(wi, wo);
        case MTSB_BSDF_ROUGHCONDUCTOR: return roughconductor_pdf(m, wi, wo);
        case MTSB_BSDF_ROUGHDIELECTRIC: return roughdielectric_pdf(m, wi, wo);
        case MTSB_BSDF_PLASTIC: return plastic_pdf(m, wi, wo);
        default: return 0.0f;
    }
}
MTSB_HD bool bsdf_sample_one(const MtsbMaterial &m, V3 wi, P2 u, float u1, BsdfSample &s) {
    switch (m.type) {
        case MTSB_BSDF_DIFFUSE: return diffuse_sample(m, wi, u, s);
        case MTSB_BSDF_ROUGHCONDUCTOR: return roughconductor_sample(m, wi, u, s);
        case MTSB_BSDF_ROUGHDIELECTRIC: return roughdielectric_sample(m, wi, u, u1, s);
        case MTSB_BSDF_DIELECTRIC: return dielectric_sample(m, wi, u, s);
        case MTSB_BSDF_CONDUCTOR: return conductor_sample(m, wi, s);
        case MTSB_BSDF_PLASTIC: return plastic_sample(m, wi, u, s);
        default: return false;
    }
}
// Resolves the TwoSided adapter: returns the material that actually scatters for this wi and
// whether the local frame has to be mirrored (twosided.cpp:111-121, 162-184).
MTSB_HD const MtsbMaterial &resolve_twosided_eval(const MtsbMaterial *mats, const MtsbMaterial &m, V3 &wi, V3 &wo) {
    if ((m.flags & MTSB_MAT_TWOSIDED) && !(cosTheta(wi) > 0)) {
        wi.z *= -1; wo.z *= -1;
        return m.back >= 0 ? mats[m.back] : m;
    }
    return m;
}
MTSB_HD V3 bsdf_eval(const MtsbMaterial *mats, int id, V3 wi, V3 wo) {
    const MtsbMaterial &m = resolve_twosided_eval(mats, mats[id], wi, wo);
    return bsdf_eval_one(m, wi, wo);
}
MTSB_HD float bsdf_pdf(const MtsbMaterial *mats, int id, V3 wi, V3 wo) {
    const MtsbMaterial &m = resolve_twosided_eval(mats, mats[id], wi, wo);
    return bsdf_pdf_one(m, wi, wo);
}
MTSB_HD bool bsdf_sample(const MtsbMaterial *mats, int id, V3 wi, P2 u, float u1, BsdfSample &s) {
    const MtsbMaterial &m = mats[id];
    if ((m.flags & MTSB_MAT_TWOSIDED) && cosTheta(wi) < 0) {
        wi.z *= -1;
        bool ok = bsdf_sample_one(m.back >= 0 ? mats[m.back] : m, wi, u, u1, s);
        if (ok && !isZero(s.weight) && s.pdf != 0) s.wo.z *= -1;
        return ok;
    }
    return bsdf_sample_one(m, wi, u, u1, s);
}
// the material whose model scatters for this wi (for queue binning by BSDF type)
MTSB_HD int bsdf_effective_material(const MtsbMaterial *mats, int id, float wi_z) {
    const MtsbMaterial &m = mats[id];
    if ((m.flags & MTSB_MAT_TWOSIDED) && wi_z < 0 && m.back >= 0) return m.back;
    return id;
}
// BSDF::getType() bits that path.cpp reads: ESmooth (path.cpp:177); ETransmission|EBackSide (records.inl:175-179)
// (a twosided wrapper reports the union of both nested types, twosided.cpp:72-99)
MTSB_HD bool bsdf_type_is_smooth(uint32_t t) {
    return t == MTSB_BSDF_DIFFUSE || t == MTSB_BSDF_ROUGHCONDUCTOR || t == MTSB_BSDF_ROUGHDIELECTRIC || t == MTSB_BSDF_PLASTIC;
}
MTSB_HD bool bsdf_type_is_delta(uint32_t t) { return t == MTSB_BSDF_DIELECTRIC || t == MTSB_BSDF_CONDUCTOR; }
MTSB_HD bool bsdf_is_smooth(const MtsbMaterial *mats, const MtsbMaterial &m) {
    if (bsdf_type_is_smooth(m.type)) return true;
    return (m.flags & MTSB_MAT_TWOSIDED) && m.back >= 0 && bsdf_type_is_smooth(mats[m.back].type);
}
MTSB_HD bool bsdf_has_back_or_transmission(const MtsbMaterial &m) {
    return m.type == MTSB_BSDF_ROUGHDIELECTRIC || m.type == MTSB_BSDF_DIELECTRIC || (m.flags & MTSB_MAT_TWOSIDED);
}

}  // namespace mtsb
