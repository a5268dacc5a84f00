// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// CPU restatement of MicrofacetDistribution (Beckmann / GGX) and of the three BSDFs on the
// `path` hot path, plus the TwoSided adapter.  Reference lines are cited per function.
#pragma once
#include "orc_math.h"
#include "../include/mtsb200.h"

namespace orc {

// ---- MicrofacetDistribution: src/bsdfs/microfacet.h -----------------------
struct Microfacet {
    int type;            // MTSB_DISTR_*
    float au, av;
    bool visible;
    // microfacet.h:86-93 (alpha clamp >= 1e-4)
    Microfacet(int t, float alphaU, float alphaV, bool sampleVisible)
        : type(t), au(std::max(alphaU, 1e-4f)), av(std::max(alphaV, 1e-4f)), visible(sampleVisible) {}
    bool isotropic() const { return au == av; }
    void scaleAlpha(float v) { au *= v; av *= v; }                         // microfacet.h:178-183

    // microfacet.h:191-234
    float eval(V3 m) const {
        if (cosTheta(m) <= 0) return 0.0f;
        float ct2 = cosTheta2(m);
        float bexp = ((m.x * m.x) / (au * au) + (m.y * m.y) / (av * av)) / ct2;
        float result;
        if (type == MTSB_DISTR_BECKMANN) {
            result = std::exp(-bexp) / (kPi * au * av * ct2 * ct2);
        } else {
            float root = (1.0f + bexp) * ct2;
            result = 1.0f / (kPi * au * av * root * root);
        }
        if (result * cosTheta(m) < 1e-20f) result = 0;
        return result;
    }

    // microfacet.h:549-559
    float projectRoughness(V3 v) const {
        float invSinTheta2 = 1 / sinTheta2(v);
        if (isotropic() || invSinTheta2 <= 0) return au;
        float cosPhi2 = v.x * v.x * invSinTheta2, sinPhi2 = v.y * v.y * invSinTheta2;
        return std::sqrt(cosPhi2 * au * au + sinPhi2 * av * av);
    }

    // microfacet.h:477-514
    float smithG1(V3 v, V3 m) const {
        if (dot(v, m) * cosTheta(v) <= 0) return 0.0f;
        float tanT = std::fabs(tanTheta(v));
        if (tanT == 0.0f) return 1.0f;
        float alpha = projectRoughness(v);
        if (type == MTSB_DISTR_BECKMANN) {
            float a = 1.0f / (alpha * tanT);
            if (a >= 1.6f) return 1.0f;
            float aSqr = a * a;
            return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
        } else {
            float root = alpha * tanT;
            return 2.0f / (1.0f + hypot2(1.0f, root));
        }
    }
    float G(V3 wi, V3 wo, V3 m) const { return smithG1(wi, m) * smithG1(wo, m); }   // :527-530

    // microfacet.h:287-395 (Beckmann / GGX branches)
    V3 sampleAll(P2 u, float &pdf) const {
        float cosThetaM = 0.0f, sinPhiM, cosPhiM, alphaSqr;
        if (isotropic()) {
            sincosf_((2.0f * kPi) * u.y, &sinPhiM, &cosPhiM);
            alphaSqr = au * au;
        } else {
            float phiM = std::atan(av / au * std::tan(kPi + 2 * kPi * u.y)) + kPi * std::floor(2 * u.y + 0.5f);
            sincosf_(phiM, &sinPhiM, &cosPhiM);
            float cosSc = cosPhiM / au, sinSc = sinPhiM / av;
            alphaSqr = 1.0f / (cosSc * cosSc + sinSc * sinSc);
        }
        if (type == MTSB_DISTR_BECKMANN) {
            float tanThetaMSqr = alphaSqr * -std::log(1.0f - u.x);
            cosThetaM = 1.0f / std::sqrt(1.0f + tanThetaMSqr);
            pdf = (1.0f - u.x) / (kPi * au * av * cosThetaM * cosThetaM * cosThetaM);
        } else {
            float tanThetaMSqr = alphaSqr * u.x / (1.0f - u.x);
            cosThetaM = 1.0f / std::sqrt(1.0f + tanThetaMSqr);
            float temp = 1 + tanThetaMSqr / alphaSqr;
            pdf = kInvPi / (au * av * cosThetaM * cosThetaM * cosThetaM * temp * temp);
        }
        if (pdf < 1e-20f) pdf = 0;
        float sinThetaM = std::sqrt(std::max(0.0f, 1 - cosThetaM * cosThetaM));
        return V3(sinThetaM * cosPhiM, sinThetaM * sinPhiM, cosThetaM);
    }
    float pdfAll(V3 m) const { return eval(m) * cosTheta(m); }                        // :404-407

    // microfacet.h:581-705
    void sampleVisible11(float thetaI, P2 u, float &sx, float &sy) const {
        const float SQRT_PI_INV = 1 / std::sqrt(kPi);
        if (type == MTSB_DISTR_BECKMANN) {
            if (thetaI < 1e-4f) {
                float s, c, r = std::sqrt(-std::log(1.0f - u.x));
                sincosf_(2 * kPi * u.y, &s, &c);
                sx = r * c; sy = r * s; return;
            }
            float tanThetaI = std::tan(thetaI), cotThetaI = 1 / tanThetaI;
            float a = -1, c = erf_as(cotThetaI);
            float sample_x = std::max(u.x, 1e-6f);
            float fit = 1 + thetaI * (-0.876f + thetaI * (0.4265f - 0.0594f * thetaI));
            float b = c - (1 + c) * std::pow(1 - sample_x, fit);
            float normalization = 1 / (1 + c + SQRT_PI_INV * tanThetaI * std::exp(-cotThetaI * cotThetaI));
            int it = 0;
            while (++it < 10) {
                if (!(b >= a && b <= c)) b = 0.5f * (a + c);
                float invErf = erfinv(b);
                float value = normalization * (1 + b + SQRT_PI_INV * tanThetaI * std::exp(-invErf * invErf)) - sample_x;
                float derivative = normalization * (1 - invErf * tanThetaI);
                if (std::fabs(value) < 1e-5f) break;
                if (value > 0) c = b; else a = b;
                b -= value / derivative;
            }
            sx = erfinv(b);
            sy = erfinv(2.0f * std::max(u.y, 1e-6f) - 1.0f);
        } else {
            if (thetaI < 1e-4f) {
                float s, c, r = safe_sqrt(u.x / (1 - u.x));
                sincosf_(2 * kPi * u.y, &s, &c);
                sx = r * c; sy = r * s; return;
            }
            float tanThetaI = std::tan(thetaI);
            float a = 1 / tanThetaI;
            float G1 = 2.0f / (1.0f + safe_sqrt(1.0f + 1.0f / (a * a)));
            float A = 2.0f * u.x / G1 - 1.0f;
            if (std::fabs(A) == 1) A -= signum(A) * kEpsilon;
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
            sy = S * z * std::sqrt(1.0f + sx * sx);
        }
    }

    // microfacet.h:421-459
    V3 sampleVisible(V3 _wi, P2 u) const {
        V3 wi = normalize(V3(au * _wi.x, av * _wi.y, _wi.z));
        float theta = 0, phi = 0;
        if (wi.z < 0.99999f) { theta = std::acos(wi.z); phi = std::atan2(wi.y, wi.x); }
        float sinPhi, cosPhi; sincosf_(phi, &sinPhi, &cosPhi);
        float sx, sy; sampleVisible11(theta, u, sx, sy);
        float rx = cosPhi * sx - sinPhi * sy, ry = sinPhi * sx + cosPhi * sy;
        rx *= au; ry *= av;
        float normalization = 1.0f / std::sqrt(rx * rx + ry * ry + 1.0f);
        return V3(-rx * normalization, -ry * normalization, normalization);
    }
    // microfacet.h:462-466
    float pdfVisible(V3 wi, V3 m) const {
        if (cosTheta(wi) == 0) return 0.0f;
        return smithG1(wi, m) * absDot(wi, m) * eval(m) / std::fabs(cosTheta(wi));
    }
    // microfacet.h:240-249, 270-275
    V3 sample(V3 wi, P2 u, float &pdf) const {
        if (visible) { V3 m = sampleVisible(wi, u); pdf = pdfVisible(wi, m); return m; }
        return sampleAll(u, pdf);
    }
    float pdf(V3 wi, V3 m) const { return visible ? pdfVisible(wi, m) : pdfAll(m); }
};

// ---- BSDF records ----------------------------------------------------------
struct BsdfSample { V3 wo; V3 weight; float pdf; float eta; bool delta; };

inline V3 rgb(const float *p) { return V3(p[0], p[1], p[2]); }

// ---- SmoothDiffuse: src/bsdfs/diffuse.cpp:110-165 --------------------------
inline V3 diffuse_eval(const MtsbMaterial &m, V3 wi, V3 wo) {
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return V3(0.0f);
    return rgb(m.reflectance) * (kInvPi * cosTheta(wo));
}
inline float diffuse_pdf(const MtsbMaterial &, V3 wi, V3 wo) {
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
    return squareToCosineHemispherePdf(wo);
}
inline bool diffuse_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {
    if (cosTheta(wi) <= 0) return false;
    s.wo = squareToCosineHemisphere(u);
    s.eta = 1.0f; s.delta = false;
    s.pdf = squareToCosineHemispherePdf(s.wo);
    s.weight = rgb(m.reflectance);
    return true;
}

// ---- RoughConductor: src/bsdfs/roughconductor.cpp:259-417 -------------------
inline Microfacet distr_of(const MtsbMaterial &m) {
    return Microfacet((int) m.distribution, m.alpha_u, m.alpha_v, (m.flags & MTSB_MAT_SAMPLE_VISIBLE) != 0);
}
inline V3 roughconductor_eval(const MtsbMaterial &m, V3 wi, V3 wo) {                  // :259-296
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return V3(0.0f);
    V3 H = normalize(wo + wi);
    Microfacet distr = distr_of(m);
    const float D = distr.eval(H);
    if (D == 0) return V3(0.0f);
    const V3 F = fresnelConductorExact(dot(wi, H), rgb(m.cond_eta), rgb(m.cond_k)) * rgb(m.reflectance);
    const float G = distr.G(wi, wo, H);
    float model = D * G / (4.0f * cosTheta(wi));
    return F * model;
}
inline float roughconductor_pdf(const MtsbMaterial &m, V3 wi, V3 wo) {                // :298-323
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
    V3 H = normalize(wo + wi);
    Microfacet distr = distr_of(m);
    if (distr.visible) return distr.eval(H) * distr.smithG1(wi, H) / (4.0f * cosTheta(wi));
    return distr.pdf(wi, H) / (4 * absDot(wo, H));
}
inline bool roughconductor_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) { // :371-417 (sample with pdf)
    if (cosTheta(wi) < 0) return false;                                                // `< 0`, not `<= 0` (:372)
    Microfacet distr = distr_of(m);
    float pdf;
    V3 mm = distr.sample(wi, u, pdf);
    if (pdf == 0) return false;
    s.wo = reflect(wi, mm);
    s.eta = 1.0f; s.delta = false;
    if (cosTheta(s.wo) <= 0) return false;
    V3 F = fresnelConductorExact(dot(wi, mm), rgb(m.cond_eta), rgb(m.cond_k)) * rgb(m.reflectance);
    float weight;
    if (distr.visible) weight = distr.smithG1(s.wo, mm);
    else weight = distr.eval(mm) * distr.G(wi, s.wo, mm) * dot(wi, mm) / (pdf * cosTheta(wi));
    pdf /= 4.0f * dot(s.wo, mm);
    s.pdf = pdf;
    s.weight = F * weight;
    return true;
}

// ---- RoughDielectric: src/bsdfs/roughdielectric.cpp:270-615 (mode == ERadiance) ----
inline V3 roughdielectric_eval(const MtsbMaterial &m, V3 wi, V3 wo) {                  // :270-348
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
    const float F = fresnelDielectricExt(dot(wi, H), m_eta);
    const float G = distr.G(wi, wo, H);
    if (refl) {
        float value = F * D * G / (4.0f * std::fabs(cosTheta(wi)));
        return rgb(m.reflectance) * value;
    } else {
        float eta = cosTheta(wi) > 0.0f ? m_eta : m_invEta;
        float sqrtDenom = dot(wi, H) + eta * dot(wo, H);
        float value = ((1 - F) * D * G * eta * eta * dot(wi, H) * dot(wo, H)) / (cosTheta(wi) * sqrtDenom * sqrtDenom);
        float factor = cosTheta(wi) > 0 ? m_invEta : m_eta;
        return rgb(m.transmittance) * std::fabs(value * factor * factor);
    }
}
inline float roughdielectric_pdf(const MtsbMaterial &m, V3 wi, V3 wo) {                // :350-422
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
    if (!sampleDistr.visible) sampleDistr.scaleAlpha(1.2f - 0.2f * std::sqrt(std::fabs(cosTheta(wi))));
    float prob = sampleDistr.pdf(signum(cosTheta(wi)) * wi, H);
    float F = fresnelDielectricExt(dot(wi, H), m_eta);
    prob *= refl ? F : (1 - F);
    return std::fabs(prob * dwh_dwo);
}
// :513-615; `u1` is the extra sampler->next1D() drawn at :554
inline bool roughdielectric_sample(const MtsbMaterial &m, V3 wi, P2 u, float u1, BsdfSample &s) {
    const float m_eta = m.eta, m_invEta = 1 / m.eta;
    Microfacet distr = distr_of(m);
    Microfacet sampleDistr = distr;
    if (!distr.visible) sampleDistr.scaleAlpha(1.2f - 0.2f * std::sqrt(std::fabs(cosTheta(wi))));
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
        weight *= rgb(m.reflectance);
        dwh_dwo = 1.0f / (4.0f * dot(s.wo, mm));
    } else {
        if (cosThetaT == 0) return false;
        s.wo = refract(wi, mm, m_eta, cosThetaT);
        s.eta = cosThetaT < 0 ? m_eta : m_invEta;
        if (cosTheta(wi) * cosTheta(s.wo) >= 0) return false;
        float factor = cosThetaT < 0 ? m_invEta : m_eta;
        weight *= rgb(m.transmittance) * (factor * factor);
        float sqrtDenom = dot(wi, mm) + s.eta * dot(s.wo, mm);
        dwh_dwo = (s.eta * s.eta * dot(s.wo, mm)) / (sqrtDenom * sqrtDenom);
    }
    if (distr.visible) weight *= distr.smithG1(s.wo, mm);
    else weight *= std::fabs(distr.eval(mm) * distr.G(wi, s.wo, mm) * dot(wi, mm) / (microfacetPDF * cosTheta(wi)));
    pdf *= std::fabs(dwh_dwo);
    s.pdf = pdf;
    s.weight = weight;
    return true;
}

// ---- SmoothDielectric: src/bsdfs/dielectric.cpp:216-330 (mode == ERadiance) --------
// Both lobes are Dirac deltas: eval()/pdf() under the solid-angle measure that `path`
// uses are zero (:224-226, :256-259); only sample() does work.
inline bool dielectric_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {      // :277-301
    const float m_eta = m.eta, m_invEta = 1 / m.eta;
    float cosThetaT;
    float F = fresnelDielectricExt(cosTheta(wi), cosThetaT, m_eta);
    s.delta = true;
    if (u.x <= F) {
        s.wo = V3(-wi.x, -wi.y, wi.z);                                                  // reflect(wi) :206-208
        s.eta = 1.0f;
        s.pdf = F;
        s.weight = rgb(m.reflectance);
    } else {
        float scale = -(cosThetaT < 0 ? m_invEta : m_eta);                              // refract(wi, cosThetaT) :211-214
        s.wo = V3(scale * wi.x, scale * wi.y, cosThetaT);
        s.eta = cosThetaT < 0 ? m_eta : m_invEta;
        s.pdf = 1 - F;
        float factor = cosThetaT < 0 ? m_invEta : m_eta;
        s.weight = rgb(m.transmittance) * (factor * factor);
    }
    return true;
}

// ---- SmoothConductor: src/bsdfs/conductor.cpp:224-285 --------------------------------
inline bool conductor_sample(const MtsbMaterial &m, V3 wi, P2, BsdfSample &s) {          // :268-285
    if (cosTheta(wi) <= 0) return false;
    s.delta = true;
    s.wo = V3(-wi.x, -wi.y, wi.z);
    s.eta = 1.0f;
    s.pdf = 1.0f;
    s.weight = rgb(m.reflectance) * fresnelConductorExact(cosTheta(wi), rgb(m.cond_eta), rgb(m.cond_k));
    return true;
}

// ---- SmoothPlastic: src/bsdfs/plastic.cpp:245-440 ---------------------------------------
// A Dirac reflection lobe (specularReflectance, stored in `transmittance`) over a diffuse base
// (diffuseReflectance, stored in `reflectance`).  Under the solid-angle measure eval()/pdf() see only
// the diffuse component (:250-252, :300-302).
inline float plastic_prob_specular(const MtsbMaterial &m, float Fi) {                  // :288-291
    return (Fi * m.spec_weight) / (Fi * m.spec_weight + (1 - Fi) * (1 - m.spec_weight));
}
inline V3 plastic_diffuse_term(const MtsbMaterial &m) {                                  // :270-275
    V3 diff = rgb(m.reflectance);
    if (m.flags & MTSB_MAT_NONLINEAR) diff = diff / (V3(1.0f) - diff * m.fdr_int);
    else diff = diff / (1 - m.fdr_int);
    return diff;
}
inline V3 plastic_eval(const MtsbMaterial &m, V3 wi, V3 wo) {                            // :245-280
    if (cosTheta(wo) <= 0 || cosTheta(wi) <= 0) return V3(0.0f);
    float Fi = fresnelDielectricExt(cosTheta(wi), m.eta);
    float Fo = fresnelDielectricExt(cosTheta(wo), m.eta);
    const float invEta2 = 1 / (m.eta * m.eta);
    return plastic_diffuse_term(m) * (squareToCosineHemispherePdf(wo) * invEta2 * (1 - Fi) * (1 - Fo));
}
inline float plastic_pdf(const MtsbMaterial &m, V3 wi, V3 wo) {                          // :282-310
    if (cosTheta(wo) <= 0 || cosTheta(wi) <= 0) return 0.0f;
    float Fi = fresnelDielectricExt(cosTheta(wi), m.eta);
    float probSpecular = plastic_prob_specular(m, Fi);
    return squareToCosineHemispherePdf(wo) * (1 - probSpecular);
}
inline bool plastic_sample(const MtsbMaterial &m, V3 wi, P2 u, BsdfSample &s) {          // :373-412
    if (cosTheta(wi) <= 0) return false;
    float Fi = fresnelDielectricExt(cosTheta(wi), m.eta);
    s.eta = 1.0f;
    float probSpecular = plastic_prob_specular(m, Fi);
    if (u.x < probSpecular) {
        s.delta = true;
        s.wo = V3(-wi.x, -wi.y, wi.z);
        s.pdf = probSpecular;
        s.weight = rgb(m.transmittance) * Fi / probSpecular;
    } else {
        s.delta = false;
        s.wo = squareToCosineHemisphere(P2{(u.x - probSpecular) / (1 - probSpecular), u.y});
        float Fo = fresnelDielectricExt(cosTheta(s.wo), m.eta);
        const float invEta2 = 1 / (m.eta * m.eta);
        s.pdf = (1 - probSpecular) * squareToCosineHemispherePdf(s.wo);
        s.weight = plastic_diffuse_term(m) * (invEta2 * (1 - Fi) * (1 - Fo) / (1 - probSpecular));
    }
    return true;
}

// ---- dispatch incl. TwoSided (src/bsdfs/twosided.cpp:108-184) -------------------
inline V3 bsdf_eval_one(const MtsbMaterial &m, V3 wi, V3 wo) {
    switch (m.type) {
        case MTSB_BSDF_DIFFUSE: return diffuse_eval(m, wi, wo);
        case MTSB_BSDF_ROUGHCONDUCTOR: return roughconductor_eval(m, wi, wo);
        case MTSB_BSDF_ROUGHDIELECTRIC: return roughdielectric_eval(m, wi, wo);
        case MTSB_BSDF_PLASTIC: return plastic_eval(m, wi, wo);
        default: return V3(0.0f);
    }
}
inline float bsdf_pdf_one(const MtsbMaterial &m, V3 wi, V3 wo) {
    switch (m.type) {
        case MTSB_BSDF_DIFFUSE: return diffuse_pdf(m, wi, wo);
        case MTSB_BSDF_ROUGHCONDUCTOR: return roughconductor_pdf(m, wi, wo);
        case MTSB_BSDF_ROUGHDIELECTRIC: return roughdielectric_pdf(m, wi, wo);
        case MTSB_BSDF_PLASTIC: return plastic_pdf(m, wi, wo);
        default: return 0.0f;
    }
}
inline bool bsdf_sample_one(const MtsbMaterial &m, V3 wi, P2 u, float u1, BsdfSample &s) {
    switch (m.type) {
        case MTSB_BSDF_DIFFUSE: return diffuse_sample(m, wi, u, s);
        case MTSB_BSDF_ROUGHCONDUCTOR: return roughconductor_sample(m, wi, u, s);
        case MTSB_BSDF_ROUGHDIELECTRIC: return roughdielectric_sample(m, wi, u, u1, s);
        case MTSB_BSDF_DIELECTRIC: return dielectric_sample(m, wi, u, s);
        case MTSB_BSDF_CONDUCTOR: return conductor_sample(m, wi, u, s);
        case MTSB_BSDF_PLASTIC: return plastic_sample(m, wi, u, s);
        default: return false;
    }
}
inline const MtsbMaterial &back_of(const MtsbMaterial *mats, const MtsbMaterial &m) {
    return m.back >= 0 ? mats[m.back] : m;
}
inline V3 bsdf_eval(const MtsbMaterial *mats, int id, V3 wi, V3 wo) {
    const MtsbMaterial &m = mats[id];
    if ((m.flags & MTSB_MAT_TWOSIDED) && !(cosTheta(wi) > 0)) {                        // twosided.cpp:111-121
        wi.z *= -1; wo.z *= -1;
        return bsdf_eval_one(back_of(mats, m), wi, wo);
    }
    return bsdf_eval_one(m, wi, wo);
}
inline float bsdf_pdf(const MtsbMaterial *mats, int id, V3 wi, V3 wo) {
    const MtsbMaterial &m = mats[id];
    if ((m.flags & MTSB_MAT_TWOSIDED) && !(wi.z > 0)) {                                 // twosided.cpp:123-135
        wi.z *= -1; wo.z *= -1;
        return bsdf_pdf_one(back_of(mats, m), wi, wo);
    }
    return bsdf_pdf_one(m, wi, wo);
}
// twosided.cpp:162-184: flip only when cosTheta(wi) < 0; un-flip wo iff result non-zero and pdf != 0
inline bool bsdf_sample(const MtsbMaterial *mats, int id, V3 wi, P2 u, float u1, BsdfSample &s) {
    const MtsbMaterial &m = mats[id];
    if ((m.flags & MTSB_MAT_TWOSIDED) && cosTheta(wi) < 0) {
        wi.z *= -1;
        bool ok = bsdf_sample_one(back_of(mats, m), wi, u, u1, s);
        if (ok && !isZero(s.weight) && s.pdf != 0) s.wo.z *= -1;
        return ok;
    }
    return bsdf_sample_one(m, wi, u, u1, s);
}
// BSDF::getType() bits that path.cpp looks at: ESmooth (path.cpp:177) and
// ETransmission|EBackSide (records.inl:175-179).
// A twosided wrapper reports the union of both nested types (twosided.cpp:72-99).
inline bool bsdf_type_is_smooth(uint32_t t) {
    return t == MTSB_BSDF_DIFFUSE || t == MTSB_BSDF_ROUGHCONDUCTOR || t == MTSB_BSDF_ROUGHDIELECTRIC || t == MTSB_BSDF_PLASTIC;
}
inline bool bsdf_is_smooth(const MtsbMaterial *mats, const MtsbMaterial &m) {
    if (bsdf_type_is_smooth(m.type)) return true;
    return (m.flags & MTSB_MAT_TWOSIDED) && m.back >= 0 && bsdf_type_is_smooth(mats[m.back].type);
}
inline bool bsdf_has_back_or_transmission(const MtsbMaterial &m) {
    return m.type == MTSB_BSDF_ROUGHDIELECTRIC || m.type == MTSB_BSDF_DIELECTRIC || (m.flags & MTSB_MAT_TWOSIDED);
}

}  // namespace orc
