// texture.h -- 2D textures on material spectra (reference: src/textures/{bitmap,checkerboard,gridtexture,scale}.cpp,
// src/librender/texture.cpp, include/mitsuba/render/mipmap.h, src/librender/intersection.cpp).
//   Texture2D::eval(its, filter)        texture.cpp:112-121   uv' = uv * uvScale + uvOffset; filtered iff its.hasUVPartials
//   Checkerboard::eval                  checkerboard.cpp:61-69
//   GridTexture::eval                   gridtexture.cpp:66-79
//   BitmapTexture::eval(uv)             bitmap.cpp:431-453    level-0 box (nearest) or bilinear lookup
//   BitmapTexture::eval(uv, d0, d1)     bitmap.cpp:488-501 -> TMIPMap::eval mipmap.h:628-716 (trilinear / EWA)
//   TMIPMap::evalTexel/evalBox/evalBilinear/evalEWA   mipmap.h:503-596, 760-840
//   Intersection::computePartials       intersection.cpp:5-98
// Texels are half precision on the device, as in the reference (bitmap.cpp:172-177: TMIPMap<Color3, Color3h>).
#pragma once
#include "envmap.h"

namespace mtsb {

#define MTSB_TEX_MAX_LEVELS 17          // 65535 texels -> 17 levels (mipmap.h:184-193)
#define MTSB_MIPMAP_LUT_SIZE 64         // mipmap.h:37

struct TextureView {
    uint32_t type, filter, wrap_u, wrap_v;
    float uv_scale[2], uv_offset[2];
    float color0[3], color1[3];
    float line_width;
    float scale[3];
    float max_anisotropy;
    int levels;
    int lw[MTSB_TEX_MAX_LEVELS], lh[MTSB_TEX_MAX_LEVELS];
    uint32_t loff[MTSB_TEX_MAX_LEVELS];     // first texel of each level in `texels`
    const HalfTexel *texels;                // half4 per texel; a luminance image is replicated into r, g, b
    const float *weight_lut;                // MTSB_MIPMAP_LUT_SIZE Gaussian weights (mipmap.h:296-302), EWA only
};

MTSB_HD int tex_modulo(int a, int b) { const int r = a % b; return r < 0 ? r + b : r; }     // math::modulo

// boundary condition on one axis (mipmap.h:506-530); returns 0 = coordinate valid, 1 = value zero, 2 = value one
MTSB_HD int tex_wrap(int &x, int size, uint32_t mode) {
    if (x >= 0 && x < size) return 0;
    switch (mode) {
        case MTSB_WRAP_REPEAT: x = tex_modulo(x, size); return 0;
        case MTSB_WRAP_CLAMP: x = x < 0 ? 0 : size - 1; return 0;
        case MTSB_WRAP_MIRROR: x = tex_modulo(x, 2 * size); if (x >= size) x = 2 * size - x - 1; return 0;
        case MTSB_WRAP_ZERO: return 1;
        default: return 2;
    }
}

MTSB_HD V3 tex_texel(const TextureView &tv, int level, int x, int y) {
    const int w = tv.lw[level], h = tv.lh[level];
    int r = tex_wrap(x, w, tv.wrap_u);
    if (r) return V3(r == 1 ? 0.0f : 1.0f);
    r = tex_wrap(y, h, tv.wrap_v);
    if (r) return V3(r == 1 ? 0.0f : 1.0f);
    const size_t idx = (size_t) tv.loff[level] + (size_t) y * (size_t) w + (size_t) x;
#if MTSB_DEVICE_CODE
    const uint2 t = __ldg(reinterpret_cast<const uint2 *>(tv.texels) + idx);
#else
    const HalfTexel t = tv.texels[idx];
#endif
    return V3(half_bits_to_float(t.x & 0xffffu), half_bits_to_float(t.x >> 16), half_bits_to_float(t.y & 0xffffu));
}

MTSB_HD V3 tex_eval_box(const TextureView &tv, int level, P2 uv) {                         // mipmap.h:565-568
    return tex_texel(tv, level, (int) floorf(uv.x * (float) tv.lw[level]), (int) floorf(uv.y * (float) tv.lh[level]));
}

MTSB_HD V3 tex_eval_bilinear(const TextureView &tv, int level, P2 uv) {                    // mipmap.h:574-596
    if (!isfinite(uv.x) || !isfinite(uv.y)) return V3(0.0f);
    if (level >= tv.levels) level = tv.levels - 1;
    const float u = uv.x * (float) tv.lw[level] - 0.5f, v = uv.y * (float) tv.lh[level] - 0.5f;
    const int xPos = (int) floorf(u), yPos = (int) floorf(v);
    const float dx1 = u - (float) xPos, dx2 = 1.0f - dx1, dy1 = v - (float) yPos, dy2 = 1.0f - dy1;
    return tex_texel(tv, level, xPos, yPos) * dx2 * dy2 + tex_texel(tv, level, xPos, yPos + 1) * dx2 * dy1
         + tex_texel(tv, level, xPos + 1, yPos) * dx1 * dy2 + tex_texel(tv, level, xPos + 1, yPos + 1) * dx1 * dy1;
}

// elliptically weighted average on one level (mipmap.h:760-840)
MTSB_HD V3 tex_eval_ewa(const TextureView &tv, int level, P2 uv, float A, float B, float C) {
    if (!isfinite(A + B + C + uv.x + uv.y)) return V3(0.0f);
    if (level >= tv.levels) return tex_eval_box(tv, tv.levels - 1, uv);
    const float u = uv.x * (float) tv.lw[level] - 0.5f, v = uv.y * (float) tv.lh[level] - 0.5f;
    const float rx = (float) tv.lw[level] / (float) tv.lw[0], ry = (float) tv.lh[level] / (float) tv.lh[0];      // m_sizeRatio
    A /= rx * rx; B /= rx * ry; C /= ry * ry;
    const float invDet = 1.0f / (-B * B + 4.0f * A * C), deltaU = 2.0f * sqrtf(C * invDet), deltaV = 2.0f * sqrtf(A * invDet);
    const int u0 = (int) ceilf(u - deltaU), u1 = (int) floorf(u + deltaU);
    const int v0 = (int) ceilf(v - deltaV), v1 = (int) floorf(v + deltaV);
    const float As = A * MTSB_MIPMAP_LUT_SIZE, Bs = B * MTSB_MIPMAP_LUT_SIZE, Cs = C * MTSB_MIPMAP_LUT_SIZE;
    V3 result(0.0f);
    float denominator = 0.0f;
    const float ddq = 2 * As, uu0 = (float) u0 - u;
    for (int vt = v0; vt <= v1; ++vt) {
        const float vv = (float) vt - v;
        float q = As * uu0 * uu0 + (Bs * uu0 + Cs * vv) * vv;
        float dq = As * (2 * uu0 + 1) + Bs * vv;
        for (int ut = u0; ut <= u1; ++ut) {
            if (q < (float) MTSB_MIPMAP_LUT_SIZE) {
                const uint32_t qi = (uint32_t) q;
                if (qi < MTSB_MIPMAP_LUT_SIZE) {
                    const float weight = tv.weight_lut[(int) q];
                    result += tex_texel(tv, level, ut, vt) * weight;
                    denominator += weight;
                }
            }
            q += dq;
            dq += ddq;
        }
    }
    if (denominator == 0) return tex_eval_bilinear(tv, level, uv);
    return result / denominator;
}

MTSB_HD float tex_hypot2(float a, float b) {                                               // math.cpp:74-86
    float r;
    if (fabsf(a) > fabsf(b)) { r = b / a; r = fabsf(a) * sqrtf(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = fabsf(b) * sqrtf(1.0f + r * r); }
    else r = 0.0f;
    return r;
}
// math::log2 (math.cpp:103-106) over math::fastlog, which on Linux/x86-64 is the double-precision log rounded to float
// (math.h:175-199); 0.6931471824645996f = std::log(2.0f)
MTSB_HD float tex_log2(float v) { const float invLn2 = 1.0f / 0.6931471824645996f; return (float) log((double) v) * invLn2; }

// TMIPMap::eval(uv, d0, d1) (mipmap.h:628-716): d0 = (du/dx, dv/dx), d1 = (du/dy, dv/dy) in uv units
MTSB_HD V3 tex_eval_filtered(const TextureView &tv, P2 uv, float d0x, float d0y, float d1x, float d1y) {
    if (tv.filter == MTSB_TEXFILTER_NEAREST) return tex_eval_box(tv, 0, uv);
    if (tv.filter == MTSB_TEXFILTER_BILINEAR) return tex_eval_bilinear(tv, 0, uv);
    const float du0 = d0x * (float) tv.lw[0], dv0 = d0y * (float) tv.lh[0], du1 = d1x * (float) tv.lw[0], dv1 = d1y * (float) tv.lh[0];
    float A = dv0 * dv0 + dv1 * dv1, B = -2.0f * (du0 * dv0 + du1 * dv1), C = du0 * du0 + du1 * du1, F = A * C - B * B * 0.25f;
    const float root = tex_hypot2(A - C, B), Aprime = 0.5f * (A + C - root), Cprime = 0.5f * (A + C + root);
    float majorRadius = Aprime != 0 ? sqrtf(F / Aprime) : 0.0f, minorRadius = Cprime != 0 ? sqrtf(F / Cprime) : 0.0f;
    if (tv.filter == MTSB_TEXFILTER_TRILINEAR || !(minorRadius > 0) || !(majorRadius > 0) || F < 0) {
        const float level = tex_log2(fmaxf(majorRadius, MTSB_EPSILON));
        const int ilevel = (int) floorf(level);
        if (ilevel < 0) return tex_eval_bilinear(tv, 0, uv);
        const float a = level - (float) ilevel;
        return tex_eval_bilinear(tv, ilevel, uv) * (1.0f - a) + tex_eval_bilinear(tv, ilevel + 1, uv) * a;
    }
    if (minorRadius * tv.max_anisotropy < majorRadius) {
        minorRadius = majorRadius / tv.max_anisotropy;
        const float theta = 0.5f * atanf(B / (A - C));
        float sinTheta, cosTheta;
        sincosf(theta, &sinTheta, &cosTheta);
        const float a2 = majorRadius * majorRadius, b2 = minorRadius * minorRadius, sinTheta2 = sinTheta * sinTheta,
                    cosTheta2 = cosTheta * cosTheta, sin2Theta = 2 * sinTheta * cosTheta;
        A = a2 * cosTheta2 + b2 * sinTheta2;
        B = (a2 - b2) * sin2Theta;
        C = a2 * sinTheta2 + b2 * cosTheta2;
        F = a2 * b2;
    }
    const float sc = 1.0f / F;
    A *= sc; B *= sc; C *= sc;
    const float level = fmaxf(0.0f, tex_log2(minorRadius));
    const int ilevel = (int) level;
    const float a = level - (float) ilevel;
    if (majorRadius < 1 || !(A > 0 && C > 0)) return tex_eval_bilinear(tv, ilevel, uv);
    return tex_eval_ewa(tv, ilevel, uv, A, B, C) * (1.0f - a) + tex_eval_ewa(tv, ilevel + 1, uv, A, B, C) * a;
}

// EnvironmentMap::evalEnvironment for a ray that carries differentials (envmap.cpp:380-410): d = ray direction,
// rxd / ryd = directions of its x / y offset rays (all in world space)
MTSB_HD V3 envmap_eval_filtered(const EnvmapView &ev, V3 d, V3 rxd, V3 ryd) {
    const V3 v = envmap_xform(ev.to_local, d);
    P2 uv; uv.x = atan2f(v.x, -v.z) * MTSB_INV_TWOPI; uv.y = envmap_safe_acos(v.y) * MTSB_INV_PI;
    const V3 dvdx = envmap_xform(ev.to_local, rxd) - v, dvdy = envmap_xform(ev.to_local, ryd) - v;
    const float t1 = MTSB_INV_TWOPI / (v.x * v.x + v.z * v.z);
    const float t2 = -MTSB_INV_PI / fmaxf(sqrtf(fmaxf(0.0f, 1.0f - v.y * v.y)), MTSB_EPSILON);
    const float dudx_x = t1 * (dvdx.z * v.x - dvdx.x * v.z), dudx_y = t2 * dvdx.y;
    const float dudy_x = t1 * (dvdy.z * v.x - dvdy.x * v.z), dudy_y = t2 * dvdy.y;
    return tex_eval_filtered(*ev.mip, uv, dudx_x, dudx_y, dudy_x, dudy_y) * ev.scale;
}

MTSB_HD V3 tex_procedural(const TextureView &tv, P2 uv) {
    if (tv.type == MTSB_TEX_CHECKERBOARD) {                                                // checkerboard.cpp:61-69
        const int x = 2 * tex_modulo((int) (uv.x * 2), 2) - 1, y = 2 * tex_modulo((int) (uv.y * 2), 2) - 1;
        return x * y == 1 ? V3(tv.color0[0], tv.color0[1], tv.color0[2]) : V3(tv.color1[0], tv.color1[1], tv.color1[2]);
    }
    float x = uv.x - (float) (int) floorf(uv.x), y = uv.y - (float) (int) floorf(uv.y);     // gridtexture.cpp:66-79
    if (x > .5) x -= 1;
    if (y > .5) y -= 1;
    return (fabsf(x) < tv.line_width || fabsf(y) < tv.line_width) ? V3(tv.color1[0], tv.color1[1], tv.color1[2])
                                                                  : V3(tv.color0[0], tv.color0[1], tv.color0[2]);
}

// uv partials of a hit (filled by compute_partials); `has` mirrors Intersection::hasUVPartials
struct UVPartials { bool has; float dudx, dvdx, dudy, dvdy; };

// Texture2D::eval(its, filter = true) (texture.cpp:112-121) followed by the folded ScalingTexture factors (scale.cpp:70)
MTSB_HD V3 texture_eval(const TextureView &tv, P2 uv_its, const UVPartials &pt) {
    P2 uv; uv.x = uv_its.x * tv.uv_scale[0] + tv.uv_offset[0]; uv.y = uv_its.y * tv.uv_scale[1] + tv.uv_offset[1];
    V3 value;
    if (tv.type != MTSB_TEX_BITMAP) value = tex_procedural(tv, uv);
    else if (pt.has) value = tex_eval_filtered(tv, uv, pt.dudx * tv.uv_scale[0], pt.dvdx * tv.uv_scale[1],
                                               pt.dudy * tv.uv_scale[0], pt.dvdy * tv.uv_scale[1]);
    else value = tv.filter == MTSB_TEXFILTER_NEAREST ? tex_eval_box(tv, 0, uv) : tex_eval_bilinear(tv, 0, uv);
    return V3(value.x * tv.scale[0], value.y * tv.scale[1], value.z * tv.scale[2]);
}

// solveLinearSystem2x2 (util.cpp:428-448)
MTSB_HD bool tex_solve2x2(const float a[2][2], const float b[2], float x[2]) {
    const float det = a[0][0] * a[1][1] - a[0][1] * a[1][0];
    if (fabsf(det) <= MTSB_RCPOVERFLOW) return false;
    const float inverse = 1.0f / det;
    x[0] = (a[1][1] * b[0] - a[0][1] * b[1]) * inverse;
    x[1] = (a[0][0] * b[1] - a[1][0] * b[0]) * inverse;
    return true;
}

// Intersection::computePartials (intersection.cpp:5-98).  n = base-frame normal (the interpolated shading normal, or the
// face normal of a mesh without vertex normals: skdtree.h:428-431); (rxd, ryd) = offset-ray directions, both from `o`.
MTSB_HD UVPartials compute_partials(V3 p, V3 n, V3 dpdu, V3 dpdv, V3 o, V3 rxd, V3 ryd) {
    UVPartials r; r.has = true; r.dudx = r.dvdx = r.dudy = r.dvdy = 0.0f;
    if (isZero(dpdu) && isZero(dpdv)) return r;
    const float pp = dot(n, p), pox = dot(n, o), poy = pox, prx = dot(n, rxd), pry = dot(n, ryd);
    if (prx == 0 || pry == 0) return r;
    const float tx = (pp - pox) / prx, ty = (pp - poy) / pry;
    const float absX = fabsf(n.x), absY = fabsf(n.y), absZ = fabsf(n.z);
    int a0, a1;
    if (absX > absY && absX > absZ) { a0 = 1; a1 = 2; } else if (absY > absZ) { a0 = 0; a1 = 2; } else { a0 = 0; a1 = 1; }
    float A[2][2], Bx[2], By[2], x[2];
    A[0][0] = comp(dpdu, a0); A[0][1] = comp(dpdv, a0); A[1][0] = comp(dpdu, a1); A[1][1] = comp(dpdv, a1);
    const V3 px = o + rxd * tx, py = o + ryd * ty;
    Bx[0] = comp(px, a0) - comp(p, a0); Bx[1] = comp(px, a1) - comp(p, a1);
    By[0] = comp(py, a0) - comp(p, a0); By[1] = comp(py, a1) - comp(p, a1);
    if (tex_solve2x2(A, Bx, x)) { r.dudx = x[0]; r.dvdx = x[1]; } else { r.dudx = 1; r.dvdx = 0; }
    if (tex_solve2x2(A, By, x)) { r.dudy = x[0]; r.dvdy = x[1]; } else { r.dudy = 1; }      // intersection.cpp:96 assigns dudy twice
    return r;
}

}  // namespace mtsb
