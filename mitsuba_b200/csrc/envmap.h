// envmap.h -- latitude-longitude environment emitter (reference: src/emitters/envmap.cpp).
//   evalEnvironment            :380-410   bilinear level-0 lookup (MIPMap::evalBilinear, include/mitsuba/render/mipmap.h:575-596)
//   internalSampleDirection    :567-600   marginal / conditional CDF over luminance x sin(theta), tent offset
//   internalPdfDirection       :603-633
//   sampleDirect / pdfDirect   :516-556
// The texels are kept in half precision, as the reference keeps them (envmap.cpp:101-103).
// Filtered lookups: the reference sends rays that carry differentials (the camera rays; path.cpp drops them after the
// first hit) through MIPMap::eval -- EWA on the Lanczos pyramid with maxAnisotropy 10, repeat in u, clamp in v
// (envmap.cpp:140-178, 393-407).  envmap_eval_filtered (texture.h, next to the pyramid code it shares with the bitmap
// texture) does the same for the first segment of a path; every other lookup is the level-0 bilinear one.
#pragma once
#include "hd.h"
#include "../../include/mtsb200.h"
#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

namespace mtsb {

#if defined(__CUDACC__)
struct __align__(8) HalfTexel { uint32_t x, y; };      // half4: (r | g << 16, b)
#else
struct alignas(8) HalfTexel { uint32_t x, y; };
#endif

struct TextureView;                 // texture.h

struct EnvmapView {
    int width, height;
    const TextureView *mip;         // the map's MIP pyramid (EWA, repeat / clamp), for lookups of rays with differentials
    const HalfTexel *texels;           // half4 per texel: (r, g, b, unused)
    const float *cdf_rows;          // height + 1
    const float *cdf_cols;          // (width + 1) * height
    const float *row_weights;       // height: sin((y + 0.5) pi / height)
    float normalization;            // envmap.cpp:318-319
    float pixel_size_x, pixel_size_y;   // :322
    float scale;
    float to_world[9], to_local[9]; // linear parts, row-major
};

#if MTSB_DEVICE_CODE
MTSB_HD float half_bits_to_float(uint32_t h) { return __half2float(__ushort_as_half((unsigned short) h)); }
#else
inline float half_bits_to_float(uint32_t h) {
    const uint32_t s = (h & 0x8000u) << 16, e = (h >> 10) & 0x1fu, m = h & 0x3ffu;
    if (e == 0) { float f = (float) m * 5.9604644775390625e-8f; return s ? -f : f; }           // zero / denormal: m * 2^-24
    if (e == 31) return u2f(s | 0x7f800000u | (m << 13));
    return u2f(s | ((e + 112u) << 23) | (m << 13));
}
#endif

// MIPMap::evalTexel (mipmap.h:503-563) with ERepeat in u and EClamp in v (envmap.cpp:176-178)
MTSB_HD V3 envmap_texel(const EnvmapView &ev, int x, int y) {
    if (x < 0 || x >= ev.width) { x %= ev.width; if (x < 0) x += ev.width; }
    if (y < 0) y = 0; else if (y >= ev.height) y = ev.height - 1;
#if MTSB_DEVICE_CODE
    const uint2 t = __ldg(reinterpret_cast<const uint2 *>(ev.texels) + (size_t) y * ev.width + x);
#else
    const HalfTexel t = ev.texels[(size_t) y * ev.width + x];
#endif
    return V3(half_bits_to_float(t.x & 0xffffu), half_bits_to_float(t.x >> 16), half_bits_to_float(t.y & 0xffffu));
}

MTSB_HD V3 envmap_xform(const float *m, V3 v) {
    return V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z);
}
MTSB_HD float envmap_safe_acos(float v) { return acosf(fminf(1.0f, fmaxf(-1.0f, v))); }
MTSB_HD float envmap_luminance(V3 c) { return c.x * 0.212671f + c.y * 0.715160f + c.z * 0.072169f; }

// the two bilinear row sums around fractional texel position (u, v) (shared by eval, sample and pdf)
MTSB_HD void envmap_rows(const EnvmapView &ev, float u, float v, V3 &value1, V3 &value2, int &yPos) {
    const int xPos = (int) floorf(u); yPos = (int) floorf(v);
    const float dx1 = u - (float) xPos, dx2 = 1.0f - dx1, dy1 = v - (float) yPos, dy2 = 1.0f - dy1;
    value1 = envmap_texel(ev, xPos, yPos) * dx2 * dy2 + envmap_texel(ev, xPos + 1, yPos) * dx1 * dy2;
    value2 = envmap_texel(ev, xPos, yPos + 1) * dx2 * dy1 + envmap_texel(ev, xPos + 1, yPos + 1) * dx1 * dy1;
}

// EnvironmentMap::evalEnvironment without differentials (envmap.cpp:380-392, 408-409); d = ray direction (world)
MTSB_HD V3 envmap_eval(const EnvmapView &ev, V3 d) {
    const V3 v = envmap_xform(ev.to_local, d);
    const float uvx = atan2f(v.x, -v.z) * MTSB_INV_TWOPI, uvy = envmap_safe_acos(v.y) * MTSB_INV_PI;
    if (!isfinite(uvx) || !isfinite(uvy)) return V3(0.0f);                       // mipmap.h:576-578
    const float u = uvx * (float) ev.width - 0.5f, w = uvy * (float) ev.height - 0.5f;
    const int xPos = (int) floorf(u), yPos = (int) floorf(w);
    const float dx1 = u - (float) xPos, dx2 = 1.0f - dx1, dy1 = w - (float) yPos, dy2 = 1.0f - dy1;
    const V3 value = envmap_texel(ev, xPos, yPos) * dx2 * dy2 + envmap_texel(ev, xPos, yPos + 1) * dx2 * dy1
                   + envmap_texel(ev, xPos + 1, yPos) * dx1 * dy2 + envmap_texel(ev, xPos + 1, yPos + 1) * dx1 * dy1;   // mipmap.h:592-595
    return value * ev.scale;
}

// EnvironmentMap::sampleReuse (envmap.cpp:655-661): std::lower_bound on a float CDF
MTSB_HD uint32_t envmap_sample_reuse(const float *cdf, uint32_t size, float &sample) {
    uint32_t lo = 0, n = size + 1;
    while (n > 0) {                                  // first entry that is not < sample
        const uint32_t half = n >> 1;
        if (cdf[lo + half] < sample) { lo += half + 1; n -= half + 1; } else n = half;
    }
    const int e = (int) lo - 1;
    uint32_t index = (uint32_t) (e < 0 ? 0 : e);
    if (index > size - 1) index = size - 1;
    sample = (sample - cdf[index]) / (cdf[index + 1] - cdf[index]);
    return index;
}
MTSB_HD float interval_to_tent(float sample) {       // warp.cpp:157-169
    float sign;
    if (sample < 0.5f) { sign = 1; sample *= 2; } else { sign = -1; sample = 2 * (sample - 0.5f); }
    return sign * (1 - sqrtf(sample));
}

// envmap.cpp:567-600; d is in the emitter's local frame
MTSB_HD void envmap_sample_direction(const EnvmapView &ev, float sx, float sy, V3 &d, V3 &value, float &pdf) {
    const uint32_t row = envmap_sample_reuse(ev.cdf_rows, (uint32_t) ev.height, sy);
    const uint32_t col = envmap_sample_reuse(ev.cdf_cols + (size_t) row * (size_t) (ev.width + 1), (uint32_t) ev.width, sx);
    const float posx = (float) col + interval_to_tent(sx), posy = (float) row + interval_to_tent(sy);
    V3 value1, value2; int yPos;
    envmap_rows(ev, posx, posy, value1, value2, yPos);
    value = (value1 + value2) * ev.scale;
    const int y0 = yPos < 0 ? 0 : (yPos > ev.height - 1 ? ev.height - 1 : yPos);
    const int y1 = yPos + 1 < 0 ? 0 : (yPos + 1 > ev.height - 1 ? ev.height - 1 : yPos + 1);
    pdf = (envmap_luminance(value1) * ev.row_weights[y0] + envmap_luminance(value2) * ev.row_weights[y1]) * ev.normalization;
    float sinPhi, cosPhi, sinTheta, cosTheta;
    sincosf(ev.pixel_size_x * (posx + 0.5f), &sinPhi, &cosPhi);
    sincosf(ev.pixel_size_y * (posy + 0.5f), &sinTheta, &cosTheta);
    d = V3(sinPhi * sinTheta, cosTheta, -cosPhi * sinTheta);
    pdf /= fmaxf(fabsf(sinTheta), MTSB_EPSILON);
}

// envmap.cpp:603-633; d in the local frame
MTSB_HD float envmap_pdf_direction(const EnvmapView &ev, V3 d) {
    const float uvx = atan2f(d.x, -d.z) * MTSB_INV_TWOPI, uvy = envmap_safe_acos(d.y) * MTSB_INV_PI;
    if (!isfinite(uvx) || !isfinite(uvy)) return 0.0f;
    const float u = uvx * (float) ev.width - 0.5f, v = uvy * (float) ev.height - 0.5f;
    V3 value1, value2; int yPos;
    envmap_rows(ev, u, v, value1, value2, yPos);
    const int y0 = yPos < 0 ? 0 : (yPos > ev.height - 1 ? ev.height - 1 : yPos);
    const int y1 = yPos + 1 < 0 ? 0 : (yPos + 1 > ev.height - 1 ? ev.height - 1 : yPos + 1);
    const float sinTheta = sqrtf(fmaxf(0.0f, 1 - d.y * d.y));                      // math::safe_sqrt
    return (envmap_luminance(value1) * ev.row_weights[y0] + envmap_luminance(value2) * ev.row_weights[y1])
           * ev.normalization / fmaxf(fabsf(sinTheta), MTSB_EPSILON);
}

}  // namespace mtsb
