// texture_host.h -- scene-creation side of bitmap textures (host code of api.cu): the MIP pyramid the reference builds in
// TMIPMap's constructor (include/mitsuba/render/mipmap.h:155-304) by repeated separable resampling with a 2-lobed Lanczos
// filter (src/textures/bitmap.cpp:280-287, src/libcore/bitmap.cpp:2230-2330, Resampler in include/mitsuba/core/rfilter.h:107-190,
// clamped to [0, 1] per pass: mipmap.h:165 maxValue = 1), each level stored in half precision.
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>
#include <cstring>

#include "texture.h"

namespace mtsb {

// tap table of one resampling direction: for target sample i the taps start[i] .. start[i] + taps - 1 with normalised weights
struct LanczosTaps {
    int taps;
    std::vector<int> start;
    std::vector<float> weight;
    LanczosTaps(int sourceRes, int targetRes) : start((size_t) targetRes) {
        const float lobes = 2.0f;
        float radius = lobes, invScale = 1.0f;
        if (targetRes < sourceRes) { const float scale = (float) sourceRes / (float) targetRes; invScale = 1 / scale; radius *= scale; }
        taps = (int) std::ceil(radius * 2);
        weight.resize((size_t) taps * targetRes);
        for (int i = 0; i < targetRes; ++i) {
            const float center = ((float) i + 0.5f) / (float) targetRes * (float) sourceRes;
            start[i] = (int) std::floor(center - radius + 0.5f);
            float *w = &weight[(size_t) i * taps];
            float sum = 0;
            for (int j = 0; j < taps; ++j) {
                const float pos = (float) (start[i] + j) + 0.5f - center;
                float x = std::fabs(pos * invScale), value;
                if (x < MTSB_EPSILON) value = 1.0f;                         // lanczos.cpp:46-58
                else if (x > lobes) value = 0.0f;
                else {
                    const float x1 = (float) (3.14159265358979323846 * (double) x), x2 = x1 / lobes;
                    value = (std::sin(x1) * std::sin(x2)) / (x1 * x2);
                }
                w[j] = value;
                sum += value;
            }
            const float normalization = 1.0f / sum;
            for (int j = 0; j < taps; ++j) w[j] = w[j] * normalization;
        }
    }
};

// one pass: `lines` lines of `sourceRes` samples -> `targetRes` samples; sample k of line l sits at base(l) + k * stride
static inline void lanczos_pass(const LanczosTaps &lt, uint32_t bc, int sourceRes, int targetRes, int lines, int channels,
                                const float *src, size_t srcLine, size_t srcStride, float *dst, size_t dstLine, size_t dstStride) {
    for (int l = 0; l < lines; ++l)
        for (int i = 0; i < targetRes; ++i)
            for (int ch = 0; ch < channels; ++ch) {
                float result = 0;
                for (int j = 0; j < lt.taps; ++j) {
                    int pos = lt.start[i] + j;
                    const int outside = tex_wrap(pos, sourceRes, bc);
                    const float v = outside ? (outside == 1 ? 0.0f : 1.0f) : src[(size_t) l * srcLine + (size_t) pos * srcStride + ch];
                    result += v * lt.weight[(size_t) i * lt.taps + j];
                }
                dst[(size_t) l * dstLine + (size_t) i * dstStride + ch] = std::min(1.0f, std::max(0.0f, result));
            }
}

// float -> half, round to nearest even, overflow to infinity (the conversion of half.cpp:86-204 / IEEE 754)
static inline uint32_t float_to_half_bits(float f) {
    uint32_t x; std::memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    x &= 0x7FFFFFFFu;
    if (x >= 0x7F800000u) return sign | (x > 0x7F800000u ? 0x7E00u : 0x7C00u);          // NaN / infinity
    if (x >= 0x477FF000u) return sign | 0x7C00u;                                         // rounds to >= 2^16
    if (x < 0x33000001u) return sign;                                                    // <= 2^-25: rounds to zero
    if (x < 0x38800000u) {                                                               // half denormal
        const uint32_t shift = 126u - (x >> 23), mant = (x & 0x7FFFFFu) | 0x800000u;     // value = mant * 2^-(shift + 23 - ...)
        const uint32_t q = mant >> shift, rem = mant & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
        return sign | (q + ((rem > halfway || (rem == halfway && (q & 1u))) ? 1u : 0u));
    }
    const uint32_t e = x - 0x38000000u;                                                  // rebias 127 -> 15
    const uint32_t q = e >> 13, rem = e & 0x1FFFu;
    return sign | (q + ((rem > 0x1000u || (rem == 0x1000u && (q & 1u))) ? 1u : 0u));
}

static inline HalfTexel pack_half_texel(const float *p, int channels) {
    const uint32_t r = float_to_half_bits(p[0]);
    const uint32_t g = channels == 3 ? float_to_half_bits(p[1]) : r;
    const uint32_t b = channels == 3 ? float_to_half_bits(p[2]) : r;
    HalfTexel t; t.x = r | (g << 16); t.y = b;
    return t;
}

// Fills tv.levels / lw / lh / loff and appends all levels to `texels` (level 0 first).
static inline void build_texture_levels(const MtsbTexture &t, TextureView &tv, std::vector<HalfTexel> &texels) {
    const int ch = t.channels;
    std::vector<float> cur(t.texels, t.texels + (size_t) t.width * t.height * ch);
    for (float &v : cur) if (v < 0) v = 0;                                   // clampNegative, mipmap.h:228-236
    int w = t.width, h = t.height, level = 0;
    const bool pyramid = t.filter != MTSB_TEXFILTER_NEAREST && t.filter != MTSB_TEXFILTER_BILINEAR;
    for (;;) {
        tv.lw[level] = w; tv.lh[level] = h; tv.loff[level] = (uint32_t) texels.size();
        for (size_t i = 0; i < (size_t) w * h; ++i) texels.push_back(pack_half_texel(&cur[i * ch], ch));
        ++level;
        if (!pyramid || (w <= 1 && h <= 1)) break;
        const int nw = std::max(1, (w + 1) / 2), nh = std::max(1, (h + 1) / 2);
        if (nw != w) {                                                       // X pass into a temporary of nw x h
            std::vector<float> tmp((size_t) nw * h * ch);
            lanczos_pass(LanczosTaps(w, nw), t.wrap_u, w, nw, h, ch, cur.data(), (size_t) w * ch, (size_t) ch, tmp.data(), (size_t) nw * ch, (size_t) ch);
            cur.swap(tmp);
        }
        if (nh != h) {                                                       // Y pass: one "line" per column
            std::vector<float> out((size_t) nw * nh * ch);
            lanczos_pass(LanczosTaps(h, nh), t.wrap_v, h, nh, nw, ch, cur.data(), (size_t) ch, (size_t) nw * ch, out.data(), (size_t) ch, (size_t) nw * ch);
            cur.swap(out);
        }
        w = nw; h = nh;
    }
    tv.levels = level;
}

}  // namespace mtsb
