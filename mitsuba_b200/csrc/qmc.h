// qmc.h -- quasi-Monte Carlo sample generators of the `path` integrator: `halton` and `ldsampler`
// (reference: src/samplers/halton.cpp, src/samplers/ldsampler.cpp, src/samplers/faure.cpp,
// src/libcore/qmc.cpp:86-196, include/mitsuba/core/qmc.h:41-121).
//
// The reference samplers are stateful objects (generate(pixel) / advance() / next1D() / next2D()); a
// wavefront needs random access, so every draw here is a pure function of
// (pixel, sample index, dimension).  `halton` is deterministic: the values are the reference's bit for
// bit (same integer digit reversal, same float products).  `ldsampler` shuffles its per-pixel point
// sets with the worker's random stream in the reference; here the scramble words and the shuffle are
// keyed by (seed, pixel, dimension) through Philox and a keyed bijection, so parity is statistical.
#pragma once
#include "hd.h"
#include "rng.h"
#include "../../include/mtsb200.h"

namespace mtsb {

#define MTSB_PRIME_TABLE_SIZE 1024            /* include/mitsuba/core/qmc.h:38 */
#define MTSB_HALTON_MAX_RESOLUTION 128        /* halton.cpp:29 */
#define MTSB_ONE_MINUS_EPS 0.999999940395355225f   /* constants.h:51 */

// Device-side description of the sample generator (part of SceneView; pointers are device pointers,
// or host pointers in the CPU emulation).
struct SamplerView {
    uint32_t type;                  // MTSB_SAMPLER_*
    uint32_t sample_count;          // sampler sampleCount (ldsampler: a power of two)
    uint32_t log2_count;            // ldsampler: log2(sample_count)
    uint32_t ld_dimension;          // ldsampler: number of low-discrepancy 1D and 2D dimensions
    // halton: space partition of the first two dimensions (halton.cpp:240-275)
    uint32_t prime_pow[2], prime_exp[2];
    uint64_t stride, mult_inv[2];
    const uint16_t *primes;         // MTSB_PRIME_TABLE_SIZE
    const uint32_t *perm_offset;    // per dimension: offset of its permutation in `perms`
    const uint16_t *perms;          // digit permutations (faure.cpp), NULL when scramble == 0
    uint16_t inv_perm2[2], inv_perm3[3], _pad;   // inverses of the first two permutations (faure.cpp:77-83)
    // sobol (sobol.cpp:146-158, sobolseq.h:99-134)
    uint32_t sobol_log_res;         // m: the film is covered by 2^m x 2^m elementary squares
    float sobol_res;                // 2^m as a float
    uint64_t sobol_scramble;        // sobol.cpp:89-102
    const uint32_t *sobol_mat;      // generator matrices, MTSB_SOBOL_SIZE columns per dimension
    const uint64_t *sobol_vdc;      // MTSB_SOBOL_SIZE entries: pixel bits flipped by sample-index bit c (sobolseq.h:110-113)
    const uint64_t *sobol_vdc_inv;  // MTSB_SOBOL_SIZE entries: index bits that set pixel bit c (sobolseq.h:126-128)
};
#define MTSB_SOBOL_SIZE 52                    /* sobolseq.h:31 */
#define MTSB_SOBOL_NUM_DIMENSIONS 1024        /* sobolseq.h:30 */

MTSB_HD float u64_to_float(uint64_t v) {
#if MTSB_DEVICE_CODE
    return __ull2float_rn(v);
#else
    return (float) v;
#endif
}
MTSB_HD uint32_t bit_reverse32(uint32_t n) {
#if MTSB_DEVICE_CODE
    return __brev(n);
#else
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ffu) << 8) | ((n & 0xff00ff00u) >> 8);
    n = ((n & 0x0f0f0f0fu) << 4) | ((n & 0xf0f0f0f0u) >> 4);
    n = ((n & 0x33333333u) << 2) | ((n & 0xccccccccu) >> 2);
    n = ((n & 0x55555555u) << 1) | ((n & 0xaaaaaaaau) >> 1);
    return n;
#endif
}

// radicalInverseFast / scrambledRadicalInverseFast (qmc.cpp:134-196): digits reversed as an integer,
// one float product per digit for the scale, one final integer->float conversion and product.
MTSB_HD float radical_inverse(uint32_t base, uint64_t index, const uint16_t *perm) {
    const float radical = fdiv(1.0f, (float) base);
    uint64_t value = 0;
    float factor = 1.0f;
    if (index >> 32) {
        while (index >> 32) {
            const uint64_t next = index / base, digit = index - next * base;
            value = value * base + (perm ? (uint64_t) perm[digit] : digit);
            factor = fmul(factor, radical);
            index = next;
        }
    }
    uint32_t idx = (uint32_t) index;       // 32-bit divisions for the common case
    while (idx) {
        const uint32_t next = idx / base, digit = idx - next * base;
        value = value * base + (perm ? (uint64_t) perm[digit] : (uint64_t) digit);
        factor = fmul(factor, radical);
        idx = next;
    }
    return fminf(fmul(u64_to_float(value), factor), MTSB_ONE_MINUS_EPS);
}

// HaltonSampler::inverseScrambledRadicalInverse (halton.cpp:193-209)
MTSB_HD uint64_t inverse_radical_inverse(uint32_t base, uint64_t inverse, uint32_t digits, const uint16_t *inv_perm) {
    uint64_t index = 0;
    while (digits) {
        uint64_t digit = inverse % base;
        if (inv_perm) digit = inv_perm[digit];
        inverse /= base;
        index = index * base + digit;
        --digits;
    }
    return index;
}

// HaltonSampler::generate (halton.cpp:277-296): first sequence index that falls into the pixel; the
// samples of the pixel are offset + stride * sampleIndex.  (px, py) are absolute film coordinates.
MTSB_HD uint64_t halton_pixel_offset(const SamplerView &sv, int px, int py, int pix_mod[2]) {
    pix_mod[0] = px % MTSB_HALTON_MAX_RESOLUTION; pix_mod[1] = py % MTSB_HALTON_MAX_RESOLUTION;
    if (sv.stride <= 1) return 0;
    uint64_t offset = 0;
    for (int i = 0; i < 2; ++i) {
        const uint16_t *inv = sv.perms ? (i == 0 ? sv.inv_perm2 : sv.inv_perm3) : (const uint16_t *) 0;
        const uint64_t o = inverse_radical_inverse(i == 0 ? 2u : 3u, (uint64_t) pix_mod[i], sv.prime_exp[i], inv);
        offset += o * (sv.stride / sv.prime_pow[i]) * sv.mult_inv[i];
    }
    return offset % sv.stride;
}

// HaltonSampler::nextFloat (halton.cpp:349-356) for an explicit dimension
MTSB_HD float halton_value(const SamplerView &sv, uint32_t dim, uint64_t index) {
    if (dim >= MTSB_PRIME_TABLE_SIZE) dim = MTSB_PRIME_TABLE_SIZE - 1;    // the reference raises an error here (halton.cpp:362-364)
    return radical_inverse(sv.primes[dim], index, sv.perms ? sv.perms + sv.perm_offset[dim] : (const uint16_t *) 0);
}

// sobol::sampleSingle (sobolseq.h:43-59)
MTSB_HD float sobol_value(const SamplerView &sv, uint32_t dim, uint64_t index) {
    if (dim >= MTSB_SOBOL_NUM_DIMENSIONS) dim = MTSB_SOBOL_NUM_DIMENSIONS - 1;     // the reference raises an error (sobol.cpp:228-230)
    uint32_t result = (uint32_t) sv.sobol_scramble;
    for (const uint32_t *col = sv.sobol_mat + (size_t) dim * MTSB_SOBOL_SIZE; index; index >>= 1, ++col)
        if (index & 1) result ^= *col;
    return fminf(fmul((float) result, 2.3283064365386963e-10f), MTSB_ONE_MINUS_EPS);
}
// sobol::look_up (sobolseq.h:99-134, single precision): index of the frame-th point inside pixel (px, py)
MTSB_HD uint64_t sobol_look_up(const SamplerView &sv, uint32_t frame, uint32_t px, uint32_t py) {
    const uint32_t m = sv.sobol_log_res;
    uint64_t index = (uint64_t) frame << (m << 1);
    uint64_t delta = 0;
    for (uint32_t c = 0; frame; frame >>= 1, ++c)
        if (frame & 1) delta ^= sv.sobol_vdc[c];
    const uint64_t scramble = (sv.sobol_scramble & 0xFFFFFFFFull) >> (32 - m);
    uint64_t b = ((((uint64_t) px ^ scramble) << m) | ((uint64_t) py ^ scramble)) ^ delta;
    for (uint32_t c = 0; b; b >>= 1, ++c)
        if (b & 1) index ^= sv.sobol_vdc_inv[c];
    return index;
}
// SobolSampler::setSampleIndex (sobol.cpp:204-216)
MTSB_HD uint64_t sobol_sample_index(const SamplerView &sv, uint32_t sample_index, int px, int py) {
    if (sv.sobol_log_res > 1 && px >= 0) return sobol_look_up(sv, sample_index, (uint32_t) px, (uint32_t) py);
    return (uint64_t) sample_index;
}

// radicalInverse2Single / sobol2Single / sample02Single (include/mitsuba/core/qmc.h:41-59, 79-121)
MTSB_HD float radical_inverse2_single(uint32_t n, uint32_t scramble) {
    n = bit_reverse32(n);
    n = (n >> (32 - 24)) ^ (scramble & 0x00FFFFFFu);
    return fdiv((float) n, 16777216.0f);
}
MTSB_HD float sobol2_single(uint32_t n, uint32_t scramble) {
    for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
        if (n & 1) scramble ^= v;
    return fdiv((float) scramble, 4294967296.0f);
}

// Keyed bijection on [0, 2^bits): the shuffle of LowDiscrepancySampler::generate1D/2D (ldsampler.cpp:152-183)
// with random access.  Every step (xor, odd multiply, xor-shift right, add; all masked) is invertible.
MTSB_HD uint32_t permute_pow2(uint32_t i, uint32_t bits, uint32_t ka, uint32_t kb) {
    if (bits == 0) return 0;
    const uint32_t mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
    const uint32_t s1 = (bits + 1) / 2, s2 = bits > 2 ? bits / 3 : 1;
    i = (i ^ ka) & mask;
    i = (i * (kb | 1u)) & mask;          i ^= i >> s1;
    i = (i * 0x9E3779B1u) & mask;        i ^= i >> s2;
    i = (i + (ka >> 16)) & mask;
    i = (i * ((kb >> 8) | 1u)) & mask;   i ^= i >> s1;
    i = (i * 0x85EBCA6Bu) & mask;        i ^= i >> s2;
    i = (i + (kb >> 20)) & mask;
    return i;
}

// Scramble words and shuffle keys of one (pixel, kind, dimension) of the ldsampler.  The Philox counter
// word 3 is 0xFFFFFFFF, which no per-sample stream uses (rng.h: sub-block 0 or 1).
MTSB_HD U32x4 ld_pixel_keys(uint64_t seed, uint64_t pixel, uint32_t kind, uint32_t dim) {
    U32x4 c; c.x = (uint32_t) pixel; c.y = (uint32_t) (pixel >> 32); c.z = (kind << 16) | dim; c.w = 0xFFFFFFFFu;
    return philox4x32_10(c, (uint32_t) seed, (uint32_t) (seed >> 32));
}
MTSB_HD float ld_value_1d(const SamplerView &sv, uint64_t seed, uint64_t pixel, uint32_t dim, uint32_t sample_index) {
    const U32x4 k = ld_pixel_keys(seed, pixel, 1, dim);
    return radical_inverse2_single(permute_pow2(sample_index, sv.log2_count, k.z, k.w), k.x);
}
MTSB_HD void ld_value_2d(const SamplerView &sv, uint64_t seed, uint64_t pixel, uint32_t dim, uint32_t sample_index, float out[2]) {
    const U32x4 k = ld_pixel_keys(seed, pixel, 2, dim);
    const uint32_t n = permute_pow2(sample_index, sv.log2_count, k.z, k.w);
    out[0] = radical_inverse2_single(n, k.x);
    out[1] = sobol2_single(n, k.y);
}

}  // namespace mtsb

