// rng.h -- counter-based replacement of IndependentSampler / Random::nextFloat
// (reference: src/samplers/independent.cpp:97-105, src/libcore/random.cpp:630-640).
// The reference draws from one SFMT19937 stream per worker thread; a wavefront renderer
// needs random access instead, so each path sample owns a Philox4x32-10 stream keyed by
// (seed, global sample id) and numbered by dimension.  The float construction is the
// reference's: 23 random mantissa bits OR 0x3f800000, minus 1  ->  [0,1).
#pragma once
#include "hd.h"

namespace mtsb {

struct U32x4 { uint32_t x, y, z, w; };

MTSB_HD void mulhilo(uint32_t a, uint32_t b, uint32_t &hi, uint32_t &lo) {
#if MTSB_DEVICE_CODE
    lo = a * b; hi = __umulhi(a, b);
#else
    uint64_t p = (uint64_t) a * b; lo = (uint32_t) p; hi = (uint32_t) (p >> 32);
#endif
}

MTSB_HD U32x4 philox4x32_10(U32x4 ctr, uint32_t k0, uint32_t k1) {
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        mulhilo(0xD2511F53u, ctr.x, hi0, lo0);
        mulhilo(0xCD9E8D57u, ctr.z, hi1, lo1);
        U32x4 n;
        n.x = hi1 ^ ctr.y ^ k0; n.y = lo1; n.z = hi0 ^ ctr.w ^ k1; n.w = lo0;
        ctr = n;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return ctr;
}

MTSB_HD float u32_to_unit_float(uint32_t x) { return u2f((x >> 9) | 0x3f800000u) - 1.0f; }

// One stream per path sample.  `block` selects a group of four 32-bit outputs; the path
// integrator uses block 0 for the pixel offset and one block per bounce afterwards.
struct SampleStream {
    uint32_t k0, k1, id_lo, id_hi;
    MTSB_HD SampleStream(uint64_t seed, uint64_t sample_id)
        : k0((uint32_t) seed), k1((uint32_t) (seed >> 32)), id_lo((uint32_t) sample_id), id_hi((uint32_t) (sample_id >> 32)) {}
    MTSB_HD void block4(uint32_t block, uint32_t sub, float out[4]) const {
        U32x4 c; c.x = id_lo; c.y = id_hi; c.z = block; c.w = sub;
        U32x4 r = philox4x32_10(c, k0, k1);
        out[0] = u32_to_unit_float(r.x); out[1] = u32_to_unit_float(r.y);
        out[2] = u32_to_unit_float(r.z); out[3] = u32_to_unit_float(r.w);
    }
};

}  // namespace mtsb
