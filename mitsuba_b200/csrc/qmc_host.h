// qmc_host.h -- host-side tables of the quasi-Monte Carlo samplers (prime table, Faure / random digit
// permutations, the space partition of the Halton sequence).  Reference: src/libcore/qmc.cpp:23-84,
// src/samplers/faure.cpp, src/samplers/halton.cpp:211-275.
#pragma once
#include "qmc.h"
#include <vector>
#include <map>
namespace mtsb {

#include "sobol_joe_kuo.inc"

// the first 1024 primes (qmc.cpp:23-84 holds them as a literal table)
inline std::vector<uint16_t> qmc_prime_table() {
    std::vector<uint16_t> primes;
    std::vector<uint8_t> composite(8200, 0);
    for (uint32_t i = 2; i < composite.size() && primes.size() < MTSB_PRIME_TABLE_SIZE; ++i) {
        if (composite[i]) continue;
        primes.push_back((uint16_t) i);
        for (uint32_t j = i * i; j < composite.size(); j += i) composite[j] = 1;
    }
    return primes;
}

// Faure's permutation of base b (faure.cpp:96-128): the reference fills a triangular table for every
// base up to 8161; the recurrence only reaches b-1 (odd b) or b/2 (even b), so memoised recursion
// touches O(log b) bases per prime.
inline const std::vector<uint16_t> &faure_permutation(uint32_t b, std::map<uint32_t, std::vector<uint16_t>> &memo) {
    auto it = memo.find(b);
    if (it != memo.end()) return it->second;
    std::vector<uint16_t> p(b);
    if (b == 1) p[0] = 0;
    else if (b == 2) { p[0] = 0; p[1] = 1; }
    else if (b & 1) {
        const std::vector<uint16_t> prev = faure_permutation(b - 1, memo);
        const uint16_t c = (uint16_t) ((b - 1) / 2);
        for (uint32_t i = 0; i < b; ++i) {
            if (i == c) p[i] = c;
            else { const uint16_t f = prev[i - (i > c ? 1 : 0)]; p[i] = (uint16_t) (f + (f >= c ? 1 : 0)); }
        }
    } else {
        const std::vector<uint16_t> half = faure_permutation(b / 2, memo);
        const uint32_t c = b / 2;
        for (uint32_t i = 0; i < b; ++i) p[i] = (uint16_t) (i < c ? 2 * half[i] : 2 * half[i - c] + 1);
    }
    return memo.emplace(b, std::move(p)).first->second;
}

// sampleTEA (include/mitsuba/core/qmc.h:139-149)
inline uint64_t sample_tea(uint32_t v0, uint32_t v1, int rounds = 4) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xA341316Cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xC8013EA4u);
        v1 += ((v0 << 4) + 0xAD90777Du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7E95761Eu);
    }
    return ((uint64_t) v1 << 32) + v0;
}

// PermutationStorage (faure.cpp:22-84): permutations of all 1024 prime bases, concatenated.
// scramble == -1: Faure; otherwise random permutations from the TEA generator seeded by `scramble`.
inline void qmc_permutations(int scramble, const std::vector<uint16_t> &primes, std::vector<uint16_t> &perms, std::vector<uint32_t> &offsets) {
    size_t total = 0;
    offsets.resize(primes.size());
    for (size_t i = 0; i < primes.size(); ++i) { offsets[i] = (uint32_t) total; total += primes[i]; }
    perms.resize(total);
    if (scramble == -1) {
        std::map<uint32_t, std::vector<uint16_t>> memo;
        for (size_t i = 0; i < primes.size(); ++i) {
            const std::vector<uint16_t> &p = faure_permutation(primes[i], memo);
            std::copy(p.begin(), p.end(), perms.begin() + offsets[i]);
        }
    } else {
        uint32_t rng_index = 0;
        auto sample_uint = [&](uint32_t n) -> uint32_t {          // faure.h:62-76
            uint64_t bitmask = n;
            bitmask |= bitmask >> 1; bitmask |= bitmask >> 2; bitmask |= bitmask >> 4; bitmask |= bitmask >> 8; bitmask |= bitmask >> 16;
            uint32_t result;
            while ((result = (uint32_t) (sample_tea((uint32_t) scramble, rng_index++) & bitmask)) >= n) { }
            return result;
        };
        for (size_t i = 0; i < primes.size(); ++i) {
            uint16_t *ptr = perms.data() + offsets[i];
            const int prime = primes[i];
            for (int j = 0; j < prime; ++j) ptr[j] = (uint16_t) j;
            for (uint16_t *it = ptr + prime - 1; it > ptr; --it) std::swap(*it, ptr[sample_uint((uint32_t) (it - ptr))]);   // faure.h:56-59
        }
    }
}

// Generator matrices of the Sobol' sequence from the Joe-Kuo seeds: column c of dimension d is the direction
// integer m_c placed below the binary point, v_c = m_c / 2^(c+1), kept to 52 bits and cut to the leading 32
// (the reference's matrices32 table, sobolseq.cpp:33-53283, is matrices64 >> 20).  Bratley-Fox recurrence:
// m_c = 2 a_1 m_{c-1} ^ 4 a_2 m_{c-2} ^ ... ^ 2^s m_{c-s} ^ m_{c-s}.
inline std::vector<uint32_t> sobol_matrices32() {
    std::vector<uint32_t> out((size_t) MTSB_SOBOL_NUM_DIMENSIONS * MTSB_SOBOL_SIZE);
    for (int d = 0; d < MTSB_SOBOL_NUM_DIMENSIONS; ++d) {
        const uint32_t poly = kSobolPoly[d];
        int s = 0; while ((poly >> (s + 1)) != 0) ++s;
        uint64_t mc[MTSB_SOBOL_SIZE];
        for (int c = 0; c < MTSB_SOBOL_SIZE; ++c) {
            if (s == 0) mc[c] = 1;                                 // dimension 0: van der Corput
            else if (c < s) mc[c] = kSobolInit[d][c];
            else {
                uint64_t v = mc[c - s] ^ (mc[c - s] << s);
                for (int k = 1; k < s; ++k) if ((poly >> (s - k)) & 1u) v ^= mc[c - k] << k;
                mc[c] = v;
            }
            out[(size_t) d * MTSB_SOBOL_SIZE + c] = (uint32_t) ((mc[c] << (MTSB_SOBOL_SIZE - 1 - c)) >> (MTSB_SOBOL_SIZE - 32));
        }
    }
    return out;
}

// The tables behind sobol::look_up for one resolution 2^m (the reference stores them for every m,
// sobolseq.cpp:106537-107997).  Dimension 0 is the bit reversal and dimension 1 the classic Sobol' matrix, so the
// leading m bits of both (= the pixel) are a linear map A of the low 2m index bits; index bit 2m + c flips the
// pixel's y bits by the leading m bits of column 2m + c of dimension 1 (vdc[c]); vdc_inv[c] = A^-1 e_c with
// pixel bit c counted from the least significant bit of (px << m | py).
inline bool sobol_vdc_tables(uint32_t m, const std::vector<uint32_t> &mat32, std::vector<uint64_t> &vdc, std::vector<uint64_t> &vdc_inv) {
    vdc.assign(MTSB_SOBOL_SIZE, 0); vdc_inv.assign(MTSB_SOBOL_SIZE, 0);
    if (m == 0 || 2 * m > MTSB_SOBOL_SIZE) return m == 0;
    // 52-bit columns of dimension 1 are not kept; its matrix is Pascal's triangle mod 2 and m <= 26 <= 32 leading bits suffice
    auto top = [&](uint32_t dim, uint32_t col) -> uint64_t { return (uint64_t) (mat32[(size_t) dim * MTSB_SOBOL_SIZE + col] >> (32 - m)); };
    for (uint32_t c = 0; 2 * m + c < MTSB_SOBOL_SIZE; ++c) vdc[c] = top(1, 2 * m + c);
    // A: column j (index bit j) -> pixel word (px << m | py)
    const uint32_t n = 2 * m;
    std::vector<uint64_t> a(n), inv(n);
    for (uint32_t j = 0; j < n; ++j) { a[j] = (top(0, j) << m) | top(1, j); inv[j] = 1ull << j; }
    // Gauss-Jordan over GF(2) on columns: reduce [A | I] so that A becomes the identity; inv then holds A^-1 by columns
    for (uint32_t r = 0; r < n; ++r) {
        uint32_t piv = r;
        while (piv < n && !((a[piv] >> r) & 1ull)) ++piv;
        if (piv == n) return false;
        std::swap(a[piv], a[r]); std::swap(inv[piv], inv[r]);
        for (uint32_t j = 0; j < n; ++j) if (j != r && ((a[j] >> r) & 1ull)) { a[j] ^= a[r]; inv[j] ^= inv[r]; }
    }
    for (uint32_t c = 0; c < n; ++c) vdc_inv[c] = inv[c];
    return true;
}

inline void extended_gcd(uint64_t a, uint64_t b, int64_t &x, int64_t &y) {      // halton.cpp:217-227
    if (b == 0) { x = 1; y = 0; return; }
    int64_t d = (int64_t) (a / b), x_, y_;
    extended_gcd(b, a % b, x_, y_);
    x = y_; y = x_ - d * y_;
}

// Fills the POD part of a SamplerView from the scene description (halton.cpp:240-275 with blocked = true,
// as Integrator::configureSampler asks for a SamplingIntegrator, integrator.cpp:38-42).  The table
// pointers are set by the caller.  Returns false (with a message) on an invalid description.
inline bool sampler_view_init(const MtsbSampler &sd, int sample_count, int crop_w, int crop_h, SamplerView &sv, const char **err) {
    sv = SamplerView();
    sv.type = sd.type;
    sv.sample_count = (uint32_t) (sample_count > 0 ? sample_count : 1);
    sv.stride = 1; sv.prime_pow[0] = sv.prime_pow[1] = 1;
    if (sd.type == MTSB_SAMPLER_INDEPENDENT) return true;
    if (sd.type == MTSB_SAMPLER_LDSAMPLER) {
        if (sv.sample_count & (sv.sample_count - 1)) { *err = "ldsampler: sampleCount must be a power of two (the host plugin rounds it up)"; return false; }
        if (sd.dimension < 0 || sd.dimension > 63) { *err = "ldsampler: dimension must be in [0, 63]"; return false; }
        uint32_t l = 0; while ((1u << l) < sv.sample_count) ++l;
        sv.log2_count = l; sv.ld_dimension = (uint32_t) sd.dimension;
        return true;
    }
    if (sd.type == MTSB_SAMPLER_HALTON) {
        const int res[2] = {crop_w, crop_h}, prime[2] = {2, 3};
        for (int i = 0; i < 2; ++i) {
            int value = 1, exp = 0;
            while (value < (res[i] < MTSB_HALTON_MAX_RESOLUTION ? res[i] : MTSB_HALTON_MAX_RESOLUTION)) { value *= prime[i]; ++exp; }
            sv.prime_pow[i] = (uint32_t) value; sv.prime_exp[i] = (uint32_t) exp; sv.stride *= (uint64_t) value;
        }
        for (int i = 0; i < 2; ++i) {
            int64_t x, y;
            const int64_t a = sv.prime_pow[1 - i], n = sv.prime_pow[i];
            extended_gcd((uint64_t) a, (uint64_t) n, x, y);
            int64_t r = x % n; if (r < 0) r += n;
            sv.mult_inv[i] = (uint64_t) r;
        }
        return true;
    }
    if (sd.type == MTSB_SAMPLER_SOBOL) {                // sobol.cpp:89-102, 146-158
        uint64_t scramble = (uint64_t) (uint32_t) sd.scramble;          // props.getSize("scramble", 0)
        if (sd.scramble < 0) { *err = "sobol: 'scramble' must not be negative"; return false; }
        if (scramble) scramble = sample_tea((uint32_t) scramble, (uint32_t) (scramble >> 32));
        sv.sobol_scramble = scramble;
        uint32_t res = (uint32_t) (crop_w > crop_h ? crop_w : crop_h), r = 1, l = 0;
        while (r < res) { r <<= 1; ++l; }                               // math::roundToPowerOfTwo / log2i
        sv.sobol_log_res = l; sv.sobol_res = (float) r;
        if (2 * l > MTSB_SOBOL_SIZE) { *err = "sobol: film resolution too large"; return false; }
        return true;
    }
    *err = "unknown sampler type";
    return false;
}

}  // namespace mtsb
