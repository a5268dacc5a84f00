// ORACLE (test infrastructure, not product code): CPU restatement of the reference's quasi-Monte Carlo
// sample generators as the stateful objects the reference has -- generate(pixel) once per pixel,
// advance() once per sample, next1D()/next2D() in call order.
//   HaltonSampler            src/samplers/halton.cpp:121-395   (+ PermutationStorage, src/samplers/faure.cpp:22-135)
//   LowDiscrepancySampler    src/samplers/ldsampler.cpp:72-215
//   radical inverses         src/libcore/qmc.cpp:86-196, include/mitsuba/core/qmc.h:41-121
// Pinned by tests/golden/qmc.npz (values produced by the reference's own qmc.cpp / faure.cpp compiled into
// oracle/_ref) and by the known-answer table of src/tests/test_samplers.cpp:37-56.
#pragma once
#include <cstdint>
#include <vector>
#include <algorithm>
#include <stdexcept>

namespace orc {

static const int kPrimeTableSize = 1024;            // qmc.h:38
static const float kOneMinusEps = 0.999999940395355225f;   // constants.h:51

// the first 1024 primes by trial division (the reference keeps a literal table, qmc.cpp:23-84)
inline const std::vector<int> &primeTable() {
    static std::vector<int> t;
    if (t.empty()) {
        for (int c = 2; (int) t.size() < kPrimeTableSize; ++c) {
            bool prime = true;
            for (size_t k = 0; k < t.size() && t[k] * t[k] <= c; ++k) if (c % t[k] == 0) { prime = false; break; }
            if (prime) t.push_back(c);
        }
    }
    return t;
}

// qmc.cpp:86-97 (generic radicalInverse, used by the incremental test of test_samplers.cpp:79-86)
inline float radicalInverse(int base, uint64_t i) {
    float radical = 1.0f / (float) base, inverse = 0.0f, digit = radical;
    while (i) { inverse += digit * (float) (i % base); digit *= radical; i /= base; }
    return std::min(inverse, kOneMinusEps);
}
// qmc.cpp:112-130
inline float radicalInverseIncremental(int base, float x) {
    float radical = 1.0f / (float) base, h, hh, r = 1.0f - x - 1e-10f;
    if (radical < r) x += radical;
    else { h = radical; do { hh = h; h *= radical; } while (h >= r); x += hh + h - 1.0f; }
    return x;
}
// qmc.cpp:139-165,198-: the "fast" variants reverse the digits as an integer and scale once
inline float radicalInverseFast(int baseIndex, uint64_t index, const uint16_t *perm = nullptr) {
    const uint64_t base = (uint64_t) primeTable()[baseIndex];
    const float radical = 1.0f / (float) base;
    uint64_t value = 0; float factor = 1.0f;
    while (index) {
        uint64_t next = index / base, digit = index - next * base;
        value = value * base + (perm ? perm[digit] : digit);
        factor *= radical;
        index = next;
    }
    return std::min((float) value * factor, kOneMinusEps);
}

// include/mitsuba/core/qmc.h:41-58
inline float radicalInverse2Single(uint32_t n, uint32_t scramble = 0) {
    n = __builtin_bswap32(n);
    n = ((n & 0x0f0f0f0f) << 4) | ((n & 0xf0f0f0f0) >> 4);
    n = ((n & 0x33333333) << 2) | ((n & 0xcccccccc) >> 2);
    n = ((n & 0x55555555) << 1) | ((n & 0xaaaaaaaa) >> 1);
    n = (n >> (32 - 24)) ^ (scramble & ~-(1 << 24));
    return (float) n / (float) (1U << 24);
}
// qmc.h:79-85
inline float sobol2Single(uint32_t n, uint32_t scramble = 0) {
    for (uint32_t v = 1U << 31; n != 0; n >>= 1, v ^= v >> 1) if (n & 1) scramble ^= v;
    return (float) scramble / (float) (1ULL << 32);
}
// qmc.h:139-149
inline uint64_t sampleTEA(uint32_t v0, uint32_t v1, int rounds = 4) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xA341316C) ^ (v1 + sum) ^ ((v1 >> 5) + 0xC8013EA4);
        v1 += ((v0 << 4) + 0xAD90777D) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7E95761E);
    }
    return ((uint64_t) v1 << 32) + v0;
}

// PermutationStorage (faure.cpp:22-135): the triangular dynamic programme of the reference, kept as it is
struct Permutations {
    std::vector<std::vector<uint16_t>> perm;      // per prime index
    std::vector<uint16_t> inv[2];
    explicit Permutations(int scramble) {
        const std::vector<int> &pt = primeTable();
        perm.resize(kPrimeTableSize);
        if (scramble == -1) {
            const int maxBase = pt[kPrimeTableSize - 1];
            std::vector<std::vector<uint16_t>> tri(maxBase + 1);
            for (int b = 1; b <= maxBase; ++b) tri[b].resize(b);
            tri[1][0] = 0; tri[2][0] = 0; tri[2][1] = 1;
            for (int b = 2; b <= maxBase; ++b) {                  // faure.cpp:106-127
                if (b & 1) {
                    uint16_t c = (uint16_t) ((b - 1) / 2);
                    for (uint16_t i = 0; i < b; ++i) {
                        if (i == c) tri[b][i] = c;
                        else { uint16_t f = tri[b - 1][i - (int) (i > c)]; tri[b][i] = (uint16_t) (f + (int) (f >= c)); }
                    }
                } else {
                    uint16_t c = (uint16_t) (b / 2);
                    for (uint16_t i = 0; i < b; ++i) tri[b][i] = (uint16_t) (i < c ? 2 * tri[c][i] : 2 * tri[c][i - c] + 1);
                }
            }
            for (int i = 0; i < kPrimeTableSize; ++i) perm[i] = tri[pt[i]];
        } else {
            uint32_t rngIndex = 0;
            auto sampleUInt = [&](uint32_t n) {                   // faure.h:62-76
                uint64_t bitmask = n;
                bitmask |= bitmask >> 1; bitmask |= bitmask >> 2; bitmask |= bitmask >> 4; bitmask |= bitmask >> 8; bitmask |= bitmask >> 16;
                uint32_t result;
                while ((result = (uint32_t) (sampleTEA((uint32_t) scramble, rngIndex++) & bitmask)) >= n) { }
                return result;
            };
            for (int i = 0; i < kPrimeTableSize; ++i) {
                std::vector<uint16_t> &p = perm[i];
                p.resize(pt[i]);
                for (int j = 0; j < pt[i]; ++j) p[j] = (uint16_t) j;
                for (int it = pt[i] - 1; it > 0; --it) std::swap(p[it], p[sampleUInt((uint32_t) it)]);
            }
        }
        for (int i = 0; i < 2; ++i) { inv[i].resize(pt[i]); for (int j = 0; j < pt[i]; ++j) inv[i][perm[i][j]] = (uint16_t) j; }
    }
};

struct SeqRandom {           // what the samplers need from Random: 64 random bits and nextFloat
    virtual uint64_t nextULong() = 0;
    virtual float nextFloat() = 0;
    virtual ~SeqRandom() {}
};

struct QmcSampler {
    virtual void generate(int px, int py) = 0;
    virtual void setSampleIndex(size_t i) = 0;
    virtual float next1D() = 0;
    virtual void next2D(float out[2]) = 0;
    virtual ~QmcSampler() {}
};

// halton.cpp
struct HaltonSampler : QmcSampler {
    static const int MAX_RESOLUTION = 128;
    const Permutations *perms;       // null: scramble == 0
    uint64_t offset = 0, stride = 1, multInverse[2] = {0, 0};
    int primePowers[2] = {1, 1}, primeExponents[2] = {0, 0}, pixelPosition[2] = {0, 0};
    uint32_t dimension = 0; size_t sampleIndex = 0;
    explicit HaltonSampler(const Permutations *p) : perms(p) {}
    static void extendedGCD(uint64_t a, uint64_t b, int64_t &x, int64_t &y) {          // :217-227
        if (b == 0) { x = 1; y = 0; return; }
        int64_t d = (int64_t) (a / b), x_, y_;
        extendedGCD(b, a % b, x_, y_);
        x = y_; y = x_ - (int64_t) (d * y_);
    }
    static uint64_t multiplicativeInverse(int64_t a, int64_t n) { int64_t x, y; extendedGCD((uint64_t) a, (uint64_t) n, x, y); int64_t r = x % n; return (uint64_t) (r < 0 ? r + n : r); }
    void setFilmResolution(int resX, int resY, bool blocked) {                          // :240-275
        const int res[2] = {resX, resY};
        if (blocked) {
            stride = 1;
            for (int i = 0; i < 2; ++i) {
                int prime = primeTable()[i], value = 1, exp = 0;
                while (value < std::min(res[i], MAX_RESOLUTION)) { value *= prime; ++exp; }
                primePowers[i] = value; primeExponents[i] = exp; stride *= (uint64_t) value;
            }
            multInverse[0] = multiplicativeInverse(primePowers[1], primePowers[0]);
            multInverse[1] = multiplicativeInverse(primePowers[0], primePowers[1]);
        } else {
            primeExponents[0] = primeExponents[1] = 0; primePowers[0] = primePowers[1] = 1; multInverse[0] = multInverse[1] = 0; stride = 1;
        }
        pixelPosition[0] = pixelPosition[1] = 0; offset = 0;
    }
    static uint64_t inverseScrambledRadicalInverse(int base, uint64_t inverse, uint64_t digits, const uint16_t *invPerm) {   // :193-209
        uint64_t index = 0;
        while (digits) {
            uint64_t digit = inverse % base;
            if (invPerm) digit = (uint64_t) invPerm[digit];
            inverse /= base; index = index * base + digit; --digits;
        }
        return index;
    }
    void generate(int px, int py) override {                                            // :277-296 (no sample arrays on this path)
        const int pos[2] = {px, py};
        offset = 0;
        if (stride > 1) {
            for (int i = 0; i < 2; ++i) {
                pixelPosition[i] = pos[i] % MAX_RESOLUTION;
                uint64_t o = inverseScrambledRadicalInverse(primeTable()[i], (uint64_t) pixelPosition[i], (uint64_t) primeExponents[i],
                                                            perms ? perms->inv[i].data() : nullptr);
                offset += o * (stride / primePowers[i]) * multInverse[i];
            }
            offset %= stride;
        }
        setSampleIndex(0);
    }
    void setSampleIndex(size_t i) override { dimension = 0; sampleIndex = i; }
    float nextFloat(uint64_t idx) {                                                     // :349-356
        uint32_t dim = dimension++;
        if (dim >= (uint32_t) kPrimeTableSize) throw std::runtime_error("Lookup dimension exceeds the prime number table size!");
        return radicalInverseFast((int) dim, idx, perms ? perms->perm[dim].data() : nullptr);
    }
    float next1D() override { return nextFloat(offset + stride * sampleIndex); }
    void next2D(float out[2]) override {                                                // :372-395
        uint64_t index = offset + stride * sampleIndex;
        if (dimension == 0) {
            out[0] = nextFloat(index) * primePowers[0] - pixelPosition[0];
            out[1] = nextFloat(index) * primePowers[1] - pixelPosition[1];
        } else { out[0] = nextFloat(index); out[1] = nextFloat(index); }
    }
};

// ldsampler.cpp
struct LowDiscrepancySampler : QmcSampler {
    size_t sampleCount, maxDimension, sampleIndex = 0, dimension1D = 0, dimension2D = 0;
    SeqRandom *random;
    std::vector<std::vector<float>> samples1D;
    std::vector<std::vector<float>> samples2D;       // interleaved x,y
    LowDiscrepancySampler(size_t count, size_t maxDim, SeqRandom *r) : sampleCount(count), maxDimension(maxDim), random(r) {
        size_t p = 1; while (p < sampleCount) p <<= 1; sampleCount = p;          // :82-86
        samples1D.assign(maxDimension, std::vector<float>(sampleCount));
        samples2D.assign(maxDimension, std::vector<float>(2 * sampleCount));
    }
    // Random::nextSize (random.cpp:604-614): rejection on a bit mask, result in [0, n)
    size_t nextSize(size_t n) {
        uint64_t bitmask = n; bitmask |= bitmask >> 1; bitmask |= bitmask >> 2; bitmask |= bitmask >> 4; bitmask |= bitmask >> 8; bitmask |= bitmask >> 16; bitmask |= bitmask >> 32;
        uint64_t r;
        while ((r = random->nextULong() & bitmask) >= n) { }
        return (size_t) r;
    }
    void generate(int, int) override {                                                  // :185-197
        for (size_t d = 0; d < maxDimension; ++d) {
            uint32_t scramble = (uint32_t) (random->nextULong() & 0xFFFFFFFFull);                     // generate1D :152-163
            for (size_t i = 0; i < sampleCount; ++i) samples1D[d][i] = radicalInverse2Single((uint32_t) i, scramble);
            for (size_t it = sampleCount - 1; it > 0; --it) std::swap(samples1D[d][it], samples1D[d][nextSize(it)]);   // Random::shuffle, random.h:149-152
            union { uint64_t qword; uint32_t dword[2]; } s; s.qword = random->nextULong();             // generate2D :165-183
            for (size_t i = 0; i < sampleCount; ++i) {
                samples2D[d][2 * i] = radicalInverse2Single((uint32_t) i, s.dword[0]);
                samples2D[d][2 * i + 1] = sobol2Single((uint32_t) i, s.dword[1]);
            }
            for (size_t it = sampleCount - 1; it > 0; --it) {
                size_t o = nextSize(it);
                std::swap(samples2D[d][2 * it], samples2D[d][2 * o]); std::swap(samples2D[d][2 * it + 1], samples2D[d][2 * o + 1]);
            }
        }
        sampleIndex = 0; dimension1D = dimension2D = 0;
    }
    void setSampleIndex(size_t i) override { sampleIndex = i; dimension1D = dimension2D = 0; }
    float next1D() override {                                                           // :199-205
        if (sampleIndex >= sampleCount) throw std::runtime_error("ldsampler: sample index out of range");
        if (dimension1D < maxDimension) return samples1D[dimension1D++][sampleIndex];
        return random->nextFloat();
    }
    void next2D(float out[2]) override {                                                // :207-213
        if (sampleIndex >= sampleCount) throw std::runtime_error("ldsampler: sample index out of range");
        if (dimension2D < maxDimension) { out[0] = samples2D[dimension2D][2 * sampleIndex]; out[1] = samples2D[dimension2D][2 * sampleIndex + 1]; ++dimension2D; }
        else { out[0] = random->nextFloat(); out[1] = random->nextFloat(); }
    }
};

// ---- Sobol' sequence: sobolseq.h:29-134 (sampleSingle, look_up) and SobolSampler, sobol.cpp:85-252 ----
// The reference carries the expanded generator matrices and the per-resolution pixel tables as literal data
// (sobolseq.cpp, 108 000 lines).  They are functions of the Joe-Kuo seed numbers, so they are rebuilt here from the
// seeds (../mitsuba_b200/csrc/sobol_joe_kuo.inc holds only those seeds, generated from the public data set) with the
// textbook direction-number recurrence v_c = v_{c-s} ^ (v_{c-s} >> s) ^ sum_k a_k v_{c-k}; tests/golden/qmc_sobol.npz
// pins every entry of the result against the reference's own tables.
namespace sobol_data {
#include "../mitsuba_b200/csrc/sobol_joe_kuo.inc"
}
struct SobolTables {
    static const uint32_t numDimensions = 1024, size = 52;
    std::vector<uint64_t> matrices64;                 // direction numbers scaled by 2^52
    std::vector<uint32_t> matrices32;                 // their leading 32 bits
    SobolTables() : matrices64((size_t) numDimensions * size), matrices32((size_t) numDimensions * size) {
        for (uint32_t d = 0; d < numDimensions; ++d) {
            uint64_t *v = &matrices64[(size_t) d * size];
            const uint32_t poly = sobol_data::kSobolPoly[d];
            uint32_t s = 0; while (poly >> (s + 1)) ++s;
            for (uint32_t c = 0; c < size; ++c) {
                if (s == 0) v[c] = 1ull << (size - 1 - c);
                else if (c < s) v[c] = (uint64_t) sobol_data::kSobolInit[d][c] << (size - 1 - c);
                else {
                    v[c] = v[c - s] ^ (v[c - s] >> s);
                    for (uint32_t k = 1; k < s; ++k) if ((poly >> (s - k)) & 1u) v[c] ^= v[c - k];
                }
                matrices32[(size_t) d * size + c] = (uint32_t) (v[c] >> (size - 32));
            }
        }
    }
    // pixel tables of one resolution 2^m: vdc[c] = y bits flipped by index bit 2m+c; inv[c] = low index bits that
    // set bit c of (px << m | py).  Built by solving the 2m x 2m system bit by bit (row elimination).
    void pixelTables(uint32_t m, std::vector<uint64_t> &vdc, std::vector<uint64_t> &inv) const {
        vdc.assign(size, 0); inv.assign(size, 0);
        if (m == 0) return;
        const uint32_t n = 2 * m;
        auto word = [&](uint32_t col) { return ((matrices64[col] >> (size - m)) << m) | (matrices64[size + col] >> (size - m)); };
        for (uint32_t c = 0; n + c < size; ++c) vdc[c] = matrices64[size + n + c] >> (size - m);
        // rows[r]: bit j = coefficient of index bit j in output bit r; augmented with the identity in the high half
        std::vector<unsigned __int128> rows(n);
        for (uint32_t r = 0; r < n; ++r) {
            unsigned __int128 row = (unsigned __int128) 1 << (64 + r);
            for (uint32_t j = 0; j < n; ++j) if ((word(j) >> r) & 1ull) row |= (unsigned __int128) 1 << j;
            rows[r] = row;
        }
        for (uint32_t j = 0; j < n; ++j) {
            uint32_t p = j; while (p < n && !((rows[p] >> j) & 1)) ++p;
            if (p == n) throw std::runtime_error("sobol: singular pixel map");
            std::swap(rows[p], rows[j]);
            for (uint32_t r = 0; r < n; ++r) if (r != j && ((rows[r] >> j) & 1)) rows[r] ^= rows[j];
        }
        // now index bit j = sum over output bits r of rows[j].high[r] * out_r  =>  column c of the inverse
        for (uint32_t c = 0; c < n; ++c) { uint64_t col = 0; for (uint32_t j = 0; j < n; ++j) if ((rows[j] >> (64 + c)) & 1) col |= 1ull << j; inv[c] = col; }
    }
    static const SobolTables &get() { static SobolTables t; return t; }
};
// sobolseq.h:43-59
inline float sobolSampleSingle(uint64_t index, uint32_t dimension, uint32_t scramble = 0) {
    const SobolTables &t = SobolTables::get();
    uint32_t result = scramble;
    for (uint32_t i = dimension * SobolTables::size; index; index >>= 1, ++i) if (index & 1) result ^= t.matrices32[i];
    return std::min(result * (1.0f / (1ULL << 32)), 0.999999940395355225f);
}
// sobolseq.h:99-134 (single precision branch)
inline uint64_t sobolLookUp(uint32_t m, uint32_t frame, uint32_t px, uint32_t py, uint64_t scramble, const std::vector<uint64_t> &vdc, const std::vector<uint64_t> &inv) {
    const uint32_t m2 = m << 1;
    uint64_t index = uint64_t(frame) << m2, delta = 0;
    for (uint32_t c = 0; frame; frame >>= 1, ++c) if (frame & 1) delta ^= vdc[c];
    scramble = (scramble & 0xFFFFFFFF) >> (32 - m);
    uint64_t b = (((uint64_t) (px ^ scramble) << m) | (py ^ scramble)) ^ delta;
    for (uint32_t c = 0; b; b >>= 1, ++c) if (b & 1) index ^= inv[c];
    return index;
}
struct SobolSampler : QmcSampler {
    uint32_t dimension = 0, logResolution = 0; uint64_t scramble, sobolSampleIndex = 0; float resolution = 1;
    int pixelPosition[2] = {0, 0}; size_t sampleIndex = 0;
    std::vector<uint64_t> vdc, inv;
    explicit SobolSampler(uint64_t scrambleProp) : scramble(scrambleProp) {             // sobol.cpp:85-107
        if (scramble) { union { uint64_t ui64; uint32_t v[2]; } u = {scramble}; scramble = sampleTEA(u.v[0], u.v[1]); }
    }
    void setFilmResolution(int resX, int resY, bool bucketed) {                         // :146-158
        if (!bucketed) { resolution = 1; logResolution = 0; }
        else {
            uint32_t r = 1, l = 0; while (r < (uint32_t) std::max(resX, resY)) { r <<= 1; ++l; }
            resolution = (float) r; logResolution = l;
        }
        SobolTables::get().pixelTables(logResolution, vdc, inv);
    }
    void generate(int px, int py) override { pixelPosition[0] = px; pixelPosition[1] = py; setSampleIndex(0); }   // :172-175
    void setSampleIndex(size_t i) override {                                            // :204-216
        dimension = 0; sampleIndex = i;
        if (logResolution > 1 && pixelPosition[0] >= 0)
            sobolSampleIndex = sobolLookUp(logResolution, (uint32_t) sampleIndex, (uint32_t) pixelPosition[0], (uint32_t) pixelPosition[1], scramble, vdc, inv);
        else sobolSampleIndex = (uint64_t) sampleIndex;
    }
    float draw() {
        if (dimension >= SobolTables::numDimensions) throw std::runtime_error("Lookup dimension exceeds the direction number table size!");
        return sobolSampleSingle(sobolSampleIndex, dimension++, (uint32_t) scramble);
    }
    float next1D() override { return draw(); }                                          // :218-229
    void next2D(float out[2]) override {                                                // :231-252
        if (dimension == 0 && sobolSampleIndex != (uint64_t) sampleIndex) {
            out[0] = draw() * resolution - pixelPosition[0];
            out[1] = draw() * resolution - pixelPosition[1];
        } else { out[0] = draw(); out[1] = draw(); }
    }
};

}  // namespace orc
