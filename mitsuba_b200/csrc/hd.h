// hd.h -- host/device portability layer of the kernel bodies.
//
// Every kernel of this library is written as a plain per-thread function
// `body(tid, args)` marked MTSB_HD plus a thin `__global__` wrapper in a .cu file.
// nvcc builds the product (sm_100a only).  The same bodies also compile with g++
// when MTSB_EMU is defined: tests/emu builds that variant to check the BVH builder
// and the traversal logic on the CPU-only CI box.  The emulation is a TEST target; the
// product library (libmtsb200.so) contains no host execution path at all.
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define MTSB_HD __host__ __device__ __forceinline__
#define MTSB_D __device__ __forceinline__
#else
#define MTSB_HD inline
#define MTSB_D inline
#endif

#if defined(__CUDA_ARCH__)
#define MTSB_DEVICE_CODE 1
#else
#define MTSB_DEVICE_CODE 0
#endif

namespace mtsb {

// ---- exactly rounded, never-contracted arithmetic -------------------------------
// The triangle test must reproduce the reference's x86 SSE evaluation bit for bit
// (no FMA contraction), so it is spelled with these.
#if MTSB_DEVICE_CODE
MTSB_HD float fmul(float a, float b) { return __fmul_rn(a, b); }
MTSB_HD float fadd(float a, float b) { return __fadd_rn(a, b); }
MTSB_HD float fsub(float a, float b) { return __fsub_rn(a, b); }
MTSB_HD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
MTSB_HD float fsqrt(float a) { return __fsqrt_rn(a); }
MTSB_HD float frcp_fast(float a) { return __frcp_rn(a); }
MTSB_HD float fdiv_approx(float a, float b) { return __fdividef(a, b); }
MTSB_HD uint32_t f2u(float f) { return __float_as_uint(f); }
MTSB_HD float u2f(uint32_t u) { return __uint_as_float(u); }
MTSB_HD int bfind(uint32_t v) { return 31 - __clz(v); }
MTSB_HD int popc(uint32_t v) { return __popc(v); }
MTSB_HD int clz64(uint64_t v) { return __clzll((long long) v); }
MTSB_HD int clz32(uint32_t v) { return __clz(v); }
// 32768 + byte J of w as a float, by one byte permute (PRMT, ALU pipe) instead of an integer->float
// conversion (I2F, quarter-rate XU pipe): 0x4700bb00 has exponent 2^15 and the byte in mantissa bits 8..15.
MTSB_HD float byte_bias_float(uint32_t w, int j, uint32_t bias) { return __uint_as_float(__byte_perm(w, bias, 0x7604u | ((uint32_t) j << 4))); }
#else
// g++ builds of this header are compiled with -ffp-contract=off
MTSB_HD float fmul(float a, float b) { return a * b; }
MTSB_HD float fadd(float a, float b) { return a + b; }
MTSB_HD float fsub(float a, float b) { return a - b; }
MTSB_HD float fdiv(float a, float b) { return a / b; }
MTSB_HD float fsqrt(float a) { return sqrtf(a); }
MTSB_HD float frcp_fast(float a) { return 1.0f / a; }
MTSB_HD float fdiv_approx(float a, float b) { return (a / b) * 1.0000002f; }   // deliberately not the exact quotient
MTSB_HD uint32_t f2u(float f) { union { float f; uint32_t u; } x; x.f = f; return x.u; }
MTSB_HD float u2f(uint32_t u) { union { float f; uint32_t u; } x; x.u = u; return x.f; }
MTSB_HD int bfind(uint32_t v) { return v ? 31 - __builtin_clz(v) : -1; }
MTSB_HD int popc(uint32_t v) { return __builtin_popcount(v); }
MTSB_HD int clz64(uint64_t v) { return v ? __builtin_clzll(v) : 64; }
MTSB_HD int clz32(uint32_t v) { return v ? __builtin_clz(v) : 32; }
MTSB_HD float byte_bias_float(uint32_t w, int j, uint32_t bias) { return u2f(bias | (((w >> (8 * j)) & 0xffu) << 8)); }
#endif

// ---- small vector type (kept independent of CUDA's float3 so g++ sees the same code) ----
struct V3 {
    float x, y, z;
    MTSB_HD V3() : x(0), y(0), z(0) {}
    MTSB_HD V3(float a, float b, float c) : x(a), y(b), z(c) {}
    MTSB_HD explicit V3(float a) : x(a), y(a), z(a) {}
    MTSB_HD float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    MTSB_HD void set(int i, float v) { if (i == 0) x = v; else if (i == 1) y = v; else z = v; }
};
MTSB_HD V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
MTSB_HD V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
MTSB_HD V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
MTSB_HD V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
MTSB_HD V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
MTSB_HD V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
// TVector3::operator/(Float) multiplies by the reciprocal (include/mitsuba/core/vector.h)
MTSB_HD V3 operator/(V3 a, float s) { float r = fdiv(1.0f, s); return V3(a.x * r, a.y * r, a.z * r); }
MTSB_HD V3 operator/(V3 a, V3 b) { return V3(fdiv(a.x, b.x), fdiv(a.y, b.y), fdiv(a.z, b.z)); }
MTSB_HD V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
MTSB_HD V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
MTSB_HD V3 &operator*=(V3 &a, float s) { a = a * s; return a; }
MTSB_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
MTSB_HD float absDot(V3 a, V3 b) { return fabsf(dot(a, b)); }
MTSB_HD V3 cross(V3 a, V3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
MTSB_HD float lengthSquared(V3 a) { return dot(a, a); }
MTSB_HD float length(V3 a) { return fsqrt(dot(a, a)); }
MTSB_HD V3 normalize(V3 a) { return a / length(a); }
MTSB_HD bool isZero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }
MTSB_HD float comp(V3 a, int i) { return i == 0 ? a.x : i == 1 ? a.y : a.z; }
MTSB_HD float maxComp(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
MTSB_HD V3 vmin(V3 a, V3 b) { return V3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
MTSB_HD V3 vmax(V3 a, V3 b) { return V3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
MTSB_HD bool finite3(V3 a) { return isfinite(a.x) && isfinite(a.y) && isfinite(a.z); }
MTSB_HD V3 ld3(const float *p) { return V3(p[0], p[1], p[2]); }
MTSB_HD void st3(float *p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

struct P2 { float x, y; };

// include/mitsuba/core/constants.h:28-31,62-68 (single precision build of the reference)
#define MTSB_EPSILON 1e-4f
#define MTSB_RCPOVERFLOW 2.93873587705571876e-39f   /* constants.h:53 */
#define MTSB_SHADOW_EPSILON 1e-3f
#define MTSB_PI 3.14159265358979323846f
#define MTSB_INV_PI 0.31830988618379067154f
#define MTSB_INV_TWOPI 0.15915494309189533577f
#define MTSB_INV_FOURPI 0.07957747154594766788f
#if MTSB_DEVICE_CODE
#define MTSB_INF __int_as_float(0x7f800000)
#else
#define MTSB_INF INFINITY
#endif

}  // namespace mtsb
