// TEST TARGET ONLY -- CPU emulation of the kernel bodies of mitsuba_b200/csrc (bvh.h, traverse.h).
//
// The product library has no host execution path.  This file compiles the SAME per-thread
// kernel bodies with g++ (MTSB_EMU) and runs them in sequential loops, so that the BVH
// builder and the traversal logic can be checked against the oracle on the GPU-less CI box
// (`pytest -m "not gpu"`).  Nothing under mitsuba_b200/ loads it.
#define MTSB_EMU 1
#include "../../mitsuba_b200/csrc/bvh.h"
#include "../../mitsuba_b200/csrc/traverse.h"
#include "../../mitsuba_b200/csrc/sensor.h"
#include "../../mitsuba_b200/csrc/rng.h"
#include "../../mitsuba_b200/csrc/qmc.h"
#include "../../mitsuba_b200/csrc/qmc_host.h"
#include "../../mitsuba_b200/csrc/texture_host.h"
#include <vector>
#include <algorithm>
#include <numeric>
#include <cstring>
#include <cmath>

using namespace mtsb;

struct EmuScene {
    std::vector<Node8> nodes;
    std::vector<TriRec> tris;
    float tight_lo[3], tight_hi[3], box_lo[3], box_hi[3];
    uint32_t n_tris;
};

extern "C" void *emu_build(const float *positions, const uint32_t *indices, uint32_t n) {
    EmuScene *s = new EmuScene();
    s->n_tris = n;
    for (int k = 0; k < 3; ++k) { s->tight_lo[k] = s->tight_hi[k] = s->box_lo[k] = s->box_hi[k] = 0; }
    if (n == 0) return s;
    std::vector<TriRec> tri_tmp(n);
    std::vector<float> prim_lo(3 * (size_t) n), prim_hi(3 * (size_t) n);
    float tlo[3] = {INFINITY, INFINITY, INFINITY}, thi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = 0; i < n; ++i) {
        tri_setup(i, positions, indices, &tri_tmp[i], &prim_lo[3 * (size_t) i], &prim_hi[3 * (size_t) i]);
        for (int k = 0; k < 3; ++k) { tlo[k] = std::min(tlo[k], prim_lo[3 * (size_t) i + k]); thi[k] = std::max(thi[k], prim_hi[3 * (size_t) i + k]); }
    }
    for (int k = 0; k < 3; ++k) {
        volatile float lo = tlo[k] - ((thi[k] - tlo[k]) * 1e-3f + 1e-3f);
        volatile float hi = thi[k] + ((thi[k] - lo) * 1e-3f + 1e-3f);
        s->box_lo[k] = lo; s->box_hi[k] = hi; s->tight_lo[k] = tlo[k]; s->tight_hi[k] = thi[k];
    }
    V3 slo(tlo[0], tlo[1], tlo[2]);
    V3 sinv(thi[0] > tlo[0] ? 1.0f / (thi[0] - tlo[0]) : 0.0f, thi[1] > tlo[1] ? 1.0f / (thi[1] - tlo[1]) : 0.0f,
            thi[2] > tlo[2] ? 1.0f / (thi[2] - tlo[2]) : 0.0f);
    float diag = std::sqrt((thi[0] - tlo[0]) * (thi[0] - tlo[0]) + (thi[1] - tlo[1]) * (thi[1] - tlo[1]) + (thi[2] - tlo[2]) * (thi[2] - tlo[2]));
    float amax = 0; for (int k = 0; k < 3; ++k) amax = std::max(amax, std::max(std::fabs(tlo[k]), std::fabs(thi[k])));
    const float leaf_pad = 4e-6f * std::max(diag, amax) + 1e-30f;

    std::vector<uint64_t> keys(n); std::vector<uint32_t> vals(n);
    for (uint32_t i = 0; i < n; ++i) keys[i] = morton_key(&prim_lo[3 * (size_t) i], &prim_hi[3 * (size_t) i], slo, sinv);
    std::iota(vals.begin(), vals.end(), 0u);
    std::stable_sort(vals.begin(), vals.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> skeys(n);
    for (uint32_t i = 0; i < n; ++i) skeys[i] = keys[vals[i]];

    std::vector<Node8> nodes(std::max<uint32_t>(1, n));
    std::vector<TriRec> tri_out(n);
    std::vector<uint32_t> wbin(std::max<uint32_t>(1, n), 0), left(n), right(n), first(n), last(n), parent(2 * (size_t) n, 0), flags(n, 0);
    std::vector<float> node_lo(6 * (size_t) n), node_hi(6 * (size_t) n);
    uint32_t counters[2] = {1, 0};
    WideBuildArgs a;
    a.n = (int) n; a.left = left.data(); a.right = right.data(); a.first = first.data(); a.last = last.data();
    a.node_lo = node_lo.data(); a.node_hi = node_hi.data(); a.sorted_prim = vals.data(); a.tri_in = tri_tmp.data();
    a.nodes = nodes.data(); a.wbin = wbin.data(); a.tri_out = tri_out.data(); a.counters = counters; a.leaf_pad = leaf_pad;
    uint32_t n_nodes = 1;
    if (n <= MTSB_LEAF_MAX) {
        for (int k = 0; k < 3; ++k) { node_lo[k] = tlo[k]; node_hi[k] = thi[k]; }
        wide_node_build(0, a, true);
    } else {
        for (int i = 0; i < (int) n - 1; ++i) radix_tree_node(i, (int) n, skeys.data(), left.data(), right.data(), parent.data(), first.data(), last.data());
        for (int j = 0; j < (int) n; ++j)
            fit_from_leaf(j, (int) n, vals.data(), prim_lo.data(), prim_hi.data(), left.data(), right.data(), parent.data(), node_lo.data(), node_hi.data(), flags.data());
        uint32_t begin = 0, end = 1;
        while (begin < end) {
            for (uint32_t w = begin; w < end; ++w) wide_node_build(w, a, false);
            begin = end; end = counters[0];
        }
        n_nodes = end;
    }
    nodes.resize(n_nodes);
    s->nodes.swap(nodes); s->tris.swap(tri_out);
    return s;
}

extern "C" void emu_nodes(void *h, Node8 *out) {
    EmuScene *s = static_cast<EmuScene *>(h);
    std::memcpy(out, s->nodes.data(), s->nodes.size() * sizeof(Node8));
}
extern "C" void emu_destroy(void *h) { delete static_cast<EmuScene *>(h); }

extern "C" void emu_info(void *h, uint32_t *n_nodes, float *tight6, float *enlarged6) {
    EmuScene *s = static_cast<EmuScene *>(h);
    *n_nodes = (uint32_t) s->nodes.size();
    for (int k = 0; k < 3; ++k) { tight6[k] = s->tight_lo[k]; tight6[3 + k] = s->tight_hi[k]; enlarged6[k] = s->box_lo[k]; enlarged6[3 + k] = s->box_hi[k]; }
}

static uint64_t g_tri_tests = 0, g_sparse_steps = 0;
extern "C" uint64_t emu_sparse_steps(void) { uint64_t v = g_sparse_steps; g_sparse_steps = 0; return v; }
extern "C" uint64_t emu_tri_tests(void) { uint64_t v = g_tri_tests; g_tri_tests = 0; return v; }
// returns the total number of node visits (traversal statistics)
extern "C" uint64_t emu_trace(void *h, const RayIn *rays, size_t n, int mode, HitOut *hits) {
    EmuScene *s = static_cast<EmuScene *>(h);
    uint64_t steps = 0;
    const U4 *nodes = reinterpret_cast<const U4 *>(s->nodes.data()), *tris = reinterpret_cast<const U4 *>(s->tris.data());
    for (size_t i = 0; i < n; ++i) {
        HitOut out; out.prim = MTSB_NO_HIT_ID; out.u = out.v = 0; out.t = INFINITY;
        float mint, maxt;
        if (!s->nodes.empty() && ray_interval(rays[i], s->box_lo, s->box_hi, mode == 1, mint, maxt)) {
            Traversal<LocalStack<64>> tr;
            tr.begin(rays[i], mint, maxt);
            if (mode == 1) { while (!tr.step<true>(nodes, tris)) {} }
            else { while (!tr.step<false>(nodes, tris)) {} }
            out = tr.result(tris); steps += tr.steps; g_tri_tests += tr.tri_tests; g_sparse_steps += tr.sparse_steps;
        }
        hits[i] = out;
    }
    return steps;
}

// copies of the triangle records in leaf order (for structural checks)
extern "C" void emu_tris(void *h, TriRec *out) {
    EmuScene *s = static_cast<EmuScene *>(h);
    std::memcpy(out, s->tris.data(), s->tris.size() * sizeof(TriRec));
}

extern "C" void emu_sensor_rays(const MtsbSensor *se, float near_x, float near_y, const float *pos, size_t n, RayIn *rays) {
    for (size_t i = 0; i < n; ++i) sensor_ray(*se, near_x, near_y, pos[2 * i], pos[2 * i + 1], rays[i]);
}

extern "C" void emu_philox(const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4) {
    U32x4 c; c.x = ctr4[0]; c.y = ctr4[1]; c.z = ctr4[2]; c.w = ctr4[3];
    U32x4 r = philox4x32_10(c, key2[0], key2[1]);
    out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}
extern "C" void emu_sample_block(uint64_t seed, uint64_t sample_id, uint32_t block, uint32_t sub, float *out4) {
    SampleStream(seed, sample_id).block4(block, sub, out4);
}

// ---- quasi-Monte Carlo draws of the kernels (qmc.h), driven in the order k_raygen / k_shade use them ----
// Halton: for each sample n2D two-dimensional draws (the first one addresses the pixel) followed by n1D one-dimensional ones.
extern "C" int emu_halton_sequence(int resX, int resY, int scramble, int px, int py, int firstSample, int nSamples, int n2D, int n1D, float *out) {
    MtsbSampler sd; sd.type = MTSB_SAMPLER_HALTON; sd.dimension = 0; sd.scramble = scramble; sd._pad = 0;
    SamplerView sv; const char *err = "";
    if (!sampler_view_init(sd, 1, resX, resY, sv, &err)) return 1;
    const std::vector<uint16_t> primes = qmc_prime_table();
    std::vector<uint16_t> perms; std::vector<uint32_t> offsets;
    sv.primes = primes.data();
    if (scramble != 0) {
        qmc_permutations(scramble, primes, perms, offsets);
        sv.perms = perms.data(); sv.perm_offset = offsets.data();
        for (int j = 0; j < 2; ++j) sv.inv_perm2[perms[offsets[0] + j]] = (uint16_t) j;
        for (int j = 0; j < 3; ++j) sv.inv_perm3[perms[offsets[1] + j]] = (uint16_t) j;
    }
    for (int j = 0; j < nSamples; ++j) {
        int pm[2];
        const uint64_t index = halton_pixel_offset(sv, px, py, pm) + sv.stride * (uint64_t) (firstSample + j);
        uint32_t dim = 0;
        for (int k = 0; k < n2D; ++k) {
            float a = halton_value(sv, dim++, index), b = halton_value(sv, dim++, index);
            if (dim == 2) { a = fsub(fmul(a, (float) sv.prime_pow[0]), (float) pm[0]); b = fsub(fmul(b, (float) sv.prime_pow[1]), (float) pm[1]); }
            *out++ = a; *out++ = b;
        }
        for (int k = 0; k < n1D; ++k) *out++ = halton_value(sv, dim++, index);
    }
    return 0;
}
extern "C" int emu_sobol_sequence(int resX, int resY, int scramble, int px, int py, int firstSample, int nSamples, int n2D, int n1D, float *out) {
    MtsbSampler sd; sd.type = MTSB_SAMPLER_SOBOL; sd.dimension = 0; sd.scramble = scramble; sd._pad = 0;
    SamplerView sv; const char *err = "";
    if (!sampler_view_init(sd, 1, resX, resY, sv, &err)) return 1;
    static const std::vector<uint32_t> mat = sobol_matrices32();
    std::vector<uint64_t> vdc, vdc_inv;
    if (!sobol_vdc_tables(sv.sobol_log_res, mat, vdc, vdc_inv)) return 2;
    sv.sobol_mat = mat.data(); sv.sobol_vdc = vdc.data(); sv.sobol_vdc_inv = vdc_inv.data();
    for (int j = 0; j < nSamples; ++j) {
        const uint32_t s_global = (uint32_t) (firstSample + j);
        const uint64_t index = sobol_sample_index(sv, s_global, px, py);
        uint32_t dim = 0;
        for (int k = 0; k < n2D; ++k) {
            float a = sobol_value(sv, dim++, index), b = sobol_value(sv, dim++, index);
            if (dim == 2 && index != (uint64_t) s_global) { a = fsub(fmul(a, sv.sobol_res), (float) px); b = fsub(fmul(b, sv.sobol_res), (float) py); }
            *out++ = a; *out++ = b;
        }
        for (int k = 0; k < n1D; ++k) *out++ = sobol_value(sv, dim++, index);
    }
    return 0;
}
extern "C" uint32_t emu_sobol_matrix32(uint32_t dimension, uint32_t column) { static const std::vector<uint32_t> mat = sobol_matrices32(); return mat[(size_t) dimension * MTSB_SOBOL_SIZE + column]; }
extern "C" int emu_sobol_vdc(uint32_t m, uint64_t *vdc52, uint64_t *inv52) {
    static const std::vector<uint32_t> mat = sobol_matrices32();
    std::vector<uint64_t> a, b;
    if (!sobol_vdc_tables(m, mat, a, b)) return 1;
    std::copy(a.begin(), a.end(), vdc52); std::copy(b.begin(), b.end(), inv52);
    return 0;
}
extern "C" float emu_radical_inverse(uint32_t base, uint64_t index, const uint16_t *perm) { return radical_inverse(base, index, perm); }
extern "C" int emu_permutation(int scramble, int baseIndex, uint16_t *out) {
    static int cached = 0; static std::vector<uint16_t> perms; static std::vector<uint32_t> offsets;
    const std::vector<uint16_t> primes = qmc_prime_table();
    if (perms.empty() || cached != scramble) { qmc_permutations(scramble, primes, perms, offsets); cached = scramble; }
    std::copy(perms.begin() + offsets[baseIndex], perms.begin() + offsets[baseIndex] + primes[baseIndex], out);
    return primes[baseIndex];
}
extern "C" int emu_prime(int i) { return qmc_prime_table()[i]; }
extern "C" void emu_sample02(uint32_t n, const uint32_t *scramble2, float *out2) { out2[0] = radical_inverse2_single(n, scramble2[0]); out2[1] = sobol2_single(n, scramble2[1]); }
extern "C" uint32_t emu_permute_pow2(uint32_t i, uint32_t bits, uint32_t ka, uint32_t kb) { return permute_pow2(i, bits, ka, kb); }
// ldsampler point sets of one pixel as the kernels see them: out1[dim][count], out2[dim][count][2] (indexed by sample index)
extern "C" int emu_ldsampler_sets(int sampleCount, int dimension, uint64_t seed, int px, int py, float *out1, float *out2) {
    MtsbSampler sd; sd.type = MTSB_SAMPLER_LDSAMPLER; sd.dimension = dimension; sd.scramble = 0; sd._pad = 0;
    SamplerView sv; const char *err = "";
    if (!sampler_view_init(sd, sampleCount, 1, 1, sv, &err)) return 1;
    const uint64_t pixel = ((uint64_t) (uint32_t) py << 32) | (uint64_t) (uint32_t) px;
    for (int d = 0; d < dimension; ++d)
        for (uint32_t i = 0; i < sv.sample_count; ++i) {
            out1[(size_t) d * sv.sample_count + i] = ld_value_1d(sv, seed, pixel, (uint32_t) d, i);
            ld_value_2d(sv, seed, pixel, (uint32_t) d, i, out2 + ((size_t) d * sv.sample_count + i) * 2);
        }
    return 0;
}

// ---- experiment support: the same wide collapse fed by a top-down binned-SAH binary tree ----
// (measures how much of the node-visit count is due to the Morton-order builder)
namespace {
struct SahBuilder {
    const float *lo, *hi; std::vector<uint32_t> &order; std::vector<uint32_t> &left, &right, &first, &last; uint32_t n; uint32_t next_internal = 0;
    SahBuilder(const float *l, const float *h, std::vector<uint32_t> &o, std::vector<uint32_t> &L, std::vector<uint32_t> &R, std::vector<uint32_t> &F, std::vector<uint32_t> &La, uint32_t nn)
        : lo(l), hi(h), order(o), left(L), right(R), first(F), last(La), n(nn) {}
    float cen(uint32_t p, int ax) const { return 0.5f * (lo[3 * (size_t) p + ax] + hi[3 * (size_t) p + ax]); }
    static float area(const float *a, const float *b) { float x = b[0] - a[0], y = b[1] - a[1], z = b[2] - a[2]; return x * y + y * z + z * x; }
    // returns unified node id of the subtree over order[b, e)
    uint32_t build(uint32_t b, uint32_t e) {
        if (e - b == 1) return n - 1 + b;
        uint32_t id = next_internal++;
        float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = b; i < e; ++i) for (int k = 0; k < 3; ++k) { float c = cen(order[i], k); clo[k] = std::min(clo[k], c); chi[k] = std::max(chi[k], c); }
        const int NB = 32; int bestAx = -1, bestSplit = -1; float bestCost = INFINITY;
        for (int ax = 0; ax < 3; ++ax) {
            if (!(chi[ax] > clo[ax])) continue;
            float blo[NB][3], bhi[NB][3]; uint32_t cnt[NB];
            for (int q = 0; q < NB; ++q) { cnt[q] = 0; for (int k = 0; k < 3; ++k) { blo[q][k] = INFINITY; bhi[q][k] = -INFINITY; } }
            float sc = NB / (chi[ax] - clo[ax]);
            for (uint32_t i = b; i < e; ++i) {
                uint32_t p = order[i]; int q = std::min(NB - 1, (int) ((cen(p, ax) - clo[ax]) * sc));
                cnt[q]++; for (int k = 0; k < 3; ++k) { blo[q][k] = std::min(blo[q][k], lo[3 * (size_t) p + k]); bhi[q][k] = std::max(bhi[q][k], hi[3 * (size_t) p + k]); }
            }
            float ra[NB]; uint32_t rc[NB]; float alo[3] = {INFINITY, INFINITY, INFINITY}, ahi[3] = {-INFINITY, -INFINITY, -INFINITY}; uint32_t c = 0;
            for (int q = NB - 1; q > 0; --q) { c += cnt[q]; for (int k = 0; k < 3; ++k) { alo[k] = std::min(alo[k], blo[q][k]); ahi[k] = std::max(ahi[k], bhi[q][k]); } ra[q] = c ? area(alo, ahi) : 0; rc[q] = c; }
            for (int k = 0; k < 3; ++k) { alo[k] = INFINITY; ahi[k] = -INFINITY; } c = 0;
            for (int q = 0; q < NB - 1; ++q) {
                c += cnt[q]; for (int k = 0; k < 3; ++k) { alo[k] = std::min(alo[k], blo[q][k]); ahi[k] = std::max(ahi[k], bhi[q][k]); }
                if (!c || !rc[q + 1]) continue;
                float cost = area(alo, ahi) * c + ra[q + 1] * rc[q + 1];
                if (cost < bestCost) { bestCost = cost; bestAx = ax; bestSplit = q; }
            }
        }
        uint32_t mid;
        if (bestAx < 0) mid = b + (e - b) / 2;
        else {
            float sc = NB / (chi[bestAx] - clo[bestAx]); float c0 = clo[bestAx]; int ax = bestAx, sp = bestSplit;
            auto it = std::partition(order.begin() + b, order.begin() + e, [&](uint32_t p) { return std::min(NB - 1, (int) ((cen(p, ax) - c0) * sc)) <= sp; });
            mid = (uint32_t) (it - order.begin());
            if (mid == b || mid == e) mid = b + (e - b) / 2;
        }
        first[id] = b; last[id] = e - 1;
        uint32_t l = build(b, mid), r = build(mid, e);
        left[id] = l; right[id] = r;
        return id;
    }
};
}

extern "C" void *emu_build_sah(const float *positions, const uint32_t *indices, uint32_t n) {
    EmuScene *s = new EmuScene();
    s->n_tris = n;
    std::vector<TriRec> tri_tmp(n);
    std::vector<float> prim_lo(3 * (size_t) n), prim_hi(3 * (size_t) n);
    float tlo[3] = {INFINITY, INFINITY, INFINITY}, thi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = 0; i < n; ++i) {
        tri_setup(i, positions, indices, &tri_tmp[i], &prim_lo[3 * (size_t) i], &prim_hi[3 * (size_t) i]);
        for (int k = 0; k < 3; ++k) { tlo[k] = std::min(tlo[k], prim_lo[3 * (size_t) i + k]); thi[k] = std::max(thi[k], prim_hi[3 * (size_t) i + k]); }
    }
    for (int k = 0; k < 3; ++k) {
        volatile float lo = tlo[k] - ((thi[k] - tlo[k]) * 1e-3f + 1e-3f);
        volatile float hi = thi[k] + ((thi[k] - lo) * 1e-3f + 1e-3f);
        s->box_lo[k] = lo; s->box_hi[k] = hi; s->tight_lo[k] = tlo[k]; s->tight_hi[k] = thi[k];
    }
    float diag = std::sqrt((thi[0] - tlo[0]) * (thi[0] - tlo[0]) + (thi[1] - tlo[1]) * (thi[1] - tlo[1]) + (thi[2] - tlo[2]) * (thi[2] - tlo[2]));
    float amax = 0; for (int k = 0; k < 3; ++k) amax = std::max(amax, std::max(std::fabs(tlo[k]), std::fabs(thi[k])));
    std::vector<uint32_t> vals(n); std::iota(vals.begin(), vals.end(), 0u);
    std::vector<uint32_t> left(n), right(n), first(n), last(n), parent(2 * (size_t) n, 0), flags(n, 0), wbin(n, 0);
    SahBuilder sb(prim_lo.data(), prim_hi.data(), vals, left, right, first, last, n);
    uint32_t root = sb.build(0, n);
    (void) root;   // == 0 (pre-order numbering)
    for (uint32_t i = 0; i + 1 < n; ++i) { parent[left[i]] = i; parent[right[i]] = i; }
    parent[0] = 0xFFFFFFFFu;
    std::vector<float> node_lo(6 * (size_t) n), node_hi(6 * (size_t) n);
    for (int j = 0; j < (int) n; ++j)
        fit_from_leaf(j, (int) n, vals.data(), prim_lo.data(), prim_hi.data(), left.data(), right.data(), parent.data(), node_lo.data(), node_hi.data(), flags.data());
    std::vector<Node8> nodes(n); std::vector<TriRec> tri_out(n);
    uint32_t counters[2] = {1, 0};
    WideBuildArgs a;
    a.n = (int) n; a.left = left.data(); a.right = right.data(); a.first = first.data(); a.last = last.data();
    a.node_lo = node_lo.data(); a.node_hi = node_hi.data(); a.sorted_prim = vals.data(); a.tri_in = tri_tmp.data();
    a.nodes = nodes.data(); a.wbin = wbin.data(); a.tri_out = tri_out.data(); a.counters = counters;
    a.leaf_pad = 4e-6f * std::max(diag, amax) + 1e-30f;
    uint32_t begin = 0, end = 1;
    while (begin < end) { for (uint32_t w = begin; w < end; ++w) wide_node_build(w, a, false); begin = end; end = counters[0]; }
    nodes.resize(end);
    s->nodes.swap(nodes); s->tris.swap(tri_out);
    return s;
}

// Per-ray traversal profile for the warp-scheduling model in tools/warp_sim.py: seq[i*max_len + k] is the
// number of triangle tests that followed node visit k of ray i (lens[i] = number of node visits, capped).
extern "C" void emu_trace_profile(void *h, const RayIn *rays, size_t n, int mode, uint8_t *seq, uint32_t max_len, uint32_t *lens) {
    EmuScene *s = static_cast<EmuScene *>(h);
    const U4 *nodes = reinterpret_cast<const U4 *>(s->nodes.data()), *tris = reinterpret_cast<const U4 *>(s->tris.data());
    for (size_t i = 0; i < n; ++i) {
        uint32_t len = 0;
        float mint, maxt;
        if (!s->nodes.empty() && ray_interval(rays[i], s->box_lo, s->box_hi, mode == 1, mint, maxt)) {
            Traversal<LocalStack<64>> tr;
            tr.begin(rays[i], mint, maxt);
            for (;;) {
                if (tr.wants_node()) {
                    tr.visit_node(nodes);
                    if (len < max_len) seq[i * (size_t) max_len + len] = (uint8_t) __builtin_popcount(tr.last_batch);
                    ++len;
                }
                bool done = false;
                if (tr.has_triangles()) done = mode == 1 ? tr.test_triangle<true>(tris) : tr.test_triangle<false>(tris);
                if (done || tr.finished()) break;
            }
        }
        lens[i] = len < max_len ? len : max_len;
    }
}

// ---- textures: csrc/texture.h evaluated on the host, pyramid from csrc/texture_host.h ---------------------------------
struct EmuTexture { TextureView tv; std::vector<HalfTexel> texels; float lut[MTSB_MIPMAP_LUT_SIZE]; };
extern "C" void *emu_texture_create(const MtsbTexture *t) {
    EmuTexture *e = new EmuTexture();
    TextureView &tv = e->tv;
    std::memset(&tv, 0, sizeof(tv));
    tv.type = t->type; tv.filter = t->filter; tv.wrap_u = t->wrap_u; tv.wrap_v = t->wrap_v;
    for (int k = 0; k < 2; ++k) { tv.uv_scale[k] = t->uv_scale[k]; tv.uv_offset[k] = t->uv_offset[k]; }
    for (int k = 0; k < 3; ++k) { tv.color0[k] = t->color0[k]; tv.color1[k] = t->color1[k]; tv.scale[k] = t->scale[k]; }
    tv.line_width = t->line_width;
    tv.max_anisotropy = t->filter == MTSB_TEXFILTER_EWA ? t->max_anisotropy : 1.0f;
    for (int i = 0; i < MTSB_MIPMAP_LUT_SIZE; ++i) { const float r2 = (float) i / (float) (MTSB_MIPMAP_LUT_SIZE - 1); e->lut[i] = (float) std::exp((double) (-2.0f * r2)) - (float) std::exp((double) -2.0f); }   // math::fastexp, math.h:175-199
    tv.weight_lut = e->lut;
    if (t->type == MTSB_TEX_BITMAP) build_texture_levels(*t, tv, e->texels);
    tv.texels = e->texels.data();
    return e;
}
extern "C" void emu_texture_destroy(void *h) { delete static_cast<EmuTexture *>(h); }
extern "C" int emu_texture_levels(void *h) { return static_cast<EmuTexture *>(h)->tv.levels; }
extern "C" void emu_texture_level_size(void *h, int level, int *wh) { const TextureView &tv = static_cast<EmuTexture *>(h)->tv; wh[0] = tv.lw[level]; wh[1] = tv.lh[level]; }
extern "C" void emu_texture_level(void *h, int level, float *out) {
    const TextureView &tv = static_cast<EmuTexture *>(h)->tv;
    for (int y = 0; y < tv.lh[level]; ++y)
        for (int x = 0; x < tv.lw[level]; ++x) {
            const V3 v = tex_texel(tv, level, x, y);
            float *o = out + 3 * ((size_t) y * tv.lw[level] + x); o[0] = v.x; o[1] = v.y; o[2] = v.z;
        }
}
extern "C" void emu_texture_eval(void *h, size_t n, const float *uv, const float *partials, float *rgb) {
    const TextureView &tv = static_cast<EmuTexture *>(h)->tv;
    for (size_t i = 0; i < n; ++i) {
        UVPartials pt; pt.has = partials != nullptr; pt.dudx = pt.dvdx = pt.dudy = pt.dvdy = 0.0f;
        if (partials) { pt.dudx = partials[4 * i]; pt.dvdx = partials[4 * i + 1]; pt.dudy = partials[4 * i + 2]; pt.dvdy = partials[4 * i + 3]; }
        P2 q; q.x = uv[2 * i]; q.y = uv[2 * i + 1];
        const V3 v = texture_eval(tv, q, pt);
        rgb[3 * i] = v.x; rgb[3 * i + 1] = v.y; rgb[3 * i + 2] = v.z;
    }
}
extern "C" void emu_compute_partials(size_t n, const float *in, float *out) {
    for (size_t i = 0; i < n; ++i) {
        const float *q = in + 21 * i;
        auto v = [&](int k) { return V3(q[3 * k], q[3 * k + 1], q[3 * k + 2]); };
        const UVPartials r = compute_partials(v(0), v(1), v(2), v(3), v(4), v(5), v(6));
        out[4 * i] = r.dudx; out[4 * i + 1] = r.dvdx; out[4 * i + 2] = r.dudy; out[4 * i + 3] = r.dvdy;
    }
}
extern "C" void emu_sensor_differentials(const MtsbSensor *se, float near_x, float near_y, int sample_count, size_t n, const float *pos, float *out6) {
    for (size_t i = 0; i < n; ++i) {
        V3 rx, ry; sensor_ray_differentials(*se, near_x, near_y, pos[2 * i], pos[2 * i + 1], sample_count, rx, ry);
        float *o = out6 + 6 * i; o[0] = rx.x; o[1] = rx.y; o[2] = rx.z; o[3] = ry.x; o[4] = ry.y; o[5] = ry.z;
    }
}
extern "C" uint32_t emu_float_to_half(float f) { return float_to_half_bits(f); }
