// bvh.h -- acceleration structure of the B200 path: layout + per-thread build steps.
//
// Replaces ShapeKDTree::build / GenericKDTree::buildInternal (reference:
// src/librender/skdtree.cpp:68-110, include/mitsuba/render/gkdtree.h:958-1244).  The
// structure is different by design: an 8-wide BVH with 80-byte
// quantised nodes built on the GPU (Morton order -> binary radix tree -> bottom-up fit ->
// greedy surface-area collapse to 8 children -> octant slot assignment), while the
// per-triangle record keeps the reference's projection-test coefficients
// (include/mitsuba/render/triaccel.h:37-100) so that t,u,v come out bit-identical.
//
// HBM layout
//   nodes : 5 x 16 B per node (Node8)              -- fetched as five 128-bit loads
//   tris  : 3 x 16 B per triangle (TriRec), stored in LEAF ORDER: the <=24 triangles
//           referenced by one node are contiguous, so a leaf fetch is 1-3 adjacent
//           48-B records.
#pragma once
#include "hd.h"

namespace mtsb {

#if defined(__CUDACC__)
typedef uint4 U4;
#else
struct alignas(16) U4 { uint32_t x, y, z, w; };
#endif

// 128-bit read-only load
#if MTSB_DEVICE_CODE
MTSB_HD U4 ld16(const U4 *p) { return __ldg(p); }
#else
MTSB_HD U4 ld16(const U4 *p) { return *p; }
#endif

// 48-byte triangle record (Wald projection coefficients, triaccel.h:37-62)
struct alignas(16) TriRec {
    uint32_t k; float n_u, n_v, n_d;
    float a_u, a_v, b_nu, b_nv;
    float c_nu, c_nv; uint32_t prim; uint32_t pad;
};

// 80-byte wide node.  Child boxes are quantised to 8 bits relative to (p, 2^e).
//   imask bit s            : slot s holds an inner child; its node = child_base + rank of s in imask
//   counts bits 2s, 2s+1   : slot s holds a leaf with 1..3 triangles (0: empty or inner); the leaf's triangles start at
//                            tri_base + (sum of the counts of the slots below s), i.e. they are stored in slot order
// Slot s is the octant (sign pattern) of the child relative to the node centre, so a ray visits the children of a
// node in the order of decreasing s ^ octinv (traverse.h) without looking at the boxes.
struct alignas(16) Node8 {
    float px, py, pz;
    uint8_t ex, ey, ez, imask;
    uint32_t child_base, tri_base;
    uint16_t counts, pad0;
    uint32_t pad1;
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");
static_assert(sizeof(TriRec) == 48, "TriRec must be 48 bytes");

#define MTSB_LEAF_MAX 3            /* triangles per leaf slot */
#define MTSB_BIN_LEAF 0x80000000u  /* binary-tree child reference: leaf flag */

// ordered-uint encoding of floats for atomic min/max
MTSB_HD uint32_t enc_f(float f) { uint32_t u = f2u(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
MTSB_HD float dec_f(uint32_t u) { return u2f((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

// ---- step A: triangle records + primitive boxes ---------------------------------
// TriAccel::load, include/mitsuba/render/triaccel.h:64-100 (A,B,C = the triangle's vertices;
// b = C-A, c = B-A, N = c x b; k = argmax |N_k| with strict '>')
MTSB_HD void tri_setup(uint32_t prim, const float *positions, const uint32_t *indices,
                       TriRec *rec_out, float *lo_out /*3*/, float *hi_out /*3*/) {
    const uint32_t i0 = indices[3 * (size_t) prim], i1 = indices[3 * (size_t) prim + 1], i2 = indices[3 * (size_t) prim + 2];
    const V3 A = ld3(positions + 3 * (size_t) i0), B = ld3(positions + 3 * (size_t) i1), C = ld3(positions + 3 * (size_t) i2);
    const V3 b(fsub(C.x, A.x), fsub(C.y, A.y), fsub(C.z, A.z));
    const V3 c(fsub(B.x, A.x), fsub(B.y, A.y), fsub(B.z, A.z));
    const V3 N(fsub(fmul(c.y, b.z), fmul(c.z, b.y)), fsub(fmul(c.z, b.x), fmul(c.x, b.z)), fsub(fmul(c.x, b.y), fmul(c.y, b.x)));
    uint32_t k = 0;
    for (int j = 0; j < 3; ++j)
        if (fabsf(N[j]) > fabsf(N[k])) k = j;
    const uint32_t u = (k + 1) % 3, v = (k + 2) % 3;
    const float n_k = N[k];
    const float denom = fsub(fmul(b[u], c[v]), fmul(b[v], c[u]));
    TriRec r;
    r.prim = prim; r.pad = 0;
    if (denom == 0.0f) {
        r.k = 3; r.n_u = r.n_v = r.n_d = r.a_u = r.a_v = r.b_nu = r.b_nv = r.c_nu = r.c_nv = 0.0f;
    } else {
        r.k = k;
        r.n_u = fdiv(N[u], n_k);
        r.n_v = fdiv(N[v], n_k);
        // dot(A, N) as the reference's TVector3 dot: x*x' + y*y' + z*z' evaluated left to right
        r.n_d = fdiv(fadd(fadd(fmul(A.x, N.x), fmul(A.y, N.y)), fmul(A.z, N.z)), n_k);
        r.b_nu = fdiv(b[u], denom);
        r.b_nv = fdiv(-b[v], denom);
        r.a_u = A[u];
        r.a_v = A[v];
        r.c_nu = fdiv(c[v], denom);
        r.c_nv = fdiv(-c[u], denom);
    }
    *rec_out = r;
    V3 lo = vmin(A, vmin(B, C)), hi = vmax(A, vmax(B, C));
    st3(lo_out, lo); st3(hi_out, hi);
}

// ---- step B: 63-bit Morton key of the box centre --------------------------------
MTSB_HD uint64_t expand21(uint64_t v) {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}
MTSB_HD uint64_t morton_key(const float *lo, const float *hi, V3 slo, V3 sinv) {
    float cx = ((lo[0] + hi[0]) * 0.5f - slo.x) * sinv.x;
    float cy = ((lo[1] + hi[1]) * 0.5f - slo.y) * sinv.y;
    float cz = ((lo[2] + hi[2]) * 0.5f - slo.z) * sinv.z;
    const float S = 2097151.0f;   // 2^21 - 1
    uint64_t ix = (uint64_t) fminf(fmaxf(cx * S, 0.0f), S);
    uint64_t iy = (uint64_t) fminf(fmaxf(cy * S, 0.0f), S);
    uint64_t iz = (uint64_t) fminf(fmaxf(cz * S, 0.0f), S);
    return (expand21(ix) << 2) | (expand21(iy) << 1) | expand21(iz);
}

// ---- step C: binary radix tree over the sorted keys (Karras 2012) ---------------
// Unified node ids: internal i in [0, n-1), leaf j -> (n-1)+j.
MTSB_HD int radix_delta(const uint64_t *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t kx = keys[i] ^ keys[j];
    if (kx == 0) return 64 + clz32((uint32_t) i ^ (uint32_t) j);
    return clz64(kx);
}
MTSB_HD void radix_tree_node(int i, int n, const uint64_t *keys, uint32_t *left, uint32_t *right,
                             uint32_t *parent, uint32_t *first, uint32_t *last) {
    int d = (radix_delta(keys, n, i, i + 1) - radix_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = radix_delta(keys, n, i, i - d);
    int lmax = 2;
    while (radix_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (radix_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = radix_delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) / 2;
        if (radix_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? d : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    uint32_t l_id = (lo == gamma) ? (uint32_t) (n - 1 + gamma) : (uint32_t) gamma;
    uint32_t r_id = (hi == gamma + 1) ? (uint32_t) (n - 1 + gamma + 1) : (uint32_t) (gamma + 1);
    left[i] = l_id; right[i] = r_id;
    parent[l_id] = (uint32_t) i; parent[r_id] = (uint32_t) i;
    first[i] = (uint32_t) lo; last[i] = (uint32_t) hi;
    if (i == 0) parent[0] = 0xFFFFFFFFu;
}

// ---- step D: bottom-up box fit (one thread per leaf, second arrival continues) ----
#if MTSB_DEVICE_CODE
#define MTSB_FENCE() __threadfence()
#define MTSB_LD_CG(p) __ldcg(p)
#define MTSB_ATOMIC_INC_FLAG(p) atomicAdd((p), 1u)
#else
#define MTSB_FENCE()
#define MTSB_LD_CG(p) (*(p))
MTSB_HD uint32_t mtsb_emu_inc(uint32_t *p) { uint32_t o = *p; *p = o + 1; return o; }
#define MTSB_ATOMIC_INC_FLAG(p) mtsb_emu_inc(p)
#endif
MTSB_HD void fit_from_leaf(int j, int n, const uint32_t *sorted_prim, const float *prim_lo, const float *prim_hi,
                           const uint32_t *left, const uint32_t *right, const uint32_t *parent,
                           float *node_lo, float *node_hi, uint32_t *flags) {
    const uint32_t leaf = (uint32_t) (n - 1 + j);
    const uint32_t p = sorted_prim[j];
    for (int k = 0; k < 3; ++k) {
        node_lo[3 * (size_t) leaf + k] = prim_lo[3 * (size_t) p + k];
        node_hi[3 * (size_t) leaf + k] = prim_hi[3 * (size_t) p + k];
    }
    if (n < 2) return;
    uint32_t cur = parent[leaf];
    while (cur != 0xFFFFFFFFu) {
        MTSB_FENCE();
        if (MTSB_ATOMIC_INC_FLAG(flags + cur) == 0) return;     // first child to arrive stops
        MTSB_FENCE();
        const uint32_t l = left[cur], r = right[cur];
        for (int k = 0; k < 3; ++k) {
            node_lo[3 * (size_t) cur + k] = fminf(MTSB_LD_CG(node_lo + 3 * (size_t) l + k), MTSB_LD_CG(node_lo + 3 * (size_t) r + k));
            node_hi[3 * (size_t) cur + k] = fmaxf(MTSB_LD_CG(node_hi + 3 * (size_t) l + k), MTSB_LD_CG(node_hi + 3 * (size_t) r + k));
        }
        cur = parent[cur];
    }
}

// ---- step E helpers: collapse one binary subtree root into a wide node ------------
struct WideBuildArgs {
    int n;                       // primitives
    const uint32_t *left, *right, *first, *last;     // binary internal nodes
    const float *node_lo, *node_hi;                  // 3 floats per unified node id
    const uint32_t *sorted_prim;                     // leaf j -> primitive id
    const TriRec *tri_in;                            // records in primitive order
    Node8 *nodes;                                    // out
    uint32_t *wbin;                                  // wide node -> binary node id it was made from
    TriRec *tri_out;                                 // out, leaf order
    uint32_t *counters;                              // [0] = wide nodes allocated, [1] = triangles emitted
    float leaf_pad;                                  // conservative growth of leaf boxes
};

MTSB_HD float box_area(const float *lo, const float *hi) {
    float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
    return ex * ey + ey * ez + ez * ex;
}
MTSB_HD uint32_t bin_size(const WideBuildArgs &a, uint32_t id) {
    return id >= (uint32_t) (a.n - 1) ? 1u : a.last[id] - a.first[id] + 1u;
}
MTSB_HD uint32_t bin_first(const WideBuildArgs &a, uint32_t id) {
    return id >= (uint32_t) (a.n - 1) ? id - (uint32_t) (a.n - 1) : a.first[id];
}

#if MTSB_DEVICE_CODE
#define MTSB_ATOMIC_ADD_U32(p, v) atomicAdd((p), (v))
#else
MTSB_HD uint32_t mtsb_emu_add(uint32_t *p, uint32_t v) { uint32_t o = *p; *p = o + v; return o; }
#define MTSB_ATOMIC_ADD_U32(p, v) mtsb_emu_add((p), (v))
#endif

// smallest e (biased, as a float exponent byte) with extent / 2^(e-127) <= 255
MTSB_HD uint8_t quant_exponent(float extent) {
    if (!(extent > 0.0f)) return 1;
    // 2^e >= extent/255  ->  take the exponent of extent/255 and round up
    float q = extent / 255.0f;
    uint32_t bits = f2u(q);
    int e = (int) ((bits >> 23) & 0xff);
    if (bits & 0x7fffffu) e += 1;
    if (e < 1) e = 1;
    if (e > 254) e = 254;
    // guard against rounding in the division above
    while (e < 254 && extent / u2f((uint32_t) e << 23) > 255.0f) ++e;
    return (uint8_t) e;
}

// `w` = index of the wide node to fill, made from binary node a.wbin[w].
// tiny == true: the whole scene (n <= MTSB_LEAF_MAX) becomes one leaf slot of the root.
MTSB_HD void wide_node_build(uint32_t w, const WideBuildArgs &a, bool tiny) {
    uint32_t cand[8]; int m;
    if (tiny) {
        m = 0;
    } else {
        const uint32_t b = a.wbin[w];
        cand[0] = a.left[b]; cand[1] = a.right[b]; m = 2;
        while (m < 8) {
            int best = -1; float bestArea = -1.0f;
            for (int i = 0; i < m; ++i) {
                if (bin_size(a, cand[i]) <= MTSB_LEAF_MAX) continue;
                float ar = box_area(a.node_lo + 3 * (size_t) cand[i], a.node_hi + 3 * (size_t) cand[i]);
                if (ar > bestArea) { bestArea = ar; best = i; }
            }
            if (best < 0) break;
            uint32_t c = cand[best];
            cand[best] = a.left[c];
            cand[m++] = a.right[c];
        }
    }
    // node box (leaf children are grown by leaf_pad, so grow the frame by it as well)
    float nlo[3], nhi[3];
    float clo[8][3], chi[8][3]; uint32_t csize[8], cfirst[8]; bool cleaf[8];
    if (tiny) {
        m = 1; csize[0] = (uint32_t) a.n; cfirst[0] = 0; cleaf[0] = true;
        // the caller stores the scene box in unified node id 0 for the tiny case
        for (int k = 0; k < 3; ++k) { clo[0][k] = a.node_lo[k] - a.leaf_pad; chi[0][k] = a.node_hi[k] + a.leaf_pad; }
    } else {
        for (int i = 0; i < m; ++i) {
            csize[i] = bin_size(a, cand[i]); cfirst[i] = bin_first(a, cand[i]);
            cleaf[i] = csize[i] <= MTSB_LEAF_MAX;
            float pad = cleaf[i] ? a.leaf_pad : a.leaf_pad;   // inner boxes contain padded leaves
            for (int k = 0; k < 3; ++k) {
                clo[i][k] = a.node_lo[3 * (size_t) cand[i] + k] - pad;
                chi[i][k] = a.node_hi[3 * (size_t) cand[i] + k] + pad;
            }
        }
    }
    for (int k = 0; k < 3; ++k) { nlo[k] = clo[0][k]; nhi[k] = chi[0][k]; }
    for (int i = 1; i < m; ++i)
        for (int k = 0; k < 3; ++k) { nlo[k] = fminf(nlo[k], clo[i][k]); nhi[k] = fmaxf(nhi[k], chi[i][k]); }

    // octant slot assignment: child c goes to the slot s whose sign vector best matches the
    // direction from the node centre to the child's centre (greedy on the 8x8 gain table)
    int slot_of[8]; bool slot_used[8]; bool child_done[8];
    for (int i = 0; i < 8; ++i) { slot_used[i] = false; child_done[i] = false; slot_of[i] = -1; }
    float ncx = 0.5f * (nlo[0] + nhi[0]), ncy = 0.5f * (nlo[1] + nhi[1]), ncz = 0.5f * (nlo[2] + nhi[2]);
    for (int it = 0; it < m; ++it) {
        float bestGain = -FLT_MAX; int bc = -1, bs = -1;
        for (int c = 0; c < m; ++c) {
            if (child_done[c]) continue;
            float dx = 0.5f * (clo[c][0] + chi[c][0]) - ncx, dy = 0.5f * (clo[c][1] + chi[c][1]) - ncy,
                  dz = 0.5f * (clo[c][2] + chi[c][2]) - ncz;
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                float g = ((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 4) ? dz : -dz);
                if (g > bestGain) { bestGain = g; bc = c; bs = s; }
            }
        }
        child_done[bc] = true; slot_used[bs] = true; slot_of[bc] = bs;
    }
    int child_in_slot[8];
    for (int s = 0; s < 8; ++s) child_in_slot[s] = -1;
    for (int c = 0; c < m; ++c) child_in_slot[slot_of[c]] = c;

    uint32_t n_inner = 0, n_tri = 0;
    for (int c = 0; c < m; ++c) { if (cleaf[c]) n_tri += csize[c]; else ++n_inner; }
    uint32_t child_base = n_inner ? MTSB_ATOMIC_ADD_U32(a.counters + 0, n_inner) : 0u;
    uint32_t tri_base = n_tri ? MTSB_ATOMIC_ADD_U32(a.counters + 1, n_tri) : 0u;

    Node8 nd;
    nd.px = nlo[0]; nd.py = nlo[1]; nd.pz = nlo[2];
    nd.ex = quant_exponent(nhi[0] - nlo[0]); nd.ey = quant_exponent(nhi[1] - nlo[1]); nd.ez = quant_exponent(nhi[2] - nlo[2]);
    const float sx = u2f((uint32_t) nd.ex << 23), sy = u2f((uint32_t) nd.ey << 23), sz = u2f((uint32_t) nd.ez << 23);
    nd.imask = 0; nd.child_base = child_base; nd.tri_base = tri_base; nd.counts = 0; nd.pad0 = 0; nd.pad1 = 0;
    uint32_t inner_rank = 0, tri_off = 0;
    for (int s = 0; s < 8; ++s) {
        int c = child_in_slot[s];
        if (c < 0) {
            nd.qlox[s] = nd.qloy[s] = nd.qloz[s] = 255; nd.qhix[s] = nd.qhiy[s] = nd.qhiz[s] = 0;
            continue;
        }
        // conservative quantisation: floor the low side, ceil the high side
        float qlx = floorf((clo[c][0] - nlo[0]) / sx), qly = floorf((clo[c][1] - nlo[1]) / sy), qlz = floorf((clo[c][2] - nlo[2]) / sz);
        float qhx = ceilf((chi[c][0] - nlo[0]) / sx), qhy = ceilf((chi[c][1] - nlo[1]) / sy), qhz = ceilf((chi[c][2] - nlo[2]) / sz);
        nd.qlox[s] = (uint8_t) fminf(fmaxf(qlx, 0.0f), 255.0f); nd.qloy[s] = (uint8_t) fminf(fmaxf(qly, 0.0f), 255.0f);
        nd.qloz[s] = (uint8_t) fminf(fmaxf(qlz, 0.0f), 255.0f);
        nd.qhix[s] = (uint8_t) fminf(fmaxf(qhx, 0.0f), 255.0f); nd.qhiy[s] = (uint8_t) fminf(fmaxf(qhy, 0.0f), 255.0f);
        nd.qhiz[s] = (uint8_t) fminf(fmaxf(qhz, 0.0f), 255.0f);
        if (cleaf[c]) {
            uint32_t cnt = csize[c];
            nd.counts |= (uint16_t) (cnt << (2 * s));
            for (uint32_t j = 0; j < cnt; ++j)
                a.tri_out[(size_t) tri_base + tri_off + j] = a.tri_in[a.sorted_prim[cfirst[c] + j]];
            tri_off += cnt;
        } else {
            nd.imask |= (uint8_t) (1u << s);
            a.wbin[child_base + inner_rank] = cand[c];
            ++inner_rank;
        }
    }
    a.nodes[w] = nd;
}

}  // namespace mtsb
