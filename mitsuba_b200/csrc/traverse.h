// traverse.h -- closest-hit / any-hit ray query against the 8-wide BVH.
//
// Replaces ShapeKDTree::rayIntersect (reference: src/librender/skdtree.cpp:112-148 closest,
// :213-232 any-hit) and SAHKDTree3D::rayIntersectHavran (include/mitsuba/render/sahkdtree3.h:178-308).
// What is kept from the reference, bit for bit:
//   * the scene-box clip with the enlarged box (include/mitsuba/core/aabb.h:308-338),
//   * the adaptive epsilon (skdtree.cpp:127-130 / :219-222),
//   * the triangle test and its acceptance rule (include/mitsuba/render/triaccel.h:102-168):
//     inclusive [mint,maxt], u>=0, v>=0, u+v<=1, an accepted hit shrinks maxt, equal t re-accepts.
// What is new: the search structure (bvh.h) and a traversal written as a resumable state
// machine, so that a persistent warp can swap finished rays for fresh ones between steps.
#pragma once
#include "bvh.h"

namespace mtsb {

struct RayIn { float ox, oy, oz, mint, dx, dy, dz, maxt; };   // == MtsbRay
struct HitOut { uint32_t prim; float u, v, t; };              // == MtsbHit
#define MTSB_NO_HIT_ID 0xFFFFFFFFu

// AABB::rayIntersect (aabb.h:308-338) on the enlarged scene box + the interval logic of
// ShapeKDTree::rayIntersect.  Returns false when the query interval is empty.
MTSB_HD bool ray_interval(const RayIn &r, const float *box_lo, const float *box_hi, bool any_hit,
                          float &mint, float &maxt) {
    float nearT = -MTSB_INF, farT = MTSB_INF;
    const float o[3] = {r.ox, r.oy, r.oz}, d[3] = {r.dx, r.dy, r.dz};
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
    for (int i = 0; i < 3; ++i) {
        const float origin = o[i], minVal = box_lo[i], maxVal = box_hi[i];
        if (d[i] == 0.0f) {
            if (origin < minVal || origin > maxVal) return false;
        } else {
            const float rcp = fdiv(1.0f, d[i]);
            float t1 = fmul(fsub(minVal, origin), rcp), t2 = fmul(fsub(maxVal, origin), rcp);
            if (t1 > t2) { float tmp = t1; t1 = t2; t2 = tmp; }
            nearT = fmaxf(t1, nearT);
            farT = fminf(t2, farT);
            if (!(nearT <= farT)) return false;
        }
    }
    float rayMinT = r.mint;
    if (rayMinT == MTSB_EPSILON) {
        float m = fmaxf(fmaxf(fabsf(r.ox), fabsf(r.oy)), fabsf(r.oz));
        if (!any_hit) m = fmaxf(m, MTSB_EPSILON);
        rayMinT = fmul(rayMinT, m);
    }
    mint = nearT; maxt = farT;
    if (rayMinT > mint) mint = rayMinT;
    if (r.maxt < maxt) maxt = r.maxt;
    return maxt > mint;
}

// TriAccel::rayIntersect, triaccel.h:102-168 (operation order and rounding preserved)
MTSB_HD bool tri_test(const U4 q0, const U4 q1, const U4 q2, float ox, float oy, float oz, float dx, float dy, float dz,
                      float mint, float maxt, float &u, float &v, float &t) {
    const uint32_t k = q0.x;
    const float n_u = u2f(q0.y), n_v = u2f(q0.z), n_d = u2f(q0.w);
    const float a_u = u2f(q1.x), a_v = u2f(q1.y), b_nu = u2f(q1.z), b_nv = u2f(q1.w);
    const float c_nu = u2f(q2.x), c_nv = u2f(q2.y);
    float o_u, o_v, o_k, d_u, d_v, d_k;
    if (k == 0) { o_u = oy; o_v = oz; o_k = ox; d_u = dy; d_v = dz; d_k = dx; }
    else if (k == 1) { o_u = oz; o_v = ox; o_k = oy; d_u = dz; d_v = dx; d_k = dy; }
    else if (k == 2) { o_u = ox; o_v = oy; o_k = oz; d_u = dx; d_v = dy; d_k = dz; }
    else return false;
    const float num = fsub(fsub(fsub(n_d, fmul(o_u, n_u)), fmul(o_v, n_v)), o_k);
    const float den = fadd(fadd(fmul(d_u, n_u), fmul(d_v, n_v)), d_k);
    {
        // Conservative early-out with an approximate quotient: a triangle is dropped here only when
        // the exactly rounded evaluation below could not accept it (margins of ~64 ulp against the
        // few ulp by which the two evaluations can differ).  Everything that survives is decided by
        // the exact arithmetic, so accepted hits and their t,u,v are unchanged.
        const float ta = fdiv_approx(num, den);
        const float et = 8e-6f * fabsf(ta);
        if (ta < mint - et || ta > maxt + et) return false;
        const float pu = ta * d_u, pv = ta * d_v;
        const float hua = (o_u + pu) - a_u, hva = (o_v + pv) - a_v;
        const float ehu = 8e-6f * (fabsf(o_u) + fabsf(pu) + fabsf(a_u)), ehv = 8e-6f * (fabsf(o_v) + fabsf(pv) + fabsf(a_v));
        const float ua = hva * b_nu + hua * b_nv, va = hua * c_nu + hva * c_nv;
        const float eu = fabsf(b_nu) * ehv + fabsf(b_nv) * ehu + 8e-6f * (fabsf(hva * b_nu) + fabsf(hua * b_nv));
        const float ev = fabsf(c_nu) * ehu + fabsf(c_nv) * ehv + 8e-6f * (fabsf(hua * c_nu) + fabsf(hva * c_nv));
        if (ua < -eu || va < -ev || ua + va > 1.0f + eu + ev + 8e-6f) return false;
    }
    t = fdiv(num, den);
    if (t < mint || t > maxt) return false;
    const float hu = fsub(fadd(o_u, fmul(t, d_u)), a_u);
    const float hv = fsub(fadd(o_v, fmul(t, d_v)), a_v);
    u = fadd(fmul(hv, b_nu), fmul(hu, b_nv));
    v = fadd(fmul(hu, c_nu), fmul(hv, c_nv));
    return u >= 0.0f && v >= 0.0f && fadd(u, v) <= 1.0f;
}

// The exactly rounded part of tri_test alone: t, u, v of a triangle that is known to be hit (same operations in the
// same order, hence the same bits).  The traversal keeps only the record index of the accepted triangle while it
// runs and evaluates the hit record once, when the ray is finished.
MTSB_HD void tri_eval(const U4 q0, const U4 q1, const U4 q2, float ox, float oy, float oz, float dx, float dy, float dz,
                      float &u, float &v, float &t) {
    const uint32_t k = q0.x;
    const float n_u = u2f(q0.y), n_v = u2f(q0.z), n_d = u2f(q0.w);
    const float a_u = u2f(q1.x), a_v = u2f(q1.y), b_nu = u2f(q1.z), b_nv = u2f(q1.w);
    const float c_nu = u2f(q2.x), c_nv = u2f(q2.y);
    float o_u, o_v, o_k, d_u, d_v, d_k;
    if (k == 0) { o_u = oy; o_v = oz; o_k = ox; d_u = dy; d_v = dz; d_k = dx; }
    else if (k == 1) { o_u = oz; o_v = ox; o_k = oy; d_u = dz; d_v = dx; d_k = dy; }
    else { o_u = ox; o_v = oy; o_k = oz; d_u = dx; d_v = dy; d_k = dz; }
    const float num = fsub(fsub(fsub(n_d, fmul(o_u, n_u)), fmul(o_v, n_v)), o_k);
    const float den = fadd(fadd(fmul(d_u, n_u), fmul(d_v, n_v)), d_k);
    t = fdiv(num, den);
    const float hu = fsub(fadd(o_u, fmul(t, d_u)), a_u);
    const float hv = fsub(fadd(o_v, fmul(t, d_v)), a_v);
    u = fadd(fmul(hv, b_nu), fmul(hu, b_nv));
    v = fadd(fmul(hu, c_nu), fmul(hv, c_nv));
}

// bit t of the result = bit (t ^ o) of the 8-bit mask m: a node's slot mask seen in the traversal order of a ray with
// octant word o (three conditional swaps of neighbouring bits / pairs / nibbles)
MTSB_HD uint32_t perm_byte(uint32_t m, uint32_t o) {
    if (o & 1u) m = ((m & 0x55u) << 1) | ((m >> 1) & 0x55u);
    if (o & 2u) m = ((m & 0x33u) << 2) | ((m >> 2) & 0x33u);
    if (o & 4u) m = ((m & 0x0fu) << 4) | ((m >> 4) & 0x0fu);
    return m;
}

// 0x3F800000 (1.0f: bits 23..29 set) when a <= b (false for NaN), else zero: one FSET.BF instead of a predicate +
// select pair.  sm_100a has no integer-mask FSET, so the hit-bit tree below works on bits 24..27 of these words.
MTSB_HD uint32_t fset_le(float a, float b) {
#if MTSB_DEVICE_CODE
    float r; asm("set.le.f32.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return __float_as_uint(r);
#else
    return a <= b ? 0x3F800000u : 0u;
#endif
}
// bitwise multiplexer (one LOP3): bits of a where c is set, bits of b elsewhere
MTSB_HD uint32_t mux(uint32_t a, uint32_t b, uint32_t c) { return (a & c) | (b & ~c); }

// Plain per-thread stack (registers / local memory).  The CUDA kernels substitute a
// shared-memory short stack with local-memory overflow (trace.cuh).
template <int N> struct LocalStack {
    uint32_t sx[N], sy[N]; int sp;
    MTSB_HD void reset() { sp = 0; }
    MTSB_HD bool empty() const { return sp == 0; }
    MTSB_HD void push(uint32_t x, uint32_t y) { if (sp < N) { sx[sp] = x; sy[sp] = y; ++sp; } }
    MTSB_HD void pop(uint32_t &x, uint32_t &y) { --sp; x = sx[sp]; y = sy[sp]; }
    MTSB_HD uint32_t perm(uint32_t m, uint32_t o) const { return perm_byte(m, o); }
};

template <class Stack> struct Traversal {
    float ox, oy, oz, dx, dy, dz, tmin, tmax;
    float idx, idy, idz;
    uint32_t octinv;
    uint32_t sel1, sel2, sel4;    // multiplexer words of the hit-bit tree (see visit_node); sel4 is all ones or zero
    uint32_t bias;                // 0x47000000 (see byte_bias_float)
    uint32_t ng_x, ng_y;          // current node group: (child base, pending inner children in traversal order << 24 | imask)
    uint32_t hit_rec;             // record index (leaf order) of the accepted triangle, MTSB_NO_HIT_ID when none
    Stack stack;
#if !MTSB_DEVICE_CODE
    uint32_t steps = 0;           // nodes visited      (emulation statistics only)
    uint32_t sparse_steps = 0;    // ... of which leaf-only nodes with <= 4 children
    uint32_t tri_tests = 0;       // triangle tests
    uint32_t last_batch = 0;      // triangles queued by the last node visit (as a mask with that many bits)
#endif

    MTSB_HD void begin(const RayIn &r, float mint, float maxt) {
        ox = r.ox; oy = r.oy; oz = r.oz; dx = r.dx; dy = r.dy; dz = r.dz; tmin = mint; tmax = maxt;
        const float ooeps = 8.271806125530277e-25f;   // 2^-80
        idx = 1.0f / (fabsf(dx) > ooeps ? dx : copysignf(ooeps, dx));
        idy = 1.0f / (fabsf(dy) > ooeps ? dy : copysignf(ooeps, dy));
        idz = 1.0f / (fabsf(dz) > ooeps ? dz : copysignf(ooeps, dz));
        octinv = (dx >= 0.0f ? 1u : 0u) | (dy >= 0.0f ? 2u : 0u) | (dz >= 0.0f ? 4u : 0u);
        sel1 = (octinv & 1u) ? 0xAAAAAAAAu : 0x55555555u;
        sel2 = (octinv & 2u) ? 0xCCCCCCCCu : 0x33333333u;
        sel4 = (octinv & 4u) ? 0u : 0xFFFFFFFFu;
#if !MTSB_DEVICE_CODE
        bias = 0x47000000u;
#endif
        ng_x = 0; ng_y = 0x80000000u;
        hit_rec = MTSB_NO_HIT_ID;
        stack.reset();
        tw = 0; tbase = 0; tcur = 0; tleft = 0; tw2 = 0; tbase2 = 0;
#if !MTSB_DEVICE_CODE
        steps = 0; sparse_steps = 0; last_batch = 0;
#endif
    }

    // The query is a small state machine, so that a warp can run its 32 rays in lock step.  Every
    // iteration each lane (a) visits one node if it has node work and a free leaf-batch slot,
    // (b) tests ONE pending triangle, (c) finishes when nothing is left.  A lane keeps up to two
    // batches of hit leaves (the one being tested and one deferred), so a lane that is working
    // through a leaf keeps descending with the rest of the warp instead of idling through the
    // (expensive) node phase; tools/warp_sim.py has the model behind this choice.
    //
    // A leaf batch is the set of hit leaf children of ONE node: word = hit leaves (8 bits, traversal order) << 24 |
    // the node's 2-bit triangle counts, plus the node's tri_base.  Leaves are expanded lazily, one at a time, when their first triangle is due: the node visit
    // itself spends no instructions on per-child triangle masks.
    uint32_t tw, tbase;           // batch whose leaves are handed out next
    uint32_t tcur, tleft;         // sequential form only: next triangle record / triangles left of the leaf being tested
    uint32_t tw2, tbase2;         // deferred batch (empty when tw2 == 0)

    MTSB_HD bool has_nodes() const { return ng_y > 0x00FFFFFFu || !stack.empty(); }
    MTSB_HD bool wants_node() const { return tw2 == 0 && has_nodes(); }
    MTSB_HD bool has_triangles() const { return tw != 0 || tleft != 0; }
    MTSB_HD bool finished() const { return tw == 0 && tleft == 0 && !has_nodes(); }

    // (a) take the nearest pending child of the current group and intersect its 8 children
    MTSB_HD void visit_node(const U4 *nodes) {
        if (ng_y <= 0x00FFFFFFu) stack.pop(ng_x, ng_y);
        const uint32_t hits_imask = ng_y;
        const int bit = bfind(hits_imask);
        const uint32_t base = ng_x;
        ng_y &= ~(1u << bit);
        if (ng_y > 0x00FFFFFFu) stack.push(ng_x, ng_y);
        const uint32_t slot = ((uint32_t) bit - 24u) ^ octinv;
        const uint32_t rel = (uint32_t) popc(hits_imask & ~(0xFFFFFFFFu << slot) & 0xFFu);
        const U4 *np = nodes + 5 * (size_t) (base + rel);
        const U4 n0 = ld16(np), n1 = ld16(np + 1), n2 = ld16(np + 2), n3 = ld16(np + 3), n4 = ld16(np + 4);
        const uint32_t e_imask = n0.w;
        const uint32_t imask = e_imask >> 24, counts = n1.z & 0xffffu;
#if !MTSB_DEVICE_CODE
        ++steps;
        {   // emulation statistics: visits of leaf-only nodes with at most four children
            int nchild = 0;
            for (int j = 0; j < 8; ++j) if ((counts >> (2 * j)) & 3u) ++nchild;
            if (imask == 0 && nchild <= 4) ++sparse_steps;
        }
#endif
        const float sdx = u2f((e_imask & 0xffu) << 23) * idx;
        const float sdy = u2f(((e_imask >> 8) & 0xffu) << 23) * idy;
        const float sdz = u2f(((e_imask >> 16) & 0xffu) << 23) * idz;
        const float ax = (u2f(n0.x) - ox) * idx, ay = (u2f(n0.y) - oy) * idy, az = (u2f(n0.z) - oz) * idz;
        // Plane bytes are turned into floats by a byte permute: byte_bias_float() = 32768 + q, so the plane
        // parameter is (32768 + q) * sd + (a - 32768 * sd).  Rounding the biased offset costs at most 2^-9 |sd|
        // (1/512 of a quantisation step); near planes are pulled in and far planes pushed out by 2^-8 |sd| so the
        // box test stays conservative (it only ever decides which nodes are visited, never a hit record).
        const float bx = ax - 32768.0f * sdx, by = ay - 32768.0f * sdy, bz = az - 32768.0f * sdz;
        const float gx = fabsf(sdx) * 0.00390625f, gy = fabsf(sdy) * 0.00390625f, gz = fabsf(sdz) * 0.00390625f;
        const float anx = bx - gx, any_ = by - gy, anz = bz - gz, afx = bx + gx, afy = by + gy, afz = bz + gz;
        // near / far planes by ray direction sign
        const bool nx = !(octinv & 1u), ny = !(octinv & 2u), nz = !(octinv & 4u);      // d < 0 (begin(): octinv bit = d >= 0)
        const uint32_t lox0 = n2.x, lox1 = n2.y, loy0 = n2.z, loy1 = n2.w, loz0 = n3.x, loz1 = n3.y;
        const uint32_t hix0 = n3.z, hix1 = n3.w, hiy0 = n4.x, hiy1 = n4.y, hiz0 = n4.z, hiz1 = n4.w;
        const uint32_t nearx[2] = {nx ? hix0 : lox0, nx ? hix1 : lox1}, farx[2] = {nx ? lox0 : hix0, nx ? lox1 : hix1};
        const uint32_t neary[2] = {ny ? hiy0 : loy0, ny ? hiy1 : loy1}, fary[2] = {ny ? loy0 : hiy0, ny ? loy1 : hiy1};
        const uint32_t nearz[2] = {nz ? hiz0 : loz0, nz ? hiz1 : loz1}, farz[2] = {nz ? loz0 : hiz0, nz ? loz1 : hiz1};
        // slab test of the eight children: one all-ones / zero word per slot (an empty slot has lo = 255 > hi = 0 and never hits)
        uint32_t m[8];
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
        for (int h = 0; h < 2; ++h) {
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
            for (int j = 0; j < 4; ++j) {
                const float tnx = byte_bias_float(nearx[h], j, bias) * sdx + anx;
                const float tny = byte_bias_float(neary[h], j, bias) * sdy + any_;
                const float tnz = byte_bias_float(nearz[h], j, bias) * sdz + anz;
                const float tfx = byte_bias_float(farx[h], j, bias) * sdx + afx;
                const float tfy = byte_bias_float(fary[h], j, bias) * sdy + afy;
                const float tfz = byte_bias_float(farz[h], j, bias) * sdz + afz;
                const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
                const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
                m[4 * h + j] = fset_le(cmin, cmax);
            }
        }
        // Hit byte in TRAVERSAL order: slot s lands on bit 24 + (s ^ octinv).  The eight words are merged by a multiplexer
        // tree whose select words depend only on the ray's octant: sel1 / sel2 put the even slot of a pair / the lower pair
        // of a quad on the low or the high side (both quads are assembled in bits 24..27), the last level puts the quads
        // in order (sel4 is all ones when the lower quad stays low).  13 logic instructions for the whole node and no
        // per-child shifts.  The node's inner-child mask is brought into the same order (table lookup on the device),
        // which splits the hits into inner children and leaves.
        const uint32_t x01 = mux(m[0], m[1], sel1), x23 = mux(m[2], m[3], sel1), x45 = mux(m[4], m[5], sel1), x67 = mux(m[6], m[7], sel1);
        const uint32_t x03 = mux(x01, x23, sel2), x47 = mux(x45, x67, sel2);
        const uint32_t qlo = mux(x03, x47, sel4), qhi = mux(x47, x03, sel4);
        const uint32_t hit24 = mux(qlo, qhi << 4, 0x0F000000u);          // qhi << 4 has no bits below 27, qlo none above 29
        const uint32_t inner24 = stack.perm(imask, octinv) << 24;
        ng_x = n1.x;
        ng_y = (hit24 & inner24) | imask;
        const uint32_t leaves24 = hit24 & ~inner24;
#if !MTSB_DEVICE_CODE
        {
            uint32_t nt = 0;
            for (int t = 0; t < 8; ++t) if (leaves24 & (1u << (24 + t))) nt += (counts >> (2 * ((uint32_t) t ^ octinv))) & 3u;
            last_batch = nt ? (0xFFFFFFFFu >> (32 - nt)) : 0u;
        }
#endif
        if (leaves24) {
            const uint32_t w = leaves24 | counts;
            if (tw == 0) { tw = w; tbase = n1.y; } else { tw2 = w; tbase2 = n1.y; }
        }
    }

    // (b) next hit leaf of the pending batch (nearest first): its first triangle record and its triangle count (1..3);
    // promotes the deferred batch when the current one has no leaves left
    MTSB_HD void next_leaf(uint32_t &rec0, uint32_t &cnt) {
        const int b = bfind(tw);                                   // 24 + position in traversal order
        tw &= ~(1u << b);
        const uint32_t s2 = (((uint32_t) b - 24u) ^ octinv) * 2u;
        const uint32_t c = tw & 0xffffu;
        const uint32_t below = c & ~(0xFFFFFFFFu << s2);
        rec0 = tbase + (uint32_t) popc(below & 0x5555u) + 2u * (uint32_t) popc(below & 0xAAAAu);
        cnt = (c >> s2) & 3u;
        if ((tw & 0xFF000000u) == 0) { tw = tw2; tbase = tbase2; tw2 = 0; }
    }

    // (b') sequential form (emulation build, trace_one): test ONE pending triangle; returns true when an any-hit query
    // is satisfied.  The leaf being tested is kept in (tcur, tleft) because next_leaf() may already have promoted the batch.
    template <bool ANY> MTSB_HD bool test_triangle(const U4 *tris) {
        if (tleft == 0) next_leaf(tcur, tleft);
#if !MTSB_DEVICE_CODE
        ++tri_tests;
#endif
        const uint32_t rec = tcur;
        const U4 *tp = tris + 3 * (size_t) rec;
        ++tcur; --tleft;
        const U4 q0 = ld16(tp), q1 = ld16(tp + 1), q2 = ld16(tp + 2);
        float u, v, t;
        if (tri_test(q0, q1, q2, ox, oy, oz, dx, dy, dz, tmin, tmax, u, v, t)) {
            hit_rec = rec;
            if (ANY) return true;
            tmax = t;
        }
        return false;
    }

    // the hit record of a finished query (closest: hit.t stays +inf when nothing was accepted, skdtree.cpp:137-140)
    MTSB_HD HitOut result(const U4 *tris) const {
        HitOut h; h.prim = MTSB_NO_HIT_ID; h.u = 0; h.v = 0; h.t = MTSB_INF;
        if (hit_rec != MTSB_NO_HIT_ID) {
            const U4 *tp = tris + 3 * (size_t) hit_rec;
            const U4 q0 = ld16(tp), q1 = ld16(tp + 1), q2 = ld16(tp + 2);
            tri_eval(q0, q1, q2, ox, oy, oz, dx, dy, dz, h.u, h.v, h.t);
            h.prim = q2.z;
        }
        return h;
    }

    // one lock-step iteration for a single ray
    template <bool ANY> MTSB_HD bool step(const U4 *nodes, const U4 *tris) {
        if (wants_node()) visit_node(nodes);
        if (has_triangles() && test_triangle<ANY>(tris)) return true;
        return finished();
    }
};

// Whole query for one ray (used by the emulation build and by small device helpers).
template <bool ANY, class Stack>
MTSB_HD void trace_one(const RayIn &r, const float *box_lo, const float *box_hi, const U4 *nodes, const U4 *tris,
                       uint32_t n_nodes, HitOut &out) {
    out.prim = MTSB_NO_HIT_ID; out.u = 0; out.v = 0; out.t = MTSB_INF;
    float mint, maxt;
    if (n_nodes == 0 || !ray_interval(r, box_lo, box_hi, ANY, mint, maxt)) return;
    Traversal<Stack> tr;
    tr.begin(r, mint, maxt);
    while (!tr.template step<ANY>(nodes, tris)) {}
    out = tr.result(tris);
}

}  // namespace mtsb
