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

#if MTSB_DEVICE_CODE
MTSB_HD U4 ld16(const U4 *p) { return __ldg(p); }
#else
MTSB_HD U4 ld16(const U4 *p) { return *p; }
#endif

// Plain per-thread stack (registers / local memory).  The CUDA kernels substitute a
// shared-memory short stack with local-memory overflow (trace.cu).
template <int N> struct LocalStack {
    uint32_t sx[N], sy[N]; int sp;
    MTSB_HD void reset() { sp = 0; }
    MTSB_HD bool empty() const { return sp == 0; }
    MTSB_HD void push(uint32_t x, uint32_t y) { if (sp < N) { sx[sp] = x; sy[sp] = y; ++sp; } }
    MTSB_HD void pop(uint32_t &x, uint32_t &y) { --sp; x = sx[sp]; y = sy[sp]; }
};

template <class Stack> struct Traversal {
    float ox, oy, oz, dx, dy, dz, tmin, tmax;
    float idx, idy, idz;
    uint32_t octinv;
    uint32_t bias;                // 0x47000000 (see byte_bias_float)
    uint32_t ng_x, ng_y;          // current node group: (child base, hit bits<<24 | imask)
    HitOut hit;
    Stack stack;
#if !MTSB_DEVICE_CODE
    uint32_t steps = 0;           // nodes visited      (emulation statistics only)
    uint32_t sparse_steps = 0;    // ... of which leaf-only nodes with <= 4 children
    uint32_t tri_tests = 0;       // triangle tests
    uint32_t last_batch = 0;      // triangles queued by the last node visit
#endif

    MTSB_HD void begin(const RayIn &r, float mint, float maxt) {
        ox = r.ox; oy = r.oy; oz = r.oz; dx = r.dx; dy = r.dy; dz = r.dz; tmin = mint; tmax = maxt;
        const float ooeps = 8.271806125530277e-25f;   // 2^-80
        idx = 1.0f / (fabsf(dx) > ooeps ? dx : copysignf(ooeps, dx));
        idy = 1.0f / (fabsf(dy) > ooeps ? dy : copysignf(ooeps, dy));
        idz = 1.0f / (fabsf(dz) > ooeps ? dz : copysignf(ooeps, dz));
        octinv = (dx >= 0.0f ? 1u : 0u) | (dy >= 0.0f ? 2u : 0u) | (dz >= 0.0f ? 4u : 0u);
#if !MTSB_DEVICE_CODE
        bias = 0x47000000u;
#endif
        ng_x = 0; ng_y = 0x80000000u;
        hit.prim = MTSB_NO_HIT_ID; hit.u = 0; hit.v = 0; hit.t = MTSB_INF;
        stack.reset();
        tg = 0; tri_base = 0; tg2 = 0; tri_base2 = 0;
#if !MTSB_DEVICE_CODE
        steps = 0; sparse_steps = 0; last_batch = 0;
#endif
    }

    // The query is a small state machine, so that a warp can run its 32 rays in lock step.  Every
    // iteration each lane (a) visits one node if it has node work and a free triangle-batch slot,
    // (b) tests ONE pending triangle, (c) finishes when nothing is left.  A lane keeps up to two
    // batches of leaf triangles (the one being tested and one deferred), so a lane that is working
    // through a leaf keeps descending with the rest of the warp instead of idling through the
    // (expensive) node phase; tools/warp_sim.py has the model behind this choice.
    uint32_t tg, tri_base;        // triangle batch being tested: pending bits, first triangle record
    uint32_t tg2, tri_base2;      // deferred batch (empty when tg2 == 0)

    MTSB_HD bool has_nodes() const { return ng_y > 0x00FFFFFFu || !stack.empty(); }
    MTSB_HD bool wants_node() const { return tg2 == 0 && has_nodes(); }
    MTSB_HD bool finished() const { return tg == 0 && !has_nodes(); }

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
#if !MTSB_DEVICE_CODE
        ++steps;
        {   // emulation statistics: visits of leaf-only nodes with at most four children
            int nchild = 0;
            for (int j = 0; j < 4; ++j) { if ((n1.z >> (8 * j)) & 0xffu) ++nchild; if ((n1.w >> (8 * j)) & 0xffu) ++nchild; }
            if ((n0.w >> 24) == 0 && nchild <= 4) ++sparse_steps;
        }
#endif

        const uint32_t e_imask = n0.w;
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
        const bool nx = dx < 0.0f, ny = dy < 0.0f, nz = dz < 0.0f;
        const uint32_t lox0 = n2.x, lox1 = n2.y, loy0 = n2.z, loy1 = n2.w, loz0 = n3.x, loz1 = n3.y;
        const uint32_t hix0 = n3.z, hix1 = n3.w, hiy0 = n4.x, hiy1 = n4.y, hiz0 = n4.z, hiz1 = n4.w;
        const uint32_t nearx[2] = {nx ? hix0 : lox0, nx ? hix1 : lox1}, farx[2] = {nx ? lox0 : hix0, nx ? lox1 : hix1};
        const uint32_t neary[2] = {ny ? hiy0 : loy0, ny ? hiy1 : loy1}, fary[2] = {ny ? loy0 : hiy0, ny ? loy1 : hiy1};
        const uint32_t nearz[2] = {nz ? hiz0 : loz0, nz ? hiz1 : loz1}, farz[2] = {nz ? loz0 : hiz0, nz ? loz1 : hiz1};
        // per-child bit position / bit pattern, four children per 32-bit word:
        //   inner child (meta = 001sssss, s >= 24): one bit at 24 + (slot ^ octinv)
        //   leaf  child (meta = uuuooooo)         : bits uuu at offset ooooo
        const uint32_t octinv4 = octinv * 0x01010101u;
        uint32_t bit_index4[2], child_bits4[2];
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
        for (int h = 0; h < 2; ++h) {
            const uint32_t meta4 = h == 0 ? n1.z : n1.w;
            const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
            const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;          // 0xff in the bytes of inner children
            bit_index4[h] = (meta4 ^ (octinv4 & inner_mask4)) & 0x1f1f1f1fu;
            child_bits4[h] = (meta4 >> 5) & 0x07070707u;
        }
        uint32_t hitmask = 0;
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
        for (int h = 0; h < 2; ++h) {
#if MTSB_DEVICE_CODE
#pragma unroll
#endif
            for (int j = 0; j < 4; ++j) {
                const int sh = 8 * j;
                const float tnx = byte_bias_float(nearx[h], j, bias) * sdx + anx;
                const float tny = byte_bias_float(neary[h], j, bias) * sdy + any_;
                const float tnz = byte_bias_float(nearz[h], j, bias) * sdz + anz;
                const float tfx = byte_bias_float(farx[h], j, bias) * sdx + afx;
                const float tfy = byte_bias_float(fary[h], j, bias) * sdy + afy;
                const float tfz = byte_bias_float(farz[h], j, bias) * sdz + afz;
                const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
                const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
                const uint32_t bits = ((child_bits4[h] >> sh) & 0xffu) << ((bit_index4[h] >> sh) & 0xffu);
                hitmask |= (cmin <= cmax) ? bits : 0u;
            }
        }
        ng_x = n1.x;
        ng_y = (hitmask & 0xFF000000u) | (e_imask >> 24);
        const uint32_t batch = hitmask & 0x00FFFFFFu;
        if (tg == 0) { tg = batch; tri_base = n1.y; } else { tg2 = batch; tri_base2 = n1.y; }
#if !MTSB_DEVICE_CODE
        last_batch = batch;
#endif
    }

    // (b) test one pending triangle; returns true when an any-hit query is satisfied
    template <bool ANY> MTSB_HD bool test_triangle(const U4 *tris) {
        const int i = bfind(tg);
        tg &= ~(1u << i);
#if !MTSB_DEVICE_CODE
        ++tri_tests;
#endif
        const U4 *tp = tris + 3 * ((size_t) tri_base + (size_t) i);
        if (tg == 0) { tg = tg2; tri_base = tri_base2; tg2 = 0; }      // promote the deferred batch
        const U4 q0 = ld16(tp), q1 = ld16(tp + 1), q2 = ld16(tp + 2);
        float u, v, t;
        if (tri_test(q0, q1, q2, ox, oy, oz, dx, dy, dz, tmin, tmax, u, v, t)) {
            hit.prim = q2.z; hit.u = u; hit.v = v; hit.t = t;
            if (ANY) return true;
            tmax = t;
        }
        return false;
    }

    // one lock-step iteration for a single ray
    template <bool ANY> MTSB_HD bool step(const U4 *nodes, const U4 *tris) {
        if (wants_node()) visit_node(nodes);
        if (tg != 0 && test_triangle<ANY>(tris)) return true;
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
    out = tr.hit;   // closest: hit.t stays +inf when nothing was accepted (skdtree.cpp:137-140)
}

}  // namespace mtsb
