// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// Ray/triangle test restated from TriAccel, the ShapeKDTree::rayIntersect interval logic,
// and a plain binned-SAH BVH2 to find candidate triangles.  The BVH is NOT the reference's
// kd-tree: it visits every triangle whose box the ray touches, so the set of accepted hits
// (and therefore the closest one) is the same as the kd-tree's on non-degenerate rays; for
// equal-t ties the kd-tree's own order is only available through oracle/_ref (ref_glue.cpp).
#pragma once
#include "orc_math.h"
#include "../include/mtsb200.h"
#include <vector>
#include <cstring>
#include <thread>
#include <atomic>

namespace orc {

// include/mitsuba/render/triaccel.h:37-62 (48-byte Wald record)
struct TriAccel {
    uint32_t k;
    float n_u, n_v, n_d;
    float a_u, a_v, b_nu, b_nv;
    float c_nu, c_nv;
    uint32_t shapeIndex, primIndex;

    // triaccel.h:64-100
    int load(V3 A, V3 B, V3 C) {
        static const int waldModulo[4] = {1, 2, 0, 1};
        V3 b = C - A, c = B - A, N = cross(c, b);
        k = 0;
        for (int j = 0; j < 3; j++)
            if (std::fabs(N[j]) > std::fabs(N[k])) k = j;
        uint32_t u = waldModulo[k], v = waldModulo[k + 1];
        const float n_k = N[k], denom = b[u] * c[v] - b[v] * c[u];
        if (denom == 0) { k = 3; return 1; }
        n_u = N[u] / n_k;
        n_v = N[v] / n_k;
        n_d = dot(A, N) / n_k;
        b_nu = b[u] / denom;
        b_nv = -b[v] / denom;
        a_u = A[u];
        a_v = A[v];
        c_nu = c[v] / denom;
        c_nv = -c[u] / denom;
        return 0;
    }

    // triaccel.h:102-168
    bool rayIntersect(const float *o, const float *d, float mint, float maxt, float &u, float &v, float &t) const {
        float o_u, o_v, o_k, d_u, d_v, d_k;
        switch (k) {
            case 0: o_u = o[1]; o_v = o[2]; o_k = o[0]; d_u = d[1]; d_v = d[2]; d_k = d[0]; break;
            case 1: o_u = o[2]; o_v = o[0]; o_k = o[1]; d_u = d[2]; d_v = d[0]; d_k = d[1]; break;
            case 2: o_u = o[0]; o_v = o[1]; o_k = o[2]; d_u = d[0]; d_v = d[1]; d_k = d[2]; break;
            default: return false;
        }
        t = (n_d - o_u * n_u - o_v * n_v - o_k) / (d_u * n_u + d_v * n_v + d_k);
        if (t < mint || t > maxt) return false;
        const float hu = o_u + t * d_u - a_u;
        const float hv = o_v + t * d_v - a_v;
        u = hv * b_nu + hu * b_nv;
        v = hu * c_nu + hv * c_nv;
        return u >= 0 && v >= 0 && u + v <= 1.0f;
    }
};

struct Box {
    V3 lo, hi;
    Box() : lo(kInf), hi(-kInf) {}
    void grow(V3 p) {
        for (int i = 0; i < 3; ++i) { lo[i] = std::min(lo[i], p[i]); hi[i] = std::max(hi[i], p[i]); }
    }
    void grow(const Box &b) { grow(b.lo); grow(b.hi); }
    float area() const { V3 e = hi - lo; return 2 * (e.x * e.y + e.y * e.z + e.z * e.x); }
};

// include/mitsuba/core/aabb.h:308-338 (uses ray.dRcp = 1/d)
inline bool aabbRayIntersect(const Box &bx, const float *o, const float *d, float &nearT, float &farT) {
    nearT = -kInf; farT = kInf;
    for (int i = 0; i < 3; i++) {
        const float origin = o[i], minVal = bx.lo[i], maxVal = bx.hi[i];
        if (d[i] == 0) {
            if (origin < minVal || origin > maxVal) return false;
        } else {
            float rcp = 1.0f / d[i];
            float t1 = (minVal - origin) * rcp, t2 = (maxVal - origin) * rcp;
            if (t1 > t2) std::swap(t1, t2);
            nearT = std::max(t1, nearT);
            farT = std::min(t2, farT);
            if (!(nearT <= farT)) return false;
        }
    }
    return true;
}

struct Accel {
    std::vector<TriAccel> tris;      // indexed by global prim id (skdtree.cpp:82-98 order)
    Box tight, enlarged;             // m_tightAABB / m_aabb (gkdtree.h:1215-1220)
    struct Node { Box box; uint32_t left, count; /* count>0: leaf [left, left+count) in order[] */ };
    std::vector<Node> nodes;
    std::vector<uint32_t> order;

    void build(const MtsbSceneDesc &sd) {
        size_t n = sd.n_tris;
        tris.resize(n);
        std::vector<Box> boxes(n);
        std::vector<V3> cent(n);
        for (uint32_t s = 0; s < sd.n_shapes; ++s) {
            const MtsbShape &sh = sd.shapes[s];
            for (uint32_t j = 0; j < sh.n_tris; ++j) {
                size_t g = (size_t) sh.first_tri + j;
                const uint32_t *ix = sd.indices + 3 * g;
                V3 p[3];
                for (int q = 0; q < 3; ++q) p[q] = V3(sd.positions[3 * ix[q]], sd.positions[3 * ix[q] + 1], sd.positions[3 * ix[q] + 2]);
                tris[g].load(p[0], p[1], p[2]);
                tris[g].shapeIndex = s;
                tris[g].primIndex = j;
                Box b; b.grow(p[0]); b.grow(p[1]); b.grow(p[2]);
                boxes[g] = b; cent[g] = (b.lo + b.hi) * 0.5f;
                tight.grow(b);
            }
        }
        // gkdtree.h:1213-1220: min -= (max-min)*1e-3 + 1e-3, then max += (max-min')*1e-3 + 1e-3
        // where min' is the already-moved minimum (two sequential statements).
        enlarged = tight;
        if (n > 0) {
            enlarged.lo = tight.lo - ((tight.hi - tight.lo) * 1e-3f + V3(1e-3f));
            enlarged.hi = tight.hi + ((tight.hi - enlarged.lo) * 1e-3f + V3(1e-3f));
        }
        order.resize(n);
        for (size_t i = 0; i < n; ++i) order[i] = (uint32_t) i;
        nodes.clear();
        nodes.reserve(2 * n / 2 + 16);
        if (n == 0) return;
        nodes.push_back(Node());
        buildRange(0, 0, (uint32_t) n, boxes, cent, 0);
    }

    void buildRange(uint32_t ni, uint32_t lo, uint32_t hi, const std::vector<Box> &boxes, const std::vector<V3> &cent, int depth) {
        Box nb, cb;
        for (uint32_t i = lo; i < hi; ++i) { nb.grow(boxes[order[i]]); cb.grow(cent[order[i]]); }
        nodes[ni].box = nb;
        uint32_t cnt = hi - lo;
        if (cnt <= 4 || depth > 60) { nodes[ni].left = lo; nodes[ni].count = cnt; return; }
        const int NB = 16;
        int bestAxis = -1, bestSplit = -1; float bestCost = kInf;
        for (int ax = 0; ax < 3; ++ax) {
            float cmin = cb.lo[ax], cmax = cb.hi[ax];
            if (!(cmax > cmin)) continue;
            Box bb[NB]; uint32_t bc[NB] = {0};
            float scale = NB / (cmax - cmin);
            for (uint32_t i = lo; i < hi; ++i) {
                int b = std::min(NB - 1, (int) ((cent[order[i]][ax] - cmin) * scale));
                bb[b].grow(boxes[order[i]]); bc[b]++;
            }
            float rightArea[NB]; uint32_t rightCnt[NB];
            Box acc; uint32_t c = 0;
            for (int b = NB - 1; b > 0; --b) { acc.grow(bb[b]); c += bc[b]; rightArea[b] = c ? acc.area() : 0; rightCnt[b] = c; }
            acc = Box(); c = 0;
            for (int b = 0; b < NB - 1; ++b) {
                acc.grow(bb[b]); c += bc[b];
                if (c == 0 || rightCnt[b + 1] == 0) continue;
                float cost = acc.area() * c + rightArea[b + 1] * rightCnt[b + 1];
                if (cost < bestCost) { bestCost = cost; bestAxis = ax; bestSplit = b; }
            }
        }
        uint32_t mid;
        if (bestAxis < 0) {
            mid = lo + cnt / 2;   // all centroids coincide: split by index
        } else {
            float cmin = cb.lo[bestAxis], cmax = cb.hi[bestAxis], scale = NB / (cmax - cmin);
            auto it = std::partition(order.begin() + lo, order.begin() + hi, [&](uint32_t id) {
                int b = std::min(NB - 1, (int) ((cent[id][bestAxis] - cmin) * scale));
                return b <= bestSplit;
            });
            mid = (uint32_t) (it - order.begin());
            if (mid == lo || mid == hi) mid = lo + cnt / 2;
        }
        uint32_t l = (uint32_t) nodes.size();
        nodes.push_back(Node()); nodes.push_back(Node());
        nodes[ni].left = l; nodes[ni].count = 0;
        buildRange(l, lo, mid, boxes, cent, depth + 1);
        buildRange(l + 1, mid, hi, boxes, cent, depth + 1);
    }

    // Conservative slab test for culling only (padded so that it can never reject a triangle
    // TriAccel would accept).
    static bool slab(const Box &b, const float *o, const float *rd, float mint, float maxt) {
        float t0 = mint, t1 = maxt;
        for (int i = 0; i < 3; ++i) {
            float ta = (b.lo[i] - o[i]) * rd[i], tb = (b.hi[i] - o[i]) * rd[i];
            if (ta > tb) std::swap(ta, tb);
            if (ta != ta || tb != tb) continue;          // 0*inf: origin on the plane of a flat axis
            t0 = std::max(t0, ta); t1 = std::min(t1, tb);
        }
        float pad = 1e-4f * std::max(1.0f, std::max(std::fabs(t0), std::fabs(t1) < kInf ? std::fabs(t1) : 0.0f));
        return t0 - pad <= t1 + pad;
    }

    // Candidate search over [mint, maxt] with TriAccel's own acceptance rule.  `anyHit`
    // returns on the first accepted triangle (sahkdtree3.h:286-288).
    bool traverse(const float *o, const float *d, float mint, float maxt, bool anyHit, MtsbHit &hit) const {
        if (nodes.empty()) return false;
        float rd[3] = {1.0f / d[0], 1.0f / d[1], 1.0f / d[2]};
        uint32_t stack[128]; int sp = 0;
        stack[sp++] = 0;
        bool found = false;
        while (sp) {
            const Node &nd = nodes[stack[--sp]];
            if (!slab(nd.box, o, rd, mint, maxt)) continue;
            if (nd.count) {
                for (uint32_t i = nd.left; i < nd.left + nd.count; ++i) {
                    uint32_t id = order[i];
                    float u, v, t;
                    if (tris[id].rayIntersect(o, d, mint, maxt, u, v, t)) {
                        if (anyHit) { hit.prim = id; hit.u = u; hit.v = v; hit.t = t; return true; }
                        // strict "<" keeps the first of two equal-t candidates; the kd-tree keeps the
                        // last one it tests (triaccel.h:157) -- equal-t rays are "degenerate" 
                        if (!found || t < hit.t) { hit.prim = id; hit.u = u; hit.v = v; hit.t = t; }
                        found = true; maxt = hit.t;
                    }
                }
            } else {
                if (sp + 2 > 128) continue;
                stack[sp++] = nd.left; stack[sp++] = nd.left + 1;
            }
        }
        return found;
    }
};

typedef int (*ExternalTraceFn)(void *user, const MtsbRay *ray, float mint, float maxt, int any, MtsbHit *hit);

// ShapeKDTree::rayIntersect(ray, its) skdtree.cpp:112-148 and rayIntersect(ray) :213-232:
// scene-box clip, adaptive epsilon, interval test; then the tree search.
inline bool sceneIntersect(const Accel &acc, const MtsbRay &ray, bool anyHit, MtsbHit &hit,
                           ExternalTraceFn ext = nullptr, void *extUser = nullptr) {
    hit.prim = MTSB_NO_HIT; hit.u = hit.v = 0; hit.t = kInf;
    float mint, maxt;
    if (!aabbRayIntersect(acc.enlarged, ray.o, ray.d, mint, maxt)) return false;
    float rayMinT = ray.mint;
    if (rayMinT == kEpsilon) {
        float m = std::max(std::max(std::fabs(ray.o[0]), std::fabs(ray.o[1])), std::fabs(ray.o[2]));
        if (!anyHit) m = std::max(m, kEpsilon);      // closest-hit variant clamps (skdtree.cpp:129-130), any-hit does not (:220-222)
        rayMinT *= m;
    }
    if (rayMinT > mint) mint = rayMinT;
    if (ray.maxt < maxt) maxt = ray.maxt;
    if (!(maxt > mint)) return false;
    bool ok;
    if (ext) ok = ext(extUser, &ray, mint, maxt, anyHit ? 1 : 0, &hit) != 0;
    else ok = acc.traverse(ray.o, ray.d, mint, maxt, anyHit, hit);
    if (!ok) { hit.prim = MTSB_NO_HIT; hit.t = kInf; return false; }
    if (!anyHit && hit.t == kInf) { hit.prim = MTSB_NO_HIT; return false; }   // skdtree.cpp:137-140
    return true;
}

}  // namespace orc
