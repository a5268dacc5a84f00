// trace.cuh -- the persistent traversal kernel template (see trace.cu for the design notes).
#pragma once
#include "scene.h"
#include "traverse.h"

namespace mtsb {

#ifndef MTSB_TRACE_BLOCK
#define MTSB_TRACE_BLOCK 128
#endif
#ifndef MTSB_TRACE_MIN_CTAS
#define MTSB_TRACE_MIN_CTAS 8     /* resident CTAs per SM the register allocation is held to (64 registers) */
#endif
#ifndef MTSB_SM_STACK
#define MTSB_SM_STACK 8        /* stack entries per thread kept in shared memory   */
#endif
#define MTSB_LOCAL_STACK 56    /* overflow entries per thread in local memory      */
#define MTSB_REFILL 26         /* refill a warp when fewer lanes than this are busy */

// Shared-memory short stack with local-memory overflow.
struct SmemStack {
    uint2 *sm;                       // this thread's column: sm[i * MTSB_TRACE_BLOCK]
    uint2 *loc;                      // overflow array (kept outside the struct so the scalar state stays in registers)
    const uint8_t *perm_tab;         // shared-memory table [8][256]: perm_byte(mask, octant)
    int sp;
    __device__ __forceinline__ uint32_t perm(uint32_t m, uint32_t o) const { return perm_tab[o * 256u + m]; }
    __device__ __forceinline__ void reset() { sp = 0; }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
    __device__ __forceinline__ void push(uint32_t x, uint32_t y) {
        if (sp < MTSB_SM_STACK) sm[sp * MTSB_TRACE_BLOCK] = make_uint2(x, y);
        else if (sp < MTSB_SM_STACK + MTSB_LOCAL_STACK) loc[sp - MTSB_SM_STACK] = make_uint2(x, y);
        ++sp;
    }
    __device__ __forceinline__ void pop(uint32_t &x, uint32_t &y) {
        --sp;
        uint2 v = sp < MTSB_SM_STACK ? sm[sp * MTSB_TRACE_BLOCK] : loc[(sp < MTSB_SM_STACK + MTSB_LOCAL_STACK ? sp : MTSB_SM_STACK + MTSB_LOCAL_STACK - 1) - MTSB_SM_STACK];
        x = v.x; y = v.y;
    }
};

struct TraceArgs {
    const U4 *nodes, *tris; uint32_t n_nodes;
    float box_lo[3], box_hi[3];
    const RayIn *rays;
    const uint32_t *d_count;     // number of rays if non-null (device), else n_max
    uint32_t n_max;
    uint32_t *work_counter;      // zeroed before the launch
    int refill;                  // refill a warp when fewer lanes than this are busy
    uint32_t prmt_bias;          // 0x47000000, passed as data so it stays a register operand of PRMT (the selector is then the immediate)
};

template <bool ANY, class Epilogue>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TRACE_MIN_CTAS) k_trace(TraceArgs a, Epilogue ep) {
    __shared__ uint2 s_stack[MTSB_SM_STACK * MTSB_TRACE_BLOCK];
    __shared__ uint8_t s_perm[8 * 256];
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += MTSB_TRACE_BLOCK) s_perm[i] = (uint8_t) perm_byte(i & 255u, i >> 8);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n = a.d_count ? min(*a.d_count, a.n_max) : a.n_max;
    uint2 overflow[MTSB_LOCAL_STACK];
    Traversal<SmemStack> tr;
    tr.stack.sm = s_stack + threadIdx.x;
    tr.stack.loc = overflow;
    tr.stack.perm_tab = s_perm;
    tr.stack.sp = 0;
    tr.bias = a.prmt_bias;
    bool active = false;
    bool exhausted = false;
    uint32_t ray_idx = 0;
    RayIn ray;
    while (true) {
        // ---- refill idle lanes ----
        const unsigned idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (idle && !exhausted) {
            const int n_idle = __popc(idle), leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int) lane == leader) base = atomicAdd(a.work_counter, (uint32_t) n_idle);
            base = __shfl_sync(0xFFFFFFFFu, base, leader);
            if (base >= n) exhausted = true;          // uniform across the warp
            if (!active) {
                const uint32_t idx = base + (uint32_t) __popc(idle & ((1u << lane) - 1u));
                if (idx < n) {
                    const float4 r0 = __ldg(reinterpret_cast<const float4 *>(a.rays) + 2 * (size_t) idx);
                    const float4 r1 = __ldg(reinterpret_cast<const float4 *>(a.rays) + 2 * (size_t) idx + 1);
                    ray.ox = r0.x; ray.oy = r0.y; ray.oz = r0.z; ray.mint = r0.w;
                    ray.dx = r1.x; ray.dy = r1.y; ray.dz = r1.z; ray.maxt = r1.w;
                    ray_idx = idx;
                    float mint, maxt;
                    if (a.n_nodes != 0 && ray_interval(ray, a.box_lo, a.box_hi, ANY, mint, maxt)) {
                        tr.begin(ray, mint, maxt);
                        active = true;
                    } else {
                        HitOut miss; miss.prim = MTSB_NO_HIT_ID; miss.u = 0; miss.v = 0; miss.t = MTSB_INF;
                        ep(idx, ray, miss);
                    }
                }
            }
        }
        if (!__any_sync(0xFFFFFFFFu, active)) {
            if (exhausted) break;
            continue;
        }
        // ---- lock-step traversal until the warp has drained below the refill threshold ----
        // Each iteration: (a) lanes without pending triangles visit one node, (b) lanes with
        // pending triangles test one, (c) pop / finish.  The three regions are warp-converged.
        unsigned live;
        do {
            if (active && tr.wants_node()) tr.visit_node(a.nodes);
            __syncwarp();
            bool done = false;
            if (active && tr.tw != 0) done = tr.template test_triangle<ANY>(a.tris);
            __syncwarp();
            if (active && (done || tr.finished())) {
                ep(ray_idx, ray, tr.result(a.tris));
                active = false;
            }
            live = __ballot_sync(0xFFFFFFFFu, active);
        } while (live != 0 && (exhausted || __popc(live) >= a.refill));
    }
}

struct StoreHit {
    HitOut *hits;
    __device__ __forceinline__ void operator()(uint32_t i, const RayIn &, const HitOut &h) const {
        reinterpret_cast<uint4 *>(hits)[i] = make_uint4(h.prim, __float_as_uint(h.u), __float_as_uint(h.v), __float_as_uint(h.t));
    }
};

uint32_t *trace_work_counter(Scene *s, cudaStream_t st);
int device_sm_count();   // api.cu: zeroed counter for the next launch
int trace_grid(Scene *s, const void *kernel);

template <bool ANY, class Epilogue>
int launch_trace_kernel(Scene *s, const RayIn *d_rays, const uint32_t *d_count, uint32_t n_max, Epilogue ep, cudaStream_t st) {
    if (n_max == 0) return MTSB_OK;
    TraceArgs a;
    a.nodes = s->v.nodes; a.tris = s->v.tris; a.n_nodes = s->v.n_nodes;
    for (int k = 0; k < 3; ++k) { a.box_lo[k] = s->v.box_lo[k]; a.box_hi[k] = s->v.box_hi[k]; }
    a.rays = d_rays; a.d_count = d_count; a.n_max = n_max;
    { static int r = 0; if (!r) { const char *e = getenv("MTSB_REFILL"); r = e ? atoi(e) : MTSB_REFILL; } a.refill = r; }
    a.prmt_bias = 0x47000000u;
    a.work_counter = trace_work_counter(s, st);
    if (!a.work_counter) return MTSB_ERR_CUDA;
    int grid = trace_grid(s, (const void *) k_trace<ANY, Epilogue>);
    if (grid <= 0) return MTSB_ERR_CUDA;
    const uint32_t need = (n_max + MTSB_TRACE_BLOCK - 1) / MTSB_TRACE_BLOCK;
    if (!d_count && need < (uint32_t) grid) grid = (int) need;
    k_trace<ANY, Epilogue><<<grid, MTSB_TRACE_BLOCK, 0, st>>>(a, ep);
    s->launches++;
    MTSB_CUDA(cudaGetLastError());
    return MTSB_OK;
}

int launch_trace(Scene *s, const RayIn *d_rays, const uint32_t *d_count, uint32_t n_max, int mode, HitOut *d_hits, cudaStream_t st);

}  // namespace mtsb
