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

// Pushes that found the stack full (MTSB_SM_STACK + MTSB_LOCAL_STACK = 64 entries; one entry per level of the wide tree at
// most, so only a pathologically deep tree gets here).  A dropped entry would silently lose a subtree, so the entry points
// read this counter after their launches and fail with MTSB_ERR_CUDA.  One copy per translation unit (no -rdc).
static __device__ unsigned int g_stack_overflow;

// ---- checked build (-DMTSB_CHECKED) ------------------------------------------------------------------------------------
// compute-sanitizer is closed on the pool this library is developed on, so the bounds a memcheck run would watch are
// checked by the kernels themselves in a checked build: every queue write against the capacity of the path buffers, every
// node / triangle-record fetch against the array sizes, the warp queue against its ring, film writes against the film.
// A failed check bumps a counter (it does not stop the kernel); mtsb_render / mtsb_trace return MTSB_ERR_CUDA when any is
// non-zero.  The shipped build compiles the checks away.  tools/checked_run.sh runs the test cases under the checked build.
enum { CHK_QUEUE = 0, CHK_NODE, CHK_TRI, CHK_WARPQ, CHK_FILM, CHK_SLOT, CHK_N };
#ifdef MTSB_CHECKED
static __device__ unsigned int g_check_fail[CHK_N];
#define MTSB_CHECK(cond, code) do { if (!(cond)) atomicAdd(&g_check_fail[code], 1u); } while (0)
#else
#define MTSB_CHECK(cond, code) do { } while (0)
#endif

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
        else atomicAdd(&g_stack_overflow, 1u);
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
    uint32_t n_tri_records;      // checked build: number of 48-B triangle records
    float box_lo[3], box_hi[3];
    const RayIn *rays;
    const uint32_t *d_count;     // number of rays if non-null (device), else n_max
    uint32_t n_max;
    uint32_t *work_counter;      // zeroed before the launch
    int refill;                  // refill a warp when fewer lanes than this are busy
    uint32_t prmt_bias;          // 0x47000000, passed as data so it stays a register operand of PRMT (the selector is then the immediate)
};

// ---- warp-cooperative triangle phase ---------------------------------------------------------------------------------
// In a lock-step loop every lane owns one ray.  Node visits keep ~25 of 32 lanes busy, but "each lane tests one of ITS
// pending triangles" ran with under 9 lanes on average while costing a quarter of all issued instructions (ncu source
// counters of the round-1 kernel, profiles/r02_trace_regions.md).  So triangles are no longer tested by the lane that owns
// the ray: every lane hands the triangles of its next hit leaf to a per-warp queue in shared memory, and when 32 tests
// are queued (or nothing else can make progress) the warp runs ONE triangle phase in which lane j tests queue entry j
// -- any ray's triangle -- against that ray's origin / direction / interval, which every lane keeps in shared memory.
// Hits go back to the owning lane through a 64-bit shared-memory minimum over (t, record, tester lane), so a ray that is
// hit by several triangles of one phase keeps the nearest, exactly as the sequential order would; the tester leaves the
// hit record (u, v, t, prim) in shared memory, where the owner picks it up.  The arithmetic per
// triangle is unchanged (tri_test with the owner's current [mint, maxt]).
#define MTSB_TQ_CAP 128            /* queue entries per warp: < 32 are left after the triangle phases, <= 96 are added per iteration */
#ifndef MTSB_TQ_FLUSH
#define MTSB_TQ_FLUSH 24           /* run a triangle phase when this many tests are queued (<= 32; 24 measured best)                                */
#endif
#ifndef MTSB_TQ_FLUSH_ANY
#define MTSB_TQ_FLUSH_ANY MTSB_TQ_FLUSH   /* the same threshold for any-hit queries (8 / 16 measured: no better)                    */
#endif
#define MTSB_TQ_OWNER_SHIFT 27     /* entry = owner lane << 27 | triangle record index (2^27 records = 134 M triangles)           */

struct WarpShared {
    uint32_t tq[MTSB_TQ_CAP];
    float4 ro[32], rd[32];         // the lanes' rays: (origin, mint), (direction, maxt); maxt follows the accepted hits
    unsigned long long best[32];   // per owner: min over (ordered t << 32 | record << 5 | tester lane) of the hits of one triangle phase
    float4 cand[32];               // per tester lane: (u, v, t, prim) of the hit it found in the current phase
    float4 hit[32];                // per owner: record of the accepted hit (copied from the winning tester's cand)
};

__device__ __forceinline__ uint32_t order_float(float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float unorder_float(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

template <bool ANY, class Epilogue>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TRACE_MIN_CTAS) k_trace(TraceArgs a, Epilogue ep) {
    __shared__ uint2 s_stack[MTSB_SM_STACK * MTSB_TRACE_BLOCK];
    __shared__ uint8_t s_perm[8 * 256];
    __shared__ WarpShared s_warp[MTSB_TRACE_BLOCK / 32];
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += MTSB_TRACE_BLOCK) s_perm[i] = (uint8_t) perm_byte(i & 255u, i >> 8);
    const uint32_t lane = threadIdx.x & 31;
    WarpShared &ws = s_warp[threadIdx.x >> 5];
    ws.best[lane] = ~0ull;
    __syncthreads();
    const uint32_t n = a.d_count ? min(*a.d_count, a.n_max) : a.n_max;
    uint2 overflow[MTSB_LOCAL_STACK];
    Traversal<SmemStack> tr;
    tr.stack.sm = s_stack + threadIdx.x;
    tr.stack.loc = overflow;
    tr.stack.perm_tab = s_perm;
    tr.stack.sp = 0;
    tr.bias = a.prmt_bias;
    tr.tw = 0; tr.tw2 = 0; tr.ng_y = 0;
    bool active = false;
    bool exhausted = false;
    uint32_t ray_idx = 0;
    uint32_t qhead = 0, qtail = 0;        // warp-uniform, monotonic: entries [qhead, qtail) of the ring are pending
    uint32_t my_last = 0;                 // queue position behind this lane's last entry: all its tests are done when qhead has passed it
    const uint32_t lt = (1u << lane) - 1u;
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
                const uint32_t idx = base + (uint32_t) __popc(idle & lt);
                if (idx < n) {
                    const float4 r0 = __ldg(reinterpret_cast<const float4 *>(a.rays) + 2 * (size_t) idx);
                    const float4 r1 = __ldg(reinterpret_cast<const float4 *>(a.rays) + 2 * (size_t) idx + 1);
                    RayIn ray;
                    ray.ox = r0.x; ray.oy = r0.y; ray.oz = r0.z; ray.mint = r0.w;
                    ray.dx = r1.x; ray.dy = r1.y; ray.dz = r1.z; ray.maxt = r1.w;
                    ray_idx = idx;
                    float mint, maxt;
                    if (a.n_nodes != 0 && ray_interval(ray, a.box_lo, a.box_hi, ANY, mint, maxt)) {
                        tr.begin(ray, mint, maxt);
                        ws.ro[lane] = make_float4(ray.ox, ray.oy, ray.oz, mint);
                        ws.rd[lane] = make_float4(ray.dx, ray.dy, ray.dz, maxt);
                        my_last = qhead;
                        active = true;
                    } else {
                        HitOut miss; miss.prim = MTSB_NO_HIT_ID; miss.u = 0; miss.v = 0; miss.t = MTSB_INF;
                        ep(idx, miss);
                    }
                }
            }
        }
        if (!__any_sync(0xFFFFFFFFu, active)) {
            if (exhausted) break;
            continue;
        }
        // ---- lock-step traversal until the warp has drained below the refill threshold ----
        unsigned live;
        do {
            // (a) node phase: lanes with node work and a free batch slot visit one node
#ifdef MTSB_CHECKED
            if (active && tr.wants_node()) {     // the node the visit is about to fetch
                uint32_t gx = tr.ng_x, gy = tr.ng_y;
                if (gy <= 0x00FFFFFFu && tr.stack.sp > 0) { const int sp = tr.stack.sp - 1; const uint2 v = sp < MTSB_SM_STACK ? tr.stack.sm[sp * MTSB_TRACE_BLOCK] : tr.stack.loc[min(sp, MTSB_SM_STACK + MTSB_LOCAL_STACK - 1) - MTSB_SM_STACK]; gx = v.x; gy = v.y; }
                const uint32_t slot = ((uint32_t) bfind(gy) - 24u) ^ tr.octinv;
                MTSB_CHECK(gx + (uint32_t) __popc(gy & ~(0xFFFFFFFFu << slot) & 0xFFu) < a.n_nodes, CHK_NODE);
            }
#endif
            if (active && tr.wants_node()) tr.visit_node(a.nodes);
            __syncwarp();
            // (b) every lane with a pending batch hands the triangles of ONE hit leaf (1..3) to the warp's queue
            uint32_t rec0 = 0, cnt = 0;
            const bool have_leaf = active && tr.tw != 0;
            if (__any_sync(0xFFFFFFFFu, have_leaf)) {
            if (have_leaf) tr.next_leaf(rec0, cnt);
            const unsigned b1 = __ballot_sync(0xFFFFFFFFu, cnt & 1u), b2 = __ballot_sync(0xFFFFFFFFu, cnt & 2u);
            if (cnt) {
                const uint32_t pos = qtail + (uint32_t) __popc(b1 & lt) + 2u * (uint32_t) __popc(b2 & lt);
                const uint32_t e = (lane << MTSB_TQ_OWNER_SHIFT) | rec0;
                ws.tq[pos & (MTSB_TQ_CAP - 1)] = e;
                if (cnt > 1) ws.tq[(pos + 1) & (MTSB_TQ_CAP - 1)] = e + 1;
                if (cnt > 2) ws.tq[(pos + 2) & (MTSB_TQ_CAP - 1)] = e + 2;
                my_last = pos + cnt;
            }
            qtail += (uint32_t) __popc(b1) + 2u * (uint32_t) __popc(b2);
            MTSB_CHECK(qtail - qhead <= (uint32_t) MTSB_TQ_CAP, CHK_WARPQ);
            }
            // (c) triangle phase: when a full warp of tests is queued, or when nothing else can make progress
            const bool other_work = __any_sync(0xFFFFFFFFu, active && (tr.tw != 0 || tr.wants_node()));
            while (qtail - qhead >= (uint32_t) (ANY ? MTSB_TQ_FLUSH_ANY : MTSB_TQ_FLUSH) || (qtail != qhead && !other_work)) {
                __syncwarp();
                const uint32_t nq = min(32u, qtail - qhead);
                if (lane < nq) {
                    const uint32_t e = ws.tq[(qhead + lane) & (MTSB_TQ_CAP - 1)];
                    const uint32_t owner = e >> MTSB_TQ_OWNER_SHIFT, rec = e & ((1u << MTSB_TQ_OWNER_SHIFT) - 1u);
                    const float4 ro = ws.ro[owner], rd = ws.rd[owner];
                    MTSB_CHECK(rec < a.n_tri_records, CHK_TRI);
                    const U4 *tp = a.tris + 3 * (size_t) rec;
                    const U4 q0 = ld16(tp), q1 = ld16(tp + 1), q2 = ld16(tp + 2);
                    float u, v, t;
                    if (tri_test(q0, q1, q2, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, ro.w, rd.w, u, v, t)) {
                        ws.cand[lane] = make_float4(u, v, t, __uint_as_float(q2.z));
                        atomicMin(&ws.best[owner], ((unsigned long long) order_float(t) << 32) | ((unsigned long long) rec << 5) | lane);
                    }
                }
                qhead += nq;
                __syncwarp();
                const unsigned long long b = ws.best[lane];
                if (b != ~0ull) {                       // this lane's ray was hit in the phase: the nearest hit is the new interval end
                    ws.best[lane] = ~0ull;
                    const float4 c = ws.cand[(uint32_t) b & 31u];                          // the winning tester's record (bit-identical to tri_eval's)
                    ws.hit[lane] = c;
                    tr.hit_rec = 0;                                                         // "has a hit"
                    if (ANY) { tr.ng_y = 0; tr.stack.reset(); tr.tw = 0; tr.tw2 = 0; }      // satisfied: only its queued tests remain to drain
                    else { tr.tmax = c.z; ws.rd[lane].w = c.z; }
                }
            }
            // (d) a ray is finished when it has no nodes, no leaves to hand out and no test left in the queue
            if (active && tr.tw == 0 && !tr.has_nodes() && (int32_t) (qhead - my_last) >= 0) {
                HitOut h; h.prim = MTSB_NO_HIT_ID; h.u = 0; h.v = 0; h.t = MTSB_INF;
                if (tr.hit_rec != MTSB_NO_HIT_ID) { const float4 c = ws.hit[lane]; h.u = c.x; h.v = c.y; h.t = c.z; h.prim = __float_as_uint(c.w); }
                ep(ray_idx, h);
                active = false;
            }
            live = __ballot_sync(0xFFFFFFFFu, active);
        } while (live != 0 && (exhausted || __popc(live) >= a.refill));
    }
}

struct StoreHit {
    HitOut *hits;
    __device__ __forceinline__ void operator()(uint32_t i, const HitOut &h) const {
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
    a.nodes = s->v.nodes; a.tris = s->v.tris; a.n_nodes = s->v.n_nodes; a.n_tri_records = (uint32_t) s->info.n_tris;
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

// reads and clears this translation unit's overflow counter (call after the stream that ran k_trace has been synchronised)
static inline int check_stack_overflow(const char *who) {
    unsigned int n = 0;
    MTSB_CUDA(cudaMemcpyFromSymbol(&n, g_stack_overflow, sizeof(n)));
    if (n) {
        const unsigned int zero = 0;
        cudaMemcpyToSymbol(g_stack_overflow, &zero, sizeof(zero));
        set_error(std::string(who) + ": traversal stack overflow (the BVH is deeper than 64 levels); results would be incomplete");
        return MTSB_ERR_CUDA;
    }
#ifdef MTSB_CHECKED
    unsigned int f[CHK_N];
    MTSB_CUDA(cudaMemcpyFromSymbol(f, g_check_fail, sizeof(f)));
    unsigned int tot = 0;
    for (int i = 0; i < CHK_N; ++i) tot += f[i];
    if (tot) {
        char buf[256];
        snprintf(buf, sizeof(buf), "%s: checked build: %u out-of-range queue writes, %u node fetches, %u triangle fetches, %u warp-queue overruns, %u film writes, %u slot indices",
                 who, f[CHK_QUEUE], f[CHK_NODE], f[CHK_TRI], f[CHK_WARPQ], f[CHK_FILM], f[CHK_SLOT]);
        const unsigned int zero[CHK_N] = {0};
        cudaMemcpyToSymbol(g_check_fail, zero, sizeof(zero));
        set_error(buf);
        return MTSB_ERR_CUDA;
    }
#endif
    return MTSB_OK;
}

int launch_trace(Scene *s, const RayIn *d_rays, const uint32_t *d_count, uint32_t n_max, int mode, HitOut *d_hits, cudaStream_t st);

}  // namespace mtsb
