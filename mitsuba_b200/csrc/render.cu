// render.cu -- wavefront formulation of MonteCarloIntegrator::renderBlock + MIPathTracer::Li
// (reference: src/librender/integrator.cpp:196-244, src/integrators/path/path.cpp:121-296).
//
// One path sample = one slot of the SoA path state.  A pass renders `spp_pass` samples of every
// pixel of a set of 32x32 tiles.  Per bounce the stages are separate kernels over queues:
//
//   k_trace<closest>   rays[cur]            -> hits                      (trace.cuh)
//   k_sort             hits                 -> one queue per BSDF model; rays that left the scene
//                                              get their environment term here and are finished
//   k_shade<MODEL>     queue[MODEL]         -> emitter hit + MIS, Russian roulette, emitter
//                                              sampling + BSDF eval (-> shadow queue),
//                                              BSDF sampling (-> rays[1-cur])
//   k_trace<any>       shadow queue         -> adds the pending contribution when unoccluded
//
// and at the end of the pass k_splat accumulates every tile in shared memory and adds it to
// the film (ImageBlock::put + the overlap-add of BlockedRenderProcess::processResult).
//
// Every kernel takes its work size from device counters (a ring of per-bounce sets, render.cuh), so the
// host enqueues the bounces of a pass back to back without waiting for the device; it only polls the
// queue sizes, asynchronously, to stop enqueueing once a pass has run out of rays.
//
// Loop restructuring with respect to path.cpp: the reference traces the BSDF ray inside
// iteration k and then finishes iteration k (emitter hit with MIS :259-266, break on miss
// :274, Russian roulette :278-288).  Here that tail runs at the head of the shading stage of
// iteration k+1, in the same order, so every sample sees the same sequence of decisions.
#include "render.cuh"
#include "trace.cuh"
#include "shading.h"
#include "sensor.h"
#include "rng.h"
#include <cooperative_groups.h>
#include <cstring>
#include <algorithm>
#include <map>
#include <mutex>

namespace cg = cooperative_groups;

namespace mtsb {

#define MTSB_TILE 32
// checked build: a film write must land inside the bordered film
#define MTSB_CHECK_FILM(sc, X, Y) MTSB_CHECK((X) >= 0 && (Y) >= 0 && (X) < (sc).film_w + 2 * (sc).film_border && (Y) < (sc).film_h + 2 * (sc).film_border, CHK_FILM)
#define MTSB_SHADE_BLOCK 128
#ifndef MTSB_SHADE_MIN_BLOCKS
#define MTSB_SHADE_MIN_BLOCKS 8      /* resident CTAs per SM the shading kernels' register allocation is held to (64 registers: the stage waits on
                                        dependent gathers, so resident warps pay more than the spills cost; 1 / 5 / 6 / 8 measured: 41.1 / 42.9 / 39.5 / 38.3 ms) */
#endif

__device__ __forceinline__ uint32_t queue_push(uint32_t *counter) {
    cg::coalesced_group g = cg::coalesced_threads();
    uint32_t base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(counter, g.size());
    return g.shfl(base, 0) + g.thread_rank();
}

__device__ __forceinline__ uint32_t compact1by1(uint32_t x) {
    x &= 0x55555555u; x = (x ^ (x >> 1)) & 0x33333333u; x = (x ^ (x >> 2)) & 0x0f0f0f0fu;
    x = (x ^ (x >> 4)) & 0x00ff00ffu; x = (x ^ (x >> 8)) & 0x0000ffffu; return x;
}

__device__ __forceinline__ void store_ray(RayIn *rays, uint32_t i, const RayIn &r) {
    reinterpret_cast<float4 *>(rays)[2 * (size_t) i] = make_float4(r.ox, r.oy, r.oz, r.mint);
    reinterpret_cast<float4 *>(rays)[2 * (size_t) i + 1] = make_float4(r.dx, r.dy, r.dz, r.maxt);
}
__device__ __forceinline__ RayIn load_ray(const RayIn *rays, uint32_t i) {
    const float4 a = reinterpret_cast<const float4 *>(rays)[2 * (size_t) i], b = reinterpret_cast<const float4 *>(rays)[2 * (size_t) i + 1];
    RayIn r; r.ox = a.x; r.oy = a.y; r.oz = a.z; r.mint = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.maxt = b.w; return r;
}

// slot -> pixel of the pass (Morton order inside the tile keeps a warp on a compact pixel patch)
__device__ __forceinline__ bool slot_pixel(const SceneView &sc, const PathBuffers &pb, const PassInfo &ps, uint32_t slot,
                                           int &px, int &py, uint32_t &s_local) {
    const uint32_t per_tile = MTSB_TILE * MTSB_TILE * ps.spp_pass;
    const uint32_t tile_local = slot / per_tile, r = slot % per_tile;
    const uint32_t pix = r / ps.spp_pass; s_local = r % ps.spp_pass;
    const int tile = pb.tiles[ps.tile_first + tile_local];
    px = (tile % (int) ps.tiles_x) * MTSB_TILE + (int) compact1by1(pix);
    py = (tile / (int) ps.tiles_x) * MTSB_TILE + (int) compact1by1(pix >> 1);
    return px < sc.film_w && py < sc.film_h;
}
__device__ __forceinline__ SampleStream slot_stream(const SceneView &sc, const PassInfo &ps, int px, int py, uint32_t s_local) {
    const uint64_t pixel = (uint64_t) (py + sc.sensor.crop_y) * (uint64_t) sc.sensor.film_width + (uint64_t) (px + sc.sensor.crop_x);
    return SampleStream(ps.seed, (pixel << 24) | (uint64_t) (ps.s_first + s_local));
}

// ---- quasi-Monte Carlo draws (halton / ldsampler) -----------------------------------------------
// The reference samplers hand out dimensions in call order (halton.cpp:358-395, ldsampler.cpp:199-213); the
// per-path dimension counter lives in the flag word.  Kept out of line: the independent sampler (the
// benchmarked configuration) never calls them and should not pay registers for them.
struct QmcDraw { float a, b; uint32_t fl; };
__device__ __noinline__ QmcDraw qmc_draw(const SamplerView *svp, uint64_t seed, int px_abs, int py_abs, uint32_t s_global,
                                         uint32_t fl, int n, float ia, float ib) {
    const SamplerView sv = *svp;
    QmcDraw r; r.a = ia; r.b = ib; r.fl = fl;
    if (sv.type == MTSB_SAMPLER_HALTON) {
        int pm[2];
        const uint64_t index = halton_pixel_offset(sv, px_abs, py_abs, pm) + sv.stride * (uint64_t) s_global;
        uint32_t dim = fl >> PF_DIM_SHIFT;
        r.a = halton_value(sv, dim++, index);
        if (n == 2) {
            r.b = halton_value(sv, dim++, index);
            if (dim == 2) {          // halton.cpp:384-386: the first two dimensions address the pixel
                r.a = fsub(fmul(r.a, (float) sv.prime_pow[0]), (float) pm[0]);
                r.b = fsub(fmul(r.b, (float) sv.prime_pow[1]), (float) pm[1]);
            }
        }
        if (dim > 0xFFFu) dim = 0xFFFu;
        r.fl = (fl & ~PF_DIM_MASK) | (dim << PF_DIM_SHIFT);
    } else if (sv.type == MTSB_SAMPLER_SOBOL) {
        const uint64_t index = sobol_sample_index(sv, s_global, px_abs, py_abs);
        uint32_t dim = fl >> PF_DIM_SHIFT;
        r.a = sobol_value(sv, dim++, index);
        if (n == 2) {
            r.b = sobol_value(sv, dim++, index);
            if (dim == 2 && index != (uint64_t) s_global) {      // sobol.cpp:243-245: the first two dimensions address the pixel
                r.a = fsub(fmul(r.a, sv.sobol_res), (float) px_abs);
                r.b = fsub(fmul(r.b, sv.sobol_res), (float) py_abs);
            }
        }
        if (dim > 0xFFFu) dim = 0xFFFu;
        r.fl = (fl & ~PF_DIM_MASK) | (dim << PF_DIM_SHIFT);
    } else if (sv.type == MTSB_SAMPLER_LDSAMPLER) {
        const uint64_t pixel = ((uint64_t) (uint32_t) py_abs << 32) | (uint64_t) (uint32_t) px_abs;
        uint32_t d1 = (fl >> PF_DIM_SHIFT) & 63u, d2 = (fl >> (PF_DIM_SHIFT + 6)) & 63u;
        if (n == 1) {
            if (d1 < sv.ld_dimension) { r.a = ld_value_1d(sv, seed, pixel, d1, s_global); ++d1; }      // else: the independent number passed in
        } else {
            if (d2 < sv.ld_dimension) { float o[2]; ld_value_2d(sv, seed, pixel, d2, s_global, o); r.a = o[0]; r.b = o[1]; ++d2; }
        }
        r.fl = (fl & ~PF_DIM_MASK) | (d1 << PF_DIM_SHIFT) | (d2 << (PF_DIM_SHIFT + 6));
    }
    return r;
}

__global__ void k_pass_begin(PathBuffers pb, uint32_t n_slots) {
    Counters *c = pb.ctr;
    for (int r = threadIdx.x; r < MTSB_CTR_RING * 16; r += blockDim.x) reinterpret_cast<uint32_t *>(c->b)[r] = 0;
    __syncthreads();
    if (threadIdx.x == 0) c->b[0].n_rays = n_slots;          // k_raygen fills queue entry `slot` for every slot
}

// integrator.cpp:218-241: pixel offset from the sampler, sensor ray, new radiance query
template <bool QMC>
__global__ void k_raygen(SceneView sc, PathBuffers pb, PassInfo ps, uint32_t n_slots) {
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n_slots; slot += gridDim.x * blockDim.x) {
        int px, py; uint32_t s_local;
        const bool valid = slot_pixel(sc, pb, ps, slot, px, py, s_local);
        pb.ray_slot[0][slot] = slot;
        if (!valid) {
            // pixel of an edge tile outside the film: an empty-interval ray (reported as a miss, ignored by k_miss / k_splat).
            // The primary queue is indexed by slot, so no queue atomics are needed here.
            pb.refn_flags[slot] = make_float4(0, 0, 0, __uint_as_float(0u));
            RayIn dead; dead.ox = dead.oy = dead.oz = 0.0f; dead.dx = dead.dy = 0.0f; dead.dz = 1.0f; dead.mint = 0.0f; dead.maxt = -1.0f;
            store_ray(pb.rays[0], slot, dead);
            continue;
        }
        float u[4];
        slot_stream(sc, ps, px, py, s_local).block4(0, 0, u);
        uint32_t fl0 = 1u | PF_FIRST | PF_VALID;
        if (QMC) {
            const QmcDraw qd = qmc_draw(sc.sampler, ps.seed, px + sc.sensor.crop_x, py + sc.sensor.crop_y, ps.s_first + s_local, fl0, 2, u[0], u[1]);
            u[0] = qd.a; u[1] = qd.b; fl0 = qd.fl;
        }
        const float sx = (float) px + u[0], sy = (float) py + u[1];
        RayIn ray;
        sensor_ray(sc.sensor, sc.near_x, sc.near_y, sx, sy, ray);
        pb.spos[slot] = make_float2(sx, sy);
        pb.thr_eta[slot] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
        pb.L_bpdf[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        pb.refn_flags[slot] = make_float4(0, 0, 0, __uint_as_float(fl0));
        store_ray(pb.rays[0], slot, ray);
    }
}

// sums the shadow rays of the last bounce that was launched
__global__ void k_pass_end(PathBuffers pb, int last_bounce) {
    Counters *c = pb.ctr;
    const uint32_t n = c->b[last_bounce % MTSB_CTR_RING].n_shadow;
    c->shadow += n; if (n) c->launches_any++;
    c->shadow_bounce[min(last_bounce, MTSB_BOUNCE_STATS - 1)] += n;
}

struct RenderConst { int max_depth, rr_depth, strict_normals, hide_emitters; };

// a ray that left the scene: environment term (path.cpp:138-145 first segment; :237-246, 259-266 afterwards); returns the
// path length of the finished sample
__device__ __forceinline__ uint32_t finish_escaped(const SceneView &sc, const PathBuffers &pb, const RenderConst &rc, int cur, uint32_t q) {
    const uint32_t slot = pb.ray_slot[cur][q];
    float4 lb = pb.L_bpdf[slot];
    const float4 rf = pb.refn_flags[slot];
    const uint32_t fl = __float_as_uint(rf.w);
    if (!(fl & PF_VALID)) return 0;                                            // dead ray of an edge tile
    if (sc.env_emitter >= 0) {
        const float4 te = pb.thr_eta[slot];
        const V3 thr(te.x, te.y, te.z);
        RayIn ray;
        if (!(fl & PF_FIRST) || sc.envmap) ray = load_ray(pb.rays[cur], q);     // a constant environment needs no direction
        else { ray.ox = ray.oy = ray.oz = ray.dx = ray.dy = 0.0f; ray.dz = 1.0f; ray.mint = ray.maxt = 0.0f; }
        V3 value;
        if ((fl & PF_FIRST) && sc.envmap) {
            // envmap.cpp:380-410: a camera ray carries differentials, so the lookup is filtered (MIPMap::eval)
            V3 rxd, ryd;
            const float2 sp = pb.spos[slot];
            sensor_ray_differentials(sc.sensor, sc.near_x, sc.near_y, sp.x, sp.y, sc.integrator.sample_count, rxd, ryd);
            value = environment_eval_filtered(sc, V3(ray.dx, ray.dy, ray.dz), rxd, ryd);
        } else value = environment_eval(sc, V3(ray.dx, ray.dy, ray.dz));        // constant.cpp:242-244 / envmap.cpp:380-410 without differentials
        if (fl & PF_FIRST) {
            if (!rc.hide_emitters) { lb.x += thr.x * value.x; lb.y += thr.y * value.y; lb.z += thr.z * value.z; }
        } else {
            // fillDirectSamplingRecord (constant.cpp:246-260): the ray must start inside the bounding sphere
            float nearT, farT;
            const V3 o(ray.ox, ray.oy, ray.oz), d(ray.dx, ray.dy, ray.dz);
            if (bsphere_intersect(sc.bsphere_c, sc.bsphere_r, o, d, nearT, farT) && !(nearT > 0 || farT < 0)) {
                DRec dRec; dRec.ref = o; dRec.refN = V3(rf.x, rf.y, rf.z); dRec.p = o; dRec.n = V3(0.0f); dRec.d = d; dRec.dist = farT;
                dRec.measure = MTSB_MEASURE_SOLID_ANGLE; dRec.emitter = sc.env_emitter;
                const float w = mi_weight(lb.w, lb.w < 0 ? 0.0f : pdf_emitter_direct(sc, dRec));
                lb.x += thr.x * value.x * w; lb.y += thr.y * value.y * w; lb.z += thr.z * value.z * w;
            }
        }
    }
    lb.w = 0.0f;
    pb.L_bpdf[slot] = lb;
    return fl & PF_DEPTH_MASK;
}

// Bins the hits of bounce `bounce` by scattering model and finishes the rays that left the scene.  Queue space is
// reserved once per BLOCK and queue (shared-memory counts, one global atomic per non-empty queue): per-warp atomics
// on a single counter run at ~1 G/s on this part and were a visible share of the stage.  The queue of a triangle
// comes from a byte table (10 MB for 10 M triangles, L2 resident) instead of the chain prim -> shape -> material.
// Block 0 also does the bookkeeping of the counter ring: ray totals of this bounce, shadow-ray total of the
// previous one, and it clears the set that bounce + MTSB_CTR_RING / 2 will use.
// n_count: rays of this bounce that count as traced (the primary queue also holds the dead rays of edge tiles), 0 = all
#define MTSB_SORT_BLOCK 512
__global__ void __launch_bounds__(MTSB_SORT_BLOCK) k_sort(SceneView sc, PathBuffers pb, RenderConst rc, int cur, int bounce, uint32_t n_count) {
    __shared__ uint32_t s_cnt[Q_COUNT], s_base[Q_COUNT];
    BounceCtr &bc = pb.ctr->b[bounce % MTSB_CTR_RING];
    const uint32_t n = bc.n_rays;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        Counters *c = pb.ctr;
        const uint32_t nc = n_count ? n_count : n;
        c->closest += nc; c->rays_bounce[min(bounce, MTSB_BOUNCE_STATS - 1)] += nc;
        if (n) c->launches_closest++;
        if (bounce > 0) {
            const uint32_t ns = c->b[(bounce - 1) % MTSB_CTR_RING].n_shadow;
            c->shadow += ns; c->shadow_bounce[min(bounce - 1, MTSB_BOUNCE_STATS - 1)] += ns;
            if (ns) c->launches_any++;
        }
        uint32_t *z = reinterpret_cast<uint32_t *>(&c->b[(bounce + MTSB_CTR_RING / 2) % MTSB_CTR_RING]);
        for (int i = 0; i < 16; ++i) z[i] = 0;
    }
    unsigned long long path_len = 0;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const uint32_t q = base + threadIdx.x;
        if (threadIdx.x < Q_COUNT) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        int which = -1;
        if (q < n) {
            const uint32_t prim = pb.hits[q].prim;
            if (prim != MTSB_NO_HIT_ID) which = (int) sc.tri_queue[prim];
            else path_len += finish_escaped(sc, pb, rc, cur, q);
        }
        // rank inside the block: warp-aggregated shared-memory atomic
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, which);
        const int leader = __ffs(peers) - 1;
        uint32_t wbase = 0;
        if (which >= 0 && (int) (threadIdx.x & 31) == leader) wbase = atomicAdd(&s_cnt[which], (uint32_t) __popc(peers));
        wbase = __shfl_sync(0xFFFFFFFFu, wbase, leader);
        const uint32_t rank = wbase + (uint32_t) __popc(peers & ((1u << (threadIdx.x & 31)) - 1u));
        __syncthreads();
        if (threadIdx.x < Q_COUNT) s_base[threadIdx.x] = s_cnt[threadIdx.x] ? atomicAdd(&bc.n_q[threadIdx.x], s_cnt[threadIdx.x]) : 0u;
        __syncthreads();
        MTSB_CHECK(which < 0 || s_base[which] + rank < pb.capacity, CHK_QUEUE);
        if (which >= 0) pb.q[which][s_base[which] + rank] = q;
    }
    for (int o = 16; o; o >>= 1) path_len += __shfl_xor_sync(0xFFFFFFFFu, path_len, o);
    if ((threadIdx.x & 31) == 0 && path_len) atomicAdd(&pb.ctr->path_len, path_len);
}

// scene creation: the packed shading record and the queue byte of every triangle (shading.h, render.cuh)
__global__ void k_pack_tris(SceneView sc, uint32_t n_tris, const uint8_t *shape_queue, U4 *out, uint8_t *tri_queue) {
    const uint32_t prim = blockIdx.x * blockDim.x + threadIdx.x;
    if (prim >= n_tris) return;
    const uint32_t s = sc.tri_shape[prim];
    const MtsbShape sh = sc.shapes[s];
    const uint32_t *ix = sc.indices + 3 * (size_t) prim;
    const uint32_t i0 = ix[0], i1 = ix[1], i2 = ix[2];
    const bool hn = sh.has_normals && sc.normals, ht = sh.has_texcoords && sc.texcoords;
    float r[32];
    for (int k = 0; k < 32; ++k) r[k] = 0.0f;
    const uint32_t iv[3] = {i0, i1, i2};
    for (int v = 0; v < 3; ++v) {
        for (int k = 0; k < 3; ++k) r[4 * v + k] = sc.positions[3 * (size_t) iv[v] + k];
        if (hn) for (int k = 0; k < 3; ++k) r[12 + 4 * v + k] = sc.normals[3 * (size_t) iv[v] + k];
    }
    if (ht) {
        const float *t0 = sc.texcoords + 2 * (size_t) i0, *t1 = sc.texcoords + 2 * (size_t) i1, *t2 = sc.texcoords + 2 * (size_t) i2;
        r[15] = t0[0]; r[19] = t0[1]; r[23] = t1[0]; r[24] = t1[1]; r[25] = t2[0]; r[26] = t2[1];
    }
    r[3] = __uint_as_float(s);
    r[7] = __uint_as_float((hn ? 1u : 0u) | (ht ? 2u : 0u));
    float4 *o = reinterpret_cast<float4 *>(out + (size_t) MTSB_SHADE_TRI_U4 * prim);
    for (int k = 0; k < 8; ++k) o[k] = make_float4(r[4 * k], r[4 * k + 1], r[4 * k + 2], r[4 * k + 3]);
    tri_queue[prim] = shape_queue[s];
}

template <int Q> __device__ __forceinline__ V3 eval_q(const MtsbMaterial *mats, int id, V3 wi, V3 wo) {
    if (Q == Q_GENERIC || Q == Q_TEXTURED) return bsdf_eval(mats, id, wi, wo);
    const MtsbMaterial &m = resolve_twosided_eval(mats, mats[id], wi, wo);
    if (Q == Q_DIFFUSE) return diffuse_eval(m, wi, wo);
    if (Q == Q_CONDUCTOR) return roughconductor_eval(m, wi, wo);
    if (Q == Q_DIELECTRIC) return roughdielectric_eval(m, wi, wo);
    return V3(0.0f);
}
template <int Q> __device__ __forceinline__ float pdf_q(const MtsbMaterial *mats, int id, V3 wi, V3 wo) {
    if (Q == Q_GENERIC || Q == Q_TEXTURED) return bsdf_pdf(mats, id, wi, wo);
    const MtsbMaterial &m = resolve_twosided_eval(mats, mats[id], wi, wo);
    if (Q == Q_DIFFUSE) return diffuse_pdf(wi, wo);
    if (Q == Q_CONDUCTOR) return roughconductor_pdf(m, wi, wo);
    if (Q == Q_DIELECTRIC) return roughdielectric_pdf(m, wi, wo);
    return 0.0f;
}
template <int Q> __device__ __forceinline__ bool sample_q(const MtsbMaterial *mats, int id, V3 wi, P2 u, float u1, BsdfSample &s) {
    if (Q == Q_GENERIC || Q == Q_TEXTURED) return bsdf_sample(mats, id, wi, u, u1, s);
    const MtsbMaterial *m = &mats[id];
    bool flip = false;
    if ((m->flags & MTSB_MAT_TWOSIDED) && cosTheta(wi) < 0) { wi.z *= -1; flip = true; if (m->back >= 0) m = &mats[m->back]; }
    bool ok;
    if (Q == Q_DIFFUSE) ok = diffuse_sample(*m, wi, u, s);
    else if (Q == Q_CONDUCTOR) ok = roughconductor_sample(*m, wi, u, s);
    else if (Q == Q_DIELECTRIC) ok = roughdielectric_sample(*m, wi, u, u1, s);
    else if (Q == Q_DELTA) ok = m->type == MTSB_BSDF_DIELECTRIC ? dielectric_sample(*m, wi, u, s) : conductor_sample(*m, wi, s);
    else ok = false;
    if (flip && ok && !isZero(s.weight) && s.pdf != 0) s.wo.z *= -1;
    return ok;
}

// Q_TEXTURED: the record(s) a hit shades with -- the shape's material (and the back side of a twosided adapter) with every
// textured spectrum replaced by the texture's value at the hit (diffuse.cpp:118 m_reflectance->eval(bRec.its), ...).
// Camera-ray hits on a BSDF that usesRayDifferentials() (a bitmap texture: bitmap.cpp:547-549) carry uv partials
// (records.inl:82-90, intersection.cpp:5-98); every later hit is looked up unfiltered.
__device__ __noinline__ void resolve_textured(const SceneView &sc, int mat, const Its &its, bool first, float2 spos, V3 cam, MtsbMaterial *lm) {
    lm[0] = sc.materials[mat];
    int n = 1;
    if ((lm[0].flags & MTSB_MAT_TWOSIDED) && lm[0].back >= 0) { lm[1] = sc.materials[lm[0].back]; lm[0].back = 1; n = 2; }
    UVPartials pt; pt.has = false; pt.dudx = pt.dvdx = pt.dudy = pt.dvdy = 0.0f;
    if (first) {
        bool bitmap = false;
        for (int i = 0; i < n; ++i) {
            if (lm[i].tex_reflectance) bitmap |= sc.textures[lm[i].tex_reflectance - 1].type == MTSB_TEX_BITMAP;
            if (lm[i].tex_transmittance) bitmap |= sc.textures[lm[i].tex_transmittance - 1].type == MTSB_TEX_BITMAP;
            if (lm[i].tex_alpha) bitmap |= sc.textures[lm[i].tex_alpha - 1].type == MTSB_TEX_BITMAP;
        }
        if (bitmap) {
            V3 rxd, ryd;
            sensor_ray_differentials(sc.sensor, sc.near_x, sc.near_y, spos.x, spos.y, sc.integrator.sample_count, rxd, ryd);
            pt = compute_partials(its.p, its.sh.n, its.dpdu, its.dpdv, cam, rxd, ryd);
        }
    }
    for (int i = 0; i < n; ++i) {
        if (lm[i].tex_reflectance) {
            const V3 c = texture_eval(sc.textures[lm[i].tex_reflectance - 1], its.uv, pt);
            lm[i].reflectance[0] = c.x; lm[i].reflectance[1] = c.y; lm[i].reflectance[2] = c.z;
        }
        if (lm[i].tex_transmittance) {
            const V3 c = texture_eval(sc.textures[lm[i].tex_transmittance - 1], its.uv, pt);
            lm[i].transmittance[0] = c.x; lm[i].transmittance[1] = c.y; lm[i].transmittance[2] = c.z;
        }
        if (lm[i].tex_alpha) {
            // roughconductor.cpp:379-384 / roughdielectric.cpp:300-305: MicrofacetDistribution(type, m_alphaU->eval(its).average(),
            // m_alphaV->eval(its).average()), whose constructor clamps both to >= 1e-4 (microfacet.h:135-136)
            const V3 c = texture_eval(sc.textures[lm[i].tex_alpha - 1], its.uv, pt);
            const float a = fmaxf((c.x + c.y + c.z) * (1.0f / 3.0f), 1e-4f);
            lm[i].alpha_u = a; lm[i].alpha_v = a;
        }
    }
}

template <int Q, bool QMC>
__global__ void __launch_bounds__(MTSB_SHADE_BLOCK, MTSB_SHADE_MIN_BLOCKS) k_shade(SceneView sc, PathBuffers pb, PassInfo ps, RenderConst rc, int cur, int bounce) {
    // Outputs (shadow ray + pending contribution, continuation ray) are staged in shared memory and queue space is
    // reserved once per block and queue at the end of the iteration: one global atomic per block instead of one per warp.
    __shared__ float4 s_sray[2 * MTSB_SHADE_BLOCK], s_contrib[MTSB_SHADE_BLOCK], s_nray[2 * MTSB_SHADE_BLOCK];
    __shared__ uint32_t s_wcnt[2][MTSB_SHADE_BLOCK / 32], s_qbase[2];
    BounceCtr &bc = pb.ctr->b[bounce % MTSB_CTR_RING];
    uint32_t *n_next = &pb.ctr->b[(bounce + 1) % MTSB_CTR_RING].n_rays;
    const uint32_t n = bc.n_q[Q];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long shade_events = 0, path_len = 0;
    for (uint32_t i0 = blockIdx.x * blockDim.x; i0 < n; i0 += gridDim.x * blockDim.x) {
        const uint32_t i = i0 + threadIdx.x;
        bool have_shadow = false;
        uint32_t spawn_slot = 0xFFFFFFFFu;
        if (i < n) {
        const uint32_t q = pb.q[Q][i];
        MTSB_CHECK(q < pb.capacity, CHK_QUEUE);
        const uint32_t slot = pb.ray_slot[cur][q];
        MTSB_CHECK(slot < pb.capacity, CHK_SLOT);
        const RayIn ray = load_ray(pb.rays[cur], q);
        const uint4 hraw = reinterpret_cast<const uint4 *>(pb.hits)[q];
        const uint32_t prim = hraw.x; const float hu = __uint_as_float(hraw.y), hv = __uint_as_float(hraw.z), ht = __uint_as_float(hraw.w);
        const float4 te = pb.thr_eta[slot], lb = pb.L_bpdf[slot], rf = pb.refn_flags[slot];
        V3 thr(te.x, te.y, te.z), L(lb.x, lb.y, lb.z);
        float eta = te.w; const float bsdfPdfPrev = lb.w;
        const V3 prevRefN(rf.x, rf.y, rf.z);
        uint32_t fl = __float_as_uint(rf.w);
        int depth = (int) (fl & PF_DEPTH_MASK);
        const bool first = (fl & PF_FIRST) != 0;
        const V3 rd(ray.dx, ray.dy, ray.dz);

        Its its;
        fill_intersection(sc, rd, prim, hu, hv, its);
        const MtsbShape &shape = sc.shapes[its.shape];
        const int mat = shape.material;
        bool alive = true;

        if (first) {
            fl |= PF_ALPHA;
            // path.cpp:150-152 (scattered is false on the first segment)
            if (shape.emitter >= 0 && !rc.hide_emitters) L += thr * area_eval(sc.emitters[shape.emitter], its.sh.n, -rd);
        } else {
            // tail of the previous iteration: emitter hit by the BSDF ray (path.cpp:230-234, 259-266)
            if (shape.emitter >= 0) {
                const V3 value = area_eval(sc.emitters[shape.emitter], its.sh.n, -rd);
                DRec dRec;
                dRec.ref = V3(ray.ox, ray.oy, ray.oz); dRec.refN = prevRefN; dRec.p = its.p; dRec.n = its.sh.n;
                dRec.d = rd; dRec.dist = ht; dRec.measure = MTSB_MEASURE_SOLID_ANGLE; dRec.emitter = shape.emitter;
                // a negative stored pdf marks a Dirac lobe: lumPdf = 0 (path.cpp:263-264)
                const float lumPdf = bsdfPdfPrev < 0 ? 0.0f : pdf_emitter_direct(sc, dRec);
                L += thr * value * mi_weight(bsdfPdfPrev, lumPdf);
            }
            // Russian roulette (path.cpp:278-288)
            const int d0 = depth++;
            if (d0 >= rc.rr_depth) {
                const float qq = fminf(maxComp(thr) * eta * eta, 0.95f);
                int px, py; uint32_t s_local; slot_pixel(sc, pb, ps, slot, px, py, s_local);
                float u[4]; slot_stream(sc, ps, px, py, s_local).block4((uint32_t) d0, 1, u);
                if (QMC) {
                    const QmcDraw qd = qmc_draw(sc.sampler, ps.seed, px + sc.sensor.crop_x, py + sc.sensor.crop_y, ps.s_first + s_local, fl, 1, u[1], 0.0f);
                    u[1] = qd.a; fl = qd.fl;
                }
                if (u[1] >= qq) alive = false;
                else thr = thr / qq;
            }
            if (alive && !(depth <= rc.max_depth || rc.max_depth < 0)) alive = false;      // while condition :137
        }
        // path.cpp:158-167
        if (alive && ((depth >= rc.max_depth && rc.max_depth > 0) ||
                      (rc.strict_normals && dot(rd, its.geoN) * cosTheta(its.wi) >= 0))) alive = false;

        bool spawned = false;
        if (alive) {
            ++shade_events;
            const bool haveMat = (Q != Q_NULL) && mat >= 0;
            DRec dRec;
            dRec.ref = its.p; dRec.refN = V3(0.0f); dRec.emitter = -1;
            float u[4] = {0, 0, 0, 0};
            if (haveMat) {
                int px, py; uint32_t s_local; slot_pixel(sc, pb, ps, slot, px, py, s_local);
                const SampleStream ss = slot_stream(sc, ps, px, py, s_local);
                ss.block4((uint32_t) depth, 0, u);
                MtsbMaterial lm[Q == Q_TEXTURED ? 2 : 1];
                const MtsbMaterial *mats = sc.materials;
                int mid = mat;
                if (Q == Q_TEXTURED) {
                    resolve_textured(sc, mat, its, first, first ? pb.spos[slot] : make_float2(0.0f, 0.0f), V3(ray.ox, ray.oy, ray.oz), lm);
                    mats = lm; mid = 0;
                }
                const MtsbMaterial &m0 = mats[mid];
                if (!bsdf_has_back_or_transmission(m0)) dRec.refN = its.sh.n;              // records.inl:175-179
                // ---- direct illumination (path.cpp:174-202): only for BSDFs with a smooth component (:176-177) ----
                const bool smooth = Q == Q_DELTA ? false : (Q == Q_GENERIC || Q == Q_TEXTURED) ? bsdf_is_smooth(mats, m0) : true;
                const bool qmc = QMC;
                if (qmc && smooth) {             // path.cpp:178: the emitter sample is drawn only for smooth BSDFs
                    const QmcDraw qd = qmc_draw(sc.sampler, ps.seed, px + sc.sensor.crop_x, py + sc.sensor.crop_y, ps.s_first + s_local, fl, 2, u[0], u[1]);
                    u[0] = qd.a; u[1] = qd.b; fl = qd.fl;
                }
                P2 ue; ue.x = u[0]; ue.y = u[1];
                V3 value(0.0f);
                dRec.pdf = 0;
                if (smooth) value = sample_emitter_direct(sc, dRec, ue);
                if (dRec.pdf != 0 && !isZero(value)) {
                    const V3 wo = its.sh.toLocal(dRec.d);
                    const V3 bsdfVal = eval_q<Q>(mats, mid, its.wi, wo);
                    if (!isZero(bsdfVal) && (!rc.strict_normals || dot(its.geoN, dRec.d) * cosTheta(wo) > 0)) {
                        const MtsbEmitter &em = sc.emitters[dRec.emitter];
                        const bool onSurface = em.type != MTSB_EMITTER_POINT;
                        const float bp = (onSurface && dRec.measure == MTSB_MEASURE_SOLID_ANGLE) ? pdf_q<Q>(mats, mid, its.wi, wo) : 0.0f;
                        const V3 contrib = thr * value * bsdfVal * mi_weight(dRec.pdf, bp);
                        // shadow segment of Scene::sampleEmitterDirect (scene.cpp:840-846)
                        RayIn sr; sr.ox = dRec.ref.x; sr.oy = dRec.ref.y; sr.oz = dRec.ref.z; sr.dx = dRec.d.x; sr.dy = dRec.d.y; sr.dz = dRec.d.z;
                        sr.mint = MTSB_EPSILON; sr.maxt = dRec.dist * (1 - MTSB_SHADOW_EPSILON);
                        s_sray[2 * threadIdx.x] = make_float4(sr.ox, sr.oy, sr.oz, sr.mint);
                        s_sray[2 * threadIdx.x + 1] = make_float4(sr.dx, sr.dy, sr.dz, sr.maxt);
                        s_contrib[threadIdx.x] = make_float4(contrib.x, contrib.y, contrib.z, __uint_as_float(slot));
                        have_shadow = true;
                    }
                }
                // ---- BSDF sampling (path.cpp:208-221) ----
                float u1 = 0.0f;
                if (qmc) {                       // path.cpp:211
                    const QmcDraw qd = qmc_draw(sc.sampler, ps.seed, px + sc.sensor.crop_x, py + sc.sensor.crop_y, ps.s_first + s_local, fl, 2, u[2], u[3]);
                    u[2] = qd.a; u[3] = qd.b; fl = qd.fl;
                }
                if (Q == Q_DIELECTRIC || Q == Q_GENERIC || Q == Q_TEXTURED) {
                    bool need = (Q == Q_DIELECTRIC);
                    if (Q == Q_GENERIC || Q == Q_TEXTURED) need = mats[bsdf_effective_material(mats, mid, its.wi.z)].type == MTSB_BSDF_ROUGHDIELECTRIC;
                    if (need) {
                        float v[4]; ss.block4((uint32_t) depth, 1, v); u1 = v[0];
                        if (qmc) {               // roughdielectric.cpp:554
                            const QmcDraw qd = qmc_draw(sc.sampler, ps.seed, px + sc.sensor.crop_x, py + sc.sensor.crop_y, ps.s_first + s_local, fl, 1, u1, 0.0f);
                            u1 = qd.a; fl = qd.fl;
                        }
                    }
                }
                BsdfSample bs;
                P2 ub; ub.x = u[2]; ub.y = u[3];
                const bool ok = sample_q<Q>(mats, mid, its.wi, ub, u1, bs);
                if (ok && !isZero(bs.weight)) {
                    fl |= PF_SCATTERED;
                    const V3 wo = its.sh.toWorld(bs.wo);
                    if (!(rc.strict_normals && dot(its.geoN, wo) * cosTheta(bs.wo) <= 0)) {
                        RayIn nr; nr.ox = its.p.x; nr.oy = its.p.y; nr.oz = its.p.z; nr.dx = wo.x; nr.dy = wo.y; nr.dz = wo.z;
                        nr.mint = MTSB_EPSILON; nr.maxt = MTSB_INF;
                        s_nray[2 * threadIdx.x] = make_float4(nr.ox, nr.oy, nr.oz, nr.mint);
                        s_nray[2 * threadIdx.x + 1] = make_float4(nr.dx, nr.dy, nr.dz, nr.maxt);
                        spawn_slot = slot;
                        thr *= bs.weight; eta *= bs.eta;
                        pb.thr_eta[slot] = make_float4(thr.x, thr.y, thr.z, eta);
                        const bool delta = (Q == Q_DELTA || Q == Q_GENERIC || Q == Q_TEXTURED) && bs.delta;
                        pb.L_bpdf[slot] = make_float4(L.x, L.y, L.z, delta ? -bs.pdf : bs.pdf);
                        fl = (fl & ~(PF_DEPTH_MASK | PF_FIRST)) | ((uint32_t) depth & PF_DEPTH_MASK);
                        pb.refn_flags[slot] = make_float4(dRec.refN.x, dRec.refN.y, dRec.refN.z, __uint_as_float(fl));
                        spawned = true;
                    }
                }
            }
        }
        if (!spawned) {
            path_len += (unsigned long long) depth;
            pb.L_bpdf[slot] = make_float4(L.x, L.y, L.z, 0.0f);
            fl = (fl & ~PF_DEPTH_MASK) | ((uint32_t) depth & PF_DEPTH_MASK);
            pb.refn_flags[slot] = make_float4(0, 0, 0, __uint_as_float(fl));
        }
        }   // i < n
        // ---- reserve queue space for the whole block, then copy the staged records out ----
        const unsigned b_s = __ballot_sync(0xFFFFFFFFu, have_shadow), b_r = __ballot_sync(0xFFFFFFFFu, spawn_slot != 0xFFFFFFFFu);
        if (lane == 0) { s_wcnt[0][warp] = (uint32_t) __popc(b_s); s_wcnt[1][warp] = (uint32_t) __popc(b_r); }
        __syncthreads();
        if (threadIdx.x < 2) {
            uint32_t tot = 0;
            for (int w = 0; w < MTSB_SHADE_BLOCK / 32; ++w) { const uint32_t c = s_wcnt[threadIdx.x][w]; s_wcnt[threadIdx.x][w] = tot; tot += c; }
            s_qbase[threadIdx.x] = tot ? atomicAdd(threadIdx.x == 0 ? &bc.n_shadow : n_next, tot) : 0u;
        }
        __syncthreads();
        const uint32_t lt = (1u << lane) - 1u;
        if (have_shadow) {
            const uint32_t sq = s_qbase[0] + s_wcnt[0][warp] + (uint32_t) __popc(b_s & lt);
            MTSB_CHECK(sq < pb.capacity, CHK_QUEUE);
            reinterpret_cast<float4 *>(pb.srays)[2 * (size_t) sq] = s_sray[2 * threadIdx.x];
            reinterpret_cast<float4 *>(pb.srays)[2 * (size_t) sq + 1] = s_sray[2 * threadIdx.x + 1];
            pb.scontrib[sq] = s_contrib[threadIdx.x];
        }
        if (spawn_slot != 0xFFFFFFFFu) {
            const uint32_t nq = s_qbase[1] + s_wcnt[1][warp] + (uint32_t) __popc(b_r & lt);
            MTSB_CHECK(nq < pb.capacity, CHK_QUEUE);
            reinterpret_cast<float4 *>(pb.rays[1 - cur])[2 * (size_t) nq] = s_nray[2 * threadIdx.x];
            reinterpret_cast<float4 *>(pb.rays[1 - cur])[2 * (size_t) nq + 1] = s_nray[2 * threadIdx.x + 1];
            pb.ray_slot[1 - cur][nq] = spawn_slot;
        }
        __syncthreads();
    }
    for (int o = 16; o; o >>= 1) { shade_events += __shfl_xor_sync(0xFFFFFFFFu, shade_events, o); path_len += __shfl_xor_sync(0xFFFFFFFFu, path_len, o); }
    if ((threadIdx.x & 31) == 0) {
        if (shade_events) atomicAdd(&pb.ctr->shade, shade_events);
        if (path_len) atomicAdd(&pb.ctr->path_len, path_len);
    }
}

// epilogue of the shadow-ray pass: unoccluded -> add the pending emitter contribution
struct ShadowResolve {
    const float4 *contrib; float4 *L_bpdf;
    __device__ __forceinline__ void operator()(uint32_t i, const HitOut &h) const {
        if (h.prim != MTSB_NO_HIT_ID) return;
        const float4 c = contrib[i];
        const uint32_t slot = __float_as_uint(c.w);
        float4 l = L_bpdf[slot];
        l.x += c.x; l.y += c.y; l.z += c.z;
        L_bpdf[slot] = l;
    }
};

// ImageBlock::put for every sample of one 32x32 tile (imageblock.h:143-206) followed by the
// overlap-add of the bordered block into the film (renderproc.cpp:149-156), written as a GATHER:
// the samples of the tile are staged in shared memory (MTSB_SPLAT_CHUNK samples per pixel at a
// time, row-major), then every thread owns bordered pixels and sums the samples whose footprint
// covers them -- same weights as the scatter form, no shared-memory atomics, and a fixed
// summation order inside the tile.  Only the final add into the film is atomic (tile borders overlap).
#define MTSB_SPLAT_CHUNK 1
__global__ void __launch_bounds__(256) k_splat(SceneView sc, PathBuffers pb, PassInfo ps, unsigned long long *bad) {
    extern __shared__ float4 s_stage[];                 // [chunk][1024] value (rgb, alpha), then [chunk][1024] float2 positions
    __shared__ float s_filter[32];                      // the discretised filter (a per-thread index into kernel parameters would serialise)
    if (threadIdx.x < 32) s_filter[threadIdx.x] = sc.filter_table[threadIdx.x];
    __syncthreads();                                    // the staging loop's scatter path reads the table
    const float f_scale = sc.filter_scale;
    const int chunk = min((int) ps.spp_pass, MTSB_SPLAT_CHUNK);
    float4 *s_val = s_stage;
    float2 *s_pos = reinterpret_cast<float2 *>(s_stage + chunk * MTSB_TILE * MTSB_TILE);
    const int b = sc.film_border;
    const int tile = pb.tiles[ps.tile_first + blockIdx.x];
    const int ox = (tile % (int) ps.tiles_x) * MTSB_TILE, oy = (tile / (int) ps.tiles_x) * MTSB_TILE;
    const int bw = min(MTSB_TILE, sc.film_w - ox), bh = min(MTSB_TILE, sc.film_h - oy);
    const int sx = bw + 2 * b, sy = bh + 2 * b;
    const uint32_t per_tile = MTSB_TILE * MTSB_TILE * ps.spp_pass;
    const uint32_t slot0 = blockIdx.x * per_tile;
    const float r = sc.filter_radius;
    const int fsx = sc.film_w + 2 * b;
    unsigned long long n_bad = 0;
    // gridDim.y CTAs share a tile: each takes every gridDim.y-th staging round of the pass (a rank of a multi-GPU
    // job renders few tiles at many samples; one CTA per tile would leave most SMs idle)
    for (uint32_t c0 = blockIdx.y * (uint32_t) chunk; c0 < ps.spp_pass; c0 += gridDim.y * (uint32_t) chunk) {
        const int nc = min(chunk, (int) (ps.spp_pass - c0));
        // ---- stage: slot (pixel in Morton order, sample) -> smem[sample][y][x] ----
        for (int i = threadIdx.x; i < nc * MTSB_TILE * MTSB_TILE; i += blockDim.x) {
            const uint32_t pix = (uint32_t) i / (uint32_t) nc, sl = (uint32_t) i % (uint32_t) nc;
            const uint32_t slot = slot0 + pix * ps.spp_pass + c0 + sl;
            const uint32_t fl = __float_as_uint(pb.refn_flags[slot].w);
            const int x = (int) compact1by1(pix), y = (int) compact1by1(pix >> 1);
            float4 val = make_float4(0, 0, 0, 0);
            float2 pos = make_float2(1e30f, 1e30f);       // skipped slots sit outside every footprint
            if (fl & PF_VALID) {
                const float4 lb = pb.L_bpdf[slot];
                if (isfinite(lb.x) && isfinite(lb.y) && isfinite(lb.z)) {               // imageblock.h:147-153
                    val = make_float4(lb.x, lb.y, lb.z, (fl & PF_ALPHA) ? 1.0f : 0.0f);
                    const float2 sp = pb.spos[slot];
                    // position inside the bordered block: pos - 0.5 - (offset - border)
                    pos = make_float2(sp.x - 0.5f - (float) (ox - b), sp.y - 0.5f - (float) (oy - b));
                    const float fx = sp.x - (float) (ox + x), fy = sp.y - (float) (oy + y);
                    if (!(fx >= 0.0f && fx < 1.0f && fy >= 0.0f && fy < 1.0f)) {
                        // A sample outside its own pixel (the reference's halton sampler with random digit permutations places
                        // some there; px + u can also round up to px + 1): the gather below assumes samples sit inside their
                        // pixel, so these take ImageBlock::put's scatter form directly (imageblock.h:155-199), clipped to the block.
                        const int x0 = max((int) ceilf(pos.x - r), 0), x1 = min((int) floorf(pos.x + r), sx - 1);
                        const int y0 = max((int) ceilf(pos.y - r), 0), y1 = min((int) floorf(pos.y + r), sy - 1);
                        for (int Y = y0; Y <= y1; ++Y)
                            for (int X = x0; X <= x1; ++X) {
                                const float w = s_filter[min((int) fabsf(((float) X - pos.x) * f_scale), 31)] *
                                                s_filter[min((int) fabsf(((float) Y - pos.y) * f_scale), 31)];
                                MTSB_CHECK_FILM(sc, ox + X, oy + Y);
                                MTSB_CHECK_FILM(sc, ox + X, oy + Y);
            float *dest = sc.film + ((size_t) (oy + Y) * fsx + (ox + X)) * 5;
                                atomicAdd(dest, w * val.x); atomicAdd(dest + 1, w * val.y); atomicAdd(dest + 2, w * val.z);
                                atomicAdd(dest + 3, w * val.w); atomicAdd(dest + 4, w);
                            }
                        val = make_float4(0, 0, 0, 0); pos = make_float2(1e30f, 1e30f);
                    }
                } else ++n_bad;
            }
            s_val[sl * MTSB_TILE * MTSB_TILE + y * MTSB_TILE + x] = val;
            s_pos[sl * MTSB_TILE * MTSB_TILE + y * MTSB_TILE + x] = pos;
        }
        __syncthreads();
        // ---- gather: one bordered pixel per thread iteration ----
        for (int p = threadIdx.x; p < sx * sy; p += blockDim.x) {
            const int X = p % sx, Y = p / sx;
            const float Xf = (float) X, Yf = (float) Y;
            // tile pixels whose samples can reach (X, Y): |X - px| <= r with px in [x' + b - 0.5, x' + b + 0.5)
            const int x_lo = max(0, (int) ceilf((float) (X - b) - r - 0.5f)), x_hi = min(bw - 1, (int) floorf((float) (X - b) + r + 0.5f));
            const int y_lo = max(0, (int) ceilf((float) (Y - b) - r - 0.5f)), y_hi = min(bh - 1, (int) floorf((float) (Y - b) + r + 0.5f));
            float acc[5] = {0, 0, 0, 0, 0};
            for (int sl = 0; sl < nc; ++sl)
                for (int yy = y_lo; yy <= y_hi; ++yy)
                    for (int xx = x_lo; xx <= x_hi; ++xx) {
                        const int j = sl * MTSB_TILE * MTSB_TILE + yy * MTSB_TILE + xx;
                        const float2 pos = s_pos[j];
                        // ImageBlock::put pixel range of this sample, ceil(pos - r) .. floor(pos + r): for an integer X,
                        // X >= ceil(y) <=> X >= y and X <= floor(z) <=> X <= z, so no rounding instruction is needed
                        if (Xf < pos.x - r || Xf > pos.x + r || Yf < pos.y - r || Yf > pos.y + r) continue;
                        const float4 val = s_val[j];
                        const float w = s_filter[min((int) fabsf((Xf - pos.x) * f_scale), 31)] *
                                        s_filter[min((int) fabsf((Yf - pos.y) * f_scale), 31)];
                        acc[0] += w * val.x; acc[1] += w * val.y; acc[2] += w * val.z; acc[3] += w * val.w; acc[4] += w;
                    }
            MTSB_CHECK_FILM(sc, ox + X, oy + Y);
            float *dest = sc.film + ((size_t) (oy + Y) * fsx + (ox + X)) * 5;
#pragma unroll
            for (int k = 0; k < 5; ++k) if (acc[k] != 0.0f) atomicAdd(dest + k, acc[k]);
        }
        __syncthreads();
    }
    for (int o = 16; o; o >>= 1) n_bad += __shfl_xor_sync(0xFFFFFFFFu, n_bad, o);
    if ((threadIdx.x & 31) == 0 && n_bad) atomicAdd(bad, n_bad);
}

// The same gather for reconstruction filters whose footprint spans at most five pixels per axis (radius <= 2.5: the
// default gaussian, box, ...).  A sample that lies inside its own pixel (px, py) only reaches columns px-2 .. px+2, so the
// staging pass evaluates its five column weights and five row weights ONCE (ImageBlock::put's separable table lookups,
// imageblock.h:170-183; zero outside the sample's pixel range) and the gather is a fixed 5x5 stencil of
// weight products without range tests or float->int conversions: about a quarter of the instructions of k_splat.
#define MTSB_SPLAT5_OWN 6          /* ceil(36 * 36 / 256) bordered pixels per thread (border <= 2) */
#define MTSB_SPLAT5_SMEM ((size_t) MTSB_TILE * MTSB_TILE * (sizeof(float4) + 10 * sizeof(float)))
__global__ void __launch_bounds__(256) k_splat5(SceneView sc, PathBuffers pb, PassInfo ps, unsigned long long *bad) {
    extern __shared__ float4 s_stage[];                 // [1024] value (rgb, alpha); then wx[5][1024], wy[5][1024]
    __shared__ float s_filter[32];
    if (threadIdx.x < 32) s_filter[threadIdx.x] = sc.filter_table[threadIdx.x];
    __syncthreads();
    const float f_scale = sc.filter_scale;
    constexpr int NPIX = MTSB_TILE * MTSB_TILE;
    float4 *s_val = s_stage;
    float *s_wx = reinterpret_cast<float *>(s_stage + NPIX);
    float *s_wy = s_wx + 5 * NPIX;
    const int b = sc.film_border;
    const int tile = pb.tiles[ps.tile_first + blockIdx.x];
    const int ox = (tile % (int) ps.tiles_x) * MTSB_TILE, oy = (tile / (int) ps.tiles_x) * MTSB_TILE;
    const int bw = min(MTSB_TILE, sc.film_w - ox), bh = min(MTSB_TILE, sc.film_h - oy);
    const int sx = bw + 2 * b, sy = bh + 2 * b;
    const uint32_t per_tile = NPIX * ps.spp_pass;
    const uint32_t slot0 = blockIdx.x * per_tile;
    const float r = sc.filter_radius;
    const int fsx = sc.film_w + 2 * b;
    unsigned long long n_bad = 0;
    float acc[MTSB_SPLAT5_OWN][5];
#pragma unroll
    for (int q = 0; q < MTSB_SPLAT5_OWN; ++q) { acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = acc[q][4] = 0.0f; }
    for (uint32_t c0 = blockIdx.y; c0 < ps.spp_pass; c0 += gridDim.y) {
        // ---- stage one sample per pixel: value + separable weights ----
        for (int pix = threadIdx.x; pix < NPIX; pix += blockDim.x) {
            const uint32_t slot = slot0 + (uint32_t) pix * ps.spp_pass + c0;
            const uint32_t fl = __float_as_uint(pb.refn_flags[slot].w);
            const int x = (int) compact1by1((uint32_t) pix), y = (int) compact1by1((uint32_t) pix >> 1);
            const int j = y * MTSB_TILE + x;
            float4 val = make_float4(0, 0, 0, 0);
            float wx[5] = {0, 0, 0, 0, 0}, wy[5] = {0, 0, 0, 0, 0};
            if (fl & PF_VALID) {
                const float4 lb = pb.L_bpdf[slot];
                if (isfinite(lb.x) && isfinite(lb.y) && isfinite(lb.z)) {               // imageblock.h:147-153
                    val = make_float4(lb.x, lb.y, lb.z, (fl & PF_ALPHA) ? 1.0f : 0.0f);
                    const float2 sp = pb.spos[slot];
                    const float posx = sp.x - 0.5f - (float) (ox - b), posy = sp.y - 0.5f - (float) (oy - b);
                    const float fx = sp.x - (float) (ox + x), fy = sp.y - (float) (oy + y);
                    if (!(fx >= 0.0f && fx < 1.0f && fy >= 0.0f && fy < 1.0f)) {
                        // outside its own pixel (see k_splat): ImageBlock::put's scatter form, clipped to the block
                        const int x0 = max((int) ceilf(posx - r), 0), x1 = min((int) floorf(posx + r), sx - 1);
                        const int y0 = max((int) ceilf(posy - r), 0), y1 = min((int) floorf(posy + r), sy - 1);
                        for (int Y = y0; Y <= y1; ++Y)
                            for (int X = x0; X <= x1; ++X) {
                                const float w = s_filter[min((int) fabsf(((float) X - posx) * f_scale), 31)] *
                                                s_filter[min((int) fabsf(((float) Y - posy) * f_scale), 31)];
                                MTSB_CHECK_FILM(sc, ox + X, oy + Y);
                                MTSB_CHECK_FILM(sc, ox + X, oy + Y);
            float *dest = sc.film + ((size_t) (oy + Y) * fsx + (ox + X)) * 5;
                                atomicAdd(dest, w * val.x); atomicAdd(dest + 1, w * val.y); atomicAdd(dest + 2, w * val.z);
                                atomicAdd(dest + 3, w * val.w); atomicAdd(dest + 4, w);
                            }
                        val = make_float4(0, 0, 0, 0);
                    } else {
#pragma unroll
                        for (int k = 0; k < 5; ++k) {
                            // bordered column / row k-2 away from the sample's pixel; inside the pixel range of the sample
                            // (ceil(pos - r) .. floor(pos + r), tested without rounding as in k_splat) its table weight
                            const float Xf = (float) (x + b + k - 2), Yf = (float) (y + b + k - 2);
                            if (!(Xf < posx - r || Xf > posx + r)) wx[k] = s_filter[min((int) fabsf((Xf - posx) * f_scale), 31)];
                            if (!(Yf < posy - r || Yf > posy + r)) wy[k] = s_filter[min((int) fabsf((Yf - posy) * f_scale), 31)];
                        }
                    }
                } else ++n_bad;
            }
            s_val[j] = val;
#pragma unroll
            for (int k = 0; k < 5; ++k) { s_wx[k * NPIX + j] = wx[k]; s_wy[k * NPIX + j] = wy[k]; }
        }
        __syncthreads();
        // ---- gather: every thread owns up to MTSB_SPLAT5_OWN bordered pixels for the whole pass and keeps their sums in
        // registers (one film atomic per pixel and CTA at the end instead of one per staging round) ----
#pragma unroll
        for (int q = 0; q < MTSB_SPLAT5_OWN; ++q) {
            const int p = (int) threadIdx.x + q * 256;
            if (p >= sx * sy) break;
            const int X = p % sx, Y = p / sx;
            const int tx = X - b, ty = Y - b;               // tile pixel under this bordered pixel (may be outside the tile)
#pragma unroll
            for (int dy = -2; dy <= 2; ++dy) {
                const int yy = ty + dy;
                if (yy < 0 || yy >= bh) continue;
#pragma unroll
                for (int dx = -2; dx <= 2; ++dx) {
                    const int xx = tx + dx;
                    if (xx < 0 || xx >= bw) continue;
                    const int j = yy * MTSB_TILE + xx;
                    // the sample of tile pixel (xx, yy) sees this pixel as its column 2 - dx and row 2 - dy
                    const float w = s_wx[(2 - dx) * NPIX + j] * s_wy[(2 - dy) * NPIX + j];
                    const float4 val = s_val[j];
                    acc[q][0] += w * val.x; acc[q][1] += w * val.y; acc[q][2] += w * val.z; acc[q][3] += w * val.w; acc[q][4] += w;
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int q = 0; q < MTSB_SPLAT5_OWN; ++q) {
        const int p = (int) threadIdx.x + q * 256;
        if (p >= sx * sy) break;
        const int X = p % sx, Y = p / sx;
        MTSB_CHECK_FILM(sc, ox + X, oy + Y);
        float *dest = sc.film + ((size_t) (oy + Y) * fsx + (ox + X)) * 5;
#pragma unroll
        for (int k = 0; k < 5; ++k) if (acc[q][k] != 0.0f) atomicAdd(dest + k, acc[q][k]);
    }
    for (int o = 16; o; o >>= 1) n_bad += __shfl_xor_sync(0xFFFFFFFFu, n_bad, o);
    if ((threadIdx.x & 31) == 0 && n_bad) atomicAdd(bad, n_bad);
}

// ---- host side ---------------------------------------------------------------------------
static void free_buffers(PathBuffers *pb) {
    if (!pb) return;
    for (int i = 0; i < 2; ++i) { cudaFree(pb->rays[i]); cudaFree(pb->ray_slot[i]); }
    cudaFree(pb->hits); cudaFree(pb->srays); cudaFree(pb->scontrib); cudaFree(pb->thr_eta); cudaFree(pb->L_bpdf);
    cudaFree(pb->refn_flags); cudaFree(pb->spos);
    for (int i = 0; i < Q_COUNT; ++i) cudaFree(pb->q[i]);
    cudaFree(pb->ctr); cudaFree(pb->tiles);
    if (pb->h_poll) cudaFreeHost(pb->h_poll);
    delete pb;
}

// Path buffers are large (about 212 B per slot) and their size does not depend on the scene, so a
// destroyed scene parks them in a per-device pool and the next scene's first render picks them up
// instead of paying cudaMalloc / cudaFree of several GB again.  The pool keeps at most
// MTSB_MAX_LANES sets per device for the life of the process.
#define MTSB_SLOT_BYTES 212ull         /* 3 x 32 (rays) + 16 (hits) + 16 (contrib) + 3 x 16 + 8 (state) + 2 x 4 (slots) + 7 x 4 (queues) */
static std::mutex g_pool_mutex;
static std::map<int, std::vector<PathBuffers *>> g_pool;

int render_free(Scene *s) {
    std::lock_guard<std::mutex> g(g_pool_mutex);
    std::vector<PathBuffers *> &pool = g_pool[s->device];
    for (int l = 0; l < MTSB_MAX_LANES; ++l) {
        if (s->paths[l]) {
            if (pool.size() < MTSB_MAX_LANES) pool.push_back(s->paths[l]);
            else free_buffers(s->paths[l]);
        }
        s->paths[l] = nullptr;
        if (s->lane_stream[l]) { cudaStreamDestroy(s->lane_stream[l]); s->lane_stream[l] = nullptr; }
    }
    return MTSB_OK;
}

// A set of path buffers is handed to the scene only when EVERY allocation succeeded: after an out-of-memory error the
// scene keeps no half-built set (whose capacity would look sufficient to the next render) and nothing stays allocated.
static int alloc_buffers(uint32_t capacity, uint32_t n_tiles, PathBuffers **out) {
    PathBuffers *pb = new PathBuffers();
    memset(pb, 0, sizeof(*pb));
    const size_t c = capacity;
    cudaError_t e = cudaSuccess;
    auto dm = [&](void **p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
    for (int i = 0; i < 2; ++i) { dm((void **) &pb->rays[i], c * sizeof(RayIn)); dm((void **) &pb->ray_slot[i], c * 4); }
    dm((void **) &pb->hits, c * sizeof(HitOut));
    dm((void **) &pb->srays, c * sizeof(RayIn));
    dm((void **) &pb->scontrib, c * sizeof(float4));
    dm((void **) &pb->thr_eta, c * sizeof(float4));
    dm((void **) &pb->L_bpdf, c * sizeof(float4));
    dm((void **) &pb->refn_flags, c * sizeof(float4));
    dm((void **) &pb->spos, c * sizeof(float2));
    for (int i = 0; i < Q_COUNT; ++i) dm((void **) &pb->q[i], c * 4);
    dm((void **) &pb->ctr, sizeof(Counters));
    dm((void **) &pb->tiles, (size_t) std::max(1u, n_tiles) * sizeof(int));
    if (e == cudaSuccess) e = cudaMallocHost((void **) &pb->h_poll, MTSB_POLL_RING * sizeof(uint32_t));
    if (e != cudaSuccess) { free_buffers(pb); return cuda_fail(e, "cudaMalloc(path buffers)", __FILE__, __LINE__); }
    pb->capacity = capacity; pb->tiles_capacity = n_tiles;
    *out = pb;
    return MTSB_OK;
}

static int ensure_buffers(Scene *s, int lane, uint32_t capacity, uint32_t n_tiles) {
    if (!s->lane_stream[lane]) MTSB_CUDA(cudaStreamCreateWithFlags(&s->lane_stream[lane], cudaStreamNonBlocking));
    if (s->paths[lane] && s->paths[lane]->capacity >= capacity && s->paths[lane]->tiles_capacity >= n_tiles) return MTSB_OK;
    {
        std::lock_guard<std::mutex> g(g_pool_mutex);
        std::vector<PathBuffers *> &pool = g_pool[s->device];
        if (s->paths[lane]) { free_buffers(s->paths[lane]); s->paths[lane] = nullptr; }
        for (size_t i = 0; i < pool.size(); ++i)
            if (pool[i]->capacity >= capacity && pool[i]->capacity <= 2 * (size_t) capacity) {
                s->paths[lane] = pool[i];
                pool.erase(pool.begin() + (long) i);
                break;
            }
        if (!s->paths[lane]) { for (PathBuffers *p : pool) free_buffers(p); pool.clear(); }    // make room before a new allocation
    }
    if (s->paths[lane]) {
        PathBuffers *pb = s->paths[lane];
        if (pb->tiles_capacity < n_tiles) {
            int *t = nullptr;
            MTSB_CUDA(cudaMalloc((void **) &t, (size_t) std::max(1u, n_tiles) * sizeof(int)));
            cudaFree(pb->tiles); pb->tiles = t; pb->tiles_capacity = n_tiles;
        }
        return MTSB_OK;
    }
    return alloc_buffers(capacity, n_tiles, &s->paths[lane]);
}

// scene creation (api.cu): packed shading records + the queue byte of every triangle
int pack_shading_records(Scene *s, uint64_t n_tris) {
    SceneView &v = s->v;
    v.shade_tris = nullptr; v.tri_queue = nullptr;
    if (n_tris == 0) return MTSB_OK;
    std::vector<uint8_t> sq(s->h_shape_material.size());
    for (size_t i = 0; i < sq.size(); ++i) sq[i] = (uint8_t) queue_of_material(s->h_materials.data(), s->h_shape_material[i]);
    const uint8_t *d_sq = nullptr;
    int rc = dev_upload(s, &d_sq, sq.data(), sq.size());
    if (rc) return rc;
    U4 *recs = nullptr; uint8_t *tq = nullptr;
    if ((rc = dev_alloc(s, &recs, (size_t) MTSB_SHADE_TRI_U4 * n_tris))) return rc;
    if ((rc = dev_alloc(s, &tq, (size_t) n_tris))) return rc;
    k_pack_tris<<<grid_for(n_tris, 256), 256, 0, s->stream>>>(v, (uint32_t) n_tris, d_sq, recs, tq);
    s->launches++;
    MTSB_CUDA(cudaGetLastError());
    v.shade_tris = recs; v.tri_queue = tq;
    return MTSB_OK;
}

struct StageTimer {
    struct Span { cudaEvent_t a, b; int stage; uint32_t n; };
    std::vector<Span> spans; std::vector<cudaEvent_t> pool; size_t next = 0; cudaStream_t st;
    explicit StageTimer(cudaStream_t s) : st(s) {}
    cudaEvent_t get() { if (next == pool.size()) { cudaEvent_t e; cudaEventCreate(&e); pool.push_back(e); } return pool[next++]; }
    void begin(int stage, uint32_t n = 0) { Span sp; sp.a = get(); sp.b = nullptr; sp.stage = stage; sp.n = n; cudaEventRecord(sp.a, st); spans.push_back(sp); }
    void end() { Span &sp = spans.back(); sp.b = get(); cudaEventRecord(sp.b, st); }
    void collect(float *out, int n_stage) {
        for (auto &sp : spans) { float ms = 0; if (sp.b && cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) out[sp.stage] += ms; }
        (void) n_stage;
        if (getenv("MTSB_PROFILE")) {       // per-launch-group times of the first passes (diagnostics)
            static const char *names[] = {"raygen", "closest", "shade", "any", "splat"};
            size_t k = 0;
            for (auto &sp : spans) { float ms = 0; if (sp.b) cudaEventElapsedTime(&ms, sp.a, sp.b); if (k++ < 60) fprintf(stderr, "[mtsb] %-8s n<=%9u %8.3f ms\n", names[sp.stage], sp.n, ms); }
        }
    }
    ~StageTimer() { for (auto e : pool) cudaEventDestroy(e); }
};
enum { ST_RAYGEN = 0, ST_CLOSEST, ST_SHADE, ST_ANY, ST_SPLAT, ST_N };

static int shade_grid(size_t n_upper, int sm_count) {
    size_t need = (n_upper + MTSB_SHADE_BLOCK - 1) / MTSB_SHADE_BLOCK;
    size_t cap = (size_t) sm_count * 16;
    return (int) std::max<size_t>(1, std::min(need, cap));
}

// One pass = a block of tiles x a block of samples, small enough for the path buffers.
struct Pass { PassInfo ps; uint32_t n_slots, n_valid; };

// The wavefront pipeline of one render call: a stream, its path buffers and the poll ring through which the host
// learns (late, and without ever waiting for it) how many rays a bounce had.
struct Pipeline {
    Scene *s = nullptr;
    cudaStream_t st = nullptr;
    PathBuffers *pb = nullptr;
    StageTimer *timer = nullptr;
    std::vector<cudaEvent_t> ev;       // MTSB_POLL_RING events: "count of bounce i is on the host"
    uint32_t queue_mask = 0;           // shading queues this scene can feed
    int sm_count = 148;
    bool qmc = false;
};

template <int Q> static void launch_shade(Pipeline &P, const Pass &p, const RenderConst &rcst, int g, int cur, int bounce) {
    if (!(P.queue_mask & (1u << Q))) return;
    if (P.qmc) k_shade<Q, true><<<g, MTSB_SHADE_BLOCK, 0, P.st>>>(P.s->v, *P.pb, p.ps, rcst, cur, bounce);
    else k_shade<Q, false><<<g, MTSB_SHADE_BLOCK, 0, P.st>>>(P.s->v, *P.pb, p.ps, rcst, cur, bounce);
    ++P.s->launches;
}

// Enqueues one bounce: closest hits -> sort (+ escaped rays) -> shading queues -> shadow rays.  n_upper bounds the
// number of rays of the bounce from above (it only sizes the grids; the kernels read the exact count on the device).
static int enqueue_bounce(Pipeline &P, const Pass &p, const RenderConst &rcst, int bounce, uint32_t n_upper) {
    Scene *s = P.s; PathBuffers &pb = *P.pb; const SceneView &v = s->v;
    const int cur = bounce & 1;
    BounceCtr *bc = &pb.ctr->b[bounce % MTSB_CTR_RING];
    P.timer->begin(ST_CLOSEST, n_upper);
    int rc = launch_trace_kernel<false>(s, pb.rays[cur], &bc->n_rays, n_upper, StoreHit{pb.hits}, P.st);
    P.timer->end();
    if (rc) return rc;
    P.timer->begin(ST_SHADE, n_upper);
    const int g = shade_grid(n_upper, P.sm_count);
    k_sort<<<std::max(1, std::min((int) ((n_upper + MTSB_SORT_BLOCK - 1) / MTSB_SORT_BLOCK), P.sm_count * 4)), MTSB_SORT_BLOCK, 0, P.st>>>(
        v, pb, rcst, cur, bounce, bounce == 0 ? p.n_valid : 0u);
    ++s->launches;
    // only the queues some material of this scene can feed are launched
    launch_shade<Q_DIFFUSE>(P, p, rcst, g, cur, bounce);
    launch_shade<Q_CONDUCTOR>(P, p, rcst, g, cur, bounce);
    launch_shade<Q_DIELECTRIC>(P, p, rcst, g, cur, bounce);
    launch_shade<Q_NULL>(P, p, rcst, g, cur, bounce);
    launch_shade<Q_TEXTURED>(P, p, rcst, g, cur, bounce);
    launch_shade<Q_GENERIC>(P, p, rcst, g, cur, bounce);
    launch_shade<Q_DELTA>(P, p, rcst, g, cur, bounce);
    P.timer->end();
    P.timer->begin(ST_ANY, n_upper);
    rc = launch_trace_kernel<true>(s, pb.srays, &bc->n_shadow, n_upper, ShadowResolve{pb.scontrib, pb.L_bpdf}, P.st);
    P.timer->end();
    if (rc) return rc;
    // the size of the next bounce travels to the host behind this bounce; nobody waits for it
    const int slot = (bounce + 1) % MTSB_POLL_RING;
    MTSB_CUDA(cudaMemcpyAsync(&pb.h_poll[slot], &pb.ctr->b[(bounce + 1) % MTSB_CTR_RING].n_rays, sizeof(uint32_t), cudaMemcpyDeviceToHost, P.st));
    MTSB_CUDA(cudaEventRecord(P.ev[(size_t) slot], P.st));
    return MTSB_OK;
}

// One pass, enqueued without waiting for the device: ray generation, then bounce after bounce until the path length
// limit -- or until the host has SEEN (in the poll ring) that a bounce ended with no rays left.  The host usually runs
// ahead of the device, so a few bounces may be enqueued that find their queues empty (a handful of microseconds each).
static int run_pass(Pipeline &P, const Pass &p, const RenderConst &rcst, size_t splat_smem) {
    Scene *s = P.s; PathBuffers &pb = *P.pb;
    P.timer->begin(ST_RAYGEN);
    k_pass_begin<<<1, 128, 0, P.st>>>(pb, p.n_slots);
    if (P.qmc) k_raygen<true><<<shade_grid(p.n_slots, P.sm_count), MTSB_SHADE_BLOCK, 0, P.st>>>(s->v, pb, p.ps, p.n_slots);
    else k_raygen<false><<<shade_grid(p.n_slots, P.sm_count), MTSB_SHADE_BLOCK, 0, P.st>>>(s->v, pb, p.ps, p.n_slots);
    s->launches += 2;
    P.timer->end();

    const int max_bounces = rcst.max_depth > 0 ? rcst.max_depth : 0x7fffffff;     // a path of depth d spawns no ray at depth >= maxDepth (path.cpp:158)
    uint32_t n_upper = p.n_slots;
    int confirmed = 0;            // bounces 0 .. confirmed have a known ray count (bounce 0: n_slots)
    int bounce = 0, last = -1;
    bool drained = false;
    std::vector<uint32_t> known(1, p.n_slots);
    while (bounce < max_bounces && !drained) {
        if (s->cancel.load()) { set_error("render cancelled"); return MTSB_ERR_CANCELLED; }
        int rc = enqueue_bounce(P, p, rcst, bounce, n_upper);
        if (rc) return rc;
        last = bounce++;
        // the poll ring holds MTSB_POLL_RING counts: never run further ahead than that
        while (bounce - confirmed >= MTSB_POLL_RING - 1) {
            MTSB_CUDA(cudaEventSynchronize(P.ev[(size_t) ((confirmed + 1) % MTSB_POLL_RING)]));
            known.push_back(pb.h_poll[(confirmed + 1) % MTSB_POLL_RING]); ++confirmed;
        }
        while (confirmed < bounce) {
            const cudaError_t q = cudaEventQuery(P.ev[(size_t) ((confirmed + 1) % MTSB_POLL_RING)]);
            if (q == cudaErrorNotReady) break;
            if (q != cudaSuccess) return cuda_fail(q, "wavefront bounce", __FILE__, __LINE__);
            known.push_back(pb.h_poll[(confirmed + 1) % MTSB_POLL_RING]); ++confirmed;
        }
        n_upper = std::min(n_upper, known.back());       // queue sizes never grow from one bounce to the next
        if (known.back() == 0) drained = true;
    }
    P.timer->begin(ST_SPLAT);
    k_pass_end<<<1, 1, 0, P.st>>>(pb, last);
    {
        // enough CTAs for ~8 per SM: split the samples of a tile over gridDim.y CTAs when the pass has few tiles
        const uint32_t rounds = (p.ps.spp_pass + MTSB_SPLAT_CHUNK - 1) / MTSB_SPLAT_CHUNK;
        const uint32_t want = (uint32_t) P.sm_count * 8u;
        uint32_t split = (want + p.ps.n_tiles - 1) / p.ps.n_tiles;
        split = std::max(1u, std::min(std::min(split, rounds), 64u));
        // filters reaching at most five pixels per axis take the precomputed-weight gather (floor(r + 0.5) <= 2)
        static const bool force_generic = getenv("MTSB_SPLAT_GENERIC") != nullptr;
        if (s->v.filter_radius < 2.5f && s->v.film_border <= 2 && !force_generic)
            k_splat5<<<dim3(p.ps.n_tiles, split), 256, MTSB_SPLAT5_SMEM, P.st>>>(s->v, pb, p.ps, s->d_film_bad);
        else
            k_splat<<<dim3(p.ps.n_tiles, split), 256, splat_smem, P.st>>>(s->v, pb, p.ps, s->d_film_bad);
    }
    s->launches += 2;
    P.timer->end();
    return MTSB_OK;
}

}  // namespace mtsb

using namespace mtsb;

extern "C" int mtsb_render(mtsb_scene_t scene, const MtsbRenderParams *params, MtsbStats *stats) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || !params) { set_error("mtsb_render: bad argument"); return MTSB_ERR_INVALID; }
    if (params->tile_count < 0 || (params->tile_count > 0 && (params->tile_index < 0 || params->tile_index >= params->tile_count)) || params->spp_offset < 0) {
        set_error("mtsb_render: bad tile partition / sample offset"); return MTSB_ERR_INVALID;
    }
    MTSB_CUDA(cudaSetDevice(s->device));
    const int sm_count = device_sm_count();
    const SceneView &v = s->v;
    const int spp = params->spp > 0 ? params->spp : v.integrator.sample_count;
    if (spp <= 0 || (uint64_t) spp + (uint64_t) params->spp_offset >= (1ull << 24)) { set_error("mtsb_render: sample count out of range"); return MTSB_ERR_INVALID; }
    const int tile_count = params->tile_count > 0 ? params->tile_count : 1, tile_index = params->tile_count > 0 ? params->tile_index : 0;
    if (v.sampler_type == MTSB_SAMPLER_LDSAMPLER) {
        // ldsampler.cpp:200,208 assert sampleIndex < sampleCount: the point sets hold exactly sampleCount entries
        int n = v.integrator.sample_count > 0 ? v.integrator.sample_count : 1;
        if (spp + params->spp_offset > n) { set_error("mtsb_render: ldsampler point sets hold sampleCount samples; spp_offset + spp exceeds it"); return MTSB_ERR_INVALID; }
    }
    const uint32_t tiles_x = (uint32_t) (v.film_w + MTSB_TILE - 1) / MTSB_TILE, tiles_y = (uint32_t) (v.film_h + MTSB_TILE - 1) / MTSB_TILE;
    std::vector<int> tiles;
    for (uint32_t t = 0; t < tiles_x * tiles_y; ++t) if ((int) (t % (uint32_t) tile_count) == tile_index) tiles.push_back((int) t);
    const uint64_t launches0 = s->launches;
    s->cancel.store(0);
    if (!(params->flags & MTSB_RENDER_KEEP_FILM)) {
        size_t nf = (size_t) (v.film_w + 2 * v.film_border) * (v.film_h + 2 * v.film_border) * 5;
        MTSB_CUDA(cudaMemsetAsync(v.film, 0, nf * sizeof(float), s->stream));
        MTSB_CUDA(cudaMemsetAsync(s->d_film_bad, 0, sizeof(unsigned long long), s->stream));
    }
    if (stats) memset(stats, 0, sizeof(*stats));
    if (stats) stats->n_devices = 1;
    if (tiles.empty()) { MTSB_CUDA(cudaStreamSynchronize(s->stream)); return MTSB_OK; }

    // pass shape: as many path slots as fit in a fifth of the free device memory, at most 64M
    // (a pass pays a fixed price for its thin late bounces, so fewer and larger passes are faster)
    const uint64_t px_per_tile = MTSB_TILE * MTSB_TILE;
    uint64_t slots_m = 64;
    {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
            uint64_t have = 0;
            for (int l = 0; l < MTSB_MAX_LANES; ++l) if (s->paths[l]) have += (uint64_t) s->paths[l]->capacity * MTSB_SLOT_BYTES;
            const uint64_t fit = ((uint64_t) free_b + have) / 5ull / MTSB_SLOT_BYTES >> 20;
            slots_m = std::max<uint64_t>(1, std::min<uint64_t>(slots_m, fit));
        }
        const char *tenv = getenv("MTSB_SLOTS_M");
        if (tenv && atoi(tenv) > 0) slots_m = (uint64_t) std::min(atoi(tenv), 2048);
    }
    const uint64_t target = slots_m << 20;
    // the samples of a pixel are split into equal passes (the tail pass is as wide as the others, so its thin late
    // bounces are not paid for a sliver of work)
    uint64_t spp_pass64 = params->spp_per_pass > 0 ? (uint64_t) params->spp_per_pass
                                                  : std::max<uint64_t>(1, std::min<uint64_t>((uint64_t) spp, target / (tiles.size() * px_per_tile)));
    spp_pass64 = std::min<uint64_t>(spp_pass64, (uint64_t) spp);
    if (params->spp_per_pass <= 0) { const uint64_t np = ((uint64_t) spp + spp_pass64 - 1) / spp_pass64; spp_pass64 = ((uint64_t) spp + np - 1) / np; }
    uint64_t tiles_per_pass = std::max<uint64_t>(1, std::min<uint64_t>(tiles.size(), target / (px_per_tile * spp_pass64)));
    { const uint64_t np = (tiles.size() + tiles_per_pass - 1) / tiles_per_pass; tiles_per_pass = (tiles.size() + np - 1) / np; }
    const uint64_t capacity64 = tiles_per_pass * px_per_tile * spp_pass64;
    if (capacity64 >= (1ull << 32)) { set_error("mtsb_render: a pass of more than 2^32 path slots (spp_per_pass / MTSB_SLOTS_M too large)"); return MTSB_ERR_INVALID; }
    const uint32_t spp_pass = (uint32_t) spp_pass64, capacity = (uint32_t) capacity64;
    std::vector<Pass> passes;
    for (uint32_t t0 = 0; t0 < tiles.size(); t0 += (uint32_t) tiles_per_pass)
        for (uint32_t s0 = 0; s0 < (uint32_t) spp; s0 += spp_pass) {
            Pass p;
            p.ps.tile_first = t0; p.ps.n_tiles = std::min<uint32_t>((uint32_t) tiles_per_pass, (uint32_t) tiles.size() - t0);
            p.ps.spp_pass = std::min<uint32_t>(spp_pass, (uint32_t) spp - s0);
            p.ps.s_first = (uint32_t) params->spp_offset + s0; p.ps.tiles_x = tiles_x; p.ps.seed = params->seed;
            p.n_slots = (uint32_t) (p.ps.n_tiles * px_per_tile * p.ps.spp_pass);
            uint64_t px_valid = 0;
            for (uint32_t t = t0; t < t0 + p.ps.n_tiles; ++t) {
                const int ox = (tiles[t] % (int) tiles_x) * MTSB_TILE, oy = (tiles[t] / (int) tiles_x) * MTSB_TILE;
                px_valid += (uint64_t) std::min(MTSB_TILE, v.film_w - ox) * (uint64_t) std::min(MTSB_TILE, v.film_h - oy);
            }
            p.n_valid = (uint32_t) (px_valid * p.ps.spp_pass);
            passes.push_back(p);
        }

    RenderConst rcst; rcst.max_depth = v.integrator.max_depth; rcst.rr_depth = v.integrator.rr_depth;
    rcst.strict_normals = v.integrator.strict_normals; rcst.hide_emitters = v.integrator.hide_emitters;
    const size_t splat_smem = (size_t) MTSB_SPLAT_CHUNK * MTSB_TILE * MTSB_TILE * (sizeof(float4) + sizeof(float2));
    MTSB_CUDA(cudaFuncSetAttribute(k_splat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) splat_smem));
    MTSB_CUDA(cudaFuncSetAttribute(k_splat5, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) MTSB_SPLAT5_SMEM));

    int status = ensure_buffers(s, 0, capacity, (uint32_t) tiles.size());
    if (status) return status;
    Pipeline P;
    P.s = s; P.st = s->lane_stream[0]; P.pb = s->paths[0]; P.sm_count = sm_count;
    P.qmc = v.sampler_type != MTSB_SAMPLER_INDEPENDENT;
    for (int32_t m : s->h_shape_material) P.queue_mask |= 1u << queue_of_material(s->h_materials.data(), m);
    P.timer = new StageTimer(P.st);
    P.ev.resize(MTSB_POLL_RING);
    for (auto &e : P.ev) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    cudaEvent_t ev0, ev1; cudaEventCreate(&ev0); cudaEventCreate(&ev1);
    cudaEventRecord(ev0, s->stream);
    cudaStreamWaitEvent(P.st, ev0, 0);                               // after the film clear
    cudaMemcpyAsync(P.pb->tiles, tiles.data(), tiles.size() * sizeof(int), cudaMemcpyHostToDevice, P.st);
    cudaMemsetAsync(P.pb->ctr, 0, sizeof(Counters), P.st);

    // The host never runs more than two passes ahead of the device: enough queued work to keep the GPU busy across the
    // pass boundary, little enough that mtsb_cancel (checked before every bounce is enqueued) takes effect within two passes.
    uint32_t n_passes = 0;
    cudaEvent_t pass_done[2] = {nullptr, nullptr};
    for (auto &e2 : pass_done) cudaEventCreateWithFlags(&e2, cudaEventDisableTiming);
    for (const Pass &p : passes) {
        if (n_passes >= 2) cudaEventSynchronize(pass_done[n_passes & 1]);
        status = run_pass(P, p, rcst, splat_smem);
        if (status) break;
        cudaEventRecord(pass_done[n_passes & 1], P.st);
        ++n_passes;
    }
    for (auto &e2 : pass_done) cudaEventDestroy(e2);
    cudaError_t e = cudaStreamSynchronize(P.st);
    if (e != cudaSuccess && status == MTSB_OK) status = cuda_fail(e, "mtsb_render", __FILE__, __LINE__);
    cudaEventRecord(ev1, P.st);
    cudaStreamWaitEvent(s->stream, ev1, 0);                          // later work on the scene's stream (develop, reduce) sees the film
    e = cudaStreamSynchronize(P.st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess && status == MTSB_OK) status = cuda_fail(e, "mtsb_render", __FILE__, __LINE__);
    if (status == MTSB_OK) { e = cudaGetLastError(); if (e != cudaSuccess) status = cuda_fail(e, "mtsb_render kernels", __FILE__, __LINE__); }
    if (status == MTSB_OK) status = check_stack_overflow("mtsb_render");
    if (stats && status == MTSB_OK) {
        Counters hc;
        cudaMemcpy(&hc, P.pb->ctr, sizeof(hc), cudaMemcpyDeviceToHost);
        float st[ST_N] = {0, 0, 0, 0, 0};
        P.timer->collect(st, ST_N);
        unsigned long long bad = 0;
        cudaMemcpy(&bad, s->d_film_bad, sizeof(bad), cudaMemcpyDeviceToHost);
        float total = 0; cudaEventElapsedTime(&total, ev0, ev1);
        // samples counts slots of whole tiles; report the ones inside the crop window
        uint64_t px_in = 0;
        for (int t : tiles) {
            int ox = (t % (int) tiles_x) * MTSB_TILE, oy = (t / (int) tiles_x) * MTSB_TILE;
            px_in += (uint64_t) std::min(MTSB_TILE, v.film_w - ox) * (uint64_t) std::min(MTSB_TILE, v.film_h - oy);
        }
        stats->samples = px_in * (uint64_t) spp;
        stats->closest_rays = hc.closest; stats->shadow_rays = hc.shadow; stats->shade_events = hc.shade;
        stats->path_length_sum = hc.path_len; stats->kernel_launches = s->launches - launches0; stats->bad_samples = bad;
        stats->ms_total = total; stats->ms_raygen = st[ST_RAYGEN]; stats->ms_trace_closest = st[ST_CLOSEST]; stats->ms_shade = st[ST_SHADE];
        stats->ms_trace_any = st[ST_ANY]; stats->ms_splat = st[ST_SPLAT]; stats->n_passes = n_passes;
        for (int b = 0; b < MTSB_BOUNCE_STATS; ++b) { stats->closest_rays_bounce[b] = hc.rays_bounce[b]; stats->shadow_rays_bounce[b] = hc.shadow_bounce[b]; }
        stats->closest_launches = (uint32_t) hc.launches_closest; stats->any_launches = (uint32_t) hc.launches_any;
    }
    delete P.timer;
    for (auto &ev : P.ev) cudaEventDestroy(ev);
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    return status;
}
