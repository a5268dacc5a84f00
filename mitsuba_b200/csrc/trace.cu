// trace.cu -- persistent-thread ray query kernels (closest hit / any hit) and the ray-set
// entry points of the C ABI (mtsb_trace, mtsb_trace_device, mtsb_gen_rays_device).
//
// Kernel shape: a grid of (SM count x resident CTAs) persistent CTAs; every warp pulls rays
// from a global counter.  When fewer than MTSB_REFILL lanes of a warp are still traversing,
// the idle lanes fetch fresh rays (warp-level compaction of live work), so divergence in
// traversal length does not leave lanes empty.  The per-thread stack lives in shared memory
// for its first MTSB_SM_STACK entries and spills to local memory beyond that.
#include "trace.cuh"
#include "sensor.h"
#include "rng.h"
#include <map>
#include <mutex>
#include <thread>
#include <cstring>
#include <algorithm>

namespace mtsb {

int launch_trace(Scene *s, const RayIn *d_rays, const uint32_t *d_count, uint32_t n_max, int mode, HitOut *d_hits, cudaStream_t st) {
    StoreHit ep{d_hits};
    if (mode == MTSB_TRACE_ANY) return launch_trace_kernel<true>(s, d_rays, d_count, n_max, ep, st);
    return launch_trace_kernel<false>(s, d_rays, d_count, n_max, ep, st);
}

// ---- synthetic ray sets of cfg 4 ----------------------------------
__global__ void k_gen_rays(SceneView v, int kind, uint64_t n, uint64_t seed, uint32_t side, float3 c, float radius, RayIn *rays) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RayIn r;
    if (kind == 0) {
        // primary rays of a side x side perspective view, row-major, pixel centres
        uint32_t x = (uint32_t) (i % side), y = (uint32_t) (i / side);
        sensor_ray_frac(v.sensor, v.near_x, v.near_y, ((float) x + 0.5f) / (float) side, ((float) y + 0.5f) / (float) side, r);
    } else {
        // chord between two uniform points on the bounding sphere (src/tests/test_kd.cpp:112-117)
        SampleStream ss(seed, i);
        float u[4]; ss.block4(0, 0, u);
        float z1 = 1.0f - 2.0f * u[1], r1 = sqrtf(fmaxf(0.0f, 1.0f - z1 * z1)), s1, c1; sincosf(2.0f * MTSB_PI * u[0], &s1, &c1);
        float z2 = 1.0f - 2.0f * u[3], r2 = sqrtf(fmaxf(0.0f, 1.0f - z2 * z2)), s2, c2; sincosf(2.0f * MTSB_PI * u[2], &s2, &c2);
        V3 p1(c.x + radius * r1 * c1, c.y + radius * r1 * s1, c.z + radius * z1);
        V3 p2(c.x + radius * r2 * c2, c.y + radius * r2 * s2, c.z + radius * z2);
        V3 d = p2 - p1; float len = length(d);
        if (!(len > 0.0f)) { d = V3(0, 0, 1); len = 1.0f; }
        d = d * (1.0f / len);
        r.ox = p1.x; r.oy = p1.y; r.oz = p1.z; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.mint = 0.0f; r.maxt = len;
    }
    reinterpret_cast<float4 *>(rays)[2 * i] = make_float4(r.ox, r.oy, r.oz, r.mint);
    reinterpret_cast<float4 *>(rays)[2 * i + 1] = make_float4(r.dx, r.dy, r.dz, r.maxt);
}

}  // namespace mtsb

using namespace mtsb;

extern "C" int mtsb_trace_device(mtsb_scene_t scene, const void *d_rays, size_t n, int mode, void *d_hits, void *stream, float *ms_kernel) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!d_rays || !d_hits)) || n > 0xFFFFFFF0ull || (mode != MTSB_TRACE_CLOSEST && mode != MTSB_TRACE_ANY)) {
        set_error("mtsb_trace_device: bad argument"); return MTSB_ERR_INVALID;
    }
    MTSB_CUDA(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t) stream : s->stream;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms_kernel) { MTSB_CUDA(cudaEventCreate(&e0)); MTSB_CUDA(cudaEventCreate(&e1)); MTSB_CUDA(cudaEventRecord(e0, st)); }
    int rc = n ? launch_trace(s, (const RayIn *) d_rays, nullptr, (uint32_t) n, mode, (HitOut *) d_hits, st) : MTSB_OK;
    if (ms_kernel) {
        cudaEventRecord(e1, st);
        cudaError_t e = cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        *ms_kernel = ms;
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (e != cudaSuccess) return cuda_fail(e, "trace kernel", __FILE__, __LINE__);
    }
    return rc;
}

// ---- mtsb_trace: host buffers in, host buffers out ---------------------------------------------------------------
// The batch moves through the GPU in chunks of MTSB_TRACE_CHUNK rays on three streams -- copy in | trace | copy out -- so that
// the PCIe transfers of neighbouring chunks overlap the kernel (one cudaMemcpyAsync per chunk and direction).  Pageable
// caller memory is staged through two pinned buffers per direction, filled / drained by a few host threads (a single
// memcpy thread moves ~10 GB/s, less than one PCIe direction); buffers the caller has pinned itself (cudaHostAlloc /
// cudaHostRegister) are used in place.  The staging memory is allocated once per device and kept.
#define MTSB_TRACE_CHUNK (1u << 20)
#define MTSB_TRACE_COPY_THREADS 4

namespace {
struct TracePipe {
    std::mutex m;
    bool ready = false;
    cudaStream_t s_in = nullptr, s_out = nullptr;
    MtsbRay *pin_rays[2] = {nullptr, nullptr}; MtsbHit *pin_hits[2] = {nullptr, nullptr};
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_k[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
};
std::mutex g_pipe_mutex;
std::map<int, TracePipe> g_pipes;

int pipe_init(TracePipe &p) {
    if (p.ready) return MTSB_OK;
    MTSB_CUDA(cudaStreamCreateWithFlags(&p.s_in, cudaStreamNonBlocking));
    MTSB_CUDA(cudaStreamCreateWithFlags(&p.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        MTSB_CUDA(cudaHostAlloc((void **) &p.pin_rays[i], (size_t) MTSB_TRACE_CHUNK * sizeof(MtsbRay), cudaHostAllocDefault));
        MTSB_CUDA(cudaHostAlloc((void **) &p.pin_hits[i], (size_t) MTSB_TRACE_CHUNK * sizeof(MtsbHit), cudaHostAllocDefault));
        MTSB_CUDA(cudaEventCreateWithFlags(&p.ev_in[i], cudaEventDisableTiming));
        MTSB_CUDA(cudaEventCreateWithFlags(&p.ev_k[i], cudaEventDisableTiming));
        MTSB_CUDA(cudaEventCreateWithFlags(&p.ev_out[i], cudaEventDisableTiming));
    }
    p.ready = true;
    return MTSB_OK;
}

bool is_pinned(const void *ptr) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

void parallel_copy(void *dst, const void *src, size_t bytes) {
    if (bytes < (4u << 20)) { std::memcpy(dst, src, bytes); return; }
    std::thread th[MTSB_TRACE_COPY_THREADS - 1];
    const size_t part = (bytes / MTSB_TRACE_COPY_THREADS + 63) & ~(size_t) 63;
    for (int t = 1; t < MTSB_TRACE_COPY_THREADS; ++t) {
        const size_t off = (size_t) t * part;
        if (off >= bytes) break;
        th[t - 1] = std::thread([=] { std::memcpy((char *) dst + off, (const char *) src + off, std::min(part, bytes - off)); });
    }
    std::memcpy(dst, src, std::min(part, bytes));
    for (auto &t : th) if (t.joinable()) t.join();
}
}  // namespace

extern "C" int mtsb_trace(mtsb_scene_t scene, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!rays || !hits)) || (mode != MTSB_TRACE_CLOSEST && mode != MTSB_TRACE_ANY)) { set_error("mtsb_trace: bad argument"); return MTSB_ERR_INVALID; }
    if (n == 0) return MTSB_OK;
    MTSB_CUDA(cudaSetDevice(s->device));
    TracePipe *pp;
    { std::lock_guard<std::mutex> g(g_pipe_mutex); pp = &g_pipes[s->device]; }      // map nodes are stable
    TracePipe &p = *pp;
    std::lock_guard<std::mutex> gd(p.m);                     // one host-buffer query per DEVICE at a time: its staging buffers are shared
    int rc = pipe_init(p);
    if (rc) return rc;
    const size_t chunk = std::min<size_t>(n, MTSB_TRACE_CHUNK);
    void *d_rays[2] = {nullptr, nullptr}, *d_hits[2] = {nullptr, nullptr};
    cudaError_t e = cudaSuccess;
    for (int i = 0; i < 2 && e == cudaSuccess; ++i) { e = dmalloc(&d_rays[i], chunk * sizeof(MtsbRay)); if (e == cudaSuccess) e = dmalloc(&d_hits[i], chunk * sizeof(MtsbHit)); }
    const bool pin_in = is_pinned(rays), pin_out = is_pinned(hits);
    const size_t n_chunks = (n + chunk - 1) / chunk;
    auto drain = [&](size_t c) {                             // hits of chunk c: pinned staging -> caller memory
        const int sl = (int) (c & 1);
        const size_t first = c * chunk, cnt = std::min(chunk, n - first);
        cudaError_t er = cudaEventSynchronize(p.ev_out[sl]);
        if (er == cudaSuccess && !pin_out) parallel_copy(hits + first, p.pin_hits[sl], cnt * sizeof(MtsbHit));
        return er;
    };
    for (size_t c = 0; c < n_chunks && e == cudaSuccess && rc == MTSB_OK; ++c) {
        const int sl = (int) (c & 1);
        const size_t first = c * chunk, cnt = std::min(chunk, n - first);
        if (c >= 2) e = drain(c - 2);                        // slot `sl` is free again once its previous hits have left
        if (e != cudaSuccess) break;
        const MtsbRay *src = rays + first;
        if (!pin_in) { parallel_copy(p.pin_rays[sl], src, cnt * sizeof(MtsbRay)); src = p.pin_rays[sl]; }
        e = cudaMemcpyAsync(d_rays[sl], src, cnt * sizeof(MtsbRay), cudaMemcpyHostToDevice, p.s_in);
        if (e == cudaSuccess) e = cudaEventRecord(p.ev_in[sl], p.s_in);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s->stream, p.ev_in[sl], 0);
        if (e != cudaSuccess) break;
        rc = launch_trace(s, (const RayIn *) d_rays[sl], nullptr, (uint32_t) cnt, mode, (HitOut *) d_hits[sl], s->stream);
        if (rc != MTSB_OK) break;
        e = cudaEventRecord(p.ev_k[sl], s->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(p.s_out, p.ev_k[sl], 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(pin_out ? (void *) (hits + first) : (void *) p.pin_hits[sl], d_hits[sl], cnt * sizeof(MtsbHit), cudaMemcpyDeviceToHost, p.s_out);
        if (e == cudaSuccess) e = cudaEventRecord(p.ev_out[sl], p.s_out);
    }
    if (e == cudaSuccess && rc == MTSB_OK) {
        if (n_chunks >= 2) e = drain(n_chunks - 2);
        if (e == cudaSuccess) e = drain(n_chunks - 1);
    }
    cudaStreamSynchronize(p.s_in); cudaStreamSynchronize(s->stream); cudaStreamSynchronize(p.s_out);     // nothing in flight when the blocks go back to the cache
    for (int i = 0; i < 2; ++i) { dfree(d_rays[i]); dfree(d_hits[i]); }
    if (rc != MTSB_OK) return rc;
    if (e != cudaSuccess) return cuda_fail(e, "mtsb_trace", __FILE__, __LINE__);
    return check_stack_overflow("mtsb_trace");
}

extern "C" int mtsb_gen_rays_device(mtsb_scene_t scene, int kind, size_t n, uint64_t seed, void *d_rays) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || !d_rays || (kind != 0 && kind != 1)) { set_error("mtsb_gen_rays_device: bad argument"); return MTSB_ERR_INVALID; }
    if (n == 0) return MTSB_OK;
    MTSB_CUDA(cudaSetDevice(s->device));
    uint32_t side = (uint32_t) ceil(sqrt((double) n));
    float3 c = make_float3(0.5f * (s->info.aabb_min[0] + s->info.aabb_max[0]), 0.5f * (s->info.aabb_min[1] + s->info.aabb_max[1]),
                           0.5f * (s->info.aabb_min[2] + s->info.aabb_max[2]));
    float dx = s->info.aabb_max[0] - c.x, dy = s->info.aabb_max[1] - c.y, dz = s->info.aabb_max[2] - c.z;
    float radius = sqrtf(dx * dx + dy * dy + dz * dz);
    k_gen_rays<<<grid_for(n, 256), 256, 0, s->stream>>>(s->v, kind, (uint64_t) n, seed, side, c, radius, (RayIn *) d_rays);
    s->launches++;
    MTSB_CUDA(cudaGetLastError());
    MTSB_CUDA(cudaStreamSynchronize(s->stream));
    return MTSB_OK;
}
