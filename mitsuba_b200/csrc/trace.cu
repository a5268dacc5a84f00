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

extern "C" int mtsb_trace(mtsb_scene_t scene, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!rays || !hits))) { set_error("mtsb_trace: bad argument"); return MTSB_ERR_INVALID; }
    if (n == 0) return MTSB_OK;
    MTSB_CUDA(cudaSetDevice(s->device));
    void *d_rays = nullptr, *d_hits = nullptr;
    MTSB_CUDA(dmalloc(&d_rays, n * sizeof(MtsbRay)));
    cudaError_t e = dmalloc(&d_hits, n * sizeof(MtsbHit));
    if (e != cudaSuccess) { dfree(d_rays); return cuda_fail(e, "cudaMalloc", __FILE__, __LINE__); }
    int rc = MTSB_OK;
    e = cudaMemcpyAsync(d_rays, rays, n * sizeof(MtsbRay), cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) rc = mtsb_trace_device(scene, d_rays, n, mode, d_hits, s->stream, nullptr);
    if (e == cudaSuccess && rc == MTSB_OK) e = cudaMemcpyAsync(hits, d_hits, n * sizeof(MtsbHit), cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess && rc == MTSB_OK) e = cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess || rc != MTSB_OK) cudaStreamSynchronize(s->stream);       // nothing in flight when the blocks go back to the cache
    dfree(d_rays); dfree(d_hits);
    if (e != cudaSuccess) return cuda_fail(e, "mtsb_trace", __FILE__, __LINE__);
    return rc;
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
