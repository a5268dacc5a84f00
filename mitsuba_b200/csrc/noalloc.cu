// noalloc.cu -- mtsb_trace_static: the host-buffer ray query of mtsb_trace for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  mtsb_trace stages the rays and the hit
// records in device blocks from the library's block cache (cudaMalloc on a miss) and takes its work counter from a ring
// that is cudaMalloc'ed on first use.  This variant runs the SAME kernel (k_trace, trace.cuh) with all of that scratch --
// ray staging, hit staging, the persistent kernel's work counter -- in ONE file-scope device array of fixed size:
//
//     MTSB_STATIC_TRACE_MAX_RAYS = 262144 rays  ->  256 B + 8 MiB (rays) + 4 MiB (hits) = 12 MiB + 256 B
//
// which every GPU that loads the library holds until the process exits.  The entry point calls no cudaMalloc, cudaFree
// or other allocating routine; it rejects n > MTSB_STATIC_TRACE_MAX_RAYS with MTSB_ERR_INVALID.  Every call shares the
// array (it keeps the last call's contents; the work counter is cleared on the stream before every launch), so calls
// that overlap on one GPU race: callers serialise them (the entry point itself synchronises the scene's stream before it
// returns, and a process-wide mutex keeps two host threads from interleaving their copies on one device).
#include "trace.cuh"
#include <mutex>

namespace mtsb {

#define MTSB_STATIC_TRACE_BYTES (256 + (size_t) MTSB_STATIC_TRACE_MAX_RAYS * (sizeof(RayIn) + sizeof(HitOut)))
static __device__ __align__(256) unsigned char g_static_trace_scratch[MTSB_STATIC_TRACE_BYTES];

// In host code the array's name is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
#define MTSB_MAX_DEVICES 64
static void *g_scratch_addr[MTSB_MAX_DEVICES];
static std::mutex g_scratch_mutex;

static int static_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= MTSB_MAX_DEVICES) { set_error("mtsb_trace_static: device index out of range"); return MTSB_ERR_INVALID; }
    if (!g_scratch_addr[device]) {
        void *p = nullptr;
        MTSB_CUDA(cudaGetSymbolAddress(&p, g_static_trace_scratch));
        if (!p) { set_error("mtsb_trace_static: cudaGetSymbolAddress returned NULL"); return MTSB_ERR_CUDA; }
        g_scratch_addr[device] = p;
    }
    *out = static_cast<unsigned char *>(g_scratch_addr[device]);
    return MTSB_OK;
}

template <bool ANY>
static int launch_static(Scene *s, const RayIn *d_rays, uint32_t n, HitOut *d_hits, uint32_t *d_counter) {
    TraceArgs a;
    a.nodes = s->v.nodes; a.tris = s->v.tris; a.n_nodes = s->v.n_nodes; a.n_tri_records = (uint32_t) s->info.n_tris;
    for (int k = 0; k < 3; ++k) { a.box_lo[k] = s->v.box_lo[k]; a.box_hi[k] = s->v.box_hi[k]; }
    a.rays = d_rays; a.d_count = nullptr; a.n_max = n;
    a.refill = MTSB_REFILL; a.prmt_bias = 0x47000000u;
    a.work_counter = d_counter;
    int grid = trace_grid(s, (const void *) k_trace<ANY, StoreHit>);        // occupancy query, cached per kernel: allocates nothing on the device
    if (grid <= 0) return MTSB_ERR_CUDA;
    const uint32_t need = (n + MTSB_TRACE_BLOCK - 1) / MTSB_TRACE_BLOCK;
    if (need < (uint32_t) grid) grid = (int) need;
    k_trace<ANY, StoreHit><<<grid, MTSB_TRACE_BLOCK, 0, s->stream>>>(a, StoreHit{d_hits});
    s->launches++;
    MTSB_CUDA(cudaGetLastError());
    return MTSB_OK;
}

}  // namespace mtsb

using namespace mtsb;

extern "C" int mtsb_trace_static_max_rays(void) { return MTSB_STATIC_TRACE_MAX_RAYS; }

extern "C" int mtsb_trace_static(mtsb_scene_t scene, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!rays || !hits)) || (mode != MTSB_TRACE_CLOSEST && mode != MTSB_TRACE_ANY)) { set_error("mtsb_trace_static: bad argument"); return MTSB_ERR_INVALID; }
    if (n > (size_t) MTSB_STATIC_TRACE_MAX_RAYS) {
        set_error("mtsb_trace_static: n exceeds MTSB_STATIC_TRACE_MAX_RAYS (262144 rays per call: the scratch is a fixed 12 MiB device array); split the batch or use mtsb_trace");
        return MTSB_ERR_INVALID;
    }
    if (n == 0) return MTSB_OK;
    std::lock_guard<std::mutex> g(g_scratch_mutex);
    MTSB_CUDA(cudaSetDevice(s->device));
    unsigned char *base = nullptr;
    int rc = static_scratch(s->device, &base);
    if (rc) return rc;
    uint32_t *d_counter = reinterpret_cast<uint32_t *>(base);
    RayIn *d_rays = reinterpret_cast<RayIn *>(base + 256);
    HitOut *d_hits = reinterpret_cast<HitOut *>(base + 256 + (size_t) MTSB_STATIC_TRACE_MAX_RAYS * sizeof(RayIn));
    MTSB_CUDA(cudaMemsetAsync(d_counter, 0, 256, s->stream));                 // the one part that must start at zero
    MTSB_CUDA(cudaMemcpyAsync(d_rays, rays, n * sizeof(MtsbRay), cudaMemcpyHostToDevice, s->stream));
    rc = mode == MTSB_TRACE_ANY ? launch_static<true>(s, d_rays, (uint32_t) n, d_hits, d_counter)
                                : launch_static<false>(s, d_rays, (uint32_t) n, d_hits, d_counter);
    cudaError_t e = cudaSuccess;
    if (rc == MTSB_OK) e = cudaMemcpyAsync(hits, d_hits, n * sizeof(MtsbHit), cudaMemcpyDeviceToHost, s->stream);
    const cudaError_t e2 = cudaStreamSynchronize(s->stream);
    if (rc != MTSB_OK) return rc;
    if (e != cudaSuccess) return cuda_fail(e, "mtsb_trace_static", __FILE__, __LINE__);
    if (e2 != cudaSuccess) return cuda_fail(e2, "mtsb_trace_static", __FILE__, __LINE__);
    return MTSB_OK;
}
