// api.cu -- C ABI: library state, scene creation / destruction (include/mtsb200.h).
// mtsb_scene_create replaces Scene::initialize (reference src/librender/scene.cpp:322-384):
// it copies the flattened scene to the device, builds the sampling tables the emitters need
// (DiscreteDistribution, include/mitsuba/core/pmf.h:101-188; TriMesh::prepareSamplingTable,
// src/librender/trimesh.cpp:388-403) and builds the BVH on the GPU.
#include "scene.h"
#include "trace.cuh"
#include "qmc_host.h"
#include "texture_host.h"
#include <map>
#include <mutex>
#include <cmath>
#include <cstring>
#include <thread>
#include <chrono>
#include <algorithm>

namespace mtsb {

static thread_local std::string g_error;
// The device a thread creates its scenes on: set per THREAD by mtsb_init (mtsb_multi_create runs one host thread per
// GPU); a thread that never called mtsb_init inherits the device of the last mtsb_init of the process.
static thread_local int g_device = -1;
static std::atomic<int> g_default_device{-1};
static int g_sm_count = 0;

void set_error(const std::string &s) { g_error = s; }
int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    char buf[512];
    snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d in %s", (int) e, cudaGetErrorString(e), file, line, what);
    g_error = buf;
    cudaGetLastError();
    return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? MTSB_ERR_NO_DEVICE : MTSB_ERR_CUDA;
}

static int ensure_device() {
    if (g_device >= 0) return MTSB_OK;
    return mtsb_init(g_default_device.load());
}

// ring of work counters for the persistent kernels
static std::mutex g_counter_mutex;
static std::map<int, std::pair<uint32_t *, uint32_t>> g_counters;   // device -> (ring, next)
static const uint32_t kCounterRing = 4096;

uint32_t *trace_work_counter(Scene *s, cudaStream_t st) {
    std::lock_guard<std::mutex> g(g_counter_mutex);
    auto &e = g_counters[s->device];
    if (!e.first) {
        if (cudaMalloc((void **) &e.first, kCounterRing * sizeof(uint32_t)) != cudaSuccess) { set_error("cudaMalloc(work counters) failed"); return nullptr; }
        e.second = 0;
    }
    uint32_t *p = e.first + (e.second++ % kCounterRing);
    if (cudaMemsetAsync(p, 0, sizeof(uint32_t), st) != cudaSuccess) { set_error("cudaMemsetAsync(work counter) failed"); return nullptr; }
    return p;
}

int device_sm_count() { return g_sm_count > 0 ? g_sm_count : 148; }

// ---- device block cache -------------------------------------------------------------------------
// cudaMalloc / cudaFree cost tens to hundreds of milliseconds per scene once the process holds a few
// GB (every call updates the page tables of the whole context and cudaFree synchronises the device).
// A frame loop that re-creates its scene per frame (the e2e path of bench.py, an animation) would pay
// that every frame, so blocks freed by mtsb_scene_destroy and the build / query temporaries are kept
// per device and size class and handed out again.  Callers synchronise their streams before dfree().
namespace {
struct BlockCache {
    std::mutex m;
    std::multimap<size_t, void *> free_blocks;       // size class -> block
    std::map<void *, size_t> live;                   // block -> size class
    size_t cached_bytes = 0;
};
std::map<int, BlockCache> g_block_cache;
std::mutex g_block_cache_mutex;
size_t cache_limit_bytes() {
    static const size_t v = [] { const char *e = getenv("MTSB_CACHE_MB"); return (size_t) (e ? std::max(0, atoi(e)) : 8192) << 20; }();
    return v;
}
size_t size_class(size_t bytes) {
    if (bytes < 16) bytes = 16;
    const size_t g = bytes <= (1u << 20) ? 512 : (size_t) 2 << 20;       // cudaMalloc granularity above 1 MiB is 2 MiB
    return (bytes + g - 1) / g * g;
}
BlockCache &cache_of_current_device() {
    int dev = 0; cudaGetDevice(&dev);
    std::lock_guard<std::mutex> g(g_block_cache_mutex);
    return g_block_cache[dev];
}
}  // namespace

cudaError_t dmalloc(void **p, size_t bytes) {
    BlockCache &c = cache_of_current_device();
    const size_t sc = size_class(bytes);
    std::lock_guard<std::mutex> g(c.m);
    auto it = c.free_blocks.find(sc);
    if (it != c.free_blocks.end()) {
        *p = it->second; c.free_blocks.erase(it); c.cached_bytes -= sc; c.live[*p] = sc;
        return cudaSuccess;
    }
    cudaError_t e = cudaMalloc(p, sc);
    if (e != cudaSuccess && !c.free_blocks.empty()) {                     // out of memory: give the cache back and retry
        cudaGetLastError();
        for (auto &b : c.free_blocks) cudaFree(b.second);
        c.free_blocks.clear(); c.cached_bytes = 0;
        e = cudaMalloc(p, sc);
    }
    if (e == cudaSuccess) c.live[*p] = sc;
    return e;
}
void dfree(void *p) {
    if (!p) return;
    BlockCache &c = cache_of_current_device();
    std::lock_guard<std::mutex> g(c.m);
    auto it = c.live.find(p);
    if (it == c.live.end()) { cudaFree(p); return; }
    const size_t sc = it->second;
    c.live.erase(it);
    if (c.cached_bytes + sc > cache_limit_bytes()) { cudaFree(p); return; }
    c.free_blocks.emplace(sc, p); c.cached_bytes += sc;
}
void dcache_release() {
    BlockCache &c = cache_of_current_device();
    std::lock_guard<std::mutex> g(c.m);
    for (auto &b : c.free_blocks) cudaFree(b.second);
    c.free_blocks.clear(); c.cached_bytes = 0;
}

// ---- staged uploads --------------------------------------------------------------------------------------------------
// cudaMemcpy from pageable memory moves ~5 GB/s (the driver stages it through one bounce buffer on one thread).  Scene
// arrays of a 10 M-triangle scene are ~300 MB, and mtsb_scene_create is inside the end-to-end step, so large arrays are
// staged here: two pinned 32 MiB buffers per device, filled by a few host threads while the previous chunk is on the bus.
namespace {
struct Stager { std::mutex m; bool ready = false; cudaStream_t st = nullptr; void *pin[2] = {nullptr, nullptr}; cudaEvent_t ev[2] = {nullptr, nullptr}; };
std::mutex g_stager_mutex;
std::map<int, Stager> g_stagers;
const size_t kStageChunk = (size_t) 32 << 20;
void stage_copy(void *dst, const void *src, size_t bytes) {
    const int nt = 4;
    if (bytes < ((size_t) 4 << 20)) { std::memcpy(dst, src, bytes); return; }
    std::thread th[nt - 1];
    const size_t part = (bytes / nt + 63) & ~(size_t) 63;
    for (int t = 1; t < nt; ++t) {
        const size_t off = (size_t) t * part;
        if (off >= bytes) break;
        th[t - 1] = std::thread([=] { std::memcpy((char *) dst + off, (const char *) src + off, std::min(part, bytes - off)); });
    }
    std::memcpy(dst, src, std::min(part, bytes));
    for (auto &t : th) if (t.joinable()) t.join();
}
}  // namespace

int staged_h2d(void *dst, const void *src, size_t bytes) {
    if (bytes < ((size_t) 1 << 20)) { MTSB_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice)); return MTSB_OK; }
    int dev = 0; MTSB_CUDA(cudaGetDevice(&dev));
    Stager *sgp;
    { std::lock_guard<std::mutex> g(g_stager_mutex); sgp = &g_stagers[dev]; }        // map nodes are stable; the copy itself is locked per device,
    Stager &sg = *sgp;                                                              // so the per-GPU threads of mtsb_multi_create upload side by side
    std::lock_guard<std::mutex> gd(sg.m);
    if (!sg.ready) {
        MTSB_CUDA(cudaStreamCreateWithFlags(&sg.st, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) { MTSB_CUDA(cudaHostAlloc(&sg.pin[i], kStageChunk, cudaHostAllocDefault)); MTSB_CUDA(cudaEventCreateWithFlags(&sg.ev[i], cudaEventDisableTiming)); }
        sg.ready = true;
    }
    size_t c = 0;
    for (size_t off = 0; off < bytes; off += kStageChunk, ++c) {
        const int sl = (int) (c & 1);
        const size_t cnt = std::min(kStageChunk, bytes - off);
        if (c >= 2) MTSB_CUDA(cudaEventSynchronize(sg.ev[sl]));
        stage_copy(sg.pin[sl], (const char *) src + off, cnt);
        MTSB_CUDA(cudaMemcpyAsync((char *) dst + off, sg.pin[sl], cnt, cudaMemcpyHostToDevice, sg.st));
        MTSB_CUDA(cudaEventRecord(sg.ev[sl], sg.st));
    }
    MTSB_CUDA(cudaStreamSynchronize(sg.st));
    return MTSB_OK;
}

int trace_grid(Scene *s, const void *kernel) {
    static std::mutex m;
    static std::map<const void *, int> cache;
    std::lock_guard<std::mutex> g(m);
    auto it = cache.find(kernel);
    if (it != cache.end()) return it->second;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, MTSB_TRACE_BLOCK, 0) != cudaSuccess || per_sm <= 0) {
        set_error("occupancy query failed"); cudaGetLastError(); return -1;
    }
    int grid = per_sm * (g_sm_count > 0 ? g_sm_count : 148);
    { const char *e = getenv("MTSB_GRID_DIV"); if (e && atoi(e) > 1) grid = std::max(1, grid / atoi(e)); }     // experiments: share the SMs between lanes
    cache[kernel] = grid;
    return grid;
}

}  // namespace mtsb

using namespace mtsb;

extern "C" int mtsb_abi_version(void) { return MTSB_ABI_VERSION; }
extern "C" const char *mtsb_last_error(void) { return g_error.c_str(); }

extern "C" int mtsb_device_count(void) {
    // a process that starts while another one is still tearing its context down has been seen to get a transient
    // failure here; ask a few times before reporting "no device"
    for (int attempt = 0; attempt < 4; ++attempt) {
        int n = 0;
        if (cudaGetDeviceCount(&n) == cudaSuccess && n > 0) return n;
        cudaGetLastError();
        if (attempt < 3) std::this_thread::sleep_for(std::chrono::milliseconds(250));
    }
    return 0;
}

extern "C" int mtsb_init(int device) {
    int n = mtsb_device_count();
    if (n <= 0) { set_error("no CUDA device present (this library has no CPU fallback)"); return MTSB_ERR_NO_DEVICE; }
    if (device < 0) {
        if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    }
    if (device >= n) { set_error("mtsb_init: device index out of range"); return MTSB_ERR_INVALID; }
    MTSB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    MTSB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) { set_error(std::string("device is not sm_100-class: ") + prop.name); return MTSB_ERR_NO_DEVICE; }
    g_device = device; g_default_device.store(device); g_sm_count = prop.multiProcessorCount;
    MTSB_CUDA(cudaFree(0));
    return MTSB_OK;
}

static int validate(const MtsbSceneDesc *d) {
    if (!d) { set_error("scene description is NULL"); return MTSB_ERR_INVALID; }
    if (d->abi_version != MTSB_ABI_VERSION) { set_error("scene description: ABI version mismatch"); return MTSB_ERR_INVALID; }
    if (d->n_tris > 0xFFFFFFF0ull / 3 || d->n_vertices > 0xFFFFFFF0ull) { set_error("scene too large for 32-bit primitive ids"); return MTSB_ERR_INVALID; }
    if ((d->n_shapes && !d->shapes) || (d->n_materials && !d->materials) || (d->n_emitters && !d->emitters) ||
        (d->n_vertices && !d->positions) || (d->n_tris && !d->indices)) { set_error("scene description: NULL array"); return MTSB_ERR_INVALID; }
    uint64_t tri_sum = 0;
    for (uint32_t i = 0; i < d->n_shapes; ++i) {
        const MtsbShape &sh = d->shapes[i];
        if ((uint64_t) sh.first_tri != tri_sum) { set_error("shapes must tile the triangle array in order"); return MTSB_ERR_INVALID; }
        tri_sum += sh.n_tris;
        if (sh.material >= (int32_t) d->n_materials || sh.emitter >= (int32_t) d->n_emitters) { set_error("shape refers to a missing material/emitter"); return MTSB_ERR_INVALID; }
        if ((uint64_t) sh.first_vertex + sh.n_vertices > d->n_vertices) { set_error("shape vertex range out of bounds"); return MTSB_ERR_INVALID; }
        if (sh.has_normals && !d->normals) { set_error("shape has_normals but normals == NULL"); return MTSB_ERR_INVALID; }
    }
    if (tri_sum != d->n_tris) { set_error("shape triangle counts do not add up to n_tris"); return MTSB_ERR_INVALID; }
    for (uint32_t i = 0; i < d->n_shapes; ++i) {
        const MtsbShape &sh = d->shapes[i];
        if (sh.emitter < 0) continue;
        if (d->emitters[sh.emitter].type != MTSB_EMITTER_AREA || d->emitters[sh.emitter].shape != (int32_t) i) { set_error("shape refers to an emitter that is not its area emitter"); return MTSB_ERR_INVALID; }
        // the area cdf of an emitter shape is built on the host from these indices: check them here (all other indices are checked on the device)
        for (uint64_t j = 0; j < 3ull * sh.n_tris; ++j) {
            const uint32_t ix = d->indices[3ull * sh.first_tri + j];
            if (ix < sh.first_vertex || ix >= sh.first_vertex + sh.n_vertices) { set_error("triangle index outside its shape's vertex range"); return MTSB_ERR_INVALID; }
        }
    }
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const MtsbMaterial &m = d->materials[i];
        if (m.type > MTSB_BSDF_PLASTIC || m.distribution > MTSB_DISTR_GGX || m.back >= (int32_t) d->n_materials) { set_error("bad material record"); return MTSB_ERR_INVALID; }
        if (m.tex_reflectance > d->n_textures || m.tex_transmittance > d->n_textures || m.tex_alpha > d->n_textures) { set_error("material refers to a missing texture"); return MTSB_ERR_INVALID; }
    }
    if (d->n_textures && !d->textures) { set_error("scene description: NULL texture array"); return MTSB_ERR_INVALID; }
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const MtsbTexture &t = d->textures[i];
        if (t.type > MTSB_TEX_GRID) { set_error("bad texture type"); return MTSB_ERR_INVALID; }
        if (t.type == MTSB_TEX_BITMAP) {
            if (t.filter > MTSB_TEXFILTER_EWA || t.wrap_u > MTSB_WRAP_ONE || t.wrap_v > MTSB_WRAP_ONE) { set_error("bitmap texture: bad filter type / wrap mode"); return MTSB_ERR_INVALID; }
            if (!t.texels || t.width <= 0 || t.height <= 0 || t.width > 0xFFFF || t.height > 0xFFFF || (t.channels != 1 && t.channels != 3)) { set_error("bitmap texture: bad image"); return MTSB_ERR_INVALID; }
        }
    }
    int n_env = 0;
    for (uint32_t i = 0; i < d->n_emitters; ++i) {
        const MtsbEmitter &e = d->emitters[i];
        if (e.type > MTSB_EMITTER_ENVMAP) { set_error("bad emitter type"); return MTSB_ERR_INVALID; }
        if (e.type == MTSB_EMITTER_ENVMAP) {
            const MtsbEnvmap *em = d->envmap;
            if (!em || !em->rgb || em->width <= 0 || em->height <= 0) { set_error("envmap emitter without image data"); return MTSB_ERR_INVALID; }
            if (em->width > 0xFFFF || em->height > 0xFFFF) { set_error("Environment maps images must be smaller than 65536 pixels in width and height"); return MTSB_ERR_INVALID; }
        }
        if (e.type == MTSB_EMITTER_AREA && (e.shape < 0 || e.shape >= (int32_t) d->n_shapes)) { set_error("area emitter without a shape"); return MTSB_ERR_INVALID; }
        if (e.type == MTSB_EMITTER_AREA && d->shapes[e.shape].emitter != (int32_t) i) { set_error("area emitter and its shape do not refer to each other"); return MTSB_ERR_INVALID; }
        if ((e.type == MTSB_EMITTER_CONSTANT || e.type == MTSB_EMITTER_ENVMAP) && ++n_env > 1) { set_error("only one environment emitter is allowed (scene.cpp:575-579)"); return MTSB_ERR_INVALID; }
    }
    const MtsbSensor &se = d->sensor;
    if (se.film_width <= 0 || se.film_height <= 0 || se.crop_width <= 0 || se.crop_height <= 0 || se.crop_x < 0 || se.crop_y < 0 ||
        se.crop_x + se.crop_width > se.film_width || se.crop_y + se.crop_height > se.film_height) { set_error("bad film / crop size"); return MTSB_ERR_INVALID; }
    if (!(se.rfilter_radius > 0) || se.rfilter_radius > 7.5f || se.rfilter > MTSB_RFILTER_BOX) { set_error("bad reconstruction filter"); return MTSB_ERR_INVALID; }
    if (d->sampler.type > MTSB_SAMPLER_SOBOL) { set_error("bad sampler type"); return MTSB_ERR_INVALID; }
    if (!(se.fov_x_deg > 0 && se.fov_x_deg < 180) || !(se.near_clip > 0) || !(se.far_clip > se.near_clip)) { set_error("bad sensor parameters"); return MTSB_ERR_INVALID; }
    return MTSB_OK;
}

// every triangle index must lie inside the vertex range of the triangle's shape (the kernels index positions / normals /
// texcoords with them); checked on the device right after the upload: one pass at HBM speed instead of a host loop
__global__ void k_check_indices(const uint32_t *indices, const uint32_t *tri_shape, const MtsbShape *shapes, uint64_t n_tris, uint32_t *bad) {
    const uint64_t prim = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (prim >= n_tris) return;
    const MtsbShape sh = shapes[tri_shape[prim]];
    bool ok = true;
    for (int k = 0; k < 3; ++k) { const uint32_t ix = indices[3 * prim + k]; ok &= ix >= sh.first_vertex && ix < sh.first_vertex + sh.n_vertices; }
    if (!ok) atomicOr(bad, 1u);
}

__global__ void k_fill_tri_shape(const MtsbShape *shapes, uint32_t n_shapes, uint64_t n_tris, uint32_t *tri_shape) {
    const uint64_t prim = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (prim >= n_tris) return;
    uint32_t lo = 0, hi = n_shapes;                 // last shape with first_tri <= prim (empty shapes share a first_tri: take the one that owns it)
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if ((uint64_t) shapes[mid].first_tri <= prim) lo = mid; else hi = mid; }
    while (lo > 0 && shapes[lo].n_tris == 0) --lo;
    tri_shape[prim] = lo;
}

static float rfilter_eval(const MtsbSensor &se, float x) {
    if (se.rfilter == MTSB_RFILTER_GAUSSIAN) {      // gaussian.cpp:50-55
        float alpha = -1.0f / (2.0f * se.rfilter_stddev * se.rfilter_stddev);
        return std::max(0.0f, (float) std::exp((double) (alpha * x * x)) - (float) std::exp((double) (alpha * se.rfilter_radius * se.rfilter_radius)));   // math::fastexp (math.h:175-199)
    }
    return std::fabs(x) <= se.rfilter_radius ? 1.0f : 0.0f;   // box.cpp:36-50
}

extern "C" int mtsb_scene_create(const MtsbSceneDesc *d, mtsb_scene_t *out) {
    if (!out) { set_error("mtsb_scene_create: out == NULL"); return MTSB_ERR_INVALID; }
    *out = nullptr;
    int rc = validate(d);
    if (rc) return rc;
    rc = ensure_device();
    if (rc) return rc;
    Scene *s = new Scene();
    s->device = g_device;
    auto fail = [&](int code) { mtsb_scene_destroy(reinterpret_cast<mtsb_scene_t>(s)); return code; };
    if (cudaSetDevice(s->device) != cudaSuccess) return fail(MTSB_ERR_CUDA);
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); return fail(MTSB_ERR_CUDA); }
    SceneView &v = s->v;
    v.n_shapes = d->n_shapes; v.n_materials = d->n_materials; v.n_emitters = d->n_emitters;
    v.sensor = d->sensor; v.integrator = d->integrator;
    if ((rc = dev_upload(s, &v.positions, d->positions, 3 * (size_t) d->n_vertices))) return fail(rc);
    if (d->normals) { if ((rc = dev_upload(s, &v.normals, d->normals, 3 * (size_t) d->n_vertices))) return fail(rc); } else v.normals = nullptr;
    if (d->texcoords) { if ((rc = dev_upload(s, &v.texcoords, d->texcoords, 2 * (size_t) d->n_vertices))) return fail(rc); } else v.texcoords = nullptr;
    if ((rc = dev_upload(s, &v.indices, d->indices, 3 * (size_t) d->n_tris))) return fail(rc);
    if ((rc = dev_upload(s, &v.shapes, d->shapes, (size_t) d->n_shapes))) return fail(rc);
    if ((rc = dev_upload(s, &v.materials, d->materials, (size_t) d->n_materials))) return fail(rc);
    s->h_materials.assign(d->materials, d->materials + d->n_materials);
    s->h_shape_material.resize(d->n_shapes);
    for (uint32_t i = 0; i < d->n_shapes; ++i) s->h_shape_material[i] = d->shapes[i].material;
    if ((rc = dev_upload(s, &v.emitters, d->emitters, (size_t) d->n_emitters))) return fail(rc);
    {   // primitive -> shape map (ShapeKDTree::findShape, skdtree.h:205-210, as a table), filled on the device: shapes tile
        // the triangle array in order (validate()), so a primitive's shape is found by bisection over first_tri
        uint32_t *ts = nullptr;
        if ((rc = dev_alloc(s, &ts, (size_t) d->n_tris))) return fail(rc);
        if (d->n_tris) {
            k_fill_tri_shape<<<grid_for(d->n_tris, 256), 256, 0, s->stream>>>(v.shapes, d->n_shapes, d->n_tris, ts);
            s->launches++;
            if (cudaGetLastError() != cudaSuccess) return fail(MTSB_ERR_CUDA);
        }
        v.tri_shape = ts;
    }
    {   // emitter selection pmf (scene.cpp:375-381) and per-mesh area pmfs (trimesh.cpp:388-403)
        std::vector<float> cdf(1, 0.0f);
        v.env_emitter = -1;
        for (uint32_t i = 0; i < d->n_emitters; ++i) {
            cdf.push_back(cdf.back() + d->emitters[i].sampling_weight);
            if (d->emitters[i].type == MTSB_EMITTER_CONSTANT || d->emitters[i].type == MTSB_EMITTER_ENVMAP) v.env_emitter = (int) i;
        }
        float sum = cdf.back();
        v.emitter_norm = 0.0f;
        if (d->n_emitters && sum > 0) {
            v.emitter_norm = 1.0f / sum;
            for (size_t i = 1; i < cdf.size(); ++i) cdf[i] *= v.emitter_norm;
            cdf.back() = 1.0f;
        }
        if ((rc = dev_upload(s, &v.emitter_cdf, cdf.data(), cdf.size()))) return fail(rc);
        std::vector<float> acdf; std::vector<uint32_t> aoff(d->n_shapes, 0xFFFFFFFFu); std::vector<float> inv_area(d->n_shapes, 0.0f);
        for (uint32_t si = 0; si < d->n_shapes; ++si) {
            const MtsbShape &sh = d->shapes[si];
            if (sh.emitter < 0) continue;
            aoff[si] = (uint32_t) acdf.size();
            size_t base = acdf.size();
            acdf.push_back(0.0f);
            for (uint32_t j = 0; j < sh.n_tris; ++j) {
                const uint32_t *ix = d->indices + 3 * ((size_t) sh.first_tri + j);
                const float *p0 = d->positions + 3 * (size_t) ix[0], *p1 = d->positions + 3 * (size_t) ix[1], *p2 = d->positions + 3 * (size_t) ix[2];
                volatile float ax = p1[0] - p0[0], ay = p1[1] - p0[1], az = p1[2] - p0[2];
                volatile float bx = p2[0] - p0[0], by = p2[1] - p0[1], bz = p2[2] - p0[2];
                volatile float cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
                volatile float area = 0.5f * std::sqrt(cx * cx + cy * cy + cz * cz);      // Triangle::surfaceArea, triangle.cpp:64-70
                acdf.push_back(acdf.back() + area);
            }
            float total = acdf.back();
            if (total > 0) {
                float norm = 1.0f / total;
                for (size_t i = base + 1; i < acdf.size(); ++i) acdf[i] *= norm;
                acdf.back() = 1.0f;
            }
            inv_area[si] = 1.0f / total;
        }
        if ((rc = dev_upload(s, &v.area_cdf, acdf.data(), acdf.size()))) return fail(rc);
        if ((rc = dev_upload(s, &v.area_cdf_offset, aoff.data(), aoff.size()))) return fail(rc);
        if ((rc = dev_upload(s, &v.inv_area, inv_area.data(), inv_area.size()))) return fail(rc);
    }
    v.envmap = nullptr;
    if (v.env_emitter >= 0 && d->emitters[v.env_emitter].type == MTSB_EMITTER_ENVMAP) {
        // EnvironmentMap::configure (envmap.cpp:260-329): half-precision texels, marginal / conditional CDFs over
        // luminance weighted by sin(theta), normalisation.  Arithmetic types follow the reference: Float sums,
        // float tables, M_PI as a double constant.
        const MtsbEnvmap &em = *d->envmap;
        const int W = em.width, H = em.height;
        std::vector<HalfTexel> tex((size_t) W * H);
        std::vector<float> cdf_cols((size_t) (W + 1) * H), cdf_rows((size_t) H + 1), row_w((size_t) H);
        size_t colPos = 0, rowPos = 0;
        volatile float rowSum = 0.0f;
        cdf_rows[rowPos++] = 0;
        for (int y = 0; y < H; ++y) {
            volatile float colSum = 0;
            cdf_cols[colPos++] = 0;
            for (int x = 0; x < W; ++x) {
                const float *p = em.rgb + 3 * ((size_t) y * W + x);
                const uint32_t hr = __half_as_ushort(__float2half_rn(p[0])), hg = __half_as_ushort(__float2half_rn(p[1])), hb = __half_as_ushort(__float2half_rn(p[2]));
                HalfTexel t; t.x = hr | (hg << 16); t.y = hb;
                tex[(size_t) y * W + x] = t;
                volatile float lum = half_bits_to_float(hr) * 0.212671f;                    // Spectrum::getLuminance, spectrum.h:724-727
                lum = lum + half_bits_to_float(hg) * 0.715160f; lum = lum + half_bits_to_float(hb) * 0.072169f;
                colSum = colSum + lum;
                cdf_cols[colPos++] = colSum;
            }
            const float normalization = 1.0f / colSum;
            for (int x = 1; x < W; ++x) cdf_cols[colPos - x - 1] *= normalization;
            cdf_cols[colPos - 1] = 1.0f;
            const float weight = (float) std::sin((double) ((float) y + 0.5f) * 3.14159265358979323846 / (double) H);
            row_w[y] = weight;
            volatile float cw = colSum * weight;
            rowSum = rowSum + cw;
            cdf_rows[rowPos++] = rowSum;
        }
        const float normalization = 1.0f / rowSum;
        for (int y = 1; y < H; ++y) cdf_rows[rowPos - y - 1] *= normalization;
        cdf_rows[rowPos - 1] = 1.0f;
        if (rowSum == 0) { set_error("The environment map is completely black -- this is not allowed."); return fail(MTSB_ERR_INVALID); }
        if (!std::isfinite(rowSum)) { set_error("The environment map contains an invalid floating point value (nan/inf) -- giving up."); return fail(MTSB_ERR_INVALID); }
        EnvmapView ev{};
        ev.width = W; ev.height = H; ev.scale = em.scale;
        ev.normalization = (float) (1.0f / ((double) rowSum * (2 * 3.14159265358979323846 / W) * (3.14159265358979323846 / H)));
        ev.pixel_size_x = (float) (2 * 3.14159265358979323846 / W); ev.pixel_size_y = (float) (3.14159265358979323846 / H);
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { ev.to_world[3 * r + c] = em.to_world[4 * r + c]; ev.to_local[3 * r + c] = em.to_local[4 * r + c]; }
        {   // the MIP pyramid of the map for filtered lookups (envmap.cpp:140-178: Lanczos-2 downsampling, repeat in u, clamp
            // in v, EWA with maxAnisotropy 10) -- the same builder and lookup code as the bitmap texture
            MtsbTexture mt;
            std::memset(&mt, 0, sizeof(mt));
            mt.type = MTSB_TEX_BITMAP; mt.filter = MTSB_TEXFILTER_EWA; mt.wrap_u = MTSB_WRAP_REPEAT; mt.wrap_v = MTSB_WRAP_CLAMP;
            mt.uv_scale[0] = mt.uv_scale[1] = 1.0f; mt.scale[0] = mt.scale[1] = mt.scale[2] = 1.0f;
            mt.max_anisotropy = 10.0f; mt.width = W; mt.height = H; mt.channels = 3; mt.texels = em.rgb;
            TextureView tv;
            std::memset(&tv, 0, sizeof(tv));
            tv.type = mt.type; tv.filter = mt.filter; tv.wrap_u = mt.wrap_u; tv.wrap_v = mt.wrap_v;
            for (int k = 0; k < 2; ++k) tv.uv_scale[k] = 1.0f;
            for (int k = 0; k < 3; ++k) tv.scale[k] = 1.0f;
            tv.max_anisotropy = mt.max_anisotropy;
            std::vector<HalfTexel> pyr;
            build_texture_levels(mt, tv, pyr);
            float lut[MTSB_MIPMAP_LUT_SIZE];
            for (int i = 0; i < MTSB_MIPMAP_LUT_SIZE; ++i) {
                const float r2 = (float) i / (float) (MTSB_MIPMAP_LUT_SIZE - 1);
                lut[i] = (float) std::exp((double) (-2.0f * r2)) - (float) std::exp((double) -2.0f);     // mipmap.h:296-302
            }
            if ((rc = dev_upload(s, &tv.weight_lut, lut, (size_t) MTSB_MIPMAP_LUT_SIZE))) return fail(rc);
            if ((rc = dev_upload(s, &tv.texels, pyr.data(), pyr.size()))) return fail(rc);
            if ((rc = dev_upload(s, &ev.mip, &tv, 1))) return fail(rc);
        }
        if ((rc = dev_upload(s, &ev.texels, tex.data(), tex.size()))) return fail(rc);
        if ((rc = dev_upload(s, &ev.cdf_rows, cdf_rows.data(), cdf_rows.size()))) return fail(rc);
        if ((rc = dev_upload(s, &ev.cdf_cols, cdf_cols.data(), cdf_cols.size()))) return fail(rc);
        if ((rc = dev_upload(s, &ev.row_weights, row_w.data(), row_w.size()))) return fail(rc);
        if ((rc = dev_upload(s, &v.envmap, &ev, 1))) return fail(rc);
    }
    v.textures = nullptr; v.n_textures = d->n_textures;
    if (d->n_textures) {
        // texture records; a bitmap brings its MIP pyramid (texture_host.h) in half precision and, for EWA, the Gaussian
        // weight table of mipmap.h:296-302 (shared by all textures)
        std::vector<TextureView> tvs(d->n_textures);
        std::vector<HalfTexel> texels;
        float lut[MTSB_MIPMAP_LUT_SIZE];
        for (int i = 0; i < MTSB_MIPMAP_LUT_SIZE; ++i) {
            const float r2 = (float) i / (float) (MTSB_MIPMAP_LUT_SIZE - 1);
            lut[i] = (float) std::exp((double) (-2.0f * r2)) - (float) std::exp((double) -2.0f);     // math::fastexp, math.h:175-199
        }
        const float *d_lut = nullptr;
        if ((rc = dev_upload(s, &d_lut, lut, (size_t) MTSB_MIPMAP_LUT_SIZE))) return fail(rc);
        for (uint32_t i = 0; i < d->n_textures; ++i) {
            const MtsbTexture &t = d->textures[i];
            TextureView &tv = tvs[i];
            std::memset(&tv, 0, sizeof(tv));
            tv.type = t.type; tv.filter = t.filter; tv.wrap_u = t.wrap_u; tv.wrap_v = t.wrap_v;
            for (int k = 0; k < 2; ++k) { tv.uv_scale[k] = t.uv_scale[k]; tv.uv_offset[k] = t.uv_offset[k]; }
            for (int k = 0; k < 3; ++k) { tv.color0[k] = t.color0[k]; tv.color1[k] = t.color1[k]; tv.scale[k] = t.scale[k]; }
            tv.line_width = t.line_width;
            tv.max_anisotropy = t.filter == MTSB_TEXFILTER_EWA ? t.max_anisotropy : 1.0f;       // bitmap.cpp:232-235
            tv.weight_lut = d_lut;
            if (t.type == MTSB_TEX_BITMAP) build_texture_levels(t, tv, texels);
        }
        const HalfTexel *d_texels = nullptr;
        if (!texels.empty() && (rc = dev_upload(s, &d_texels, texels.data(), texels.size()))) return fail(rc);
        for (TextureView &tv : tvs) tv.texels = d_texels;
        if ((rc = dev_upload(s, &v.textures, tvs.data(), tvs.size()))) return fail(rc);
    }
    {   // sample generator tables (halton: primes + digit permutations, halton.cpp:139-149, faure.cpp:22-84)
        const char *err = "";
        SamplerView sv;
        if (!sampler_view_init(d->sampler, d->integrator.sample_count, d->sensor.crop_width, d->sensor.crop_height, sv, &err)) {
            set_error(err); return fail(MTSB_ERR_INVALID);
        }
        if (d->sampler.type == MTSB_SAMPLER_HALTON) {
            const std::vector<uint16_t> primes = qmc_prime_table();
            if ((rc = dev_upload(s, &sv.primes, primes.data(), primes.size()))) return fail(rc);
            if (d->sampler.scramble != 0) {
                std::vector<uint16_t> perms; std::vector<uint32_t> offsets;
                qmc_permutations(d->sampler.scramble, primes, perms, offsets);
                if ((rc = dev_upload(s, &sv.perms, perms.data(), perms.size()))) return fail(rc);
                if ((rc = dev_upload(s, &sv.perm_offset, offsets.data(), offsets.size()))) return fail(rc);
                for (int j = 0; j < 2; ++j) sv.inv_perm2[perms[offsets[0] + j]] = (uint16_t) j;     // faure.cpp:130-135
                for (int j = 0; j < 3; ++j) sv.inv_perm3[perms[offsets[1] + j]] = (uint16_t) j;
            }
        }
        if (d->sampler.type == MTSB_SAMPLER_SOBOL) {
            const std::vector<uint32_t> mat = sobol_matrices32();
            std::vector<uint64_t> vdc, vdc_inv;
            if (!sobol_vdc_tables(sv.sobol_log_res, mat, vdc, vdc_inv)) { set_error("sobol: singular pixel map"); return fail(MTSB_ERR_INVALID); }
            if ((rc = dev_upload(s, &sv.sobol_mat, mat.data(), mat.size()))) return fail(rc);
            if ((rc = dev_upload(s, &sv.sobol_vdc, vdc.data(), vdc.size()))) return fail(rc);
            if ((rc = dev_upload(s, &sv.sobol_vdc_inv, vdc_inv.data(), vdc_inv.size()))) return fail(rc);
        }
        v.sampler_type = sv.type;
        if ((rc = dev_upload(s, &v.sampler, &sv, 1))) return fail(rc);
    }
    if (d->n_tris) {
        uint32_t *d_bad = nullptr, h_bad = 0;
        if ((rc = dev_alloc(s, &d_bad, 1))) return fail(rc);
        if (cudaMemsetAsync(d_bad, 0, sizeof(uint32_t), s->stream) != cudaSuccess) return fail(MTSB_ERR_CUDA);
        k_check_indices<<<grid_for(d->n_tris, 256), 256, 0, s->stream>>>(v.indices, v.tri_shape, v.shapes, d->n_tris, d_bad);
        s->launches++;
        if (cudaMemcpyAsync(&h_bad, d_bad, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream) != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) {
            set_error("index check failed"); cudaGetLastError(); return fail(MTSB_ERR_CUDA);
        }
        if (h_bad) { set_error("triangle index outside its shape's vertex range"); return fail(MTSB_ERR_INVALID); }
    }
    if ((rc = build_bvh(s, v.positions, v.indices, d->n_tris))) return fail(rc);
    if ((rc = pack_shading_records(s, d->n_tris))) return fail(rc);
    {   // Scene::m_aabb -> bounding sphere of the environment emitter (scene.cpp:388-414, constant.cpp:67-71)
        float lo[3], hi[3];
        for (int k = 0; k < 3; ++k) { lo[k] = d->n_tris ? v.box_lo[k] : 0.0f; hi[k] = d->n_tris ? v.box_hi[k] : 0.0f; }
        auto grow = [&](const float *p) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } };
        float cam[3] = {d->sensor.to_world[3], d->sensor.to_world[7], d->sensor.to_world[11]};
        grow(cam);
        for (uint32_t i = 0; i < d->n_emitters; ++i) if (d->emitters[i].type == MTSB_EMITTER_POINT) grow(d->emitters[i].position);
        float r2 = 0;
        for (int k = 0; k < 3; ++k) { v.bsphere_c[k] = (hi[k] + lo[k]) * 0.5f; float e = v.bsphere_c[k] - hi[k]; r2 += e * e; }
        v.bsphere_r = std::max(1e-4f, std::sqrt(r2) * 1.5f);
    }
    {   // perspective projection constants and the film / filter table (rfilter.cpp:36-55)
        const MtsbSensor &se = d->sensor;
        double aspect = (double) se.film_width / (double) se.film_height;
        double tan_half = std::tan(0.5 * (double) se.fov_x_deg * 3.14159265358979323846 / 180.0);
        v.near_x = (float) ((double) se.near_clip * tan_half);
        v.near_y = (float) ((double) se.near_clip * tan_half / aspect);
        const int RES = 31;
        float sum = 0.0f;
        for (int i = 0; i < RES; ++i) { float val = rfilter_eval(se, (se.rfilter_radius * i) / RES); v.filter_table[i] = val; sum += val; }
        v.filter_table[RES] = 0.0f;
        sum *= 2 * se.rfilter_radius / RES;
        float normalization = 1.0f / sum;
        for (int i = 0; i < RES; ++i) v.filter_table[i] *= normalization;
        v.filter_radius = se.rfilter_radius; v.filter_scale = RES / se.rfilter_radius;
        v.film_border = (int) std::ceil(se.rfilter_radius - 0.5f);
        v.film_w = se.crop_width; v.film_h = se.crop_height;
        size_t nf = (size_t) (v.film_w + 2 * v.film_border) * (v.film_h + 2 * v.film_border) * 5;
        if ((rc = dev_alloc(s, &v.film, nf))) return fail(rc);
        if (cudaMemsetAsync(v.film, 0, nf * sizeof(float), s->stream) != cudaSuccess) return fail(MTSB_ERR_CUDA);
        if ((rc = dev_alloc(s, &s->d_film_bad, 1))) return fail(rc);
        if (cudaMemsetAsync(s->d_film_bad, 0, sizeof(unsigned long long), s->stream) != cudaSuccess) return fail(MTSB_ERR_CUDA);
    }
    s->info.n_emitters = d->n_emitters;
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) { set_error("scene upload failed"); return fail(MTSB_ERR_CUDA); }
    *out = reinterpret_cast<mtsb_scene_t>(s);
    return MTSB_OK;
}

extern "C" int mtsb_scene_info(mtsb_scene_t scene, MtsbSceneInfo *info) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || !info) { set_error("mtsb_scene_info: bad argument"); return MTSB_ERR_INVALID; }
    *info = s->info;
    return MTSB_OK;
}

extern "C" void mtsb_scene_destroy(mtsb_scene_t scene) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    render_free(s);
    cudaDeviceSynchronize();                 // what cudaFree did implicitly: nothing may still read a block that goes back to the cache
    for (void *p : s->allocs) dfree(p);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

extern "C" int mtsb_cancel(mtsb_scene_t scene) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s) { set_error("mtsb_cancel: bad argument"); return MTSB_ERR_INVALID; }
    s->cancel.store(1);
    return MTSB_OK;
}
