// multi.cu -- one frame over several GPUs of one node, inside the library (include/mtsb200.h: mtsb_multi_*).
//
// Reference: the block scheduler gives the image blocks of ONE frame to its workers and merges the returned blocks
// into one film under a mutex (src/librender/renderproc.cpp:149-156, imageproc.cpp:43-78, imageblock.h:103-107).
// Here a "worker" is a GPU: the scene is replicated (one host thread per GPU uploads it and builds its BVH), device i
// renders every sample of the 32x32 tiles t with t % n == i, and the bordered per-device films are summed onto the
// first device by ONE ncclReduce over NVLink -- the same overlap-add, done once per frame instead of once per block.
// The sample streams are keyed by (seed, pixel, sample index), so the reduced film equals the single-GPU film up to
// fp32 addition order.
//
// NCCL is bound at run time (dlopen of libnccl.so.2) by mtsb_multi_create only: the library itself has no link-time
// dependency on it, and a process that already holds an NCCL (PyTorch) shares that copy.
#include "scene.h"
#include <nccl.h>
#include <dlfcn.h>
#include <thread>
#include <vector>
#include <cstring>
#include <algorithm>

namespace mtsb {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static bool load_nccl(NcclApi &api) {
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) { api.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL); if (api.handle) break; }
    if (!api.handle) { set_error(std::string("mtsb_multi_create: cannot load NCCL (libnccl.so.2): ") + (dlerror() ? dlerror() : "")); return false; }
    auto sym = [&](const char *name) { return dlsym(api.handle, name); };
    api.CommInitAll = (decltype(api.CommInitAll)) sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy)) sym("ncclCommDestroy");
    api.Reduce = (decltype(api.Reduce)) sym("ncclReduce");
    api.GroupStart = (decltype(api.GroupStart)) sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd)) sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString)) sym("ncclGetErrorString");
    if (!api.CommInitAll || !api.CommDestroy || !api.Reduce || !api.GroupStart || !api.GroupEnd || !api.GetErrorString) {
        set_error("mtsb_multi_create: libnccl.so.2 lacks a required entry point"); return false;
    }
    return true;
}

struct Multi {
    std::vector<int> devices;
    std::vector<mtsb_scene_t> scenes;
    std::vector<ncclComm_t> comms;
    NcclApi nccl;
};

}  // namespace mtsb

using namespace mtsb;

static int nccl_fail(const NcclApi &api, ncclResult_t r, const char *what) {
    set_error(std::string("NCCL error in ") + what + ": " + api.GetErrorString(r));
    return MTSB_ERR_CUDA;
}

extern "C" int mtsb_multi_create(const MtsbSceneDesc *desc, int n_devices, const int *device_ids, mtsb_multi_t *out) {
    if (!out) { set_error("mtsb_multi_create: out == NULL"); return MTSB_ERR_INVALID; }
    *out = nullptr;
    const int have = mtsb_device_count();
    if (have <= 0) { set_error("no CUDA device present (this library has no CPU fallback)"); return MTSB_ERR_NO_DEVICE; }
    if (n_devices <= 0 || n_devices > have) { set_error("mtsb_multi_create: n_devices out of range"); return MTSB_ERR_INVALID; }
    Multi *m = new Multi();
    for (int i = 0; i < n_devices; ++i) {
        const int d = device_ids ? device_ids[i] : i;
        if (d < 0 || d >= have || std::find(m->devices.begin(), m->devices.end(), d) != m->devices.end()) {
            delete m; set_error("mtsb_multi_create: bad or repeated device id"); return MTSB_ERR_INVALID;
        }
        m->devices.push_back(d);
    }
    m->scenes.assign((size_t) n_devices, nullptr);
    // one host thread per GPU: upload + BVH build run side by side (the description is read-only host memory)
    std::vector<int> rcs((size_t) n_devices, MTSB_OK);
    std::vector<std::string> errs((size_t) n_devices);
    std::vector<std::thread> th;
    for (int i = 0; i < n_devices; ++i)
        th.emplace_back([&, i] {
            int rc = mtsb_init(m->devices[(size_t) i]);
            if (!rc) rc = mtsb_scene_create(desc, &m->scenes[(size_t) i]);
            rcs[(size_t) i] = rc;
            if (rc) errs[(size_t) i] = mtsb_last_error();
        });
    for (auto &t : th) t.join();
    for (int i = 0; i < n_devices; ++i)
        if (rcs[(size_t) i]) { const int rc = rcs[(size_t) i]; const std::string e = errs[(size_t) i]; mtsb_multi_destroy(reinterpret_cast<mtsb_multi_t>(m)); set_error(e); return rc; }
    if (n_devices > 1) {
        if (!load_nccl(m->nccl)) { mtsb_multi_destroy(reinterpret_cast<mtsb_multi_t>(m)); return MTSB_ERR_CUDA; }
        m->comms.assign((size_t) n_devices, nullptr);
        const ncclResult_t r = m->nccl.CommInitAll(m->comms.data(), n_devices, m->devices.data());
        if (r != ncclSuccess) { const int rc = nccl_fail(m->nccl, r, "ncclCommInitAll"); m->comms.clear(); mtsb_multi_destroy(reinterpret_cast<mtsb_multi_t>(m)); return rc; }
    }
    *out = reinterpret_cast<mtsb_multi_t>(m);
    return MTSB_OK;
}

extern "C" int mtsb_multi_device_count(mtsb_multi_t mm) { Multi *m = reinterpret_cast<Multi *>(mm); return m ? (int) m->devices.size() : 0; }
extern "C" mtsb_scene_t mtsb_multi_scene(mtsb_multi_t mm, int i) {
    Multi *m = reinterpret_cast<Multi *>(mm);
    return (m && i >= 0 && i < (int) m->scenes.size()) ? m->scenes[(size_t) i] : nullptr;
}

extern "C" int mtsb_multi_cancel(mtsb_multi_t mm) {
    Multi *m = reinterpret_cast<Multi *>(mm);
    if (!m) { set_error("mtsb_multi_cancel: bad argument"); return MTSB_ERR_INVALID; }
    for (mtsb_scene_t s : m->scenes) if (s) mtsb_cancel(s);
    return MTSB_OK;
}

extern "C" int mtsb_multi_render(mtsb_multi_t mm, const MtsbRenderParams *params, MtsbStats *stats) {
    Multi *m = reinterpret_cast<Multi *>(mm);
    if (!m || !params) { set_error("mtsb_multi_render: bad argument"); return MTSB_ERR_INVALID; }
    if (params->tile_count > 1 || params->tile_index != 0) { set_error("mtsb_multi_render: the tile partition belongs to the library (tile_index / tile_count must be 0)"); return MTSB_ERR_INVALID; }
    const int n = (int) m->devices.size();
    std::vector<int> rcs((size_t) n, MTSB_OK);
    std::vector<std::string> errs((size_t) n);
    std::vector<MtsbStats> st((size_t) n);
    std::vector<std::thread> th;
    for (int i = 0; i < n; ++i)
        th.emplace_back([&, i] {
            MtsbRenderParams rp = *params;
            rp.tile_index = i; rp.tile_count = n;
            rcs[(size_t) i] = mtsb_render(m->scenes[(size_t) i], &rp, &st[(size_t) i]);
            if (rcs[(size_t) i]) errs[(size_t) i] = mtsb_last_error();
        });
    for (auto &t : th) t.join();
    for (int i = 0; i < n; ++i) if (rcs[(size_t) i]) { set_error(errs[(size_t) i]); return rcs[(size_t) i]; }
    float ms_reduce = 0.0f;
    if (n > 1) {
        // merge: film[0] += film[i] for every i, one reduce over NVLink (processResult's put(), renderproc.cpp:149-156)
        Scene *root = reinterpret_cast<Scene *>(m->scenes[0]);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        MTSB_CUDA(cudaSetDevice(root->device));
        MTSB_CUDA(cudaEventCreate(&e0)); MTSB_CUDA(cudaEventCreate(&e1));
        MTSB_CUDA(cudaEventRecord(e0, root->stream));
        ncclResult_t r = m->nccl.GroupStart();
        if (r != ncclSuccess) return nccl_fail(m->nccl, r, "ncclGroupStart");
        for (int i = 0; i < n && r == ncclSuccess; ++i) {
            Scene *s = reinterpret_cast<Scene *>(m->scenes[(size_t) i]);
            const size_t nf = (size_t) (s->v.film_w + 2 * s->v.film_border) * (s->v.film_h + 2 * s->v.film_border) * 5;
            cudaSetDevice(s->device);
            r = m->nccl.Reduce(s->v.film, s->v.film, nf, ncclFloat, ncclSum, 0, m->comms[(size_t) i], s->stream);
        }
        const ncclResult_t r2 = m->nccl.GroupEnd();
        if (r != ncclSuccess) return nccl_fail(m->nccl, r, "ncclReduce");
        if (r2 != ncclSuccess) return nccl_fail(m->nccl, r2, "ncclGroupEnd");
        MTSB_CUDA(cudaSetDevice(root->device));
        MTSB_CUDA(cudaEventRecord(e1, root->stream));
        for (int i = 0; i < n; ++i) {
            Scene *s = reinterpret_cast<Scene *>(m->scenes[(size_t) i]);
            MTSB_CUDA(cudaSetDevice(s->device));
            MTSB_CUDA(cudaStreamSynchronize(s->stream));
        }
        MTSB_CUDA(cudaSetDevice(root->device));
        cudaEventElapsedTime(&ms_reduce, e0, e1);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        // bad-sample counts live per device; fold them into the root so that film statistics describe the whole frame
    }
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        for (int i = 0; i < n; ++i) {
            const MtsbStats &a = st[(size_t) i];
            stats->samples += a.samples; stats->closest_rays += a.closest_rays; stats->shadow_rays += a.shadow_rays;
            stats->shade_events += a.shade_events; stats->path_length_sum += a.path_length_sum; stats->kernel_launches += a.kernel_launches;
            stats->bad_samples += a.bad_samples; stats->closest_launches += a.closest_launches; stats->any_launches += a.any_launches;
            stats->n_passes += a.n_passes;
            stats->ms_total = std::max(stats->ms_total, a.ms_total); stats->ms_trace_closest = std::max(stats->ms_trace_closest, a.ms_trace_closest);
            stats->ms_trace_any = std::max(stats->ms_trace_any, a.ms_trace_any); stats->ms_shade = std::max(stats->ms_shade, a.ms_shade);
            stats->ms_raygen = std::max(stats->ms_raygen, a.ms_raygen); stats->ms_splat = std::max(stats->ms_splat, a.ms_splat);
            for (int b = 0; b < 16; ++b) { stats->closest_rays_bounce[b] += a.closest_rays_bounce[b]; stats->shadow_rays_bounce[b] += a.shadow_rays_bounce[b]; }
        }
        stats->ms_reduce = ms_reduce; stats->ms_total += ms_reduce; stats->n_devices = (uint32_t) n;
    }
    return MTSB_OK;
}

extern "C" void mtsb_multi_destroy(mtsb_multi_t mm) {
    Multi *m = reinterpret_cast<Multi *>(mm);
    if (!m) return;
    for (ncclComm_t c : m->comms) if (c && m->nccl.CommDestroy) m->nccl.CommDestroy(c);
    for (mtsb_scene_t s : m->scenes) if (s) mtsb_scene_destroy(s);
    if (m->nccl.handle) dlclose(m->nccl.handle);
    delete m;
}
