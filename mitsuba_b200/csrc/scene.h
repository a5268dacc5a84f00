// scene.h -- host-side handle behind mtsb_scene_t: device pointers + small host tables.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include <atomic>
#include "../../include/mtsb200.h"
#include "bvh.h"
#include "scene_view.h"

namespace mtsb {

struct PathBuffers;   // render.cu
#define MTSB_MAX_LANES 4

struct Scene {
    int device = 0;
    SceneView v{};
    std::vector<void *> allocs;             // everything to cudaFree on destroy
    MtsbSceneInfo info{};
    std::atomic<int> cancel{0};
    unsigned long long *d_film_bad = nullptr;
    float *d_rgb = nullptr;                 // developed image staging (mtsb_film_develop)
    PathBuffers *paths[MTSB_MAX_LANES] = {};       // one wavefront pipeline per lane (render.cu)
    cudaStream_t lane_stream[MTSB_MAX_LANES] = {};
    std::vector<MtsbMaterial> h_materials;  // host copies for launch decisions (which shading queues can be non-empty)
    std::vector<int32_t> h_shape_material;
    uint64_t launches = 0;                  // kernels launched by this library for this scene
    cudaStream_t stream = nullptr;
};

void set_error(const std::string &s);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);
#define MTSB_CUDA(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) return mtsb::cuda_fail(e__, #x, __FILE__, __LINE__); } while (0)

// cached device blocks (api.cu): dfree() keeps a block for the next dmalloc() of the same size class
cudaError_t dmalloc(void **p, size_t bytes);
void dfree(void *p);
void dcache_release();

template <class T> int dev_alloc(Scene *s, T **p, size_t count) {
    void *q = nullptr;
    cudaError_t e = dmalloc(&q, count ? count * sizeof(T) : sizeof(T));
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc", __FILE__, __LINE__);
    if (s) s->allocs.push_back(q);
    *p = static_cast<T *>(q);
    return MTSB_OK;
}
// host -> device copy of a caller-owned (usually pageable) array; large arrays go through pinned staging buffers filled by
// several host threads, one cudaMemcpyAsync per chunk (api.cu).  Returns after the data has left the host array.
int staged_h2d(void *dst, const void *src, size_t bytes);

template <class T> int dev_upload(Scene *s, const T **p, const T *host, size_t count) {
    T *q = nullptr;
    int rc = dev_alloc(s, &q, count);
    if (rc) return rc;
    if (count && (rc = staged_h2d(q, host, count * sizeof(T)))) return rc;
    *p = q;
    return MTSB_OK;
}

int build_bvh(Scene *s, const float *d_positions, const uint32_t *d_indices, uint64_t n_tris);
int render_free(Scene *s);
int pack_shading_records(Scene *s, uint64_t n_tris);   // render.cu: per-triangle shading records + queue bytes

inline unsigned grid_for(size_t n, unsigned block) { return (unsigned) ((n + block - 1) / block); }

}  // namespace mtsb
