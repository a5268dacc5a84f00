// bvh_build.cu -- GPU construction of the 8-wide BVH (replaces ShapeKDTree::build,
// reference src/librender/skdtree.cpp:68-110; see bvh.h for the layout and the steps).
#include "scene.h"
#include <cub/device/device_radix_sort.cuh>
#include <cstring>
#include <algorithm>

namespace mtsb {

__global__ void k_tri_setup(uint32_t n, const float *positions, const uint32_t *indices, TriRec *rec,
                            float *prim_lo, float *prim_hi, uint32_t *scene_box /*6 encoded*/) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t e[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
    if (tid < n) {
        float lo[3], hi[3];
        tri_setup(tid, positions, indices, rec + tid, lo, hi);
        for (int k = 0; k < 3; ++k) {
            prim_lo[3 * (size_t) tid + k] = lo[k]; prim_hi[3 * (size_t) tid + k] = hi[k];
            e[k] = enc_f(lo[k]); e[3 + k] = enc_f(hi[k]);
        }
    }
    for (int k = 0; k < 3; ++k) {
        uint32_t mn = __reduce_min_sync(0xFFFFFFFFu, e[k]);
        uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, e[3 + k]);
        if ((threadIdx.x & 31) == 0) { atomicMin(scene_box + k, mn); atomicMax(scene_box + 3 + k, mx); }
    }
}

__global__ void k_morton(uint32_t n, const float *prim_lo, const float *prim_hi, V3 slo, V3 sinv,
                         uint64_t *keys, uint32_t *vals) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= n) return;
    keys[tid] = morton_key(prim_lo + 3 * (size_t) tid, prim_hi + 3 * (size_t) tid, slo, sinv);
    vals[tid] = tid;
}

__global__ void k_radix_tree(int n, const uint64_t *keys, uint32_t *left, uint32_t *right, uint32_t *parent,
                             uint32_t *first, uint32_t *last) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    radix_tree_node(i, n, keys, left, right, parent, first, last);
}

__global__ void k_fit(int n, const uint32_t *sorted_prim, const float *prim_lo, const float *prim_hi,
                      const uint32_t *left, const uint32_t *right, const uint32_t *parent,
                      float *node_lo, float *node_hi, uint32_t *flags) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    fit_from_leaf(j, n, sorted_prim, prim_lo, prim_hi, left, right, parent, node_lo, node_hi, flags);
}

__global__ void k_wide_level(uint32_t begin, uint32_t end, WideBuildArgs a, int tiny) {
    uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= end) return;
    wide_node_build(w, a, tiny != 0);
}

#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { rc = cuda_fail(e__, #x, __FILE__, __LINE__); goto done; } } while (0)

int build_bvh(Scene *s, const float *d_positions, const uint32_t *d_indices, uint64_t n_tris64) {
    int rc = MTSB_OK;
    const uint32_t n = (uint32_t) n_tris64;
    std::vector<void *> tmp;
    auto talloc = [&](size_t bytes) -> void * {
        void *p = nullptr;
        if (dmalloc(&p, bytes ? bytes : 16) != cudaSuccess) return nullptr;
        tmp.push_back(p);
        return p;
    };
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    s->v.nodes = nullptr; s->v.tris = nullptr; s->v.n_nodes = 0;
    s->info.n_tris = n; s->info.n_bvh_nodes = 0; s->info.bvh_bytes = 0; s->info.tri_bytes = 0; s->info.build_ms = 0;
    for (int k = 0; k < 3; ++k) { s->info.aabb_min[k] = 0; s->info.aabb_max[k] = 0; s->v.box_lo[k] = 0; s->v.box_hi[k] = 0; }
    if (n == 0) return MTSB_OK;
    {
        cudaEventCreate(&ev0); cudaEventCreate(&ev1);
        cudaEventRecord(ev0, s->stream);
        const unsigned B = 256;
        TriRec *tri_tmp = (TriRec *) talloc((size_t) n * sizeof(TriRec));
        float *prim_lo = (float *) talloc((size_t) n * 12), *prim_hi = (float *) talloc((size_t) n * 12);
        uint32_t *scene_box = (uint32_t *) talloc(6 * 4);
        uint64_t *keys = (uint64_t *) talloc((size_t) n * 8), *keys2 = (uint64_t *) talloc((size_t) n * 8);
        uint32_t *vals = (uint32_t *) talloc((size_t) n * 4), *vals2 = (uint32_t *) talloc((size_t) n * 4);
        if (!tri_tmp || !prim_lo || !prim_hi || !scene_box || !keys || !keys2 || !vals || !vals2) { set_error("out of device memory (bvh build)"); rc = MTSB_ERR_CUDA; goto done; }
        uint32_t init_box[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
        CK(cudaMemcpyAsync(scene_box, init_box, sizeof(init_box), cudaMemcpyHostToDevice, s->stream));
        k_tri_setup<<<grid_for(n, B), B, 0, s->stream>>>(n, d_positions, d_indices, tri_tmp, prim_lo, prim_hi, scene_box);
        s->launches++;
        uint32_t hbox[6];
        CK(cudaMemcpyAsync(hbox, scene_box, sizeof(hbox), cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        float tlo[3], thi[3];
        for (int k = 0; k < 3; ++k) { tlo[k] = dec_f(hbox[k]); thi[k] = dec_f(hbox[3 + k]); }
        // enlarged box: gkdtree.h:1215-1220 (two sequential statements; the second sees the moved minimum)
        for (int k = 0; k < 3; ++k) {
            volatile float lo = tlo[k] - ((thi[k] - tlo[k]) * 1e-3f + 1e-3f);
            volatile float hi = thi[k] + ((thi[k] - lo) * 1e-3f + 1e-3f);
            s->v.box_lo[k] = lo; s->v.box_hi[k] = hi;
            s->info.aabb_min[k] = tlo[k]; s->info.aabb_max[k] = thi[k];
        }
        V3 slo(tlo[0], tlo[1], tlo[2]);
        V3 sinv(thi[0] > tlo[0] ? 1.0f / (thi[0] - tlo[0]) : 0.0f, thi[1] > tlo[1] ? 1.0f / (thi[1] - tlo[1]) : 0.0f,
                thi[2] > tlo[2] ? 1.0f / (thi[2] - tlo[2]) : 0.0f);
        float diag = sqrtf((thi[0] - tlo[0]) * (thi[0] - tlo[0]) + (thi[1] - tlo[1]) * (thi[1] - tlo[1]) + (thi[2] - tlo[2]) * (thi[2] - tlo[2]));
        float amax = 0; for (int k = 0; k < 3; ++k) amax = std::max(amax, std::max(fabsf(tlo[k]), fabsf(thi[k])));
        const float leaf_pad = 4e-6f * std::max(diag, amax) + 1e-30f;

        k_morton<<<grid_for(n, B), B, 0, s->stream>>>(n, prim_lo, prim_hi, slo, sinv, keys, vals);
        s->launches++;
        {
            size_t sort_bytes = 0;
            cub::DoubleBuffer<uint64_t> dk(keys, keys2); cub::DoubleBuffer<uint32_t> dv(vals, vals2);
            CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, dk, dv, (int) n, 0, 63, s->stream));
            void *sort_tmp = talloc(sort_bytes);
            if (!sort_tmp) { set_error("out of device memory (sort)"); rc = MTSB_ERR_CUDA; goto done; }
            CK(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, dk, dv, (int) n, 0, 63, s->stream));
            keys = dk.Current(); vals = dv.Current();
        }
        // final arrays
        size_t max_nodes = std::max<size_t>(1, n);       // every wide node consumes >= 1 binary internal node
        Node8 *nodes_tmp = (Node8 *) talloc(max_nodes * sizeof(Node8));
        TriRec *tri_out = nullptr;
        rc = dev_alloc(s, &tri_out, (size_t) n);
        if (rc) goto done;
        uint32_t *wbin = (uint32_t *) talloc(max_nodes * 4);
        uint32_t *counters = (uint32_t *) talloc(8);
        float *node_lo = (float *) talloc((size_t) (2 * (size_t) n) * 12), *node_hi = (float *) talloc((size_t) (2 * (size_t) n) * 12);
        uint32_t *left = (uint32_t *) talloc((size_t) n * 4), *right = (uint32_t *) talloc((size_t) n * 4);
        uint32_t *first = (uint32_t *) talloc((size_t) n * 4), *last = (uint32_t *) talloc((size_t) n * 4);
        uint32_t *parent = (uint32_t *) talloc((size_t) (2 * (size_t) n) * 4), *flags = (uint32_t *) talloc((size_t) n * 4);
        if (!nodes_tmp || !wbin || !counters || !node_lo || !node_hi || !left || !right || !first || !last || !parent || !flags) {
            set_error("out of device memory (bvh build)"); rc = MTSB_ERR_CUDA; goto done;
        }
        WideBuildArgs a;
        a.n = (int) n; a.left = left; a.right = right; a.first = first; a.last = last; a.node_lo = node_lo; a.node_hi = node_hi;
        a.sorted_prim = vals; a.tri_in = tri_tmp; a.nodes = nodes_tmp; a.wbin = wbin; a.tri_out = tri_out; a.counters = counters;
        a.leaf_pad = leaf_pad;
        uint32_t h_counters[2] = {1, 0};
        CK(cudaMemcpyAsync(counters, h_counters, 8, cudaMemcpyHostToDevice, s->stream));
        uint32_t n_nodes = 1;
        if (n <= MTSB_LEAF_MAX) {
            // whole scene in one leaf slot: node id 0 of node_lo/hi carries the scene box
            CK(cudaMemcpyAsync(node_lo, tlo, 12, cudaMemcpyHostToDevice, s->stream));
            CK(cudaMemcpyAsync(node_hi, thi, 12, cudaMemcpyHostToDevice, s->stream));
            k_wide_level<<<1, 32, 0, s->stream>>>(0, 1, a, 1);
            s->launches++;
        } else {
            CK(cudaMemsetAsync(flags, 0, (size_t) n * 4, s->stream));
            k_radix_tree<<<grid_for(n - 1, B), B, 0, s->stream>>>((int) n, keys, left, right, parent, first, last);
            k_fit<<<grid_for(n, B), B, 0, s->stream>>>((int) n, vals, prim_lo, prim_hi, left, right, parent, node_lo, node_hi, flags);
            s->launches += 2;
            uint32_t zero = 0;
            CK(cudaMemcpyAsync(wbin, &zero, 4, cudaMemcpyHostToDevice, s->stream));
            uint32_t begin = 0, end = 1;
            while (begin < end) {
                k_wide_level<<<grid_for(end - begin, 128), 128, 0, s->stream>>>(begin, end, a, 0);
                s->launches++;
                CK(cudaMemcpyAsync(h_counters, counters, 8, cudaMemcpyDeviceToHost, s->stream));
                CK(cudaStreamSynchronize(s->stream));
                begin = end; end = h_counters[0];
                if (end > max_nodes) { set_error("bvh build: node pool overflow"); rc = MTSB_ERR_CUDA; goto done; }
            }
            n_nodes = end;
        }
        CK(cudaGetLastError());
        Node8 *nodes = nullptr;
        rc = dev_alloc(s, &nodes, (size_t) n_nodes);
        if (rc) goto done;
        CK(cudaMemcpyAsync(nodes, nodes_tmp, (size_t) n_nodes * sizeof(Node8), cudaMemcpyDeviceToDevice, s->stream));
        cudaEventRecord(ev1, s->stream);
        CK(cudaStreamSynchronize(s->stream));
        float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1);
        s->v.nodes = reinterpret_cast<const U4 *>(nodes); s->v.tris = reinterpret_cast<const U4 *>(tri_out); s->v.n_nodes = n_nodes;
        s->info.n_bvh_nodes = n_nodes; s->info.bvh_bytes = (uint64_t) n_nodes * sizeof(Node8);
        s->info.tri_bytes = (uint64_t) n * sizeof(TriRec); s->info.build_ms = ms;
    }
done:
    for (void *p : tmp) dfree(p);            // the stream was synchronised above (or the build failed)
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    return rc;
}

}  // namespace mtsb
