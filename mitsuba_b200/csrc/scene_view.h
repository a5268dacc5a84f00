// scene_view.h -- device-side view of a scene, handed to the kernels by value.
#pragma once
#include "../../include/mtsb200.h"
#include "bvh.h"
#include "qmc.h"
#include "envmap.h"
#include "texture.h"

namespace mtsb {

// Device-side view handed to the kernels by value (all pointers are device pointers).
struct SceneView {
    // acceleration structure
    const U4 *nodes; const U4 *tris; uint32_t n_nodes;
    float box_lo[3], box_hi[3];             // enlarged scene box (gkdtree.h:1215-1220)
    // geometry, indexed by global primitive id / global vertex id
    const float *positions, *normals, *texcoords; const uint32_t *indices; const uint32_t *tri_shape;
    // packed per-triangle shading records (8 x 16 B each, shading.h) and the shading queue every triangle feeds (1 B each)
    const U4 *shade_tris; const uint8_t *tri_queue;
    const MtsbShape *shapes; const MtsbMaterial *materials; const MtsbEmitter *emitters;
    uint32_t n_shapes, n_materials, n_emitters;
    // emitter sampling tables (DiscreteDistribution, include/mitsuba/core/pmf.h)
    const float *emitter_cdf;               // n_emitters + 1
    float emitter_norm;                     // 1 / sum of sampling weights
    const float *area_cdf;                  // concatenated per emitter shape
    const uint32_t *area_cdf_offset;        // per shape: offset into area_cdf (n_tris+1 entries), or 0xFFFFFFFF
    const float *inv_area;                  // per shape
    int env_emitter;                        // index of the environment emitter (`constant` or `envmap`) or -1
    const EnvmapView *envmap;               // device copy of the `envmap` tables, or NULL
    const TextureView *textures;            // device copies of the texture records (src/textures), or NULL
    uint32_t n_textures;
    float bsphere_c[3], bsphere_r;          // constant.cpp:67-71
    // sensor (perspective.cpp:127-150, 286-317)
    MtsbSensor sensor; float near_x, near_y;
    MtsbIntegrator integrator;
    uint32_t sampler_type;                  // MTSB_SAMPLER_*
    const SamplerView *sampler;             // device copy of the generator description (qmc.h); read only when sampler_type != independent
    // film (imageblock.h:124-206)
    float *film; int film_w, film_h, film_border; float filter_radius, filter_scale;
    float filter_table[32];
};

}  // namespace mtsb
