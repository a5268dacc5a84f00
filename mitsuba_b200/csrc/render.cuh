// render.cuh -- wavefront path state of the `path` integrator (SoA, one entry per path slot).
#pragma once
#include "scene.h"
#include "traverse.h"

namespace mtsb {

// Shading queues: one per scattering model so that every k_shade instance runs one code path.
// Q_CONDUCTOR / Q_DIELECTRIC are the rough (microfacet) models, Q_DELTA the smooth `dielectric` and
// `conductor` (Dirac lobes: no emitter sampling), Q_GENERIC twosided wrappers that mix two models,
// Q_TEXTURED materials whose spectra come from a texture (evaluated at the hit, then shaded like Q_GENERIC).
enum { Q_DIFFUSE = 0, Q_CONDUCTOR = 1, Q_DIELECTRIC = 2, Q_NULL = 3, Q_GENERIC = 4, Q_DELTA = 5, Q_TEXTURED = 6, Q_MISS = 7, Q_COUNT = 8 };

// path flags (stored as float bits in refn_flags.w)
#define PF_DEPTH_MASK 0xFFFFu
#define PF_FIRST      (1u << 16)   /* rRec.type still has EEmittedRadiance (first segment)  */
#define PF_SCATTERED  (1u << 17)
#define PF_ALPHA      (1u << 18)   /* first hit was valid (records.inl:136-152)            */
#define PF_VALID      (1u << 19)   /* slot maps to a pixel inside the crop window          */
/* bits 20..31: dimension counter of the quasi-Monte Carlo samplers (qmc.h).  halton: next dimension
 * (12 bits); ldsampler: next 1D dimension (bits 20..25) and next 2D dimension (bits 26..31).      */
#define PF_DIM_SHIFT  20
#define PF_DIM_MASK   0xFFF00000u

struct Counters {
    uint32_t n_rays[2];        // closest-ray queue sizes (double buffered)
    uint32_t n_shadow;
    uint32_t n_q[Q_COUNT];
    uint32_t pad[2];
    unsigned long long closest, shadow, shade, path_len, bad, samples;
};

struct PathBuffers {
    uint32_t capacity;
    RayIn *rays[2]; uint32_t *ray_slot[2];
    HitOut *hits;
    RayIn *srays; float4 *scontrib;
    float4 *thr_eta, *L_bpdf, *refn_flags; float2 *spos;
    uint32_t *q[Q_COUNT];
    Counters *ctr;             // device
    Counters *h_ctr;           // pinned host mirror
    int *tiles;                // device: tile ids rendered by this call
    uint32_t tiles_capacity;
};

struct PassInfo {
    uint32_t tile_first, n_tiles;       // range in PathBuffers::tiles
    uint32_t spp_pass, s_first;         // samples of this pass: global indices [s_first, s_first + spp_pass)
    uint32_t tiles_x;                   // tiles per row of the crop window
    uint64_t seed;
};

}  // namespace mtsb
