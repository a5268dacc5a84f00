// render.cuh -- wavefront path state of the `path` integrator (SoA, one entry per path slot).
#pragma once
#include "scene.h"
#include "traverse.h"

namespace mtsb {

// Shading queues: one per scattering model so that every k_shade instance runs one code path.
// Q_CONDUCTOR / Q_DIELECTRIC are the rough (microfacet) models, Q_DELTA the smooth `dielectric` and
// `conductor` (Dirac lobes: no emitter sampling), Q_GENERIC twosided wrappers that mix two models,
// Q_TEXTURED materials whose spectra come from a texture (evaluated at the hit, then shaded like Q_GENERIC).
// A ray that leaves the scene is finished inside k_sort (environment term), it has no queue.
enum { Q_DIFFUSE = 0, Q_CONDUCTOR = 1, Q_DIELECTRIC = 2, Q_NULL = 3, Q_GENERIC = 4, Q_DELTA = 5, Q_TEXTURED = 6, Q_COUNT = 7 };
#define Q_NONE 0xFFu               /* tri_queue entry of a triangle nothing can hit (never produced) */

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

// Queue sizes live in a ring of per-bounce sets: bounce b uses set b % MTSB_CTR_RING (its closest-ray count is written by
// the shading stage of bounce b - 1, its shading-queue and shadow-ray counts by its own stages) and clears the set that
// bounce b + MTSB_CTR_RING / 2 will use.  Nothing on the host has to reset or read a counter between two bounces, so a
// whole pass is enqueued without a round trip; the host only polls the sizes (asynchronously) to stop early.
#define MTSB_CTR_RING 8
#define MTSB_BOUNCE_STATS 16       /* per-bounce ray counts kept for the statistics (later bounces add to the last entry) */
struct BounceCtr {
    uint32_t n_rays;           // closest-hit rays of this bounce
    uint32_t n_shadow;         // shadow rays of this bounce
    uint32_t n_q[Q_COUNT];     // shading queue sizes
    uint32_t pad[16 - 2 - Q_COUNT];
};
struct Counters {
    BounceCtr b[MTSB_CTR_RING];
    unsigned long long closest, shadow, shade, path_len, bad, samples;
    unsigned long long rays_bounce[MTSB_BOUNCE_STATS], shadow_bounce[MTSB_BOUNCE_STATS];
    unsigned long long launches_closest, launches_any;     // k_trace launches that found rays in their queue
};

struct PathBuffers {
    uint32_t capacity;
    RayIn *rays[2]; uint32_t *ray_slot[2];
    HitOut *hits;
    RayIn *srays; float4 *scontrib;
    float4 *thr_eta, *L_bpdf, *refn_flags; float2 *spos;
    uint32_t *q[Q_COUNT];
    Counters *ctr;             // device
    uint32_t *h_poll;          // pinned host: closest-ray counts of the bounces of the pass in flight (MTSB_POLL_RING entries)
    int *tiles;                // device: tile ids rendered by this call
    uint32_t tiles_capacity;
};

struct PassInfo {
    uint32_t tile_first, n_tiles;       // range in PathBuffers::tiles
    uint32_t spp_pass, s_first;         // samples of this pass: global indices [s_first, s_first + spp_pass)
    uint32_t tiles_x;                   // tiles per row of the crop window
    uint64_t seed;
};

#define MTSB_POLL_RING 64

// which shading queue a material feeds (host + device)
MTSB_HD int queue_of_type(uint32_t t) {
    return t == MTSB_BSDF_DIFFUSE ? Q_DIFFUSE : t == MTSB_BSDF_ROUGHCONDUCTOR ? Q_CONDUCTOR : t == MTSB_BSDF_ROUGHDIELECTRIC ? Q_DIELECTRIC
         : (t == MTSB_BSDF_DIELECTRIC || t == MTSB_BSDF_CONDUCTOR) ? Q_DELTA : t == MTSB_BSDF_PLASTIC ? Q_GENERIC : Q_NULL;
}
MTSB_HD int queue_of_material(const MtsbMaterial *mats, int mat) {
    if (mat < 0) return Q_NULL;
    const MtsbMaterial &m = mats[mat];
    const int q = queue_of_type(m.type);
    if (m.tex_reflectance | m.tex_transmittance | m.tex_alpha) return Q_TEXTURED;
    if ((m.flags & MTSB_MAT_TWOSIDED) && m.back >= 0) {
        if (mats[m.back].tex_reflectance | mats[m.back].tex_transmittance | mats[m.back].tex_alpha) return Q_TEXTURED;
        if (queue_of_type(mats[m.back].type) != q) return Q_GENERIC;
    }
    return q;
}

}  // namespace mtsb
