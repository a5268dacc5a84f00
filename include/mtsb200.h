/*
 * mtsb200.h -- C ABI of the B200-native `path` integrator hot path.
 *
 * This is the drop-in boundary.  Everything behind it is
 * hand-written CUDA for sm_100a; everything in front of it is host code that
 * mirrors the reference's plugin surface (Integrator / BSDF / Emitter / Shape /
 * Sensor / Film) and flattens a scene into the POD `MtsbSceneDesc` below.
 *
 * Reference interfaces each entry point replaces (paths relative to the
 * reference tree):
 *   mtsb_scene_create   <- Scene::initialize            src/librender/scene.cpp:322-384
 *                          ShapeKDTree::build           src/librender/skdtree.cpp:68-110
 *   mtsb_trace          <- ShapeKDTree::rayIntersect    src/librender/skdtree.cpp:112-148 (closest)
 *                                                       src/librender/skdtree.cpp:213-232 (any-hit)
 *   mtsb_bsdf_query     <- BSDF::eval/pdf/sample        include/mitsuba/render/bsdf.h
 *                          (diffuse.cpp:110-165, roughconductor.cpp:259-417,
 *                           roughdielectric.cpp:270-615, twosided.cpp:108-184)
 *   mtsb_emitter_sample <- Scene::sampleEmitterDirect   src/librender/scene.cpp:828-855 (no visibility test)
 *   mtsb_film_put       <- ImageBlock::put              include/mitsuba/render/imageblock.h:124-206
 *   mtsb_render         <- SamplingIntegrator::render   src/librender/integrator.cpp:95-129
 *                          MonteCarloIntegrator::renderBlock  src/librender/integrator.cpp:196-244
 *                          MIPathTracer::Li             src/integrators/path/path.cpp:121-296
 *   mtsb_film_develop   <- HDRFilm::develop             src/films/hdrfilm.cpp:481-
 *   mtsb_cancel         <- Integrator::cancel           src/librender/integrator.cpp:90-93
 *   mtsb_multi_render   <- RenderJob / BlockedRenderProcess: one frame's blocks over several workers, merged film
 *                          src/librender/renderproc.cpp:149-156, src/librender/imageproc.cpp:43-78
 *
 * Conventions: plain C types only, opaque handles, `int` status (0 = ok),
 * error text through mtsb_last_error().  No exception crosses this boundary.
 * Caller owns every host buffer; the library owns the device memory behind a
 * handle.  There is NO CPU fallback: every compute entry point fails with
 * MTSB_ERR_NO_DEVICE when no sm_100-class GPU is present.
 */
#ifndef MTSB200_H
#define MTSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTSB_ABI_VERSION 6

/* ---- status codes ---------------------------------------------------- */
enum {
    MTSB_OK = 0,
    MTSB_ERR_INVALID = 1,     /* bad argument / inconsistent scene description */
    MTSB_ERR_NO_DEVICE = 2,   /* no CUDA device (there is no CPU fallback)    */
    MTSB_ERR_CUDA = 3,        /* a CUDA call or kernel failed                 */
    MTSB_ERR_CANCELLED = 4,   /* mtsb_cancel() was called during a render     */
    MTSB_ERR_IO = 5,          /* scene file / mesh file problem               */
    MTSB_ERR_PARSE = 6        /* XML / property error (reference: Log(EError) throws) */
};

/* ---- material table (reference: src/bsdfs) --------------------------- */
enum { MTSB_BSDF_DIFFUSE = 0, MTSB_BSDF_ROUGHCONDUCTOR = 1, MTSB_BSDF_ROUGHDIELECTRIC = 2,
       MTSB_BSDF_NULL = 3 /* shape without a BSDF: absorbs (emitter-only quads) */,
       MTSB_BSDF_DIELECTRIC = 4 /* smooth: dielectric.cpp, both lobes Dirac deltas */,
       MTSB_BSDF_CONDUCTOR = 5  /* smooth: conductor.cpp, one delta reflection lobe */,
       MTSB_BSDF_PLASTIC = 6    /* smooth: plastic.cpp, delta reflection over a diffuse base */ };
enum { MTSB_DISTR_BECKMANN = 0, MTSB_DISTR_GGX = 1 };           /* microfacet.h:48-57 */
enum { MTSB_MAT_TWOSIDED = 1u, MTSB_MAT_SAMPLE_VISIBLE = 2u, MTSB_MAT_NONLINEAR = 4u /* plastic `nonlinear` */ };

typedef struct MtsbMaterial {
    uint32_t type;            /* MTSB_BSDF_*                                          */
    uint32_t flags;           /* MTSB_MAT_*                                           */
    uint32_t distribution;    /* MTSB_DISTR_*                                         */
    int32_t  back;            /* twosided: material used when wi.z<0 (twosided.cpp:111-121); -1 = same */
    float    alpha_u, alpha_v;/* already clamped to >= 1e-4 (microfacet.h:135-136)     */
    float    eta;             /* roughdielectric: intIOR/extIOR (roughdielectric.cpp:199) */
    float    fdr_int;         /* plastic: fresnelDiffuseReflectance(1/eta) (plastic.cpp:196) */
    float    reflectance[3];  /* diffuse `reflectance` / `specularReflectance`; plastic: `diffuseReflectance` */
    float    transmittance[3];/* (rough)dielectric `specularTransmittance`; plastic: `specularReflectance`   */
    float    cond_eta[3];     /* roughconductor eta/extEta (roughconductor.cpp:189)   */
    float    cond_k[3];       /* roughconductor k/extEta                              */
    float    spec_weight;     /* plastic: m_specularSamplingWeight (plastic.cpp:199-204) */
    uint32_t tex_reflectance; /* 1 + index into MtsbSceneDesc::textures of the texture that replaces `reflectance`, 0 = constant */
    uint32_t tex_transmittance; /* same for `transmittance`                            */
    uint32_t tex_alpha;       /* same for the roughness `alpha` of roughconductor / roughdielectric: the texture's channel
                                 average replaces alpha_u and alpha_v at the hit (roughconductor.cpp:379-384,
                                 roughdielectric.cpp:300-305), clamped to >= 1e-4 as MicrofacetDistribution does */
} MtsbMaterial;              /* 96 bytes */

/* ---- textures (reference: src/textures/{bitmap,checkerboard,gridtexture,scale}.cpp, src/librender/texture.cpp) ----
 * A texture replaces a constant spectrum of a material and is evaluated at the hit's uv (Texture2D::eval,
 * texture.cpp:112-121): uv' = uv * uv_scale + uv_offset.  `scale` plugins are folded into `scale`.
 * bitmap: texels are kept in half precision as the reference keeps them (bitmap.cpp:172-177); filter NEAREST /
 * BILINEAR look up level 0 (mipmap.h:565-596); TRILINEAR / EWA use the Lanczos-2 MIP pyramid (mipmap.h:248-276,
 * built at scene creation) with the uv partials of the camera ray on the first hit (intersection.cpp:5-98,
 * mipmap.h:628-716) and the level-0 bilinear lookup on every later hit (bitmap.cpp:431-453).                */
enum { MTSB_TEX_BITMAP = 0, MTSB_TEX_CHECKERBOARD = 1, MTSB_TEX_GRID = 2 };
enum { MTSB_TEXFILTER_NEAREST = 0, MTSB_TEXFILTER_BILINEAR = 1, MTSB_TEXFILTER_TRILINEAR = 2, MTSB_TEXFILTER_EWA = 3 }; /* mipmap.h:54-63 */
enum { MTSB_WRAP_CLAMP = 0, MTSB_WRAP_REPEAT = 1, MTSB_WRAP_MIRROR = 2, MTSB_WRAP_ZERO = 3, MTSB_WRAP_ONE = 4 };  /* rfilter.h:53-64 EBoundaryCondition, same numbering */

typedef struct MtsbTexture {
    uint32_t type;                  /* MTSB_TEX_*                                        */
    uint32_t filter;                /* bitmap: MTSB_TEXFILTER_*                          */
    uint32_t wrap_u, wrap_v;        /* bitmap: MTSB_WRAP_*                               */
    float    uv_scale[2];           /* Texture2D `uscale`/`vscale` (texture.cpp:87-91)   */
    float    uv_offset[2];          /* `uoffset`/`voffset`                               */
    float    color0[3], color1[3];  /* checkerboard.cpp:46-47 / gridtexture.cpp:47-48    */
    float    line_width;            /* gridtexture `lineWidth`                           */
    float    scale[3];              /* product of the enclosing `scale` textures (scale.cpp:70), 1 = none */
    float    max_anisotropy;        /* bitmap `maxAnisotropy` (bitmap.cpp:232-235)       */
    int32_t  width, height;         /* bitmap: level-0 size, < 65536 each                */
    int32_t  channels;              /* bitmap: 1 (luminance) or 3 (linear RGB)           */
    const float *texels;            /* bitmap: width*height*channels linear values, row-major, row 0 = v 0 */
} MtsbTexture;                      /* 96 bytes */

/* ---- shapes: every shape is a triangle mesh (createTriMesh'd on the host) */
typedef struct MtsbShape {
    uint32_t first_vertex, n_vertices;
    uint32_t first_tri, n_tris;     /* global prim id = first_tri + local index  (skdtree.cpp:68-70) */
    int32_t  material;              /* index into materials, -1 = none            */
    int32_t  emitter;               /* index into emitters (area light), -1       */
    uint32_t has_normals;           /* vertex normals present (skdtree.h:383-396) */
    uint32_t has_texcoords;
} MtsbShape;                        /* 32 bytes */

/* ---- emitters (reference: src/emitters/{area,point,constant,envmap}.cpp) ----- */
enum { MTSB_EMITTER_AREA = 0, MTSB_EMITTER_POINT = 1, MTSB_EMITTER_CONSTANT = 2,
       MTSB_EMITTER_ENVMAP = 3 /* image data in MtsbSceneDesc::envmap; `radiance` unused */ };

typedef struct MtsbEmitter {
    uint32_t type;
    int32_t  shape;                 /* area: owning shape                              */
    float    radiance[3];           /* area/constant `radiance`, point `intensity`    */
    float    position[3];           /* point: world position                           */
    float    sampling_weight;       /* Emitter::getSamplingWeight, default 1           */
    float    _pad[3];
} MtsbEmitter;                      /* 48 bytes */

/* Latitude-longitude environment map (envmap.cpp): the image is stored in half precision on the device as the
 * reference stores it (envmap.cpp:101-103), looked up bilinearly at level 0 (repeat in u, clamp in v) and importance
 * sampled through marginal / conditional CDFs over luminance x sin(theta) (envmap.cpp:260-329, 567-600).           */
typedef struct MtsbEnvmap {
    int32_t width, height;          /* < 65536 each (envmap.cpp:164-166)                */
    const float *rgb;               /* width*height*3 linear RGB, row-major, row 0 = +Y pole */
    float to_world[16];             /* emitter `toWorld`, row-major; only its linear part acts on directions */
    float to_local[16];             /* its inverse (Transform keeps both, transform.h)  */
    float scale;                    /* `scale`, default 1 (envmap.cpp:188)              */
    uint32_t _pad[3];
} MtsbEnvmap;

/* ---- sensor + film + filter ------------------------------------------- */
enum { MTSB_RFILTER_GAUSSIAN = 0, MTSB_RFILTER_BOX = 1 };

typedef struct MtsbSensor {
    float to_world[16];             /* row-major camera-to-world (no scale: perspective.cpp:116-118) */
    float fov_x_deg;                /* resolved x field of view (sensor.cpp:289-309)   */
    float near_clip, far_clip;      /* defaults 1e-2, 1e4 (sensor.cpp:162-164)         */
    int32_t film_width, film_height;/* full film size                                  */
    int32_t crop_x, crop_y, crop_width, crop_height;  /* film.cpp:28-54                */
    uint32_t rfilter;               /* MTSB_RFILTER_*                                  */
    float rfilter_radius;           /* gaussian: 4*stddev (gaussian.cpp:41); box: 0.5  */
    float rfilter_stddev;           /* gaussian only                                   */
} MtsbSensor;

typedef struct MtsbIntegrator {
    int32_t max_depth;              /* -1 = unbounded (integrator.cpp:259)             */
    int32_t rr_depth;               /* default 5                                       */
    int32_t strict_normals;
    int32_t hide_emitters;
    int32_t sample_count;           /* sampler `sampleCount`, default 4 (independent.cpp:57) */
    uint32_t _pad;
} MtsbIntegrator;

/* ---- sample generator (reference: src/samplers/{independent,ldsampler,halton}.cpp) ----
 * independent: i.i.d. uniforms (counter-based here, independent.cpp:97-105).
 * ldsampler:   randomly scrambled (0,2)-sequences for the first `dimension` 1D and 2D draws of a
 *              sample, shuffled per pixel, independent numbers afterwards (ldsampler.cpp:152-215);
 *              sample_count must be a power of two (the host plugin rounds up, ldsampler.cpp:82-86).
 * halton:      the Halton sequence enumerated per pixel (Gruenschloss et al., halton.cpp:240-330),
 *              deterministic: scramble -1 = Faure permutations (default), 0 = none, any other value =
 *              pseudorandom permutations seeded by it (faure.cpp:34-74).
 * sobol:       the Sobol' sequence (Joe-Kuo direction numbers) enumerated per pixel (sobol.cpp:204-252,
 *              sobolseq.h:43-134), deterministic; scramble 0 = none (default), otherwise the XOR mask is
 *              derived from it by TEA (sobol.cpp:96-102).                                                */
enum { MTSB_SAMPLER_INDEPENDENT = 0, MTSB_SAMPLER_LDSAMPLER = 1, MTSB_SAMPLER_HALTON = 2, MTSB_SAMPLER_SOBOL = 3 };
typedef struct MtsbSampler {
    uint32_t type;                  /* MTSB_SAMPLER_*                                  */
    int32_t  dimension;             /* ldsampler `dimension`, default 4 (ldsampler.cpp:78) */
    int32_t  scramble;              /* halton `scramble`, default -1 (halton.cpp:126); sobol `scramble`, default 0 (sobol.cpp:91) */
    uint32_t _pad;
} MtsbSampler;

typedef struct MtsbSceneDesc {
    uint32_t abi_version;           /* MTSB_ABI_VERSION                                */
    uint32_t n_shapes, n_materials, n_emitters;
    uint64_t n_vertices, n_tris;
    const MtsbShape    *shapes;
    const MtsbMaterial *materials;
    const MtsbEmitter  *emitters;
    const float    *positions;      /* 3*n_vertices                                    */
    const float    *normals;        /* 3*n_vertices or NULL; ignored where !has_normals */
    const float    *texcoords;      /* 2*n_vertices or NULL                            */
    const uint32_t *indices;        /* 3*n_tris, GLOBAL vertex indices                 */
    MtsbSensor      sensor;
    MtsbIntegrator  integrator;
    MtsbSampler     sampler;        /* zero-initialised = independent                  */
    const MtsbEnvmap *envmap;       /* required iff an emitter has type MTSB_EMITTER_ENVMAP, else NULL */
    const MtsbTexture *textures;    /* n_textures records, NULL when there are none     */
    uint32_t n_textures;
    uint32_t _pad;
} MtsbSceneDesc;

/* ---- rays and hit records (cfg 4 and the parity tests) ---------------- */
typedef struct MtsbRay { float o[3]; float mint; float d[3]; float maxt; } MtsbRay;   /* 32 B */
/* prim = global primitive index (first_tri + local), 0xFFFFFFFF on a miss;
 * u,v = barycentrics of vertices 1 and 2 (triaccel.h:160-167); t = distance. */
typedef struct MtsbHit { uint32_t prim; float u, v, t; } MtsbHit;                      /* 16 B */
#define MTSB_NO_HIT 0xFFFFFFFFu
enum { MTSB_TRACE_CLOSEST = 0, MTSB_TRACE_ANY = 1 };

/* ---- BSDF query records (parity tests; mirrors BSDFSamplingRecord) ---- */
typedef struct MtsbBsdfQuery {
    float wi[3];                    /* local frame                                      */
    float wo[3];                    /* eval/pdf input                                   */
    float u[3];                     /* sample(): 2D sample + the extra next1D() of
                                       roughdielectric.cpp:554                          */
} MtsbBsdfQuery;
typedef struct MtsbBsdfResult {
    float eval[3];  float pdf;      /* eval(bRec)*cos, pdf(bRec) for the given wo       */
    float weight[3]; float spdf;    /* sample(bRec,pdf,u): weight, pdf                  */
    float swo[3];   float seta;     /* sampled wo, bRec.eta                              */
} MtsbBsdfResult;

/* ---- direct-illumination sampling records (parity tests) -------------- */
typedef struct MtsbDirectQuery { float ref[3]; float ref_n[3]; float u[2]; } MtsbDirectQuery;
typedef struct MtsbDirectResult {
    float value[3]; float pdf;      /* radiance/pdf incl. emitter-pick pdf, solid-angle pdf */
    float d[3]; float dist;
    float p[3]; int32_t emitter;    /* sampled point and emitter index (-1: pdf==0)         */
    float n[3]; int32_t measure;    /* 0 solid angle, 1 discrete                             */
} MtsbDirectResult;

/* ---- render parameters / statistics ----------------------------------- */
typedef struct MtsbRenderParams {
    int32_t  spp;                   /* samples per pixel to render in this call (<=0: scene sampleCount) */
    int32_t  spp_offset;            /* first global sample index (sample-split across ranks) */
    uint64_t seed;                  /* counter-based RNG key                                  */
    /* tile partition: this call renders 32x32 tiles t with t % tile_count == tile_index   */
    int32_t  tile_index, tile_count;
    int32_t  spp_per_pass;          /* <=0: automatic                                         */
    uint32_t flags;                 /* MTSB_RENDER_*                                          */
} MtsbRenderParams;
enum { MTSB_RENDER_KEEP_FILM = 1u   /* accumulate onto the device film instead of clearing */ };

typedef struct MtsbStats {
    uint64_t samples;               /* path samples started                               */
    uint64_t closest_rays;          /* == reference `raysTraced` (skdtree.cpp:46)         */
    uint64_t shadow_rays;           /* == reference `shadowRaysTraced` (skdtree.cpp:47)   */
    uint64_t shade_events;          /* surface interactions shaded                        */
    uint64_t path_length_sum;       /* sum of rRec.depth (path.cpp:292-293)               */
    uint64_t kernel_launches;       /* kernels of this library launched in the call       */
    uint64_t bad_samples;           /* non-finite samples rejected (imageblock.h:147-153) */
    float    ms_total;              /* device time of the call (CUDA events)              */
    float    ms_trace_closest, ms_trace_any, ms_shade, ms_raygen, ms_splat;
    float    ms_h2d, ms_d2h;
    uint32_t n_passes;
    uint32_t closest_launches;      /* k_trace<closest> launches that had rays (speculative empty launches excluded) */
    uint32_t any_launches;          /* k_trace<any> launches that had rays                 */
    uint32_t n_devices;             /* GPUs that took part (mtsb_multi_render; 1 otherwise) */
    float    ms_reduce;             /* mtsb_multi_render: NCCL reduce of the film (device time on device 0) */
    uint32_t _pad;
    /* rays in flight per bounce, summed over the passes of the call: [b] = closest-hit rays of bounce b
     * (b = 0: camera rays), shadow rays generated at bounce b; bounces >= 15 add to the last entry */
    uint64_t closest_rays_bounce[16], shadow_rays_bounce[16];
} MtsbStats;

typedef struct MtsbSceneInfo {
    uint64_t n_tris, n_bvh_nodes;
    uint64_t bvh_bytes, tri_bytes;
    float    aabb_min[3], aabb_max[3];
    float    build_ms;
    uint32_t n_emitters;
} MtsbSceneInfo;

typedef struct mtsb_scene_s *mtsb_scene_t;

/* ---- entry points ----------------------------------------------------- */
int  mtsb_abi_version(void);
const char *mtsb_last_error(void);
/* Bind this process to one CUDA device (one process per GPU). device<0: current. */
int  mtsb_init(int device);
int  mtsb_device_count(void);

int  mtsb_scene_create(const MtsbSceneDesc *desc, mtsb_scene_t *out);
int  mtsb_scene_info(mtsb_scene_t scene, MtsbSceneInfo *info);
void mtsb_scene_destroy(mtsb_scene_t scene);

/* Host-buffer ray query: copies rays in, traces, copies hits out. */
int  mtsb_trace(mtsb_scene_t scene, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits);
/* Device-resident variant: d_rays/d_hits are device pointers; `stream` is a cudaStream_t. The kernel is ordered after
 * the work already queued on that stream (pass the stream that produced d_rays). NULL selects the scene's own
 * non-blocking stream: then the caller must have synchronised the producer of d_rays itself.
 * ms_kernel (optional) receives the kernel time measured with CUDA events on that stream. */
int  mtsb_trace_device(mtsb_scene_t scene, const void *d_rays, size_t n, int mode, void *d_hits,
                       void *stream, float *ms_kernel);

int  mtsb_bsdf_query(mtsb_scene_t scene, int material, const MtsbBsdfQuery *q, size_t n, MtsbBsdfResult *out);
int  mtsb_emitter_sample(mtsb_scene_t scene, const MtsbDirectQuery *q, size_t n, MtsbDirectResult *out);
/* Intersection record at a hit: p, geometric normal, shading frame (n,s,t), uv: 17 floats per hit
 * (skdtree.h:343-430). */
int  mtsb_fill_intersection(mtsb_scene_t scene, const MtsbRay *rays, const MtsbHit *hits, size_t n, float *out17);
/* Primary rays for pixel samples (perspective.cpp:286-317): pos = 2 floats per sample. */
int  mtsb_sensor_rays(mtsb_scene_t scene, const float *sample_pos, size_t n, MtsbRay *rays);

/* Splat n samples (pos.xy, rgb, alpha = 6 floats each) into the scene's device film. */
int  mtsb_film_clear(mtsb_scene_t scene);
int  mtsb_film_put(mtsb_scene_t scene, const float *samples6, size_t n);
/* Raw film: (crop_h+2b) x (crop_w+2b) x 5 floats (R,G,B,alpha,weight), b = border. */
int  mtsb_film_size(mtsb_scene_t scene, int *width, int *height, int *border);
int  mtsb_film_download(mtsb_scene_t scene, float *film5);
/* Device pointer of the raw film (for the NCCL reduce in the caller's process). */
int  mtsb_film_device_ptr(mtsb_scene_t scene, void **d_film5, size_t *n_floats);
/* Develop: rgb = sum(w*L)/sum(w), cropped: crop_h x crop_w x 3 floats. */
int  mtsb_film_develop(mtsb_scene_t scene, float *rgb);

/* Render: generate, trace, shade, splat. Film stays on the device. */
int  mtsb_render(mtsb_scene_t scene, const MtsbRenderParams *params, MtsbStats *stats);
int  mtsb_cancel(mtsb_scene_t scene);

/* ---- several GPUs of one node inside the library (reference: the block scheduler hands the blocks of ONE frame to
 * its workers and merges them into one film: src/librender/renderproc.cpp:149-156, imageproc.cpp:43-78).
 * mtsb_multi_create replicates the scene on n_devices GPUs (one host thread per GPU uploads it and builds its BVH) and
 * opens an NCCL communicator over them; mtsb_multi_render renders the 32x32 tiles t with t % n_devices == i on device i
 * (every sample of those tiles; params->tile_index / tile_count must be 0) and sums the bordered per-device films onto
 * device_ids[0] with ONE ncclReduce over NVLink; the film calls below then act on that device.  NCCL is loaded at run
 * time (libnccl.so.2) by mtsb_multi_create only; with n_devices == 1 no NCCL is needed.  stats are summed over the
 * devices, times are the maximum over devices.                                                                      */
typedef struct mtsb_multi_s *mtsb_multi_t;
int  mtsb_multi_create(const MtsbSceneDesc *desc, int n_devices, const int *device_ids /* NULL: 0..n-1 */, mtsb_multi_t *out);
int  mtsb_multi_render(mtsb_multi_t m, const MtsbRenderParams *params, MtsbStats *stats);
int  mtsb_multi_cancel(mtsb_multi_t m);
mtsb_scene_t mtsb_multi_scene(mtsb_multi_t m, int i);   /* scene on the i-th device (i = 0 holds the reduced film) */
int  mtsb_multi_device_count(mtsb_multi_t m);
void mtsb_multi_destroy(mtsb_multi_t m);

/* ---- allocation-free ray query --------------------------------------------------------------------------------
 * mtsb_trace_static has NO counterpart in the reference; it was added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * inputs, outputs and results as mtsb_trace (host buffers; mode = MTSB_TRACE_CLOSEST / MTSB_TRACE_ANY), same kernel, but
 * every byte of scratch (ray staging, hit staging, the persistent kernel's work counter) lives in one file-scope device
 * array of fixed size, so the call allocates nothing: no cudaMalloc, no cudaFree, no allocating library routine.
 * Limit: n <= MTSB_STATIC_TRACE_MAX_RAYS per call (larger n: MTSB_ERR_INVALID, nothing is traced); the array is
 * 12 MiB + 256 B and every GPU that loads the library holds it until the process exits.  All calls share the array, which
 * keeps the last call's contents (the work counter is cleared before every launch): calls that overlap on one GPU
 * race -- serialise them.  The call returns after the hits are in `hits`.                                            */
#define MTSB_STATIC_TRACE_MAX_RAYS 262144
int  mtsb_trace_static(mtsb_scene_t scene, const MtsbRay *rays, size_t n, int mode, MtsbHit *hits);
int  mtsb_trace_static_max_rays(void);

/* Synthetic workload generators used by bench.py / tests (device-side ray sets of cfg 4). */
int  mtsb_gen_rays_device(mtsb_scene_t scene, int kind /*0 camera-coherent, 1 bsphere chords*/,
                          size_t n, uint64_t seed, void *d_rays);

#ifdef __cplusplus
}
#endif
#endif /* MTSB200_H */
