// query.cu -- record-level entry points of the C ABI (parity tests call the same device
// functions the wavefront kernels use): mtsb_bsdf_query, mtsb_emitter_sample,
// mtsb_fill_intersection, mtsb_sensor_rays, and the film entry points.
#include "scene.h"
#include "shading.h"
#include "sensor.h"

namespace mtsb {

__global__ void k_bsdf_query(SceneView sc, int material, const MtsbBsdfQuery *q, size_t n, MtsbBsdfResult *out) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 wi = ld3(q[i].wi), wo = ld3(q[i].wo);
    V3 e = bsdf_eval(sc.materials, material, wi, wo);
    MtsbBsdfResult r;
    st3(r.eval, e);
    r.pdf = bsdf_pdf(sc.materials, material, wi, wo);
    BsdfSample bs; bs.wo = V3(0.0f); bs.weight = V3(0.0f); bs.pdf = 0; bs.eta = 1;
    P2 u; u.x = q[i].u[0]; u.y = q[i].u[1];
    bool ok = bsdf_sample(sc.materials, material, wi, u, q[i].u[2], bs);
    if (!ok) { bs.weight = V3(0.0f); bs.pdf = 0; bs.wo = V3(0.0f); bs.eta = 1; }
    st3(r.weight, bs.weight); r.spdf = bs.pdf; st3(r.swo, bs.wo); r.seta = bs.eta;
    out[i] = r;
}

__global__ void k_emitter_sample(SceneView sc, const MtsbDirectQuery *q, size_t n, MtsbDirectResult *out) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    DRec d; d.ref = ld3(q[i].ref); d.refN = ld3(q[i].ref_n);
    d.p = d.n = d.d = V3(0.0f); d.dist = 0; d.measure = 0;
    P2 u; u.x = q[i].u[0]; u.y = q[i].u[1];
    V3 v = sample_emitter_direct(sc, d, u);
    MtsbDirectResult r;
    if (d.emitter < 0) { memset(&r, 0, sizeof(r)); r.emitter = -1; out[i] = r; return; }
    st3(r.value, v); r.pdf = d.pdf; st3(r.d, d.d); r.dist = d.dist; st3(r.p, d.p); r.emitter = d.emitter;
    st3(r.n, d.n); r.measure = d.measure;
    out[i] = r;
}

__global__ void k_fill_intersection(SceneView sc, const MtsbRay *rays, const MtsbHit *hits, size_t n, float *out17) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float *o = out17 + 17 * i;
    if (hits[i].prim == MTSB_NO_HIT) { for (int k = 0; k < 17; ++k) o[k] = 0; return; }
    Its its;
    fill_intersection(sc, ld3(rays[i].d), hits[i].prim, hits[i].u, hits[i].v, its);
    st3(o, its.p); st3(o + 3, its.geoN); st3(o + 6, its.sh.n); st3(o + 9, its.sh.s); st3(o + 12, its.sh.t);
    o[15] = its.uv.x; o[16] = its.uv.y;
}

__global__ void k_sensor_rays(SceneView sc, const float *pos, size_t n, MtsbRay *rays) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RayIn r;
    sensor_ray(sc.sensor, sc.near_x, sc.near_y, pos[2 * i], pos[2 * i + 1], r);
    MtsbRay o; o.o[0] = r.ox; o.o[1] = r.oy; o.o[2] = r.oz; o.mint = r.mint; o.d[0] = r.dx; o.d[1] = r.dy; o.d[2] = r.dz; o.maxt = r.maxt;
    rays[i] = o;
}

// ImageBlock::put for an arbitrary sample list (imageblock.h:124-206): one thread per sample,
// red.global.add into the bordered film.  The renderer itself uses the tiled kernel in render.cu.
__global__ void k_film_put(SceneView sc, const float *s6, size_t n, unsigned long long *bad) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *s = s6 + 6 * i;
    float v[5] = {s[2], s[3], s[4], s[5], 1.0f};
    for (int k = 0; k < 5; ++k) if (!isfinite(v[k])) { atomicAdd(bad, 1ull); return; }
    const int sx = sc.film_w + 2 * sc.film_border, sy = sc.film_h + 2 * sc.film_border;
    Footprint f;
    film_footprint(sc, s[0], s[1], 0, 0, sx, sy, f);
    for (int y = 0; y < f.ny; ++y)
        for (int x = 0; x < f.nx; ++x) {
            const float w = f.wx[x] * f.wy[y];
            float *dest = sc.film + ((size_t) (f.miny + y) * sx + (f.minx + x)) * 5;
            for (int k = 0; k < 5; ++k) atomicAdd(dest + k, w * v[k]);
        }
}

// HDRFilm::develop -> Bitmap::convert from ESpectrumAlphaWeight (hdrfilm.cpp:481-, bitmap.cpp:422-426)
__global__ void k_develop(SceneView sc, float *rgb) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) sc.film_w * sc.film_h) return;
    int x = (int) (i % sc.film_w), y = (int) (i / sc.film_w);
    const float *p = sc.film + ((size_t) (y + sc.film_border) * (sc.film_w + 2 * sc.film_border) + x + sc.film_border) * 5;
    float w = p[4], inv = w != 0 ? 1.0f / w : 0.0f;
    rgb[3 * i] = p[0] * inv; rgb[3 * i + 1] = p[1] * inv; rgb[3 * i + 2] = p[2] * inv;
}

// host <-> device round trip helper for the record-level entry points
template <class In, class Out, class Launch>
static int roundtrip(Scene *s, const In *in, size_t n_in, Out *out, size_t n_out, Launch launch) {
    MTSB_CUDA(cudaSetDevice(s->device));
    In *d_in = nullptr; Out *d_out = nullptr;
    MTSB_CUDA(cudaMalloc((void **) &d_in, n_in * sizeof(In)));
    cudaError_t e = cudaMalloc((void **) &d_out, n_out * sizeof(Out));
    if (e != cudaSuccess) { cudaFree(d_in); return cuda_fail(e, "cudaMalloc", __FILE__, __LINE__); }
    e = cudaMemcpyAsync(d_in, in, n_in * sizeof(In), cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) { launch(d_in, d_out); s->launches++; e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, n_out * sizeof(Out), cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) return cuda_fail(e, "record query", __FILE__, __LINE__);
    return MTSB_OK;
}

}  // namespace mtsb

using namespace mtsb;

extern "C" int mtsb_bsdf_query(mtsb_scene_t scene, int material, const MtsbBsdfQuery *q, size_t n, MtsbBsdfResult *out) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!q || !out))) { set_error("mtsb_bsdf_query: bad argument"); return MTSB_ERR_INVALID; }
    if (material < 0 || (uint32_t) material >= s->v.n_materials) { set_error("mtsb_bsdf_query: bad material"); return MTSB_ERR_INVALID; }
    if (!n) return MTSB_OK;
    SceneView v = s->v; cudaStream_t st = s->stream;
    return roundtrip(s, q, n, out, n, [&](MtsbBsdfQuery *di, MtsbBsdfResult *dout) { k_bsdf_query<<<grid_for(n, 128), 128, 0, st>>>(v, material, di, n, dout); });
}

extern "C" int mtsb_emitter_sample(mtsb_scene_t scene, const MtsbDirectQuery *q, size_t n, MtsbDirectResult *out) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!q || !out))) { set_error("mtsb_emitter_sample: bad argument"); return MTSB_ERR_INVALID; }
    if (!n) return MTSB_OK;
    SceneView v = s->v; cudaStream_t st = s->stream;
    return roundtrip(s, q, n, out, n, [&](MtsbDirectQuery *di, MtsbDirectResult *dout) { k_emitter_sample<<<grid_for(n, 128), 128, 0, st>>>(v, di, n, dout); });
}

extern "C" int mtsb_fill_intersection(mtsb_scene_t scene, const MtsbRay *rays, const MtsbHit *hits, size_t n, float *out17) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!rays || !hits || !out17))) { set_error("mtsb_fill_intersection: bad argument"); return MTSB_ERR_INVALID; }
    if (!n) return MTSB_OK;
    for (size_t i = 0; i < n; ++i)
        if (hits[i].prim != MTSB_NO_HIT && hits[i].prim >= s->info.n_tris) { set_error("mtsb_fill_intersection: primitive index out of range"); return MTSB_ERR_INVALID; }
    MTSB_CUDA(cudaSetDevice(s->device));
    MtsbHit *d_hits = nullptr;
    MTSB_CUDA(cudaMalloc((void **) &d_hits, n * sizeof(MtsbHit)));
    cudaError_t e = cudaMemcpyAsync(d_hits, hits, n * sizeof(MtsbHit), cudaMemcpyHostToDevice, s->stream);
    if (e != cudaSuccess) { cudaFree(d_hits); return cuda_fail(e, "cudaMemcpy", __FILE__, __LINE__); }
    SceneView v = s->v; cudaStream_t st = s->stream;
    int rc = roundtrip(s, rays, n, out17, 17 * n, [&](MtsbRay *dr, float *dout) { k_fill_intersection<<<grid_for(n, 128), 128, 0, st>>>(v, dr, d_hits, n, dout); });
    cudaFree(d_hits);
    return rc;
}

extern "C" int mtsb_sensor_rays(mtsb_scene_t scene, const float *sample_pos, size_t n, MtsbRay *rays) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && (!sample_pos || !rays))) { set_error("mtsb_sensor_rays: bad argument"); return MTSB_ERR_INVALID; }
    if (!n) return MTSB_OK;
    SceneView v = s->v; cudaStream_t st = s->stream;
    return roundtrip(s, sample_pos, 2 * n, rays, n, [&](float *dp, MtsbRay *dr) { k_sensor_rays<<<grid_for(n, 128), 128, 0, st>>>(v, dp, n, dr); });
}

// ---- film --------------------------------------------------------------------------------
static size_t film_floats(const Scene *s) { return (size_t) (s->v.film_w + 2 * s->v.film_border) * (s->v.film_h + 2 * s->v.film_border) * 5; }

extern "C" int mtsb_film_clear(mtsb_scene_t scene) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s) { set_error("mtsb_film_clear: bad argument"); return MTSB_ERR_INVALID; }
    MTSB_CUDA(cudaSetDevice(s->device));
    MTSB_CUDA(cudaMemsetAsync(s->v.film, 0, film_floats(s) * sizeof(float), s->stream));
    MTSB_CUDA(cudaMemsetAsync(s->d_film_bad, 0, sizeof(unsigned long long), s->stream));
    MTSB_CUDA(cudaStreamSynchronize(s->stream));
    return MTSB_OK;
}

extern "C" int mtsb_film_size(mtsb_scene_t scene, int *width, int *height, int *border) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s) { set_error("mtsb_film_size: bad argument"); return MTSB_ERR_INVALID; }
    if (width) *width = s->v.film_w;
    if (height) *height = s->v.film_h;
    if (border) *border = s->v.film_border;
    return MTSB_OK;
}

extern "C" int mtsb_film_put(mtsb_scene_t scene, const float *samples6, size_t n) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || (n && !samples6)) { set_error("mtsb_film_put: bad argument"); return MTSB_ERR_INVALID; }
    if (!n) return MTSB_OK;
    MTSB_CUDA(cudaSetDevice(s->device));
    float *d = nullptr;
    MTSB_CUDA(cudaMalloc((void **) &d, 6 * n * sizeof(float)));
    cudaError_t e = cudaMemcpyAsync(d, samples6, 6 * n * sizeof(float), cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) { k_film_put<<<grid_for(n, 128), 128, 0, s->stream>>>(s->v, d, n, s->d_film_bad); s->launches++; e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cudaFree(d);
    if (e != cudaSuccess) return cuda_fail(e, "mtsb_film_put", __FILE__, __LINE__);
    return MTSB_OK;
}

extern "C" int mtsb_film_download(mtsb_scene_t scene, float *film5) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || !film5) { set_error("mtsb_film_download: bad argument"); return MTSB_ERR_INVALID; }
    MTSB_CUDA(cudaSetDevice(s->device));
    MTSB_CUDA(cudaMemcpyAsync(film5, s->v.film, film_floats(s) * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    MTSB_CUDA(cudaStreamSynchronize(s->stream));
    return MTSB_OK;
}

extern "C" int mtsb_film_device_ptr(mtsb_scene_t scene, void **d_film5, size_t *n_floats) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || !d_film5) { set_error("mtsb_film_device_ptr: bad argument"); return MTSB_ERR_INVALID; }
    *d_film5 = s->v.film;
    if (n_floats) *n_floats = film_floats(s);
    return MTSB_OK;
}

extern "C" int mtsb_film_develop(mtsb_scene_t scene, float *rgb) {
    Scene *s = reinterpret_cast<Scene *>(scene);
    if (!s || !rgb) { set_error("mtsb_film_develop: bad argument"); return MTSB_ERR_INVALID; }
    MTSB_CUDA(cudaSetDevice(s->device));
    size_t n = (size_t) s->v.film_w * s->v.film_h;
    if (!s->d_rgb) { int rc = dev_alloc(s, &s->d_rgb, 3 * n); if (rc) return rc; }      // kept for the life of the scene
    float *d = s->d_rgb;
    k_develop<<<grid_for(n, 256), 256, 0, s->stream>>>(s->v, d);
    s->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb, d, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "mtsb_film_develop", __FILE__, __LINE__);
    return MTSB_OK;
}
