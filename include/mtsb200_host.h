/*
 * mtsb200_host.h -- C ABI of the host side in front of mtsb200.h: scene XML loading through the
 * plugin mirror (mitsuba_b200/host) and the render-job driver.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   mtsbh_load_file / mtsbh_load_string <- SceneHandler + SAX parse    src/librender/scenehandler.cpp:197-751
 *                                          (as driven by mitsuba.cpp:330-372 / Utility::loadScene)
 *   mtsbh_scene_desc                    <- Scene::initialize flattening src/librender/scene.cpp:322-384
 *   mtsbh_render                        <- RenderJob::run               src/librender/renderjob.cpp:91-150
 *                                          (preprocess -> render -> film develop)
 *   mtsbh_render_multi                  <- the same job with the scheduler's workers = the GPUs of the node
 *                                          (Scheduler / BlockedRenderProcess, src/librender/renderproc.cpp:149-156)
 *   mtsbh_write_pfm                     <- Bitmap::write(EPFM)          src/libcore/bitmap.cpp (PFM writer)
 *
 * The library (libmtsb200host.so) is plain C++ and links against libmtsb200.so; it never computes
 * radiance itself.
 */
#ifndef MTSB200_HOST_H
#define MTSB200_HOST_H

#include "mtsb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mtsbh_scene_s *mtsbh_scene_t;

const char *mtsbh_last_error(void);

/* `defines`: NULL or a NULL-terminated array of "name=value" strings (the -D options of the
 * reference's command line, mitsuba.cpp:190-199). Returns MTSB_OK, MTSB_ERR_IO or MTSB_ERR_PARSE. */
int  mtsbh_load_file(const char *path, const char *const *defines, mtsbh_scene_t *out);
int  mtsbh_load_string(const char *xml, const char *name, const char *const *defines, mtsbh_scene_t *out);
void mtsbh_scene_free(mtsbh_scene_t scene);

/* The flattened scene; pointers stay valid until mtsbh_scene_free. */
const MtsbSceneDesc *mtsbh_scene_desc(mtsbh_scene_t scene);
/* `seed` of the sampler (independent.cpp:58, default 1004). */
int  mtsbh_scene_seed(mtsbh_scene_t scene);
/* Log(EWarn) messages collected while loading (unqueried properties etc.). */
int  mtsbh_warning_count(mtsbh_scene_t scene);
const char *mtsbh_warning(mtsbh_scene_t scene, int i);

/* Render job: builds the device scene (mtsb_scene_create), renders `spp` samples per pixel
 * (<=0: the sampler's sampleCount) with mtsb_render, develops the film into
 * rgb[crop_height*crop_width*3]. `stats` may be NULL. */
int  mtsbh_render(mtsbh_scene_t scene, int spp, float *rgb, MtsbStats *stats);
/* The same job on `n_devices` GPUs of this node (device ids 0..n-1): the blocks of the frame are distributed over the
 * GPUs and merged into one film, as the reference's scheduler distributes them over its workers
 * (mtsb_multi_create / mtsb_multi_render; src/librender/renderproc.cpp:149-156).  n_devices <= 0: every GPU present. */
int  mtsbh_render_multi(mtsbh_scene_t scene, int spp, int n_devices, float *rgb, MtsbStats *stats);

/* Little-endian colour PFM, bottom row first as the format requires. */
int  mtsbh_write_pfm(const char *path, const float *rgb, int width, int height);

#ifdef __cplusplus
}
#endif
#endif /* MTSB200_HOST_H */
