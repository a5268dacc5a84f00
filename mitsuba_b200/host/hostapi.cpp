// hostapi.cpp -- extern "C" surface of the host library (include/mtsb200_host.h).
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/mtsb200_host.h"
#include "loader.h"

using namespace mtsbhost;

struct mtsbh_scene_s {
    std::shared_ptr<Scene> scene;
    FlatScene flat;
    std::vector<std::string> warnings;
};

static thread_local std::string g_error;

const char *mtsbh_last_error(void) { return g_error.c_str(); }

static int finishLoad(std::shared_ptr<Scene> scene, mtsbh_scene_t *out) {
    mtsbh_scene_s *h = new mtsbh_scene_s();
    h->scene = scene;
    scene->flatten(h->flat);
    h->warnings = warnings();
    *out = h;
    return MTSB_OK;
}

static void parseDefines(const char *const *defines, ParameterMap &params) {
    if (!defines) return;
    for (; *defines; ++defines) {
        std::string d(*defines);
        size_t eq = d.find('=');
        if (eq == std::string::npos || eq == 0) throw std::runtime_error("Parameter definition \"" + d + "\" must have the form name=value");
        params[d.substr(0, eq)] = d.substr(eq + 1);
    }
}

int mtsbh_load_file(const char *path, const char *const *defines, mtsbh_scene_t *out) {
    if (!path || !out) { g_error = "mtsbh_load_file: NULL argument"; return MTSB_ERR_INVALID; }
    warnings().clear();
    try {
        FILE *f = fopen(path, "rb");
        if (!f) { g_error = std::string("Unable to open the scene file \"") + path + "\""; return MTSB_ERR_IO; }
        fclose(f);
        ParameterMap params;
        parseDefines(defines, params);
        return finishLoad(loadSceneFile(path, params), out);
    } catch (const std::exception &e) {
        g_error = e.what();
        return MTSB_ERR_PARSE;
    }
}

int mtsbh_load_string(const char *xml, const char *name, const char *const *defines, mtsbh_scene_t *out) {
    if (!xml || !out) { g_error = "mtsbh_load_string: NULL argument"; return MTSB_ERR_INVALID; }
    warnings().clear();
    try {
        ParameterMap params;
        parseDefines(defines, params);
        return finishLoad(loadSceneString(xml, name ? name : "<string>", params), out);
    } catch (const std::exception &e) {
        g_error = e.what();
        return MTSB_ERR_PARSE;
    }
}

void mtsbh_scene_free(mtsbh_scene_t s) { delete s; }
const MtsbSceneDesc *mtsbh_scene_desc(mtsbh_scene_t s) { return s ? &s->flat.desc : nullptr; }
int mtsbh_scene_seed(mtsbh_scene_t s) { return s ? s->flat.seed : 0; }
int mtsbh_warning_count(mtsbh_scene_t s) { return s ? (int) s->warnings.size() : 0; }
const char *mtsbh_warning(mtsbh_scene_t s, int i) { return (s && i >= 0 && i < (int) s->warnings.size()) ? s->warnings[i].c_str() : nullptr; }

int mtsbh_render(mtsbh_scene_t s, int spp, float *rgb, MtsbStats *stats) {
    if (!s || !rgb) { g_error = "mtsbh_render: NULL argument"; return MTSB_ERR_INVALID; }
    int rc = mtsb_init(-1);
    if (rc) { g_error = mtsb_last_error(); return rc; }
    mtsb_scene_t dev = nullptr;
    if ((rc = mtsb_scene_create(&s->flat.desc, &dev))) { g_error = mtsb_last_error(); return rc; }
    MtsbRenderParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.spp = spp;
    rp.seed = (uint64_t) s->flat.seed;
    rp.tile_index = 0;
    rp.tile_count = 1;
    MtsbStats local;
    rc = mtsb_render(dev, &rp, stats ? stats : &local);
    if (!rc) rc = mtsb_film_develop(dev, rgb);
    if (rc) g_error = mtsb_last_error();
    mtsb_scene_destroy(dev);
    return rc;
}

int mtsbh_render_multi(mtsbh_scene_t s, int spp, int n_devices, float *rgb, MtsbStats *stats) {
    if (!s || !rgb) { g_error = "mtsbh_render_multi: NULL argument"; return MTSB_ERR_INVALID; }
    if (n_devices <= 0) n_devices = mtsb_device_count();
    if (n_devices <= 0) { g_error = "no CUDA device present (this library has no CPU fallback)"; return MTSB_ERR_NO_DEVICE; }
    if (n_devices == 1) return mtsbh_render(s, spp, rgb, stats);
    mtsb_multi_t multi = nullptr;
    int rc = mtsb_multi_create(&s->flat.desc, n_devices, nullptr, &multi);
    if (rc) { g_error = mtsb_last_error(); return rc; }
    MtsbRenderParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.spp = spp;
    rp.seed = (uint64_t) s->flat.seed;
    MtsbStats local;
    rc = mtsb_multi_render(multi, &rp, stats ? stats : &local);
    if (!rc) rc = mtsb_film_develop(mtsb_multi_scene(multi, 0), rgb);
    if (rc) g_error = mtsb_last_error();
    mtsb_multi_destroy(multi);
    return rc;
}

int mtsbh_write_pfm(const char *path, const float *rgb, int width, int height) {
    if (!path || !rgb || width <= 0 || height <= 0) { g_error = "mtsbh_write_pfm: bad argument"; return MTSB_ERR_INVALID; }
    FILE *f = fopen(path, "wb");
    if (!f) { g_error = std::string("Unable to open \"") + path + "\" for writing"; return MTSB_ERR_IO; }
    fprintf(f, "PF\n%d %d\n-1\n", width, height);
    for (int y = height - 1; y >= 0; --y)
        if (fwrite(rgb + (size_t) y * width * 3, sizeof(float), (size_t) width * 3, f) != (size_t) width * 3) {
            fclose(f);
            g_error = "short write";
            return MTSB_ERR_IO;
        }
    fclose(f);
    return MTSB_OK;
}
