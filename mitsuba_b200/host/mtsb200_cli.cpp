// mtsb200 -- command line front end, the subset of the reference's `mitsuba` binary that drives the
// path hot path (src/mitsuba/mitsuba.cpp:80-140 usage, 176-260 option parsing, 330-372 job loop):
//     mtsb200 [-D name=value]... [-o out.pfm] [-n spp] [-g gpus] [-q] scene.xml [scene2.xml ...]
// Options of the reference that configure its CPU scheduler or network (-p, -c, -s servers, -j, -r, -b,
// -x, -z) have no meaning here; -p is accepted and ignored so that existing command lines keep working.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mtsb200_host.h"

static void usage() {
    printf("mtsb200 -- B200 path tracer front end\n"
           "Usage: mtsb200 [options] <one or more scene XML files>\n"
           "   -h          Display this help text\n"
           "   -D key=val  Define a constant, which can be referenced as \"$key\" in the scene\n"
           "   -o fname    Write the output image (PFM) to the file denoted by \"fname\"\n"
           "   -n spp      Override the sampler's sampleCount\n"
           "   -g count    Render every frame on this many GPUs of the node (0 = all; default 1): the blocks of the\n"
           "               frame are distributed over the GPUs and merged with one NCCL reduce\n"
           "   -p count    Ignored (CPU core count of the reference)\n"
           "   -q          Quiet mode - do not print any log messages to stdout\n");
}

int main(int argc, char **argv) {
    std::vector<std::string> defs, scenes;
    std::string out;
    bool quiet = false;
    int spp = 0, gpus = 1;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto need = [&](const char *opt) -> const char * {
            if (i + 1 >= argc) { fprintf(stderr, "Option %s needs an argument\n", opt); exit(-1); }
            return argv[++i];
        };
        if (a == "-h") { usage(); return 0; }
        else if (a == "-q") quiet = true;
        else if (a == "-D") defs.push_back(need("-D"));
        else if (a.size() > 2 && a.compare(0, 2, "-D") == 0) defs.push_back(a.substr(2));
        else if (a == "-o") out = need("-o");
        else if (a == "-n") spp = atoi(need("-n"));
        else if (a == "-g") gpus = atoi(need("-g"));
        else if (a == "-p") need("-p");
        else if (!a.empty() && a[0] == '-') { fprintf(stderr, "Unknown option %s\n", a.c_str()); usage(); return -1; }
        else scenes.push_back(a);
    }
    if (scenes.empty()) { usage(); return 0; }
    std::vector<const char *> dptr;
    for (auto &d : defs) dptr.push_back(d.c_str());
    dptr.push_back(nullptr);

    for (const std::string &path : scenes) {
        mtsbh_scene_t scene = nullptr;
        if (mtsbh_load_file(path.c_str(), dptr.data(), &scene)) { fprintf(stderr, "%s\n", mtsbh_last_error()); return -1; }
        if (!quiet)
            for (int i = 0; i < mtsbh_warning_count(scene); ++i) fprintf(stderr, "WARN  %s\n", mtsbh_warning(scene, i));
        const MtsbSceneDesc *d = mtsbh_scene_desc(scene);
        const int w = d->sensor.crop_width, h = d->sensor.crop_height;
        std::vector<float> rgb((size_t) w * h * 3);
        MtsbStats st;
        if (gpus == 1 ? mtsbh_render(scene, spp, rgb.data(), &st) : mtsbh_render_multi(scene, spp, gpus, rgb.data(), &st)) { fprintf(stderr, "%s\n", mtsbh_last_error()); mtsbh_scene_free(scene); return -1; }
        // mitsuba.cpp:352-360: the default destination is the scene file name with the film's extension
        std::string dest = out;
        if (dest.empty()) {
            dest = path;
            size_t dot = dest.rfind('.');
            if (dot != std::string::npos) dest = dest.substr(0, dot);
            dest += ".pfm";
        }
        if (mtsbh_write_pfm(dest.c_str(), rgb.data(), w, h)) { fprintf(stderr, "%s\n", mtsbh_last_error()); mtsbh_scene_free(scene); return -1; }
        if (!quiet)
            printf("%s: %llu triangles, %dx%d, %llu samples in %.1f ms (%.1f M samples/s), %llu rays -> %s\n", path.c_str(),
                   (unsigned long long) d->n_tris, w, h, (unsigned long long) st.samples, st.ms_total,
                   st.ms_total > 0 ? st.samples / st.ms_total * 1e-3 : 0.0, (unsigned long long) (st.closest_rays + st.shadow_rays), dest.c_str());
        mtsbh_scene_free(scene);
    }
    return 0;
}
