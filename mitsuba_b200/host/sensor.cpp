// sensor.cpp -- perspective sensor, hdrfilm, reconstruction filter, independent sampler and the
// `path` integrator's property handling.
#include <cmath>
#include <cstring>

#include "plugins.h"

namespace mtsbhost {

// ---- reconstruction filters (src/rfilters/gaussian.cpp:33-42, box.cpp:36-40) ---------------------
RFilter::RFilter(const Properties &p) : ConfigurableObject(p) {
    const std::string type = lower(p.getPluginName());
    if (type == "gaussian") {
        kind = MTSB_RFILTER_GAUSSIAN;
        stddev = p.getFloat("stddev", 0.5f);
        radius = 4 * stddev;
    } else if (type == "box") {
        kind = MTSB_RFILTER_BOX;
        stddev = 0;
        radius = p.getFloat("radius", 0.5f) + 1e-5f;
    } else {
        error("Plugin \"" + type + "\" not found -- the B200 path implements the reconstruction filters gaussian and box.");
    }
}

// ---- film ------------------------------------------------------------------------------------------
Film::Film(const Properties &p) : ConfigurableObject(p) {
    const std::string type = lower(p.getPluginName());
    if (type != "hdrfilm")
        error("Plugin \"" + type + "\" not found -- the B200 path implements the film hdrfilm.");
    width = p.getInteger("width", 768);
    height = p.getInteger("height", 576);
    cropX = p.getInteger("cropOffsetX", 0);
    cropY = p.getInteger("cropOffsetY", 0);
    cropW = p.getInteger("cropWidth", width);
    cropH = p.getInteger("cropHeight", height);
    if (cropX < 0 || cropY < 0 || cropW <= 0 || cropH <= 0 || cropX + cropW > width || cropY + cropH > height)
        error("Invalid crop window specification!");
    if (p.getBoolean("highQualityEdges", false))
        error("hdrfilm: 'highQualityEdges' is not part of the B200 path.");
    // output-side properties of hdrfilm.cpp:205-260: accepted, they do not influence the radiance estimate
    p.getBoolean("banner", true);
    p.getBoolean("attachLog", true);
    p.getString("fileFormat", "openexr");
    p.getString("pixelFormat", "rgb");
    p.getString("channelNames", "");
    p.getString("componentFormat", "float16");
}

void Film::addChild(const std::string &name, const ObjRef &child) {
    if (child->getClass() == CRFilter) filter = std::static_pointer_cast<RFilter>(child);
    else ConfigurableObject::addChild(name, child);
}

void Film::configure() {
    if (!filter) {                        // film.cpp:84-92
        filter = std::make_shared<RFilter>(Properties("gaussian"));
        filter->configure();
    }
}

// ---- sampler ---------------------------------------------------------------------------------------
Sampler::Sampler(const Properties &p) : ConfigurableObject(p) {
    const std::string type = lower(p.getPluginName());
    kind = MTSB_SAMPLER_INDEPENDENT; dimension = 4; scramble = -1;
    if (type == "independent") {                     // independent.cpp:55-60
        sampleCount = p.getInteger("sampleCount", 4);
        seed = p.getInteger("seed", 1004);
    } else if (type == "ldsampler") {                // ldsampler.cpp:75-86
        kind = MTSB_SAMPLER_LDSAMPLER;
        sampleCount = p.getInteger("sampleCount", 4);
        dimension = p.getInteger("dimension", 4);
        seed = 1004;
        if (sampleCount > 0 && (sampleCount & (sampleCount - 1))) {
            int r = 1; while (r < sampleCount) r <<= 1;
            sampleCount = r;
            warn("Sample count should be a power of two -- rounding to " + std::to_string(sampleCount));
        }
        if (dimension < 0 || dimension > 63) error("ldsampler: 'dimension' must be between 0 and 63 on the B200 path");
    } else if (type == "halton") {                   // halton.cpp:121-130
        kind = MTSB_SAMPLER_HALTON;
        sampleCount = p.getInteger("sampleCount", 4);
        scramble = p.getInteger("scramble", -1);
        seed = 1004;
    } else if (type == "sobol") {                    // sobol.cpp:85-107
        kind = MTSB_SAMPLER_SOBOL;
        sampleCount = p.getInteger("sampleCount", 4);
        scramble = p.getInteger("scramble", 0);
        seed = 1004;
        if (scramble < 0) error("sobol: 'scramble' must not be negative");
    } else {
        error("Plugin \"" + type + "\" not found -- the B200 path implements the samplers independent (counter-based streams), ldsampler, halton and sobol.");
    }
    if (sampleCount <= 0) error(type + ": 'sampleCount' must be positive");
}

// ---- perspective sensor ----------------------------------------------------------------------------
static bool hasScale(const Transform &t) {
    // Transform::hasScale (transform.h): the columns of the linear part must have unit length
    for (int i = 0; i < 3; ++i) {
        float s = t.m[i] * t.m[i] + t.m[4 + i] * t.m[4 + i] + t.m[8 + i] * t.m[8 + i];
        if (std::fabs(s - 1.0f) > 1e-3f) return true;
    }
    return false;
}

Sensor::Sensor(const Properties &p) : ConfigurableObject(p) {
    const std::string type = lower(p.getPluginName());
    if (type != "perspective")
        error("Plugin \"" + type + "\" not found -- the B200 path implements the sensor perspective.");
    m_toWorld = p.getTransform("toWorld", Transform());
    if (hasScale(m_toWorld))              // perspective.cpp:116-118
        error("Scale factors in the camera-to-world transformation are not allowed!");
    m_near = p.getFloat("nearClip", 1e-2f);
    m_far = p.getFloat("farClip", 1e4f);
    p.getFloat("focusDistance", m_far);
    if (m_near <= 0) error("The 'nearClip' parameter must be greater than zero!");
    if (m_near >= m_far) error("The 'nearClip' parameter must be smaller than 'farClip'.");
    p.markQueried("fov"); p.markQueried("fovAxis"); p.markQueried("focalLength");
    if (p.hasProperty("fov") && p.hasProperty("focalLength"))
        error("Please specify either a focal length ('focalLength') or a field of view ('fov')!");
    if (p.getFloat("shutterOpen", 0.0f) != p.getFloat("shutterClose", 0.0f))
        error("perspective: a non-zero exposure interval (motion blur) is not part of the B200 path.");
}

void Sensor::addChild(const std::string &name, const ObjRef &child) {
    if (child->getClass() == CSampler) sampler = std::static_pointer_cast<Sampler>(child);
    else if (child->getClass() == CFilm) film = std::static_pointer_cast<Film>(child);
    else ConfigurableObject::addChild(name, child);
}

void Sensor::configure() {
    // sensor.cpp:84-108
    if (!film) { film = std::make_shared<Film>(Properties("hdrfilm")); film->configure(); }
    if (!sampler) {
        Properties sp("independent");
        sp.setInteger("sampleCount", 4);
        sampler = std::make_shared<Sampler>(sp);
    }
    const float aspect = film->width / (float) film->height;
    auto setX = [&](float x) {
        if (x <= 0 || x >= 180) error("The horizontal field of view must be in the interval (0, 180)!");
        m_xfov = x;
    };
    const float d2r = (float) (M_PI / 180.0), r2d = (float) (180.0 / M_PI);
    auto setDiag = [&](float dfov) {       // sensor.cpp:305-309
        float diagonal = 2 * std::tan(0.5f * (dfov * d2r));
        float width = diagonal / std::sqrt(1.0f + 1.0f / (aspect * aspect));
        setX(r2d * (2 * std::atan(width * 0.5f)));
    };
    // sensor.cpp:243-282
    if (m_props.hasProperty("fov")) {
        float fov = m_props.getFloat("fov");
        std::string axis = lower(m_props.getString("fovAxis", "x"));
        if (axis == "smaller") axis = aspect > 1 ? "y" : "x";
        else if (axis == "larger") axis = aspect > 1 ? "x" : "y";
        if (axis == "x") setX(fov);
        else if (axis == "y") setX(r2d * (2 * std::atan(std::tan(0.5f * (fov * d2r)) * aspect)));
        else if (axis == "diagonal") setDiag(fov);
        else error("The 'fovAxis' parameter must be set to one of 'smaller', 'larger', 'diagonal', 'x', or 'y'!");
    } else {
        std::string f = m_props.getString("focalLength", "50mm");
        if (f.size() >= 2 && f.compare(f.size() - 2, 2, "mm") == 0) f = f.substr(0, f.size() - 2);
        char *end = nullptr;
        float value = (float) strtod(f.c_str(), &end);
        if (*end != '\0' || f.empty())
            error("Could not parse the focal length (must be of the form <x>mm, where <x> is a positive integer)!");
        setDiag((float) (2 * 180 / M_PI * std::atan(std::sqrt((float) (36 * 36 + 24 * 24)) / (2 * value))));
    }
}

void Sensor::fill(MtsbSensor &s) const {
    std::memset(&s, 0, sizeof(s));
    std::memcpy(s.to_world, m_toWorld.m.data(), sizeof(s.to_world));
    s.fov_x_deg = m_xfov;
    s.near_clip = m_near; s.far_clip = m_far;
    s.film_width = film->width; s.film_height = film->height;
    s.crop_x = film->cropX; s.crop_y = film->cropY; s.crop_width = film->cropW; s.crop_height = film->cropH;
    s.rfilter = film->filter->kind;
    s.rfilter_radius = film->filter->radius;
    s.rfilter_stddev = film->filter->stddev;
}

// ---- path integrator (integrator.cpp:252-287) ------------------------------------------------------
PathIntegrator::PathIntegrator(const Properties &p) : ConfigurableObject(p) {
    const std::string type = lower(p.getPluginName());
    if (type != "path")
        error("Plugin \"" + type + "\" not found -- this library implements the integrator path (the unidirectional MI path tracer).");
    rrDepth = p.getInteger("rrDepth", 5);
    maxDepth = p.getInteger("maxDepth", -1);
    strictNormals = p.getBoolean("strictNormals", false);
    hideEmitters = p.getBoolean("hideEmitters", false);
    if (rrDepth <= 0) error("'rrDepth' must be set to a value greater than zero!");
    if (maxDepth <= 0 && maxDepth != -1) error("'maxDepth' must be set to -1 (infinite) or a value greater than zero!");
    if (p.getInteger("removeOutlier", -1) != -1)          // path.cpp:114 (fork-specific)
        error("path: 'removeOutlier' is not part of the B200 path.");
}

}  // namespace mtsbhost
