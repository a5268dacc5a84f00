// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// C wrapper around the reference's MIP map, COMPILED FROM /root/reference (nothing is copied):
//   TMIPMap<TSpectrum<Float,3>, TSpectrum<half,3>>     include/mitsuba/render/mipmap.h (the type behind `bitmap` textures,
//                                                       src/textures/bitmap.cpp:172-177)
//   Bitmap::resample / Resampler / LanczosSincFilter   src/libcore/bitmap.cpp, include/mitsuba/core/rfilter.h, src/rfilters/lanczos.cpp
// The input is handed over as a linear float32 bitmap (gamma 1), so Bitmap::convert returns the bitmap itself and no
// FormatConverter is needed (fmtconv.cpp needs boost::mpl, properties.cpp boost::variant: neither exists in this image).
// The few Properties / ConfigurableObject / PluginManager symbols the objects link against are stand-ins defined below.
#include <mitsuba/mitsuba.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/rfilter.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/core/plugin.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/render/mipmap.h>
#include <map>
#include <stdexcept>

MTS_NAMESPACE_BEGIN

// ---- stand-ins -------------------------------------------------------------------------------------------------
struct PropertyElement { int value; };
Properties::Properties() : m_elements(new std::map<std::string, PropertyElement>()), m_id("unnamed") {}
Properties::Properties(const std::string &pluginName) : m_elements(new std::map<std::string, PropertyElement>()), m_pluginName(pluginName), m_id("unnamed") {}
Properties::Properties(const Properties &p) : m_elements(new std::map<std::string, PropertyElement>(*p.m_elements)), m_pluginName(p.m_pluginName), m_id(p.m_id) {}
Properties::~Properties() { delete m_elements; }
void Properties::operator=(const Properties &p) { *m_elements = *p.m_elements; m_pluginName = p.m_pluginName; m_id = p.m_id; }
void Properties::setInteger(const std::string &name, const int &value, bool) { (*m_elements)[name].value = value; }
int Properties::getInteger(const std::string &name, const int &defVal) const {
    std::map<std::string, PropertyElement>::const_iterator it = m_elements->find(name);
    return it == m_elements->end() ? defVal : it->second.value;
}
void Properties::putPropertyNames(std::vector<std::string> &results) const { for (auto &kv : *m_elements) results.push_back(kv.first); }
std::string Properties::getAsString(const std::string &name) const { return std::to_string(getInteger(name, 0)); }
bool Properties::operator==(const Properties &p) const { return m_pluginName == p.m_pluginName && m_id == p.m_id; }
void Properties::merge(const Properties &) {}

ConfigurableObject::ConfigurableObject(Stream *stream, InstanceManager *manager) : SerializableObject(stream, manager) {}
void ConfigurableObject::setParent(ConfigurableObject *) {}
void ConfigurableObject::addChild(const std::string &, ConfigurableObject *) { throw std::runtime_error("[reference] addChild is not part of this build"); }
void ConfigurableObject::configure() {}
void ConfigurableObject::serialize(Stream *, InstanceManager *) const {}
MTS_IMPLEMENT_CLASS(ConfigurableObject, true, SerializableObject)

ref<PluginManager> PluginManager::m_instance;
ConfigurableObject *PluginManager::createObject(const Class *, const Properties &) { throw std::runtime_error("[reference] PluginManager is not part of this build"); }
const FormatConverter *FormatConverter::getInstance(Conversion) { throw std::runtime_error("[reference] FormatConverter is not part of this build"); }
void FormatConverter::staticInitialization() {}
void FormatConverter::staticShutdown() {}
int Thread::getID() { return 0; }
void Logger::setLogLevel(ELogLevel) {}
void Logger::logProgress(Float, const std::string &, const std::string &, const std::string &, const void *) {}
namespace stats {       // defined in src/librender/texture.cpp:25-28, which is not part of this build
StatsCounter mipStorage("Texture system", "Cumulative MIP map memory allocations", EByteCount);
StatsCounter clampedAnisotropy("Texture system", "Lookups with clamped anisotropy", EPercentage);
StatsCounter avgEWASamples("Texture system", "Average EWA samples / lookup", EAverage);
StatsCounter filteredLookups("Texture system", "Filtered texture lookups", EPercentage);
}

MTS_NAMESPACE_END

#include "src/rfilters/lanczos.cpp"

using namespace mitsuba;
typedef TSpectrum<Float, 3> RefColor3;
typedef TSpectrum<half, 3> RefColor3h;
typedef TMIPMap<RefColor3, RefColor3h> RefMIPMap3;

struct RefMip { ref<RefMIPMap3> mip; };

extern "C" {

// rgb: w*h*3 linear floats; wrap / filter numbering = MTSB_WRAP_* / MTSB_TEXFILTER_* (same order as the reference's enums)
void *ref_mipmap_create(int w, int h, const float *rgb, int wrapU, int wrapV, int filterType, float maxAnisotropy) {
    ref<Bitmap> bm = new Bitmap(Bitmap::ERGB, Bitmap::EFloat32, Vector2i(w, h));
    memcpy(bm->getFloat32Data(), rgb, sizeof(float) * 3 * (size_t) w * (size_t) h);
    Properties p("lanczos");
    p.setInteger("lobes", 2);
    ref<ReconstructionFilter> rf = new LanczosSincFilter(p);
    rf->configure();
    if (filterType != EEWA) maxAnisotropy = 1.0f;                         // bitmap.cpp:232-235
    RefMip *r = new RefMip();
    r->mip = new RefMIPMap3(bm, Bitmap::ERGB, Bitmap::EFloat, rf, (ReconstructionFilter::EBoundaryCondition) wrapU,
                            (ReconstructionFilter::EBoundaryCondition) wrapV, (EMIPFilterType) filterType, maxAnisotropy);
    return r;
}
void ref_mipmap_destroy(void *h) { delete static_cast<RefMip *>(h); }
int ref_mipmap_levels(void *h) { return static_cast<RefMip *>(h)->mip->getLevels(); }
void ref_mipmap_level_size(void *h, int level, int *wh) {
    const RefMIPMap3::Array2DType &a = static_cast<RefMip *>(h)->mip->getArray(level);
    wh[0] = a.getWidth(); wh[1] = a.getHeight();
}
void ref_mipmap_level(void *h, int level, float *out) {
    const RefMIPMap3 *m = static_cast<RefMip *>(h)->mip.get();
    const RefMIPMap3::Array2DType &a = m->getArray(level);
    for (int y = 0; y < a.getHeight(); ++y)
        for (int x = 0; x < a.getWidth(); ++x) {
            RefColor3 v = m->evalTexel(level, x, y);
            float *o = out + 3 * ((size_t) y * a.getWidth() + x); o[0] = v[0]; o[1] = v[1]; o[2] = v[2];
        }
}
// mode 0: evalBox(level 0), 1: evalBilinear(level 0), 2: eval(uv, d0, d1) with q = u v d0x d0y d1x d1y
void ref_mipmap_eval(void *h, int mode, size_t n, const float *q, float *rgb) {
    const RefMIPMap3 *m = static_cast<RefMip *>(h)->mip.get();
    for (size_t i = 0; i < n; ++i) {
        const float *p = q + 6 * i;
        RefColor3 v;
        if (mode == 0) v = m->evalBox(0, Point2(p[0], p[1]));
        else if (mode == 1) v = m->evalBilinear(0, Point2(p[0], p[1]));
        else v = m->eval(Point2(p[0], p[1]), Vector2(p[2], p[3]), Vector2(p[4], p[5]));
        rgb[3 * i] = v[0]; rgb[3 * i + 1] = v[1]; rgb[3 * i + 2] = v[2];
    }
}
void ref_mipmap_stats(void *h, float *out9) {
    const RefMIPMap3 *m = static_cast<RefMip *>(h)->mip.get();
    for (int c = 0; c < 3; ++c) { out9[c] = m->getMinimum()[c]; out9[3 + c] = m->getMaximum()[c]; out9[6 + c] = m->getAverage()[c]; }
}

}  // extern "C"
