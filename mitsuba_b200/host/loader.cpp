// loader.cpp -- scene XML -> object graph, the role of SceneHandler in the reference
// (src/librender/scenehandler.cpp:197-751).  The reference drives this from SAX start/end events;
// here the small DOM of xml.h is walked in document order, which gives the same ordering of
// <default> definitions, $parameter substitution, id registration and addChild() calls.
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>

#include "loader.h"
#include "xml.h"

namespace mtsbhost {

namespace {

struct Context {
    Context *parent = nullptr;
    std::string tag;
    std::map<std::string, std::string> attr;
    Properties props;
    std::vector<std::pair<std::string, ObjRef>> children;
};

struct Handler {
    ParameterMap &params;
    std::map<std::string, ObjRef> &named;
    std::string file, baseDir;
    bool included;
    Transform transform;
    std::shared_ptr<Scene> scene;

    [[noreturn]] void fail(const XmlNode &n, const std::string &msg) const {
        // XMLLog(EError, ...): "Error in file \"%s\" (near line %i): ..."
        throw std::runtime_error("Error in file \"" + file + "\" (near line " + std::to_string(n.line) + "): " + msg);
    }

    // scenehandler.cpp:182-195
    float parseFloat(const XmlNode &n, const std::string &name, const std::string &str, const float *dflt = nullptr) const {
        if (str.empty()) {
            if (dflt) return *dflt;
            fail(n, "Missing floating point value (in <" + name + ">)");
        }
        char *end = nullptr;
        float r = (float) strtod(str.c_str(), &end);
        if (*end != '\0') fail(n, "Invalid floating point value specified (in <" + name + ">)");
        return r;
    }
    float attrFloat(const XmlNode &n, Context &c, const char *key) const { return parseFloat(n, c.tag, c.attr[key]); }
    float attrFloat(const XmlNode &n, Context &c, const char *key, float d) const { return parseFloat(n, c.tag, c.attr[key], &d); }

    static std::vector<std::string> tokenize(const std::string &s, const char *delims) {
        std::vector<std::string> out;
        size_t pos = 0;
        while (pos < s.size()) {
            size_t b = s.find_first_not_of(delims, pos);
            if (b == std::string::npos) break;
            size_t e = s.find_first_of(delims, b);
            if (e == std::string::npos) e = s.size();
            out.push_back(s.substr(b, e - b));
            pos = e;
        }
        return out;
    }

    Vec3 parseColor(const XmlNode &n, Context &c, const char *what) const {
        // scenehandler.cpp:476-497 / 508-526: "#rrggbb", one value, or three values
        std::vector<std::string> tok = tokenize(c.attr["value"], ", ");
        Vec3 v;
        if (tok.size() == 1 && tok[0].size() == 7 && tok[0][0] == '#') {
            char *end = nullptr;
            long enc = strtol(tok[0].c_str() + 1, &end, 16);
            if (*end != '\0') fail(n, std::string("Invalid ") + what + " value specified (in <" + c.attr["name"] + ">)");
            v = Vec3{((enc & 0xFF0000) >> 16) / 255.0f, ((enc & 0x00FF00) >> 8) / 255.0f, (enc & 0x0000FF) / 255.0f};
        } else if (tok.size() == 1) {
            v.x = v.y = v.z = parseFloat(n, c.tag, tok[0]);
        } else if (tok.size() == 3) {
            v = Vec3{parseFloat(n, c.tag, tok[0]), parseFloat(n, c.tag, tok[1]), parseFloat(n, c.tag, tok[2])};
        } else fail(n, std::string("Invalid ") + what + " value specified");
        return v;
    }

    static float fromSRGB(float v) {       // spectrum.cpp:400-405
        if (v <= 0.04045f) return v * (float) (1.0 / 12.92);
        return std::pow((v + 0.055f) * (float) (1.0 / 1.055), 2.4f);
    }

    static bool classOfTag(const std::string &t, EClass &c) {
        static const std::pair<const char *, EClass> m[] = {{"scene", CScene}, {"shape", CShape}, {"sampler", CSampler}, {"film", CFilm},
            {"integrator", CIntegrator}, {"texture", CTexture}, {"sensor", CSensor}, {"emitter", CEmitter}, {"bsdf", CBsdf}, {"rfilter", CRFilter}};
        for (auto &e : m) if (t == e.first) { c = e.second; return true; }
        return false;
    }

    void element(const XmlNode &n, Context *parent);
    void finish(const XmlNode &n, Context &c);
};

void Handler::element(const XmlNode &n, Context *parent) {
    static const char *known[] = {"scene", "shape", "sampler", "film", "integrator", "texture", "sensor", "emitter", "subsurface", "medium",
        "volume", "phase", "bsdf", "rfilter", "null", "ref", "integer", "float", "boolean", "string", "translate", "rotate", "lookat", "lookAt",
        "scale", "matrix", "point", "vector", "rgb", "srgb", "blackbody", "spectrum", "transform", "animation", "include", "alias", "default"};
    bool ok = false;
    for (const char *k : known) ok |= n.tag == k;
    if (!ok) fail(n, "Unhandled tag \"" + n.tag + "\" encountered!");
    if (parent == nullptr && n.tag != "scene") fail(n, "The root element must be <scene>");

    Context c;
    c.parent = parent;
    c.tag = n.tag;
    // startElement: $parameter substitution, longest names first (reverse map order), scenehandler.cpp:207-224
    for (auto &kv : n.attr) {
        std::string v = kv.second;
        if (!v.empty() && v.find('$') != std::string::npos) {
            for (auto it = params.rbegin(); it != params.rend(); ++it) {
                std::string key = "$" + it->first;
                size_t pos = 0;
                while ((pos = v.find(key, pos)) != std::string::npos) { v.replace(pos, key.size(), it->second); ++pos; }
            }
            if (v.find('$') != std::string::npos && v.find('[') == std::string::npos)
                fail(n, "The scene referenced an undefined parameter: \"" + v + "\"");
        }
        c.attr[kv.first] = v;
    }
    if (n.tag == "scene") {
        std::string ver = c.attr["version"];
        if (ver.empty())
            fail(n, "The requested scene cannot be loaded, since it is missing version information! "
                    "The version should be specified within the 'scene' tag, e.g. <scene version=\"0.5.0\">");
        // Version::isCompatible: same major.minor as MTS_VERSION 0.5.0 (0.6 kept the 0.5 format)
        std::vector<std::string> t = tokenize(ver, ".");
        if (t.size() != 3) fail(n, "Unable to parse version number \"" + ver + "\"");
        int major = atoi(t[0].c_str()), minor = atoi(t[1].c_str());
        if (major == 0 && minor < 5)
            fail(n, "The requested scene is from an older version of Mitsuba (file version: " + ver + ", current version: 0.5.0), hence the loading process was stopped.");
        if (major > 0 || minor > 6)
            fail(n, "The requested scene is from an incompatible future version of Mitsuba (file version: " + ver + ", current version: 0.5.0). Giving up.");
    } else if (n.tag == "transform") {
        transform = Transform();
    } else if (n.tag == "animation") {
        fail(n, "<animation>: animated transformations are not part of the B200 path.");
    }
    for (auto &ch : n.children) element(*ch, &c);
    finish(n, c);
}

void Handler::finish(const XmlNode &n, Context &c) {
    const std::string &name = n.tag;
    c.props.setPluginName(lower(c.attr["type"]));
    if (c.attr.count("id")) c.props.setID(c.attr["id"]);
    Context *P = c.parent;
    ObjRef object;
    bool isNull = false;
    auto needParent = [&]() { if (!P) fail(n, "<" + name + "> needs a parent element"); };
    try {
        if (name == "null") {
            isNull = true;
        } else if (name == "ref") {
            std::string id = c.attr["id"];
            if (!named.count(id)) fail(n, "Referenced object '" + id + "' not found!");
            object = named[id];
        } else if (name == "integer") {
            needParent();
            char *end = nullptr;
            long long v = strtoll(c.attr["value"].c_str(), &end, 10);
            if (*end != '\0' || c.attr["value"].empty()) fail(n, "Invalid integer value specified (in <" + c.attr["name"] + ">)");
            P->props.setInteger(c.attr["name"], v);
        } else if (name == "float") {
            needParent();
            P->props.setFloat(c.attr["name"], attrFloat(n, c, "value"));
        } else if (name == "boolean") {
            needParent();
            const std::string &s = c.attr["value"];
            if (s != "true" && s != "false") fail(n, "Unsupported boolean constant '" + s + "' -- must be 'true' or 'false'!");
            P->props.setBoolean(c.attr["name"], s == "true");
        } else if (name == "string") {
            needParent();
            P->props.setString(c.attr["name"], c.attr["value"]);
        } else if (name == "translate") {
            transform = Transform::translate(Vec3{attrFloat(n, c, "x", 0), attrFloat(n, c, "y", 0), attrFloat(n, c, "z", 0)}) * transform;
        } else if (name == "rotate") {
            Vec3 axis{attrFloat(n, c, "x", 0), attrFloat(n, c, "y", 0), attrFloat(n, c, "z", 0)};
            transform = Transform::rotate(axis, attrFloat(n, c, "angle")) * transform;
        } else if (name == "lookat" || name == "lookAt") {
            auto vec = [&](const char *key, bool optional, Vec3 &out) -> bool {
                std::vector<std::string> t = tokenize(c.attr[key], ", ");
                if (t.size() == 0 && optional) return false;
                if (t.size() != 3) fail(n, std::string("<lookat>: invalid '") + key + "' argument");
                out = Vec3{parseFloat(n, name, t[0]), parseFloat(n, name, t[1]), parseFloat(n, name, t[2])};
                return true;
            };
            Vec3 o, t, u{0, 0, 0};
            vec("origin", false, o);
            vec("target", false, t);
            vec("up", true, u);
            if (u.x * u.x + u.y * u.y + u.z * u.z == 0) {
                // scenehandler.cpp:392-396: no 'up' -> coordinateSystem(normalize(t-o), u, unused), i.e. u = c x d with c the
                // axis-aligned orthogonal of d (core/util.cpp coordinateSystem)
                Vec3 d{t.x - o.x, t.y - o.y, t.z - o.z}, cc;
                float l = std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
                d = Vec3{d.x / l, d.y / l, d.z / l};
                if (std::fabs(d.x) > std::fabs(d.y)) { float il = 1.0f / std::sqrt(d.x * d.x + d.z * d.z); cc = Vec3{d.z * il, 0, -d.x * il}; }
                else { float il = 1.0f / std::sqrt(d.y * d.y + d.z * d.z); cc = Vec3{0, d.z * il, -d.y * il}; }
                u = Vec3{cc.y * d.z - cc.z * d.y, cc.z * d.x - cc.x * d.z, cc.x * d.y - cc.y * d.x};
            }
            transform = Transform::lookAt(o, t, u) * transform;
        } else if (name == "scale") {
            float x, y, z;
            if (c.attr.count("value")) x = y = z = attrFloat(n, c, "value");
            else { x = attrFloat(n, c, "x", 1); y = attrFloat(n, c, "y", 1); z = attrFloat(n, c, "z", 1); }
            transform = Transform::scale(Vec3{x, y, z}) * transform;
        } else if (name == "matrix") {
            std::vector<std::string> t = tokenize(c.attr["value"], ", ");
            if (t.size() != 16) fail(n, "Invalid matrix specified");
            Transform m;
            for (int i = 0; i < 16; ++i) m.m[i] = parseFloat(n, name, t[i]);
            transform = m * transform;
        } else if (name == "point") {
            needParent();
            P->props.setPoint(c.attr["name"], Vec3{attrFloat(n, c, "x"), attrFloat(n, c, "y"), attrFloat(n, c, "z")});
        } else if (name == "vector") {
            needParent();
            P->props.setVector(c.attr["name"], Vec3{attrFloat(n, c, "x"), attrFloat(n, c, "y"), attrFloat(n, c, "z")});
        } else if (name == "rgb") {
            needParent();
            if (c.attr.count("intent")) {
                std::string i = lower(c.attr["intent"]);
                if (i != "reflectance" && i != "illuminant") fail(n, "Invalid intent \"" + i + "\", must be \"reflectance\" or \"illuminant\"");
            }
            P->props.setSpectrum(c.attr["name"], parseColor(n, c, "rgb"));       // RGB build: fromLinearRGB stores the triple
        } else if (name == "srgb") {
            needParent();
            Vec3 v = parseColor(n, c, "sRGB");
            P->props.setSpectrum(c.attr["name"], Vec3{fromSRGB(v.x), fromSRGB(v.y), fromSRGB(v.z)});
        } else if (name == "spectrum") {
            needParent();
            bool hasValue = c.attr.count("value"), hasFile = c.attr.count("filename");
            if (hasValue == hasFile) fail(n, "<spectrum>: please provide one of 'value' or 'filename'");
            if (hasFile) fail(n, "<spectrum filename=...>: measured spectra are not part of the B200 path (RGB build) -- give an <rgb> value.");
            std::vector<std::string> t = tokenize(c.attr["value"], ", ");
            if (t.size() == 1 && t[0].find(':') == std::string::npos) {
                float v = parseFloat(n, name, t[0]);                               // Spectrum(v) resp. D65*v == v in an RGB build
                P->props.setSpectrum(c.attr["name"], Vec3{v, v, v});
            } else if (t.size() == 3 && t[0].find(':') == std::string::npos) {
                // SPECTRUM_SAMPLES == 3: three plain values are taken as the coefficients (scenehandler.cpp:612-622)
                P->props.setSpectrum(c.attr["name"], Vec3{parseFloat(n, name, t[0]), parseFloat(n, name, t[1]), parseFloat(n, name, t[2])});
            } else fail(n, "<spectrum>: wavelength:value lists are not part of the B200 path (RGB build) -- give an <rgb> value.");
        } else if (name == "blackbody") {
            fail(n, "<blackbody> is not part of the B200 path (RGB build) -- give an <rgb> value.");
        } else if (name == "transform") {
            needParent();
            P->props.setTransform(c.attr["name"], transform);
        } else if (name == "alias") {
            std::string id = c.attr["id"], as = c.attr["as"];
            if (!named.count(id)) fail(n, "Referenced object '" + id + "' not found!");
            if (named.count(as)) fail(n, "Duplicate ID '" + as + "' used in scene description!");
            named[as] = named[id];
        } else if (name == "include") {
            std::string fn = c.attr["filename"];
            if (!fn.empty() && fn[0] != '/' && !baseDir.empty()) fn = baseDir + "/" + fn;
            object = loadSceneFile(fn, params, &named, true);
        } else if (name == "default") {
            if (!params.count(c.attr["name"])) params[c.attr["name"]] = c.attr["value"];
        } else if (name == "subsurface" || name == "medium" || name == "volume" || name == "phase") {
            fail(n, "<" + name + ">: participating media and subsurface scattering are not part of the B200 path.");
        } else {
            EClass cls;
            if (!classOfTag(name, cls)) fail(n, "Internal error: could not instantiate an object corresponding to the tag '" + name + "'");
            try {
                object = createObject(cls, c.props, baseDir);
            } catch (const std::exception &e) {
                fail(n, std::string("Error while creating object: ") + e.what());
            }
            if (cls == CScene) scene = std::static_pointer_cast<Scene>(object);
        }
    } catch (const std::runtime_error &e) {
        if (std::strncmp(e.what(), "Error in file", 13) == 0 || std::strncmp(e.what(), "Error while parsing", 19) == 0) throw;
        fail(n, e.what());
    }

    if (object || isNull) {
        std::string id = c.attr.count("id") ? c.attr["id"] : "", nodeName = c.attr.count("name") ? c.attr["name"] : "";
        if (object) {
            try {
                if (P) P->children.push_back({nodeName, object});
                for (auto &ch : c.children) if (ch.second) object->addChild(ch.first, ch.second);
                c.children.clear();
                // an included file's Scene is configured by the including file's scene
                if (name != "include" && name != "ref" && !(included && object->getClass() == CScene)) object->configure();
            } catch (const std::runtime_error &e) {
                if (std::strncmp(e.what(), "Error in file", 13) == 0) throw;
                fail(n, e.what());
            }
        }
        if (!id.empty() && name != "ref") {
            if (named.count(id)) fail(n, "Duplicate ID '" + id + "' used in scene description!");
            named[id] = object;
        }
    }
    for (const std::string &u : c.props.getUnqueried())
        warn("Unqueried attribute \"" + u + "\" in element \"" + name + "\"");
}

}  // namespace

std::shared_ptr<Scene> loadSceneString(const std::string &xml, const std::string &fileName, ParameterMap &params,
                                       std::map<std::string, ObjRef> *namedObjects, bool included) {
    std::map<std::string, ObjRef> localNamed;
    std::string dir;
    size_t slash = fileName.find_last_of('/');
    if (slash != std::string::npos) dir = fileName.substr(0, slash);
    Handler h{params, namedObjects ? *namedObjects : localNamed, fileName, dir, included, Transform(), nullptr};
    std::unique_ptr<XmlNode> root = XmlParser(xml, fileName).parse();
    h.element(*root, nullptr);
    if (!h.scene) throw std::runtime_error("Error in file \"" + fileName + "\": no <scene> element");
    return h.scene;
}

std::shared_ptr<Scene> loadSceneFile(const std::string &path, ParameterMap &params, std::map<std::string, ObjRef> *namedObjects,
                                     bool included) {
    std::ifstream is(path, std::ios::binary);
    if (!is.good()) throw std::runtime_error("Unable to open the scene file \"" + path + "\"");
    std::stringstream ss;
    ss << is.rdbuf();
    return loadSceneString(ss.str(), path, params, namedObjects, included);
}

}  // namespace mtsbhost
