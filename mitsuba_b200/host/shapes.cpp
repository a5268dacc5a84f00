// shapes.cpp -- shape plugins of the path, all realised as triangle meshes (the GPU path intersects
// triangles only; the reference's analytic `rectangle` has the createTriMesh() form used here,
// src/shapes/rectangle.cpp:172-205).
#include <cmath>
#include <cstring>

#include "plugins.h"

namespace mtsbhost {

static inline Vec3 sub(const Vec3 &a, const Vec3 &b) { return Vec3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline Vec3 crs(const Vec3 &a, const Vec3 &b) { return Vec3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline float dot3(const Vec3 &a, const Vec3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float len(const Vec3 &a) { return std::sqrt(dot3(a, a)); }
static inline Vec3 nrm(const Vec3 &a) { float l = len(a); return Vec3{a.x / l, a.y / l, a.z / l}; }
static inline Vec3 getv(const std::vector<float> &v, size_t i) { return Vec3{v[3 * i], v[3 * i + 1], v[3 * i + 2]}; }

// unitAngle (core/util.h): numerically robust angle between unit vectors
static float unitAngle(const Vec3 &u, const Vec3 &v) {
    if (dot3(u, v) < 0) {
        Vec3 s{v.x + u.x, v.y + u.y, v.z + u.z};
        return (float) M_PI - 2 * std::asin(std::min(1.0f, 0.5f * len(s)));
    }
    return 2 * std::asin(std::min(1.0f, 0.5f * len(sub(v, u))));
}

// trimesh.cpp:608-681
void computeNormals(Mesh &m, bool faceNormals, bool flipNormals) {
    const size_t nv = m.vertexCount(), nt = m.triangleCount();
    if (faceNormals) {
        m.normals.clear();
        if (flipNormals)
            for (size_t i = 0; i < nt; ++i) std::swap(m.indices[3 * i], m.indices[3 * i + 1]);   // winding order
        return;
    }
    if (!m.normals.empty()) {
        if (flipNormals) for (auto &x : m.normals) x = -x;
        return;
    }
    // angle-weighted vertex normals (Thuermer & Wuethrich, JGT 1998)
    m.normals.assign(3 * nv, 0.0f);
    for (size_t t = 0; t < nt; ++t) {
        Vec3 n{0, 0, 0};
        for (int i = 0; i < 3; ++i) {
            uint32_t i0 = m.indices[3 * t + i], i1 = m.indices[3 * t + (i + 1) % 3], i2 = m.indices[3 * t + (i + 2) % 3];
            Vec3 a = sub(getv(m.positions, i1), getv(m.positions, i0)), b = sub(getv(m.positions, i2), getv(m.positions, i0));
            if (i == 0) {
                n = crs(a, b);
                float l = len(n);
                if (l == 0) break;
                n = Vec3{n.x / l, n.y / l, n.z / l};
            }
            float ang = unitAngle(nrm(a), nrm(b));
            m.normals[3 * i0] += n.x * ang; m.normals[3 * i0 + 1] += n.y * ang; m.normals[3 * i0 + 2] += n.z * ang;
        }
    }
    int invalid = 0;
    for (size_t i = 0; i < nv; ++i) {
        Vec3 n = getv(m.normals, i);
        float l = len(n);
        if (flipNormals) l = -l;
        if (l != 0) n = Vec3{n.x / l, n.y / l, n.z / l};
        else { ++invalid; n = Vec3{1, 0, 0}; }
        m.normals[3 * i] = n.x; m.normals[3 * i + 1] = n.y; m.normals[3 * i + 2] = n.z;
    }
    if (invalid) warn("\"" + m.name + "\": Unable to generate " + std::to_string(invalid) + " vertex normals");
}

static void push3(std::vector<float> &v, const Vec3 &a) { v.push_back(a.x); v.push_back(a.y); v.push_back(a.z); }

// rectangle.cpp:80-85, 172-205: (-1,-1,0)..(1,1,0) in object space, triangles (0,1,2),(2,3,0)
static Mesh makeRectangle(const Properties &p) {
    Transform t = p.getTransform("toWorld", Transform());
    const bool flip = p.getBoolean("flipNormals", false);
    if (flip) t = t * Transform::scale(Vec3{1, 1, -1});
    Vec3 du = t.vector(Vec3{2, 0, 0}), dv = t.vector(Vec3{0, 2, 0});
    if (std::fabs(dot3(nrm(du), nrm(dv))) > 1e-4f) error("Error: 'toWorld' transformation contains shear!");
    Vec3 n = nrm(t.normal(Vec3{0, 0, 1}));
    Mesh m;
    m.name = "rectangle";
    const float c[4][2] = {{-1, -1}, {1, -1}, {1, 1}, {-1, 1}}, uv[4][2] = {{0, 0}, {1, 0}, {1, 1}, {0, 1}};
    for (int i = 0; i < 4; ++i) {
        push3(m.positions, t.point(Vec3{c[i][0], c[i][1], 0}));
        push3(m.normals, n);
        m.texcoords.push_back(uv[i][0]); m.texcoords.push_back(uv[i][1]);
    }
    // the analytic rectangle's geometric normal is the (possibly flipped) frame normal: keep the winding consistent with it
    if (flip) m.indices = {0, 2, 1, 2, 0, 3};
    else m.indices = {0, 1, 2, 2, 3, 0};
    return m;
}

// cube.cpp:24-30, 75-104: 6 faces x 4 vertices with per-face normals and [0,1]^2 uv per face; the face
// order (-y, +y, +x, +z, -x, -z) and the two triangles (0,1,2),(3,0,2) per face are part of the prim-id contract.
static Mesh makeCube(const Properties &p) {
    Transform t = p.getTransform("toWorld", Transform());
    static const signed char face[6][4][3] = {
        {{1, -1, -1}, {1, -1, 1}, {-1, -1, 1}, {-1, -1, -1}},   {{1, 1, -1}, {-1, 1, -1}, {-1, 1, 1}, {1, 1, 1}},
        {{1, -1, -1}, {1, 1, -1}, {1, 1, 1}, {1, -1, 1}},       {{1, -1, 1}, {1, 1, 1}, {-1, 1, 1}, {-1, -1, 1}},
        {{-1, -1, 1}, {-1, 1, 1}, {-1, 1, -1}, {-1, -1, -1}},   {{1, 1, -1}, {1, -1, -1}, {-1, -1, -1}, {-1, 1, -1}}};
    static const signed char fn[6][3] = {{0, -1, 0}, {0, 1, 0}, {1, 0, 0}, {0, 0, 1}, {-1, 0, 0}, {0, 0, -1}};
    static const float uv[4][2] = {{0, 1}, {1, 1}, {1, 0}, {0, 0}};
    Mesh m;
    m.name = "cube";
    for (int f = 0; f < 6; ++f) {
        Vec3 n = nrm(t.normal(Vec3{(float) fn[f][0], (float) fn[f][1], (float) fn[f][2]}));
        for (int v = 0; v < 4; ++v) {
            push3(m.positions, t.point(Vec3{(float) face[f][v][0], (float) face[f][v][1], (float) face[f][v][2]}));
            push3(m.normals, n);
            m.texcoords.push_back(uv[v][0]); m.texcoords.push_back(uv[v][1]);
        }
        uint32_t b = 4 * f;
        for (uint32_t k : {b + 0, b + 1, b + 2, b + 3, b + 0, b + 2}) m.indices.push_back(k);
    }
    return m;
}

static std::string resolvePath(const std::string &baseDir, const std::string &f) {
    if (f.empty()) error("Property \"filename\" is empty");
    if (f[0] == '/' || baseDir.empty()) return f;
    return baseDir + "/" + f;      // FileResolver: the scene file's directory is searched first (scenehandler / mitsuba.cpp)
}

Shape::Shape(const Properties &p, const std::string &baseDir) : ConfigurableObject(p) {
    const std::string type = lower(p.getPluginName());
    if (type == "obj") {
        // obj.cpp:189-222
        m_faceNormals = p.getBoolean("faceNormals", false);
        m_flipNormals = p.getBoolean("flipNormals", false);
        bool collapse = p.getBoolean("collapse", false);
        bool flipTex = p.getBoolean("flipTexCoords", true);
        int shapeIndex = p.getInteger("shapeIndex", -1);
        Transform t = p.getTransform("toWorld", Transform());
        if (p.getBoolean("loadMaterials", true)) { /* MTL import is outside the path: the shape's own <bsdf> is used */ }
        if (p.hasProperty("maxSmoothAngle")) {
            if (m_faceNormals) error("The properties 'maxSmoothAngle' and 'faceNormals' can't be specified at the same time!");
            error("obj: 'maxSmoothAngle' (topology rebuild) is not part of the B200 path.");
        }
        loadOBJ(resolvePath(baseDir, p.getString("filename")), t, flipTex, collapse, shapeIndex, m_meshes);
    } else if (type == "ply") {
        // ply.cpp:218-240
        m_faceNormals = p.getBoolean("faceNormals", false);
        m_flipNormals = p.getBoolean("flipNormals", false);
        p.getBoolean("srgb", true);                       // vertex colours are not used by the path
        if (p.hasProperty("maxSmoothAngle")) error("ply: 'maxSmoothAngle' (topology rebuild) is not part of the B200 path.");
        loadPLY(resolvePath(baseDir, p.getString("filename")), p.getTransform("toWorld", Transform()), m_meshes);
    } else if (type == "serialized") {
        // serialized.cpp:148-210
        const int shapeIndex = p.getInteger("shapeIndex", 0);
        m_faceNormals = p.getBoolean("faceNormals", false);
        m_flipNormals = p.getBoolean("flipNormals", false);
        if (p.hasProperty("maxSmoothAngle")) {
            if (m_faceNormals) error("The properties 'maxSmoothAngle' and 'faceNormals' can't be specified at the same time!");
            error("serialized: 'maxSmoothAngle' (topology rebuild) is not part of the B200 path.");
        }
        loadSerialized(resolvePath(baseDir, p.getString("filename")), p.getTransform("toWorld", Transform()), shapeIndex, m_meshes);
    } else if (type == "rectangle") {
        m_meshes.push_back(makeRectangle(p));
    } else if (type == "cube") {
        m_faceNormals = p.getBoolean("faceNormals", false);     // TriMesh(props), trimesh.cpp:59-60
        m_flipNormals = p.getBoolean("flipNormals", false);
        m_meshes.push_back(makeCube(p));
    } else if (type == "sphere" || type == "disk" || type == "cylinder" || type == "hair" || type == "heightfield") {
        error("Shape plugin \"" + type + "\" is an analytic primitive; the B200 path intersects triangle meshes (obj, ply, serialized, rectangle, cube).");
    } else if (type == "instance" || type == "shapegroup") {
        error("Shape plugin \"" + type + "\": instancing is flattened by the exporter; it is not part of the B200 path.");
    } else {
        error("Plugin \"" + type + "\" not found -- the B200 path implements the shapes obj, ply, serialized, rectangle and cube.");
    }
}

void Shape::addChild(const std::string &name, const ObjRef &child) {
    // shape.cpp:101-133 (+ obj.cpp:721-760 for compound shapes)
    if (child->getClass() == CBsdf) {
        m_bsdf = std::static_pointer_cast<Bsdf>(child);
    } else if (child->getClass() == CEmitter) {
        auto e = std::static_pointer_cast<Emitter>(child);
        if (m_emitter) error("Tried to attach multiple emitters to a shape!");
        if (!e->isArea()) error("Only area emitters can be attached to a shape!");
        if (m_meshes.size() > 1) error("Cannot attach an emitter to an OBJ file containing multiple objects!");
        if (e->m_parents++ > 0) error("An area light cannot be parent of multiple shapes");
        m_emitter = e;
    } else {
        ConfigurableObject::addChild(name, child);
    }
}

void Shape::configure() {
    // shape.cpp:49-62: a shape without BSDF is an absorbing diffuse surface when it is an emitter, else 0.5 diffuse
    if (!m_bsdf) {
        Properties bp("diffuse");
        float r = m_emitter ? 0.0f : 0.5f;
        bp.setSpectrum("reflectance", Vec3{r, r, r});
        m_bsdf = std::make_shared<Bsdf>(bp);
        m_bsdf->configure();
    }
    for (auto &m : m_meshes) {
        if (m.triangleCount() == 0) error("Encountered an empty triangle mesh!");      // trimesh.cpp:389-392
        computeNormals(m, m_faceNormals, m_flipNormals);
    }
    m_flipNormals = false;
}

}  // namespace mtsbhost
