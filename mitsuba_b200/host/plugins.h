// plugins.h -- host-side mirror of the reference's plugin surface for the `path` hot path.
//
// One class per plugin the path touches, constructed from Properties exactly like the reference's
// (same property names, defaults and error conditions; file:line cited per class), with the
// reference's addChild()/configure() life cycle.  Nothing here computes radiance: the objects only
// validate, resolve defaults and flatten themselves into the POD records of include/mtsb200.h; the
// work happens behind mtsb_render() on the GPU.
//
//   reference interface                         here
//   ConfigurableObject (core/cobject.h)         ConfigurableObject
//   PluginManager::createObject (plugin.cpp)    createObject()
//   BSDF: diffuse/roughconductor/roughdielectric/twosided   Bsdf
//   Texture: bitmap/checkerboard/gridtexture/scale          Texture
//   Emitter: area/point/constant                Emitter
//   Shape: obj/ply/rectangle/cube/(trimesh)     Shape (always holds triangle meshes)
//   Sensor perspective + Film hdrfilm + rfilter + Sampler independent
//   Integrator path                             PathIntegrator::render() -> mtsb_render()
//   Scene                                       Scene (addChild, configure, flatten)
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "../../include/mtsb200.h"
#include "props.h"

namespace mtsbhost {

enum EClass { CScene, CIntegrator, CSensor, CFilm, CSampler, CRFilter, CShape, CBsdf, CEmitter, CTexture };
const char *className(EClass c);

class ConfigurableObject {
public:
    explicit ConfigurableObject(const Properties &p) : m_props(p) {}
    virtual ~ConfigurableObject() {}
    virtual EClass getClass() const = 0;
    // cobject.cpp:45-50: the base class rejects every child
    virtual void addChild(const std::string &name, const std::shared_ptr<ConfigurableObject> &child);
    virtual void configure() {}
    const Properties &getProperties() const { return m_props; }
    const std::string &getID() const { return m_props.getID(); }
protected:
    Properties m_props;
};
typedef std::shared_ptr<ConfigurableObject> ObjRef;

// ---- textures ----------------------------------------------------------------------------------
struct Image {              // linear RGB, row 0 on top
    int width = 0, height = 0;
    std::vector<float> rgb;
    bool luminance = false; // the file held one channel (replicated into r, g, b)
};
// imageio.cpp: PFM, Radiance RGBE, PNG (8/16 bit, sRGB -> linear as Bitmap::convert does).  `what` names the consumer in error texts.
Image readImage(const std::string &path, const char *what = "Environment map", float gammaOverride = 0.0f);

// src/textures/{bitmap,checkerboard,gridtexture,scale}.cpp + Texture2D (src/librender/texture.cpp:81-96)
class Texture : public ConfigurableObject {
public:
    Texture(const Properties &p, const std::string &baseDir);
    EClass getClass() const override { return CTexture; }
    void addChild(const std::string &name, const ObjRef &child) override;   // scale.cpp:84-89
    void configure() override;                                                // scale.cpp:79-82
    bool isConstant() const;                        // a `scale` around a constant `value` (scale.cpp:55-57, 112-114)
    Vec3 constantValue() const;                     // its value, scales applied
    Vec3 getMaximum() const;                        // component-wise, scales applied (ensureEnergyConservation's input)
    Vec3 getAverage() const;
    // the leaf record with the product of the enclosing scales (and `extra`) folded into MtsbTexture::scale;
    // `texels` points into this object
    MtsbTexture record(const Vec3 &extra) const;
private:
    enum Kind { KBitmap, KChecker, KGrid, KScale } m_kind;
    MtsbTexture m_rec;
    Image m_image; std::vector<float> m_lum;        // bitmap payload (m_lum: one channel per texel for luminance images)
    Vec3 m_scale{1, 1, 1};                          // KScale
    bool m_hasValue = false; Vec3 m_value{0, 0, 0}; // KScale with a constant `value`
    Vec3 m_max{0, 0, 0}, m_avg{0, 0, 0};            // leaf statistics
    std::shared_ptr<Texture> m_nested;
};

// ---- BSDFs -------------------------------------------------------------------------------------
class Bsdf : public ConfigurableObject {
public:
    explicit Bsdf(const Properties &p);
    EClass getClass() const override { return CBsdf; }
    void addChild(const std::string &name, const ObjRef &child) override;   // twosided.cpp:186-196
    void configure() override;
    bool hasTransmission() const { return m_rec.type == MTSB_BSDF_ROUGHDIELECTRIC || m_rec.type == MTSB_BSDF_DIELECTRIC; }
    bool isTwosided() const { return m_twosided; }
    // appends this BSDF's record(s) to `out` (twosided with two children: back first), returns its index;
    // textures the records refer to are appended to `tex`
    int flatten(std::vector<MtsbMaterial> &out, std::vector<MtsbTexture> &tex) const;
private:
    MtsbMaterial record(std::vector<MtsbTexture> &tex) const;
    MtsbMaterial m_rec;
    // textures given as children: slot 0 replaces MtsbMaterial::reflectance, slot 1 ::transmittance
    std::shared_ptr<Texture> m_tex[3];                         // reflectance-like, transmittance-like, roughness `alpha`
    Vec3 m_texScale[3] = {Vec3{1, 1, 1}, Vec3{1, 1, 1}, Vec3{1, 1, 1}};      // ensureEnergyConservation's factor (slots 0, 1)
    bool m_twosided = false;
    std::vector<std::shared_ptr<Bsdf>> m_nested;
};

// ---- emitters ----------------------------------------------------------------------------------
class Emitter : public ConfigurableObject {
public:
    explicit Emitter(const Properties &p, const std::string &baseDir = std::string());
    EClass getClass() const override { return CEmitter; }
    bool isArea() const { return m_rec.type == MTSB_EMITTER_AREA; }
    const MtsbEmitter &record() const { return m_rec; }
    int m_parents = 0;      // area.cpp:192-201: exactly one parent shape
    bool isEnvmap() const { return m_rec.type == MTSB_EMITTER_ENVMAP; }
    // envmap.cpp:105-189: image + toWorld + scale; `rgb` of the returned record points into this object
    MtsbEnvmap envmapRecord() const;
private:
    MtsbEmitter m_rec;
    Image m_image; Transform m_toWorld; float m_scale = 1.0f;
};

// ---- shapes ------------------------------------------------------------------------------------
struct Mesh {
    std::string name;
    std::vector<float> positions, normals, texcoords;   // 3,3,2 per vertex; normals/texcoords may be empty
    std::vector<uint32_t> indices;                       // 3 per triangle, local
    size_t vertexCount() const { return positions.size() / 3; }
    size_t triangleCount() const { return indices.size() / 3; }
};

class Shape : public ConfigurableObject {
public:
    explicit Shape(const Properties &p, const std::string &baseDir);
    EClass getClass() const override { return CShape; }
    void addChild(const std::string &name, const ObjRef &child) override;   // shape.cpp:101-133
    void configure() override;                                                // shape.cpp:49-62 + trimesh.cpp:362-386
    const std::vector<Mesh> &meshes() const { return m_meshes; }
    std::shared_ptr<Bsdf> bsdf() const { return m_bsdf; }
    std::shared_ptr<Emitter> emitter() const { return m_emitter; }
private:
    std::vector<Mesh> m_meshes;
    std::shared_ptr<Bsdf> m_bsdf;
    std::shared_ptr<Emitter> m_emitter;
    bool m_faceNormals = false, m_flipNormals = false;
};

// mesh readers (meshio.cpp).  `flipTexCoords` as in obj.cpp:210.
void loadOBJ(const std::string &path, const Transform &toWorld, bool flipTexCoords, bool collapse, int shapeIndex,
             std::vector<Mesh> &out);
void loadPLY(const std::string &path, const Transform &toWorld, std::vector<Mesh> &out);
void loadSerialized(const std::string &path, const Transform &toWorld, int shapeIndex, std::vector<Mesh> &out);   // serializedio.cpp
// trimesh.cpp:608-681
void computeNormals(Mesh &m, bool faceNormals, bool flipNormals);

// ---- sensor side -------------------------------------------------------------------------------
class RFilter : public ConfigurableObject {
public:
    explicit RFilter(const Properties &p);
    EClass getClass() const override { return CRFilter; }
    uint32_t kind; float radius, stddev;
};

class Film : public ConfigurableObject {
public:
    explicit Film(const Properties &p);          // film.cpp:24-54, hdrfilm.cpp:205-260
    EClass getClass() const override { return CFilm; }
    void addChild(const std::string &name, const ObjRef &child) override;   // film.cpp:72-82
    void configure() override;                   // film.cpp:84-92: default gaussian filter
    int width, height, cropX, cropY, cropW, cropH;
    std::shared_ptr<RFilter> filter;
};

class Sampler : public ConfigurableObject {
public:
    explicit Sampler(const Properties &p);       // independent.cpp:55-60, ldsampler.cpp:75-86, halton.cpp:121-130
    EClass getClass() const override { return CSampler; }
    int sampleCount; int seed;
    uint32_t kind;                               // MTSB_SAMPLER_*
    int dimension, scramble;                     // ldsampler `dimension`, halton `scramble`
};

class Sensor : public ConfigurableObject {
public:
    explicit Sensor(const Properties &p);        // sensor.cpp:157-175, 224-233; perspective.cpp:108-120
    EClass getClass() const override { return CSensor; }
    void addChild(const std::string &name, const ObjRef &child) override;   // sensor.cpp:149-157
    void configure() override;                   // sensor.cpp:84-108, 243-282
    void fill(MtsbSensor &s) const;
    std::shared_ptr<Film> film;
    std::shared_ptr<Sampler> sampler;
private:
    Transform m_toWorld;
    float m_near, m_far, m_xfov = 0;
};

// ---- integrator --------------------------------------------------------------------------------
class PathIntegrator : public ConfigurableObject {
public:
    explicit PathIntegrator(const Properties &p);   // integrator.cpp:252-287, path.cpp:110-116
    EClass getClass() const override { return CIntegrator; }
    int maxDepth, rrDepth; bool strictNormals, hideEmitters;
};

// ---- scene -------------------------------------------------------------------------------------
struct FlatScene {          // owns the buffers MtsbSceneDesc points into
    std::vector<MtsbShape> shapes;
    std::vector<MtsbMaterial> materials;
    std::vector<MtsbEmitter> emitters;
    std::vector<MtsbTexture> textures;          // `texels` point into the Texture objects of the Scene
    std::vector<float> positions, normals, texcoords;
    std::vector<uint32_t> indices;
    MtsbEnvmap envmap;          // valid when an emitter has type MTSB_EMITTER_ENVMAP (points into the Emitter object)
    MtsbSceneDesc desc;
    int seed = 1004;
};

class Scene : public ConfigurableObject {
public:
    explicit Scene(const Properties &p) : ConfigurableObject(p) {}
    EClass getClass() const override { return CScene; }
    void addChild(const std::string &name, const ObjRef &child) override;   // scene.cpp:558-617
    void configure() override;                                                // scene.cpp:240-320
    // Scene::initialize's flattening (scene.cpp:322-384): shapes in insertion order, prim ids by offset
    void flatten(FlatScene &out) const;
    std::shared_ptr<PathIntegrator> integrator;
    std::shared_ptr<Sensor> sensor;
    std::vector<std::shared_ptr<Shape>> shapes;
    std::vector<std::shared_ptr<Emitter>> emitters;     // non-area emitters, in insertion order
};

// PluginManager::createObject: instantiate by (expected class, props.getPluginName()); throws like the
// reference when the plugin is unknown ("Plugin \"x\" not found") or of the wrong class.
ObjRef createObject(EClass expected, const Properties &props, const std::string &baseDir);

std::vector<std::string> &warnings();     // Log(EWarn) sink (unqueried properties, clamped values)

inline std::string lower(std::string s) { for (auto &c : s) c = (char) tolower((unsigned char) c); return s; }
[[noreturn]] inline void error(const std::string &msg) { throw std::runtime_error(msg); }   // Log(EError) throws
inline void warn(const std::string &m) { warnings().push_back(m); }

}  // namespace mtsbhost
