// scene.cpp -- Scene object: child bookkeeping, defaults and the flattening into MtsbSceneDesc
// (reference: src/librender/scene.cpp:240-384, 558-617).
#include <cstring>
#include <map>

#include "plugins.h"

namespace mtsbhost {

ObjRef createObject(EClass expected, const Properties &props, const std::string &baseDir) {
    // plugin.cpp:196-216: createObject(classType, props) verifies the class of the instantiated plugin
    switch (expected) {
        case CScene: return std::make_shared<Scene>(props);
        case CIntegrator: return std::make_shared<PathIntegrator>(props);
        case CSensor: return std::make_shared<Sensor>(props);
        case CFilm: return std::make_shared<Film>(props);
        case CSampler: return std::make_shared<Sampler>(props);
        case CRFilter: return std::make_shared<RFilter>(props);
        case CShape: return std::make_shared<Shape>(props, baseDir);
        case CBsdf: return std::make_shared<Bsdf>(props);
        case CEmitter: return std::make_shared<Emitter>(props, baseDir);
        case CTexture: return std::make_shared<Texture>(props, baseDir);
    }
    error("createObject: unknown class");
}

void Scene::addChild(const std::string &name, const ObjRef &child) {
    switch (child->getClass()) {
        case CIntegrator: integrator = std::static_pointer_cast<PathIntegrator>(child); break;
        case CSensor: sensor = std::static_pointer_cast<Sensor>(child); break;          // scene.cpp:566-570: the last one wins
        case CShape: {
            auto s = std::static_pointer_cast<Shape>(child);
            shapes.push_back(s);
            if (s->emitter()) emitters.push_back(s->emitter());       // Scene::addShape, scene.cpp:596-597
            break;
        }
        case CEmitter: {
            auto e = std::static_pointer_cast<Emitter>(child);
            if (e->isArea()) error("An area light must be child of a shape instance");   // area.cpp:198-201
            emitters.push_back(e);
            break;
        }
        case CScene: {
            // an <include>d scene: take over its contents (scene.cpp:522-545), emitters, sensor, then shapes
            auto inc = std::static_pointer_cast<Scene>(child);
            for (auto &e : inc->emitters) emitters.push_back(e);
            if (inc->sensor) sensor = inc->sensor;
            if (inc->integrator && !integrator) integrator = inc->integrator;
            for (auto &s : inc->shapes) shapes.push_back(s);
            break;
        }
        case CTexture:          // a named top-level texture that BSDFs <ref>erence
        case CBsdf: break;      // a named top-level BSDF that shapes <ref>erence; the scene itself ignores it (scene.cpp:612)
        default: ConfigurableObject::addChild(name, child);
    }
}

void Scene::configure() {
    // scene.cpp:272-277 installs the `direct` integrator when none is given; that integrator is not this path
    if (!integrator)
        error("The scene has no <integrator>: the reference would install `direct`; this library runs `path` -- add <integrator type=\"path\"/>.");
    if (!sensor)
        error("The scene has no <sensor>: add a perspective sensor (the reference's automatic camera placement is not implemented).");
    int nEnv = 0;
    for (auto &e : emitters) if (e->record().type == MTSB_EMITTER_CONSTANT || e->record().type == MTSB_EMITTER_ENVMAP) ++nEnv;
    if (nEnv > 1) error("The scene may only contain one environment emitter");      // scene.cpp:339-341
}

void Scene::flatten(FlatScene &f) const {
    std::map<const Bsdf *, int> matIndex;
    std::map<const Emitter *, int> emIndex;
    for (auto &e : emitters) {
        emIndex[e.get()] = (int) f.emitters.size();
        f.emitters.push_back(e->record());
    }
    bool anyNormals = false, anyUV = false;
    for (auto &s : shapes) for (auto &m : s->meshes()) { anyNormals |= !m.normals.empty(); anyUV |= !m.texcoords.empty(); }
    for (auto &s : shapes) {
        const Bsdf *b = s->bsdf().get();
        auto it = matIndex.find(b);
        int mat = it != matIndex.end() ? it->second : (matIndex[b] = b->flatten(f.materials, f.textures));
        for (auto &m : s->meshes()) {
            MtsbShape rec;
            rec.first_vertex = (uint32_t) (f.positions.size() / 3);
            rec.n_vertices = (uint32_t) m.vertexCount();
            rec.first_tri = (uint32_t) (f.indices.size() / 3);
            rec.n_tris = (uint32_t) m.triangleCount();
            rec.material = mat;
            rec.emitter = -1;
            rec.has_normals = !m.normals.empty();
            rec.has_texcoords = !m.texcoords.empty();
            if (s->emitter()) {
                rec.emitter = emIndex[s->emitter().get()];
                f.emitters[rec.emitter].shape = (int32_t) f.shapes.size();
            }
            f.positions.insert(f.positions.end(), m.positions.begin(), m.positions.end());
            if (anyNormals) {
                if (m.normals.empty()) f.normals.insert(f.normals.end(), 3 * m.vertexCount(), 0.0f);
                else f.normals.insert(f.normals.end(), m.normals.begin(), m.normals.end());
            }
            if (anyUV) {
                if (m.texcoords.empty()) f.texcoords.insert(f.texcoords.end(), 2 * m.vertexCount(), 0.0f);
                else f.texcoords.insert(f.texcoords.end(), m.texcoords.begin(), m.texcoords.end());
            }
            for (uint32_t i : m.indices) f.indices.push_back(i + rec.first_vertex);
            f.shapes.push_back(rec);
        }
    }
    if (f.positions.size() / 3 > 0xFFFFFFF0ull || f.indices.size() / 3 > 0xFFFFFFF0ull) error("Scene too large for 32-bit primitive ids");

    MtsbSceneDesc &d = f.desc;
    std::memset(&d, 0, sizeof(d));
    d.abi_version = MTSB_ABI_VERSION;
    d.n_shapes = (uint32_t) f.shapes.size();
    d.n_materials = (uint32_t) f.materials.size();
    d.n_emitters = (uint32_t) f.emitters.size();
    d.n_vertices = f.positions.size() / 3;
    d.n_tris = f.indices.size() / 3;
    d.shapes = f.shapes.data();
    d.materials = f.materials.data();
    d.emitters = f.emitters.data();
    d.positions = f.positions.data();
    d.normals = anyNormals ? f.normals.data() : nullptr;
    d.texcoords = anyUV ? f.texcoords.data() : nullptr;
    d.indices = f.indices.data();
    sensor->fill(d.sensor);
    d.integrator.max_depth = integrator->maxDepth;
    d.integrator.rr_depth = integrator->rrDepth;
    d.integrator.strict_normals = integrator->strictNormals;
    d.integrator.hide_emitters = integrator->hideEmitters;
    d.integrator.sample_count = sensor->sampler->sampleCount;
    d.envmap = nullptr;
    for (auto &e : emitters) if (e->isEnvmap()) { f.envmap = e->envmapRecord(); d.envmap = &f.envmap; }
    d.textures = f.textures.empty() ? nullptr : f.textures.data();
    d.n_textures = (uint32_t) f.textures.size();
    d.sampler.type = sensor->sampler->kind; d.sampler.dimension = sensor->sampler->dimension; d.sampler.scramble = sensor->sampler->scramble; d.sampler._pad = 0;
    f.seed = sensor->sampler->seed;
}

}  // namespace mtsbhost
