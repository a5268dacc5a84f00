// meshio.cpp -- Wavefront OBJ and PLY readers producing the vertex / triangle order of the
// reference's loaders, which the global primitive ids depend on (src/shapes/obj.cpp:189-350, 608-715;
// src/shapes/ply.cpp).
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

#include "plugins.h"

namespace mtsbhost {

namespace {

struct ObjTri { int p[3] = {0, 0, 0}, n[3] = {0, 0, 0}, uv[3] = {0, 0, 0}; };
struct ObjVertex {
    float v[8];      // p(3) n(3) uv(2)
    bool operator<(const ObjVertex &o) const { return std::memcmp(v, o.v, sizeof(v)) < 0; }   // any strict order keyed on all 8 values
};

std::string trim(const std::string &s) {
    size_t b = s.find_first_not_of(" \t\r\n"), e = s.find_last_not_of(" \t\r\n");
    return b == std::string::npos ? "" : s.substr(b, e - b + 1);
}

// obj.cpp:367-388: "p", "p/uv", "p//n", "p/uv/n"
void parseCorner(ObjTri &t, int i, const std::string &tok) {
    size_t pos = 0;
    bool doubleSlash = tok.find("//") != std::string::npos;
    std::vector<std::string> parts;
    while (pos <= tok.size()) {
        size_t e = tok.find('/', pos);
        if (e == std::string::npos) e = tok.size();
        if (e > pos) parts.push_back(tok.substr(pos, e - pos));
        pos = e + 1;
    }
    if (parts.size() == 1) t.p[i] = atoi(parts[0].c_str());
    else if (parts.size() == 2) {
        t.p[i] = atoi(parts[0].c_str());
        if (doubleSlash) t.n[i] = atoi(parts[1].c_str()); else t.uv[i] = atoi(parts[1].c_str());
    } else if (parts.size() == 3) {
        t.p[i] = atoi(parts[0].c_str()); t.uv[i] = atoi(parts[1].c_str()); t.n[i] = atoi(parts[2].c_str());
    } else error("Invalid OBJ face format!");
}

// obj.cpp:608-715: vertices are created in order of first use, merged when position, normal and uv all match
void createMesh(const std::string &name, const std::vector<Vec3> &vertices, const std::vector<Vec3> &normals,
                const std::vector<float> &texcoords, const std::vector<ObjTri> &tris, const Transform &toWorld,
                std::vector<Mesh> &out) {
    if (tris.empty()) return;
    std::map<ObjVertex, uint32_t> vmap;
    std::vector<ObjVertex> vbuf;
    bool hasN = false, hasUV = false;
    Mesh m;
    m.name = name;
    m.indices.reserve(3 * tris.size());
    // the normal transform needs the inverse once, not per vertex
    Transform inv = toWorld.inverse();
    for (const ObjTri &t : tris) {
        for (int j = 0; j < 3; ++j) {
            int vi = t.p[j], ni = t.n[j], ti = t.uv[j];
            if (vi < 0) vi += (int) vertices.size() + 1;
            if (ni < 0) ni += (int) normals.size() + 1;
            if (ti < 0) ti += (int) (texcoords.size() / 2) + 1;
            if (vi > (int) vertices.size() || vi <= 0)
                error("Out of bounds: tried to access vertex " + std::to_string(vi) + " (max: " + std::to_string(vertices.size()) + ")");
            ObjVertex v;
            std::memset(&v, 0, sizeof(v));
            Vec3 p = toWorld.point(vertices[vi - 1]);
            v.v[0] = p.x; v.v[1] = p.y; v.v[2] = p.z;
            if (ni != 0) {
                if (ni > (int) normals.size() || ni < 0)
                    error("Out of bounds: tried to access normal " + std::to_string(ni) + " (max: " + std::to_string(normals.size()) + ")");
                const Vec3 &n0 = normals[ni - 1];
                Vec3 n{inv.m[0] * n0.x + inv.m[4] * n0.y + inv.m[8] * n0.z, inv.m[1] * n0.x + inv.m[5] * n0.y + inv.m[9] * n0.z,
                       inv.m[2] * n0.x + inv.m[6] * n0.y + inv.m[10] * n0.z};
                float l = std::sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
                if (l != 0) n = Vec3{n.x / l, n.y / l, n.z / l};
                v.v[3] = n.x; v.v[4] = n.y; v.v[5] = n.z;
                hasN = true;
            }
            if (ti != 0) {
                if (ti > (int) (texcoords.size() / 2) || ti < 0)
                    error("Out of bounds: tried to access uv " + std::to_string(ti) + " (max: " + std::to_string(texcoords.size() / 2) + ")");
                v.v[6] = texcoords[2 * (ti - 1)]; v.v[7] = texcoords[2 * (ti - 1) + 1];
                hasUV = true;
            }
            for (float &f : v.v) if (f == 0.0f) f = 0.0f;       // -0 and +0 compare equal in the reference's key order
            auto it = vmap.find(v);
            uint32_t key;
            if (it != vmap.end()) key = it->second;
            else { key = (uint32_t) vbuf.size(); vmap[v] = key; vbuf.push_back(v); }
            m.indices.push_back(key);
        }
    }
    m.positions.reserve(3 * vbuf.size());
    for (const ObjVertex &v : vbuf) {
        m.positions.insert(m.positions.end(), v.v, v.v + 3);
        if (hasN) m.normals.insert(m.normals.end(), v.v + 3, v.v + 6);
        if (hasUV) m.texcoords.insert(m.texcoords.end(), v.v + 6, v.v + 8);
    }
    out.push_back(std::move(m));
}

}  // namespace

void loadOBJ(const std::string &path, const Transform &toWorld, bool flipTexCoords, bool collapse, int shapeIndex,
             std::vector<Mesh> &out) {
    std::ifstream is(path);
    if (!is.good()) error("Wavefront OBJ file '" + path + "' not found!");
    std::string stem = path.substr(path.find_last_of('/') == std::string::npos ? 0 : path.find_last_of('/') + 1);
    if (stem.rfind('.') != std::string::npos) stem = stem.substr(0, stem.rfind('.'));

    std::vector<Vec3> vertices, normals;
    std::vector<float> texcoords;
    std::vector<ObjTri> tris;
    std::string name = stem, line, buf;
    int geomIndex = 0;
    bool nameBeforeGeometry = false;
    auto flush = [&](const std::string &n) {
        geomIndex += 1;
        if (shapeIndex < 0 || geomIndex - 1 == shapeIndex) createMesh(n, vertices, normals, texcoords, tris, toWorld, out);
        tris.clear();
    };
    while (std::getline(is, line)) {
        // obj.cpp:172-186: a trailing backslash continues the line
        while (!line.empty() && (line.back() == '\r' || line.back() == ' ')) line.pop_back();
        while (!line.empty() && line.back() == '\\') {
            std::string next;
            line.pop_back();
            if (!std::getline(is, next)) break;
            line += next;
        }
        std::istringstream iss(line);
        if (!(iss >> buf)) continue;
        if (buf == "v") { Vec3 p; iss >> p.x >> p.y >> p.z; vertices.push_back(p); }
        else if (buf == "vn") { Vec3 n; iss >> n.x >> n.y >> n.z; normals.push_back(n); }
        else if (buf == "vt") {
            float u = 0, v = 0;
            iss >> u >> v;
            if (flipTexCoords) v = 1 - v;
            texcoords.push_back(u); texcoords.push_back(v);
        } else if (buf == "g" && !collapse) {
            std::string newName = trim(line.substr(1));
            std::string target = nameBeforeGeometry ? name : newName;     // two naming conventions, obj.cpp:261-270
            if (!tris.empty()) flush(target); else nameBeforeGeometry = true;
            name = newName;
        } else if (buf == "usemtl") {
            if (!tris.empty() && !collapse) { flush(name); name = stem; }
        } else if (buf == "f") {
            std::string tok;
            ObjTri t;
            for (int i = 0; i < 3; ++i) { if (!(iss >> tok)) error("Invalid OBJ face format!"); parseCorner(t, i, tok); }
            tris.push_back(t);
            while (iss >> tok) {            // n-gons as a fan (convex assumption, obj.cpp:312-319)
                t.p[1] = t.p[2]; t.uv[1] = t.uv[2]; t.n[1] = t.n[2];
                parseCorner(t, 2, tok);
                tris.push_back(t);
            }
        }
    }
    flush(name);
    if (out.empty()) error("Wavefront OBJ file '" + path + "' contains no geometry" + (shapeIndex >= 0 ? " at the requested shapeIndex" : ""));
}

}  // namespace mtsbhost

