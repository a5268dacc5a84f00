// plyio.cpp -- PLY reader (ascii, binary_little_endian, binary_big_endian) for the `ply` shape
// (src/shapes/ply.cpp uses the rply library; the element/property handling below follows its
// callbacks: vertex x y z [nx ny nz] [u v | s t], face vertex_indices|vertex_index lists, quads
// split as (0,1,2),(0,2,3), ply.cpp:343-377).
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>

#include "plugins.h"

namespace mtsbhost {

namespace {

enum PlyType { I8, U8, I16, U16, I32, U32, F32, F64, TNone };
struct PlyProp { std::string name; PlyType type = TNone, countType = TNone; bool list = false; };
struct PlyElem { std::string name; size_t count = 0; std::vector<PlyProp> props; };

PlyType parseType(const std::string &s) {
    if (s == "char" || s == "int8") return I8;
    if (s == "uchar" || s == "uint8") return U8;
    if (s == "short" || s == "int16") return I16;
    if (s == "ushort" || s == "uint16") return U16;
    if (s == "int" || s == "int32") return I32;
    if (s == "uint" || s == "uint32") return U32;
    if (s == "float" || s == "float32") return F32;
    if (s == "double" || s == "float64") return F64;
    error("PLY: unknown property type \"" + s + "\"");
}
size_t typeSize(PlyType t) { static const size_t s[] = {1, 1, 2, 2, 4, 4, 4, 8, 0}; return s[t]; }

struct Reader {
    std::istream &is;
    int format;       // 0 ascii, 1 little endian, 2 big endian
    double next(PlyType t) {
        if (format == 0) {
            double v;
            if (!(is >> v)) error("PLY: unexpected end of file");
            return v;
        }
        unsigned char b[8];
        size_t n = typeSize(t);
        is.read((char *) b, (std::streamsize) n);
        if ((size_t) is.gcount() != n) error("PLY: unexpected end of file");
        if (format == 2) for (size_t i = 0; i < n / 2; ++i) std::swap(b[i], b[n - 1 - i]);
        switch (t) {
            case I8: return (double) *(int8_t *) b;
            case U8: return (double) *(uint8_t *) b;
            case I16: { int16_t v; std::memcpy(&v, b, 2); return v; }
            case U16: { uint16_t v; std::memcpy(&v, b, 2); return v; }
            case I32: { int32_t v; std::memcpy(&v, b, 4); return v; }
            case U32: { uint32_t v; std::memcpy(&v, b, 4); return v; }
            case F32: { float v; std::memcpy(&v, b, 4); return v; }
            case F64: { double v; std::memcpy(&v, b, 8); return v; }
            default: return 0;
        }
    }
};

}  // namespace

void loadPLY(const std::string &path, const Transform &toWorld, std::vector<Mesh> &out) {
    std::ifstream is(path, std::ios::binary);
    if (!is.good()) error("PLY file \"" + path + "\" could not be found!");
    std::string line, tok;
    if (!std::getline(is, line) || line.substr(0, 3) != "ply") error("\"" + path + "\" is not a PLY file");
    int format = -1;
    std::vector<PlyElem> elems;
    while (std::getline(is, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ls(line);
        if (!(ls >> tok)) continue;
        if (tok == "format") {
            ls >> tok;
            format = tok == "ascii" ? 0 : tok == "binary_little_endian" ? 1 : tok == "binary_big_endian" ? 2 : -1;
            if (format < 0) error("PLY: unknown format \"" + tok + "\"");
        } else if (tok == "element") {
            PlyElem e;
            ls >> e.name >> e.count;
            elems.push_back(e);
        } else if (tok == "property") {
            if (elems.empty()) error("PLY: property before element");
            PlyProp p;
            ls >> tok;
            if (tok == "list") {
                std::string ct, it;
                ls >> ct >> it >> p.name;
                p.list = true; p.countType = parseType(ct); p.type = parseType(it);
            } else { p.type = parseType(tok); ls >> p.name; }
            elems.back().props.push_back(p);
        } else if (tok == "end_header") break;
    }
    if (format < 0) error("PLY: missing format line");

    Reader rd{is, format};
    Mesh m;
    m.name = path.substr(path.find_last_of('/') == std::string::npos ? 0 : path.find_last_of('/') + 1);
    Transform inv = toWorld.inverse();
    for (const PlyElem &e : elems) {
        if (e.name == "vertex") {
            int ix = -1, iy = -1, iz = -1, inx = -1, iny = -1, inz = -1, iu = -1, iv = -1;
            for (size_t k = 0; k < e.props.size(); ++k) {
                const std::string &n = e.props[k].name;
                if (e.props[k].list) error("PLY: list property in the vertex element");
                if (n == "x") ix = (int) k; else if (n == "y") iy = (int) k; else if (n == "z") iz = (int) k;
                else if (n == "nx") inx = (int) k; else if (n == "ny") iny = (int) k; else if (n == "nz") inz = (int) k;
                else if (n == "u" || n == "s") iu = (int) k; else if (n == "v" || n == "t") iv = (int) k;
            }
            if (ix < 0 || iy < 0 || iz < 0) error("PLY: vertex element without x/y/z");
            const bool hasN = inx >= 0 && iny >= 0 && inz >= 0, hasUV = iu >= 0 && iv >= 0;
            m.positions.reserve(3 * e.count);
            std::vector<double> row(e.props.size());
            for (size_t i = 0; i < e.count; ++i) {
                for (size_t k = 0; k < e.props.size(); ++k) row[k] = rd.next(e.props[k].type);
                Vec3 p = toWorld.point(Vec3{(float) row[ix], (float) row[iy], (float) row[iz]});
                m.positions.push_back(p.x); m.positions.push_back(p.y); m.positions.push_back(p.z);
                if (hasN) {
                    Vec3 n0{(float) row[inx], (float) row[iny], (float) row[inz]};
                    Vec3 n{inv.m[0] * n0.x + inv.m[4] * n0.y + inv.m[8] * n0.z, inv.m[1] * n0.x + inv.m[5] * n0.y + inv.m[9] * n0.z,
                           inv.m[2] * n0.x + inv.m[6] * n0.y + inv.m[10] * n0.z};
                    float l = std::sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
                    if (l != 0) n = Vec3{n.x / l, n.y / l, n.z / l};
                    m.normals.push_back(n.x); m.normals.push_back(n.y); m.normals.push_back(n.z);
                }
                if (hasUV) { m.texcoords.push_back((float) row[iu]); m.texcoords.push_back((float) row[iv]); }
            }
        } else if (e.name == "face") {
            for (size_t i = 0; i < e.count; ++i) {
                for (const PlyProp &p : e.props) {
                    if (!p.list) { rd.next(p.type); continue; }
                    size_t cnt = (size_t) rd.next(p.countType);
                    std::vector<uint32_t> idx(cnt);
                    for (size_t k = 0; k < cnt; ++k) idx[k] = (uint32_t) rd.next(p.type);
                    if (p.name != "vertex_indices" && p.name != "vertex_index") continue;
                    if (cnt != 3 && cnt != 4) error("PLY: only triangle and quad faces are supported!");     // ply.cpp:349-351
                    for (uint32_t k : idx) if (k >= m.vertexCount()) error("PLY: face index out of range");
                    m.indices.insert(m.indices.end(), {idx[0], idx[1], idx[2]});
                    if (cnt == 4) m.indices.insert(m.indices.end(), {idx[0], idx[2], idx[3]});
                }
            }
        } else {
            // skip unknown elements
            for (size_t i = 0; i < e.count; ++i)
                for (const PlyProp &p : e.props) {
                    if (!p.list) { rd.next(p.type); continue; }
                    size_t cnt = (size_t) rd.next(p.countType);
                    for (size_t k = 0; k < cnt; ++k) rd.next(p.type);
                }
        }
    }
    if (m.indices.empty()) error("PLY file \"" + path + "\" contains no faces");
    out.push_back(std::move(m));
}

}  // namespace mtsbhost
