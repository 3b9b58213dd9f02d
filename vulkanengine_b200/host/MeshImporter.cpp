// Binary FBX (7.x) reader for the submesh table — see MeshImporter.h for what is and is not reproduced.
#include "MeshImporter.h"

#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>

namespace VulkanCore {

namespace {

// ---- FBX binary records -----------------------------------------------------------------------------
// file = 27-byte header ("Kaydara FBX Binary  \0", 0x1A 0x00, uint32 version) + node records.
// node = endOffset, numProperties, propertyListLen (uint32 each; uint64 from version 7500), uint8 nameLen, name,
//        properties, nested nodes up to endOffset (closed by an all-zero record).
// property = type char + payload: Y int16, C bool, I int32, F float, D double, L int64,
//            f/d/l/i/b arrays {uint32 count, uint32 encoding (1 = zlib), uint32 byteLen, data}, S/R {uint32 len, bytes}.
struct Prop {
    char type = 0;
    int64_t i = 0;
    double d = 0.0;
    std::string s;
    std::vector<double> dv;
    std::vector<int64_t> iv;
    double number() const { return (type == 'D' || type == 'F') ? d : (double)i; }
};

struct Rec {
    std::string name;
    std::vector<Prop> props;
    std::vector<Rec> kids;
    const Rec* child(const char* n) const
    {
        for (const Rec& k : kids)
            if (k.name == n) return &k;
        return nullptr;
    }
};

struct Cursor {
    const std::vector<unsigned char>& b;
    size_t pos;
    uint32_t version;
    template <class T>
    T get()
    {
        if (pos + sizeof(T) > b.size()) throw std::runtime_error("fbx: truncated");
        T v;
        std::memcpy(&v, &b[pos], sizeof(T));
        pos += sizeof(T);
        return v;
    }
    uint64_t word() { return version >= 7500 ? get<uint64_t>() : get<uint32_t>(); }
};

std::vector<unsigned char> inflate_bytes(const unsigned char* src, size_t n, size_t out_bytes)
{
    std::vector<unsigned char> out(out_bytes);
    uLongf len = (uLongf)out_bytes;
    if (uncompress(out.data(), &len, src, (uLong)n) != Z_OK || len != out_bytes) throw std::runtime_error("fbx: zlib array");
    return out;
}

Prop read_prop(Cursor& c)
{
    Prop p;
    p.type = (char)c.get<uint8_t>();
    switch (p.type) {
    case 'Y': p.i = c.get<int16_t>(); break;
    case 'C': p.i = c.get<uint8_t>(); break;
    case 'I': p.i = c.get<int32_t>(); break;
    case 'L': p.i = c.get<int64_t>(); break;
    case 'F': p.d = c.get<float>(); break;
    case 'D': p.d = c.get<double>(); break;
    case 'S':
    case 'R': {
        const uint32_t n = c.get<uint32_t>();
        if (c.pos + n > c.b.size()) throw std::runtime_error("fbx: truncated string");
        p.s.assign(reinterpret_cast<const char*>(&c.b[c.pos]), n);
        c.pos += n;
        break;
    }
    case 'f': case 'd': case 'l': case 'i': case 'b': {
        const uint32_t count = c.get<uint32_t>(), enc = c.get<uint32_t>(), bytes = c.get<uint32_t>();
        if (c.pos + bytes > c.b.size()) throw std::runtime_error("fbx: truncated array");
        const size_t elem = p.type == 'f' || p.type == 'i' ? 4 : p.type == 'b' ? 1 : 8;
        // the element count is a bare uint32 from the file: bound the allocation by what zlib can expand
        // the stored bytes to (deflate tops out near 1032:1)
        if (enc == 1 && (uint64_t)count * elem > (uint64_t)bytes * 1032 + 64) throw std::runtime_error("fbx: array count");
        std::vector<unsigned char> tmp;
        const unsigned char* data = &c.b[c.pos];
        if (enc == 1) {
            tmp = inflate_bytes(data, bytes, (size_t)count * elem);
            data = tmp.data();
        } else if ((size_t)count * elem != bytes) {
            throw std::runtime_error("fbx: array length");
        }
        c.pos += bytes;
        for (uint32_t k = 0; k < count; ++k) {
            if (p.type == 'd') { double v; std::memcpy(&v, data + k * 8, 8); p.dv.push_back(v); }
            else if (p.type == 'f') { float v; std::memcpy(&v, data + k * 4, 4); p.dv.push_back(v); }
            else if (p.type == 'l') { int64_t v; std::memcpy(&v, data + k * 8, 8); p.iv.push_back(v); }
            else if (p.type == 'i') { int32_t v; std::memcpy(&v, data + k * 4, 4); p.iv.push_back(v); }
            else p.iv.push_back((int8_t)data[k]);
        }
        break;
    }
    default: throw std::runtime_error(std::string("fbx: property type '") + p.type + "'");
    }
    return p;
}

// keep the arrays of the few records the submesh table needs; skip the bulk (normals, UVs, ...) undecoded
bool wanted_array_owner(const std::string& name)
{
    return name == "Vertices" || name == "PolygonVertexIndex" || name == "Materials";
}

bool read_rec(Cursor& c, Rec& out)
{
    const size_t rec_start = c.pos;
    const uint64_t end = c.word(), nprops = c.word(), plen = c.word();
    const uint8_t nlen = c.get<uint8_t>();
    if (c.pos + nlen > c.b.size()) throw std::runtime_error("fbx: truncated name");
    out.name.assign(reinterpret_cast<const char*>(&c.b[c.pos]), nlen);
    c.pos += nlen;
    if (end == 0) return false;  // the all-zero record closing a list
    if (end > c.b.size()) throw std::runtime_error("fbx: record past end of file");
    // offsets come from the file: a record must end after it starts and its properties must end inside it,
    // or the parent's loop would walk backwards / past the record
    if (end <= rec_start || plen > end - c.pos) throw std::runtime_error("fbx: record offsets out of order");
    const size_t props_end = c.pos + plen;
    const bool bulky = out.name.rfind("LayerElement", 0) != 0 && !wanted_array_owner(out.name) &&
                       (out.name == "Normals" || out.name == "UV" || out.name == "UVIndex" || out.name == "NormalsW" ||
                        out.name == "Tangents" || out.name == "Binormals" || out.name == "Colors" || out.name == "ColorIndex" ||
                        out.name == "Edges" || out.name == "KeyTime" || out.name == "KeyValueFloat" || out.name == "Content");
    if (bulky) {
        c.pos = props_end;
    } else {
        for (uint64_t k = 0; k < nprops; ++k) out.props.push_back(read_prop(c));
        if (c.pos != props_end) throw std::runtime_error("fbx: property list length");
    }
    while (c.pos < end) {
        Rec kid;
        if (!read_rec(c, kid)) break;
        out.kids.push_back(std::move(kid));
    }
    c.pos = end;
    return true;
}

// ---- double 4x4 (column-major like mat4) -------------------------------------------------------------
struct M4 {
    double m[4][4];
    static M4 id()
    {
        M4 r;
        std::memset(&r, 0, sizeof r);
        r.m[0][0] = r.m[1][1] = r.m[2][2] = r.m[3][3] = 1.0;
        return r;
    }
};
M4 mul(const M4& a, const M4& b)  // a * b, columns of b transformed by a
{
    M4 r;
    for (int col = 0; col < 4; ++col)
        for (int row = 0; row < 4; ++row) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += a.m[k][row] * b.m[col][k];
            r.m[col][row] = s;
        }
    return r;
}
M4 rot_axis(int axis, double deg)
{
    const double a = deg * 3.14159265358979323846 / 180.0, cs = std::cos(a), sn = std::sin(a);
    M4 r = M4::id();
    const int u = (axis + 1) % 3, v = (axis + 2) % 3;  // rotation in the (u, v) plane, right-handed
    r.m[u][u] = cs; r.m[v][v] = cs;
    r.m[u][v] = sn;   // column u, row v
    r.m[v][u] = -sn;  // column v, row u
    return r;
}

struct Model {
    int64_t id = 0;
    std::string name, subtype;
    double t[3] = {0, 0, 0}, r[3] = {0, 0, 0}, s[3] = {1, 1, 1};
};

void read_vec3(const Rec& p, double out[3])
{
    // P: name, type, label, flags, x, y, z
    if (p.props.size() >= 7)
        for (int k = 0; k < 3; ++k) out[k] = p.props[4 + k].number();
}

}  // namespace

std::shared_ptr<MeshSource> MeshImporter::ImportAssimpMesh(const std::string& filepath, std::string* error)
{
    auto fail = [&](const std::string& why) -> std::shared_ptr<MeshSource> {
        if (error) *error = filepath + ": " + why;
        return nullptr;
    };
    FILE* f = fopen(filepath.c_str(), "rb");
    if (!f) return fail("cannot open");
    std::vector<unsigned char> bytes;
    unsigned char buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, f)) > 0) bytes.insert(bytes.end(), buf, buf + got);
    fclose(f);
    if (bytes.size() < 27 || std::memcmp(bytes.data(), "Kaydara FBX Binary  ", 20) != 0) return fail("not a binary FBX file");

    try {
        uint32_t version;
        std::memcpy(&version, &bytes[23], 4);
        Cursor c{bytes, 27, version};
        std::vector<Rec> top;
        while (c.pos + (version >= 7500 ? 25 : 13) <= bytes.size()) {
            Rec r;
            if (!read_rec(c, r)) break;
            top.push_back(std::move(r));
        }
        const Rec *objects = nullptr, *connections = nullptr, *settings = nullptr;
        for (const Rec& r : top) {
            if (r.name == "Objects") objects = &r;
            if (r.name == "Connections") connections = &r;
            if (r.name == "GlobalSettings") settings = &r;
        }
        if (!objects || !connections) return fail("no Objects / Connections");

        double unit_scale = 1.0;
        if (settings)
            if (const Rec* p70 = settings->child("Properties70"))
                for (const Rec& p : p70->kids)
                    if (!p.props.empty() && p.props[0].s == "UnitScaleFactor" && p.props.size() >= 5) unit_scale = p.props[4].number();
        const double file_scale = unit_scale * 0.01;  // FBXImporter.cpp:171-179, applied by aiProcess_GlobalScale

        std::map<int64_t, Model> models;
        std::map<int64_t, const Rec*> geometries;
        std::map<int64_t, bool> materials;
        for (const Rec& o : objects->kids) {
            if (o.props.size() < 3) continue;
            const int64_t id = o.props[0].i;
            if (o.name == "Model") {
                Model m;
                m.id = id;
                m.name = o.props[1].s.substr(0, o.props[1].s.find('\0'));
                m.subtype = o.props[2].s;
                if (const Rec* p70 = o.child("Properties70"))
                    for (const Rec& p : p70->kids) {
                        if (p.props.empty()) continue;
                        if (p.props[0].s == "Lcl Translation") read_vec3(p, m.t);
                        if (p.props[0].s == "Lcl Rotation") read_vec3(p, m.r);
                        if (p.props[0].s == "Lcl Scaling") read_vec3(p, m.s);
                    }
                models[id] = m;
            } else if (o.name == "Geometry" && o.props[2].s == "Mesh") {
                geometries[id] = &o;
            } else if (o.name == "Material") {
                materials[id] = true;
            }
        }
        // object-object connections, in file order: child -> parent
        std::vector<std::pair<int64_t, int64_t>> links;
        for (const Rec& k : connections->kids)
            if (k.name == "C" && k.props.size() >= 3 && k.props[0].s == "OO") links.emplace_back(k.props[1].i, k.props[2].i);

        auto source = std::make_shared<MeshSource>();
        std::map<int64_t, int> material_index;  // scene-wide material numbers, in order of first use
        uint32_t base_vertex = 0, base_index = 0;

        // depth-first over the model hierarchy from the root (id 0), children in connection order
        std::vector<int64_t> stack{0};
        std::vector<int64_t> order;
        while (!stack.empty()) {
            const int64_t parent = stack.back();
            stack.pop_back();
            if (parent) order.push_back(parent);
            std::vector<int64_t> kids;
            for (const auto& l : links)
                if (l.second == parent && models.count(l.first)) kids.push_back(l.first);
            for (size_t k = kids.size(); k-- > 0;) stack.push_back(kids[k]);
        }
        for (int64_t mid : order) {
            const Model& m = models[mid];
            if (m.subtype != "Mesh") continue;
            std::vector<int64_t> slots;  // the model's material list
            for (const auto& l : links)
                if (l.second == mid && materials.count(l.first)) slots.push_back(l.first);
            M4 T = M4::id(), S = M4::id();
            for (int k = 0; k < 3; ++k) {
                T.m[3][k] = m.t[k] * file_scale;
                S.m[k][k] = m.s[k];
            }
            const M4 R = mul(rot_axis(2, m.r[2]), mul(rot_axis(1, m.r[1]), rot_axis(0, m.r[0])));  // eEulerXYZ
            const M4 local = mul(T, mul(R, S));
            mat4 local_f;
            for (int col = 0; col < 4; ++col)
                for (int row = 0; row < 4; ++row) local_f.c[col][row] = (float)local.m[col][row];

            for (const auto& l : links) {
                if (l.second != mid || !geometries.count(l.first)) continue;
                const Rec& g = *geometries[l.first];
                const Rec *vr = g.child("Vertices"), *pr = g.child("PolygonVertexIndex");
                if (!vr || !pr || vr->props.empty() || pr->props.empty()) continue;
                const std::vector<double>& verts = vr->props[0].dv;
                const std::vector<int64_t>& pvi = pr->props[0].iv;
                std::vector<int64_t> per_poly;  // material slot of each polygon; empty = all slot `all_same`
                int64_t all_same = 0;
                if (const Rec* lm = g.child("LayerElementMaterial")) {
                    const Rec *map = lm->child("MappingInformationType"), *mats = lm->child("Materials");
                    if (mats && !mats->props.empty()) {
                        const std::vector<int64_t>& mv = mats->props[0].iv;
                        if (map && !map->props.empty() && map->props[0].s == "ByPolygon") per_poly = mv;
                        else if (!mv.empty()) all_same = mv[0];
                    }
                }
                struct Acc {
                    uint32_t corners = 0, indices = 0;
                    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
                    std::vector<vec3> positions;    // one vertex per polygon corner, polygon by polygon
                    std::vector<uint32_t> fan;      // triangle fans over them, relative to the submesh's first vertex
                };
                std::map<int64_t, Acc> by_slot;
                size_t poly = 0, start = 0;
                for (size_t k = 0; k < pvi.size(); ++k) {
                    if (pvi[k] >= 0) continue;
                    const size_t corners = k - start + 1;
                    const int64_t slot = per_poly.empty() ? all_same : (poly < per_poly.size() ? per_poly[poly] : 0);
                    if (corners >= 3) {
                        Acc& a = by_slot[slot];
                        const uint32_t v0 = a.corners;
                        for (uint32_t f = 1; f + 1 < corners; ++f) {  // aiProcess_Triangulate on a convex polygon: a fan
                            a.fan.push_back(v0);
                            a.fan.push_back(v0 + f);
                            a.fan.push_back(v0 + f + 1);
                        }
                        a.corners += (uint32_t)corners;
                        a.indices += (uint32_t)(corners - 2) * 3u;
                        for (size_t q = start; q <= k; ++q) {
                            const int64_t vi = pvi[q] < 0 ? ~pvi[q] : pvi[q];
                            if ((size_t)vi * 3 + 2 >= verts.size()) throw std::runtime_error("fbx: vertex index out of range");
                            float p[3];
                            for (int ax = 0; ax < 3; ++ax) {
                                p[ax] = (float)verts[(size_t)vi * 3 + ax] * (float)file_scale;  // aiVector3D is float
                                a.lo[ax] = std::min(a.lo[ax], p[ax]);
                                a.hi[ax] = std::max(a.hi[ax], p[ax]);
                            }
                            a.positions.push_back({p[0], p[1], p[2]});
                        }
                    }
                    start = k + 1;
                    ++poly;
                }
                for (const auto& kv : by_slot) {
                    const Acc& a = kv.second;
                    Submesh sm;
                    sm.BaseVertex = base_vertex;
                    sm.BaseIndex = base_index;
                    sm.VertexCount = a.corners;
                    sm.IndexCount = a.indices;
                    int mat = 0;
                    if (kv.first >= 0 && (size_t)kv.first < slots.size()) {
                        auto it = material_index.find(slots[(size_t)kv.first]);
                        if (it == material_index.end()) it = material_index.emplace(slots[(size_t)kv.first], (int)material_index.size()).first;
                        mat = it->second;
                    }
                    sm.MaterialIndex = mat;
                    sm.LocalTransform = local_f;
                    sm.BoundsMin = {a.lo[0], a.lo[1], a.lo[2]};
                    sm.BoundsMax = {a.hi[0], a.hi[1], a.hi[2]};
                    source->GetSubmeshes().push_back(sm);
                    source->GetPositions().insert(source->GetPositions().end(), a.positions.begin(), a.positions.end());
                    source->GetIndices().insert(source->GetIndices().end(), a.fan.begin(), a.fan.end());
                    base_vertex += a.corners;
                    base_index += a.indices;
                }
            }
        }
        if (source->GetSubmeshes().empty()) return fail("no mesh geometry");
        return source;
    } catch (const std::exception& e) {
        return fail(e.what());
    }
}

}  // namespace VulkanCore
