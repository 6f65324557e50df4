// rt_resource.cpp — model loaders, mirroring /root/reference/src/core/resource.rs.
//
// Upstream leans on the tobj 3.2.5 and gltf 1.1.0 crates for parsing; neither is vendored
// in the reference tree, so the parsing itself is written here against the published
// formats (Wavefront OBJ; glTF 2.0 JSON + GLB container), keeping exactly what the
// reference does with the parsed data: one Mesh record per OBJ model / glTF primitive,
// pos = (mesh_count*3, 0, 0), colour (0.2,0.2,1,1), specular 0.5, triangulated and
// single-indexed OBJ (resource.rs:57-63), node transforms ignored for .glb (:228-276),
// positions scaled by the running mesh count for .gltf (:180), normals required.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <tuple>
#include <zlib.h>
#include "rt_host.hpp"

namespace rtb {

namespace {

// texture extension: when set, the loaders keep texture coordinates (load_model_uv)
thread_local bool g_keep_uv = false;

std::string read_file(const std::string& path, bool binary)
{
    std::ifstream f(path, binary ? std::ios::binary : std::ios::in);
    if (!f) throw std::runtime_error((binary ? "Binary file does not exist: " : "Text file does not exist: ") + path);  // resource.rs:8-12,19-23
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

bool ends_with(const std::string& s, const char* suf)
{
    size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

std::string sibling(const std::string& path, const std::string& name)
{
    size_t p = path.find_last_of('/');
    return p == std::string::npos ? name : path.substr(0, p + 1) + name;
}

// RFC 4648 base64 (padding optional, whitespace skipped); anything else throws
std::string base64_decode(const char* s, size_t n)
{
    std::string out; out.reserve(n / 4 * 3);
    uint32_t acc = 0; int bits = 0;
    for (size_t i = 0; i < n; i++) {
        const unsigned char c = (unsigned char)s[i];
        int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A';
        else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
        else if (c >= '0' && c <= '9') v = c - '0' + 52;
        else if (c == '+' || c == '-') v = 62;
        else if (c == '/' || c == '_') v = 63;
        else if (c == '=' ) break;
        else if (c == ' ' || c == '\n' || c == '\r' || c == '\t') continue;
        else throw std::runtime_error("glTF: bad character in a base64 data: URI");
        acc = (acc << 6) | (uint32_t)v; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((char)((acc >> bits) & 0xffu)); }
    }
    return out;
}

// Scene-graph import only (extension): a glTF `uri` is a sibling file or an RFC 2397 `data:` URI with a base64
// payload.  Upstream's loaders treat every uri as a file name (resource.rs:129-132, 211-214) and so do ours.
std::string resolve_uri(const std::string& path, const std::string& uri)
{
    if (uri.compare(0, 5, "data:") != 0) return read_file(sibling(path, uri), true);
    const size_t comma = uri.find(',');
    if (comma == std::string::npos) throw std::runtime_error("glTF: data: URI without a payload");
    if (uri.compare(comma >= 7 ? comma - 7 : 0, 7, ";base64") != 0) throw std::runtime_error("glTF: data: URI is not base64");
    return base64_decode(uri.data() + comma + 1, uri.size() - comma - 1);
}

// ---- minimal JSON (enough for glTF) ----------------------------------------------------
struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    double num = 0; bool b = false; std::string str;
    std::vector<Json> arr; std::vector<std::pair<std::string, Json>> obj;
    const Json* get(const char* k) const { for (auto& p : obj) if (p.first == k) return &p.second; return nullptr; }
    const Json& at(const char* k) const { const Json* j = get(k); if (!j) throw std::runtime_error(std::string("glTF: missing key ") + k); return *j; }
    // numbers as integers, saturating (a cast of an out-of-range double is undefined behaviour): NaN -> 0, clamp to +-2^62
    static long long to_ll(double v) { if (!(v == v)) return 0; if (v >= 4.6e18) return 1ll << 62; if (v <= -4.6e18) return -(1ll << 62); return (long long)v; }
    long long integer(const char* k, long long dflt) const { const Json* j = get(k); return j ? to_ll(j->num) : dflt; }
    // an array index: anything negative or absurd becomes SIZE_MAX, which every bounds test and .at() rejects
    size_t index() const { const long long v = to_ll(num); return v < 0 ? (size_t)-1 : (size_t)v; }
};

struct JsonParser {
    const char* p; const char* e;
    void ws() { while (p < e && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++; }
    [[noreturn]] void bad(const char* m) { throw std::runtime_error(std::string("JSON: ") + m); }
    Json value()
    {
        ws();
        if (p >= e) bad("unexpected end");
        Json j;
        if (*p == '{') {
            j.kind = Json::Obj; p++; ws();
            if (p < e && *p == '}') { p++; return j; }
            for (;;) {
                ws(); Json k = value(); if (k.kind != Json::Str) bad("object key");
                ws(); if (p >= e || *p != ':') bad("':'"); p++;
                j.obj.emplace_back(k.str, value());
                ws(); if (p < e && *p == ',') { p++; continue; }
                if (p < e && *p == '}') { p++; return j; }
                bad("'}'");
            }
        } else if (*p == '[') {
            j.kind = Json::Arr; p++; ws();
            if (p < e && *p == ']') { p++; return j; }
            for (;;) {
                j.arr.push_back(value());
                ws(); if (p < e && *p == ',') { p++; continue; }
                if (p < e && *p == ']') { p++; return j; }
                bad("']'");
            }
        } else if (*p == '"') {
            j.kind = Json::Str; p++;
            while (p < e && *p != '"') {
                if (*p == '\\' && p + 1 < e) {
                    p++;
                    switch (*p) {
                    case 'n': j.str += '\n'; break; case 't': j.str += '\t'; break; case 'r': j.str += '\r'; break;
                    case 'b': j.str += '\b'; break; case 'f': j.str += '\f'; break;
                    case 'u': { if (p + 4 >= e) bad("\\u"); unsigned c = (unsigned)strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16); j.str += (c < 128) ? (char)c : '?'; p += 4; break; }
                    default: j.str += *p;
                    }
                    p++;
                } else j.str += *p++;
            }
            if (p >= e) bad("unterminated string");
            p++;
            return j;
        } else if (!strncmp(p, "true", 4)) { j.kind = Json::Bool; j.b = true; p += 4; return j; }
        else if (!strncmp(p, "false", 5)) { j.kind = Json::Bool; p += 5; return j; }
        else if (!strncmp(p, "null", 4)) { p += 4; return j; }
        else {
            char* end = nullptr;
            j.kind = Json::Num; j.num = strtod(p, &end);
            if (end == p) bad("value");
            p = end;
            return j;
        }
    }
};

Json parse_json(const std::string& s) { JsonParser jp{s.data(), s.data() + s.size()}; return jp.value(); }

// ---- glTF accessors --------------------------------------------------------------------
struct Accessor { const unsigned char* base; size_t stride, count; int comp, ncomp; };

Accessor accessor(const Json& g, const std::vector<std::string>& buffers, size_t k)
{
    const Json& a = g.at("accessors").arr.at(k);
    const long long j_view = a.integer("bufferView", 0);
    const Json& bv = g.at("bufferViews").arr.at(j_view < 0 ? (size_t)-1 : (size_t)j_view);
    int comp = (int)a.integer("componentType", 0);
    const std::string& type = a.at("type").str;
    int nc = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : type == "MAT4" ? 16 : 0;
    size_t sz = (comp == 5120 || comp == 5121) ? 1 : (comp == 5122 || comp == 5123) ? 2 : (comp == 5125 || comp == 5126) ? 4 : 0;
    if (!nc || !sz) throw std::runtime_error("glTF: unsupported accessor type");
    const long long j_stride = bv.integer("byteStride", 0), j_bo = bv.integer("byteOffset", 0), j_ao = a.integer("byteOffset", 0),
                    j_count = a.integer("count", 0), j_buf = bv.integer("buffer", 0);
    if (j_stride < 0 || j_bo < 0 || j_ao < 0 || j_count < 0 || j_buf < 0 || j_stride > (1 << 20)) throw std::runtime_error("glTF: negative or absurd accessor field");
    const size_t elem = sz * (size_t)nc;
    size_t stride = (size_t)j_stride;
    if (!stride) stride = elem;
    const size_t count = (size_t)j_count;
    const std::string& buf = buffers.at((size_t)j_buf);
    // overflow-safe: off <= size, elem <= size - off, (count - 1) * stride <= size - off - elem
    if ((unsigned long long)j_bo > buf.size() || (unsigned long long)j_ao > buf.size() - (size_t)j_bo) throw std::runtime_error("glTF: accessor outside its buffer");
    const size_t off = (size_t)j_bo + (size_t)j_ao;
    if (count && (elem > buf.size() - off || (count - 1) > (buf.size() - off - elem) / stride)) throw std::runtime_error("glTF: accessor outside its buffer");
    return Accessor{(const unsigned char*)buf.data() + off, stride, count, comp, nc};
}

void append_primitive(const Json& g, const std::vector<std::string>& buffers, const Json& prim, bool scale_by_mesh_count,
                      std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes)
{
    const Json& attrs = prim.at("attributes");
    const Json* jp = attrs.get("POSITION"); const Json* jn = attrs.get("NORMAL");
    if (!jp || !jn) throw std::runtime_error("glTF: primitive without POSITION/NORMAL (the reference unwraps both)");  // resource.rs:149-150
    Accessor ap = accessor(g, buffers, jp->index()), an = accessor(g, buffers, jn->index());
    if (ap.comp != 5126 || an.comp != 5126 || ap.ncomp != 3 || an.ncomp != 3) throw std::runtime_error("glTF: POSITION/NORMAL must be f32 VEC3");
    std::vector<uint32_t> idx;
    if (const Json* ji = prim.get("indices")) {
        Accessor ai = accessor(g, buffers, ji->index());
        idx.resize(ai.count);
        for (size_t r = 0; r < ai.count; r++) {
            const unsigned char* q = ai.base + r * ai.stride;
            if (ai.comp == 5121) idx[r] = *q;
            else if (ai.comp == 5123) { uint16_t v; memcpy(&v, q, 2); idx[r] = v; }
            else if (ai.comp == 5125) { uint32_t v; memcpy(&v, q, 4); idx[r] = v; }
            else throw std::runtime_error("glTF: index component type");
        }
    } else {
        idx.resize(ap.count);
        for (size_t r = 0; r < ap.count; r++) idx[r] = (uint32_t)r;
    }
    meshes.push_back(loader_mesh((uint32_t)meshes.size(), (uint32_t)idx.size(), (uint32_t)indices.size(), (uint32_t)vertices.size()));
    const float scale = (float)meshes.size();                     // resource.rs:180 (after the push)
    size_t n = ap.count < an.count ? ap.count : an.count;         // positions.zip(normals)
    Accessor at{nullptr, 0, 0, 0, 0};
    if (g_keep_uv)
        if (const Json* jt = attrs.get("TEXCOORD_0")) {
            at = accessor(g, buffers, jt->index());
            if (at.comp != 5126 || at.ncomp != 2) at.base = nullptr;     // only f32 VEC2 coordinates are kept
        }
    for (size_t r = 0; r < n; r++) {
        rt_vertex v{};
        memcpy(v.pos, ap.base + r * ap.stride, 12);
        memcpy(v.normal, an.base + r * an.stride, 12);
        if (at.base && r < at.count) { memcpy(&v.tu, at.base + r * at.stride, 4); memcpy(&v.tv, at.base + r * at.stride + 4, 4); }
        if (scale_by_mesh_count) { v.pos[0] *= scale; v.pos[1] *= scale; v.pos[2] *= scale; }
        vertices.push_back(v);
    }
    indices.insert(indices.end(), idx.begin(), idx.end());
}

} // namespace

rt_mesh loader_mesh(uint32_t mesh_count, uint32_t n_indices, uint32_t first, uint32_t offset)
{
    rt_mesh m{};
    m.triangles = n_indices / 3; m.first = first; m.offset = offset;
    m.pos[0] = (float)mesh_count * 3.0f;
    m.color[0] = 0.2f; m.color[1] = 0.2f; m.color[2] = 1.0f; m.color[3] = 1.0f;
    m.specular = 0.5f;
    return m;
}

void load_model(const std::string& path, std::vector<rt_vertex>& v, std::vector<uint32_t>& i, std::vector<rt_mesh>& m)
{
    if (ends_with(path, ".obj")) load_model_obj(path, v, i, m);
    else if (ends_with(path, ".gltf")) load_model_gltf(path, v, i, m);
    else if (ends_with(path, ".glb")) load_model_glb(path, v, i, m);
    else throw std::runtime_error("Unsupported model format");     // resource.rs:43
}

namespace {
// Wavefront MTL, the statements the scene import uses (extension: upstream never opens the .mtl, resource.rs:57-69
// drops tobj's materials): Kd, Ke, Ns, map_Kd, map_Bump / bump / norm.
struct ObjMaterial { std::string name; float kd[3] = {0.8f, 0.8f, 0.8f}, ke[3] = {0, 0, 0}, ns = 0.0f; std::string map_kd, map_bump; };
std::vector<ObjMaterial> parse_mtl(const std::string& path)
{
    std::vector<ObjMaterial> out;
    std::istringstream in(read_file(path, false));
    std::string line;
    while (std::getline(in, line)) {
        while (!line.empty() && (line.back() == '\r' || line.back() == ' ')) line.pop_back();
        std::istringstream ls(line);
        std::string tag;
        if (!(ls >> tag) || tag[0] == '#') continue;
        if (tag == "newmtl") { out.emplace_back(); ls >> out.back().name; continue; }
        if (out.empty()) continue;
        ObjMaterial& m = out.back();
        if (tag == "Kd") ls >> m.kd[0] >> m.kd[1] >> m.kd[2];
        else if (tag == "Ke") ls >> m.ke[0] >> m.ke[1] >> m.ke[2];
        else if (tag == "Ns") ls >> m.ns;
        else if (tag == "map_Kd" || tag == "map_Bump" || tag == "map_bump" || tag == "bump" || tag == "norm") {
            std::string tok, last;                                    // options (-bm 1.0, -s ...) come first, the file name last
            while (ls >> tok) last = tok;
            (tag == "map_Kd" ? m.map_kd : m.map_bump) = last;
        }
    }
    return out;
}
struct ObjScene { std::vector<ObjMaterial> materials; std::vector<std::string> images; };   // images in handle order (handle = index + 1)

void load_obj(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes, ObjScene* scene)
{
    std::istringstream in(read_file(path, false));
    std::vector<float> pos, nrm, tex;
    struct Model { std::map<std::tuple<long, long, long>, uint32_t> map; std::vector<rt_vertex> v; std::vector<uint32_t> idx; std::string mtl; };
    std::vector<Model> models;
    std::unique_ptr<Model> cur;
    std::string cur_mtl;
    auto flush = [&]() { if (cur && !cur->idx.empty()) models.push_back(std::move(*cur)); cur.reset(); };
    auto fix = [](long i, size_t n) -> long { return i > 0 ? i - 1 : (long)n + i; };
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        std::string tag;
        if (!(ls >> tag) || tag[0] == '#') continue;
        if (tag == "v") { float x, y, z; ls >> x >> y >> z; pos.insert(pos.end(), {x, y, z}); }
        else if (tag == "vn") { float x, y, z; ls >> x >> y >> z; nrm.insert(nrm.end(), {x, y, z}); }
        else if (tag == "vt") { float x = 0, y = 0; ls >> x >> y; tex.insert(tex.end(), {x, y}); }
        else if (tag == "o" || tag == "g") { flush(); cur.reset(new Model()); cur->mtl = cur_mtl; }
        else if (scene && tag == "mtllib") { std::string f; ls >> f; for (ObjMaterial& m : parse_mtl(sibling(path, f))) scene->materials.push_back(std::move(m)); }
        else if (scene && tag == "usemtl") { ls >> cur_mtl; if (cur && !cur->idx.empty()) { flush(); } if (!cur) cur.reset(new Model()); cur->mtl = cur_mtl; }
        else if (tag == "f") {
            if (!cur) { cur.reset(new Model()); cur->mtl = cur_mtl; }
            std::vector<uint32_t> corner;
            std::string c;
            while (ls >> c) {
                long vi = 0, ti = 0, ni = 0; bool ht = false, hn = false;
                size_t s1 = c.find('/');
                vi = atol(c.substr(0, s1).c_str());
                if (s1 != std::string::npos) {
                    size_t s2 = c.find('/', s1 + 1);
                    std::string t = c.substr(s1 + 1, s2 == std::string::npos ? std::string::npos : s2 - s1 - 1);
                    if (!t.empty()) { ti = atol(t.c_str()); ht = true; }
                    if (s2 != std::string::npos && s2 + 1 < c.size()) { ni = atol(c.substr(s2 + 1).c_str()); hn = true; }
                }
                long v0 = fix(vi, pos.size() / 3), t0 = ht ? fix(ti, tex.size() / 2) : -1, n0 = hn ? fix(ni, nrm.size() / 3) : -1;
                if (v0 < 0 || (size_t)v0 >= pos.size() / 3) throw std::runtime_error("OBJ: vertex index out of range");
                if (n0 < 0 || (size_t)n0 >= nrm.size() / 3) throw std::runtime_error("OBJ: face without a valid normal (the reference indexes normals unconditionally)");  // resource.rs:96-100
                auto key = std::make_tuple(v0, t0, n0);
                auto it = cur->map.find(key);
                if (it == cur->map.end()) {
                    rt_vertex vv{};
                    memcpy(vv.pos, &pos[(size_t)v0 * 3], 12); memcpy(vv.normal, &nrm[(size_t)n0 * 3], 12);
                    if (g_keep_uv) {                             // extension: (u, 1 - v), image row 0 = top row
                        bool ok = t0 >= 0 && (size_t)t0 < tex.size() / 2;
                        vv.tu = ok ? tex[(size_t)t0 * 2] : 0.0f;
                        vv.tv = 1.0f - (ok ? tex[(size_t)t0 * 2 + 1] : 1.0f);
                    }
                    it = cur->map.emplace(key, (uint32_t)cur->v.size()).first;
                    cur->v.push_back(vv);
                }
                corner.push_back(it->second);
            }
            for (size_t k = 1; k + 1 < corner.size(); k++) { cur->idx.push_back(corner[0]); cur->idx.push_back(corner[k]); cur->idx.push_back(corner[k + 1]); }
        }
    }
    flush();
    auto image_handle = [&](const std::string& file) -> uint32_t {
        if (file.empty()) return 0u;
        for (size_t k = 0; k < scene->images.size(); k++) if (scene->images[k] == file) return (uint32_t)k + 1u;
        scene->images.push_back(file);
        return (uint32_t)scene->images.size();
    };
    for (auto& mdl : models) {
        rt_mesh m = loader_mesh((uint32_t)meshes.size(), (uint32_t)mdl.idx.size(), (uint32_t)indices.size(), (uint32_t)vertices.size());
        if (scene) {                                                 // material import: in place (pos 0), colours and textures from the .mtl
            m.pos[0] = 0.0f;
            const ObjMaterial* mat = nullptr;
            for (const ObjMaterial& c : scene->materials) if (c.name == mdl.mtl) mat = &c;
            if (mat) {
                m.color[0] = mat->kd[0]; m.color[1] = mat->kd[1]; m.color[2] = mat->kd[2]; m.color[3] = 1.0f;
                const float e = std::max(mat->ke[0], std::max(mat->ke[1], mat->ke[2]));
                if (e > 0.0f) { m.emission_color[0] = mat->ke[0]; m.emission_color[1] = mat->ke[1]; m.emission_color[2] = mat->ke[2]; m.emission_color[3] = 1.0f; m.emission_strength = 1.0f; }
                const float sm = std::sqrt(std::max(mat->ns, 0.0f) / 900.0f);          // Blender writes Ns = 900 (1 - roughness)^2
                m.specular = sm > 1.0f ? 1.0f : sm;
                m.diffuse_tex = image_handle(mat->map_kd); m.normal_tex = image_handle(mat->map_bump);
            }
        }
        meshes.push_back(m);
        vertices.insert(vertices.end(), mdl.v.begin(), mdl.v.end());
        indices.insert(indices.end(), mdl.idx.begin(), mdl.idx.end());
    }
}
} // namespace

void load_model_obj(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes)
{
    load_obj(path, vertices, indices, meshes, nullptr);
}

namespace {
// GLB container: JSON chunk + first BIN chunk; buffers[] resolved like resource.rs:203-216
void read_glb(const std::string& path, Json& g, std::vector<std::string>& buffers)
{
    std::string data = read_file(path, true);
    if (data.size() < 20 || memcmp(data.data(), "glTF", 4) != 0) throw std::runtime_error("GLB: bad magic");
    std::string json, blob; bool have_blob = false;
    size_t off = 12;
    while (off + 8 <= data.size()) {
        uint32_t clen, ctype; memcpy(&clen, &data[off], 4); memcpy(&ctype, &data[off + 4], 4);
        if (off + 8 + clen > data.size()) throw std::runtime_error("GLB: truncated chunk");
        if (ctype == 0x4E4F534Au) json.assign(data, off + 8, clen);
        else if (ctype == 0x004E4942u && !have_blob) { blob.assign(data, off + 8, clen); have_blob = true; }
        off += 8 + (size_t)clen;
    }
    g = parse_json(json);
    if (const Json* jb = g.get("buffers"))
        for (const Json& b : jb->arr) {
            if (const Json* uri = b.get("uri")) buffers.push_back(read_file(sibling(path, uri->str), true));
            else { if (!have_blob) throw std::runtime_error("Missing blob"); buffers.push_back(blob); }   // resource.rs:211-216
        }
}
} // namespace

void load_model_glb(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes)
{
    Json g; std::vector<std::string> buffers;
    read_glb(path, g, buffers);
    if (const Json* jm = g.get("meshes"))
        for (const Json& mesh : jm->arr)
            for (const Json& prim : mesh.at("primitives").arr) append_primitive(g, buffers, prim, false, vertices, indices, meshes);
}

// ---- scene-graph import (extension; upstream ignores node transforms and materials, resource.rs:228-276) ----
namespace {
struct Mat4 { float m[16]; };                                     // column-major, as glTF stores matrices
Mat4 mat_identity() { Mat4 r{}; r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f; return r; }
Mat4 mat_mul(const Mat4& a, const Mat4& b)
{
    Mat4 r{};
    for (int c = 0; c < 4; c++)
        for (int rw = 0; rw < 4; rw++) {
            float acc = 0.0f;
            for (int k = 0; k < 4; k++) acc += a.m[k * 4 + rw] * b.m[c * 4 + k];
            r.m[c * 4 + rw] = acc;
        }
    return r;
}
float num_or(const Json* j, size_t k, float d) { return (j && k < j->arr.size()) ? (float)j->arr[k].num : d; }
Mat4 node_local(const Json& node)
{
    if (const Json* jm = node.get("matrix")) {
        Mat4 r{};
        for (size_t k = 0; k < 16; k++) r.m[k] = num_or(jm, k, (k % 5 == 0) ? 1.0f : 0.0f);
        return r;
    }
    const Json *jt = node.get("translation"), *jr = node.get("rotation"), *js = node.get("scale");
    const float tx = num_or(jt, 0, 0.f), ty = num_or(jt, 1, 0.f), tz = num_or(jt, 2, 0.f);
    const float x = num_or(jr, 0, 0.f), y = num_or(jr, 1, 0.f), z = num_or(jr, 2, 0.f), w = num_or(jr, 3, 1.f);
    const float sx = num_or(js, 0, 1.f), sy = num_or(js, 1, 1.f), sz = num_or(js, 2, 1.f);
    Mat4 r{};                                                      // T * R * S
    r.m[0] = (1.f - 2.f * (y * y + z * z)) * sx; r.m[1] = (2.f * (x * y + z * w)) * sx;       r.m[2] = (2.f * (x * z - y * w)) * sx;
    r.m[4] = (2.f * (x * y - z * w)) * sy;       r.m[5] = (1.f - 2.f * (x * x + z * z)) * sy; r.m[6] = (2.f * (y * z + x * w)) * sy;
    r.m[8] = (2.f * (x * z + y * w)) * sz;       r.m[9] = (2.f * (y * z - x * w)) * sz;       r.m[10] = (1.f - 2.f * (x * x + y * y)) * sz;
    r.m[12] = tx; r.m[13] = ty; r.m[14] = tz; r.m[15] = 1.f;
    return r;
}
// material -> Mesh record: base colour factor, smoothness = 1 - roughness, emissive factor, texture handles
// (handle = image index + 1, the order load_glb_images returns them in)
void apply_material(const Json& g, const Json& prim, rt_mesh& m)
{
    const Json* jmi = prim.get("material");
    const Json* mats = g.get("materials");
    if (!jmi || !mats || jmi->index() >= mats->arr.size()) return;
    const Json& mat = mats->arr[jmi->index()];
    auto image_handle = [&](const Json* texref) -> uint32_t {
        if (!texref) return 0u;
        const Json* ji = texref->get("index"); const Json* texs = g.get("textures");
        if (!ji || !texs || ji->index() >= texs->arr.size()) return 0u;
        const Json* src = texs->arr[ji->index()].get("source");
        return (src && src->index() < 0xffffffu) ? (uint32_t)src->index() + 1u : 0u;
    };
    if (const Json* pbr = mat.get("pbrMetallicRoughness")) {
        if (const Json* bc = pbr->get("baseColorFactor")) for (size_t k = 0; k < 4; k++) m.color[k] = num_or(bc, k, 1.0f);
        else { m.color[0] = m.color[1] = m.color[2] = m.color[3] = 1.0f; }
        const Json* jr = pbr->get("roughnessFactor");
        float smooth = 1.0f - (jr ? (float)jr->num : 1.0f);
        m.specular = smooth < 0.0f ? 0.0f : (smooth > 1.0f ? 1.0f : smooth);
        m.diffuse_tex = image_handle(pbr->get("baseColorTexture"));
    }
    m.normal_tex = image_handle(mat.get("normalTexture"));
    if (const Json* je = mat.get("emissiveFactor")) {
        const float e0 = num_or(je, 0, 0.f), e1 = num_or(je, 1, 0.f), e2 = num_or(je, 2, 0.f);
        if (e0 != 0.f || e1 != 0.f || e2 != 0.f) { m.emission_color[0] = e0; m.emission_color[1] = e1; m.emission_color[2] = e2; m.emission_color[3] = 1.0f; m.emission_strength = 1.0f; }
    }
}
void visit_node(const Json& g, const std::vector<std::string>& buffers, size_t ni, const Mat4& parent, int depth,
                std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes)
{
    const Json& nodes = g.at("nodes");
    if (ni >= nodes.arr.size() || depth > 256) throw std::runtime_error("GLB: bad node reference");
    const Json& node = nodes.arr[ni];
    const Mat4 world = mat_mul(parent, node_local(node));
    if (const Json* jm = node.get("mesh")) {
        // inverse transpose of the upper 3x3 for normals = cofactor matrix / det; the normalisation removes the scale
        const float* a = world.m;
        const float c00 = a[5] * a[10] - a[6] * a[9], c01 = a[6] * a[8] - a[4] * a[10], c02 = a[4] * a[9] - a[5] * a[8];
        const float c10 = a[2] * a[9] - a[1] * a[10], c11 = a[0] * a[10] - a[2] * a[8], c12 = a[1] * a[8] - a[0] * a[9];
        const float c20 = a[1] * a[6] - a[2] * a[5], c21 = a[2] * a[4] - a[0] * a[6], c22 = a[0] * a[5] - a[1] * a[4];
        const float det = a[0] * c00 + a[1] * c01 + a[2] * c02;
        for (const Json& prim : g.at("meshes").arr.at(jm->index()).at("primitives").arr) {
            const size_t v0 = vertices.size();
            append_primitive(g, buffers, prim, false, vertices, indices, meshes);
            rt_mesh& m = meshes.back();
            m.pos[0] = m.pos[1] = m.pos[2] = 0.0f;                 // the node transform places the mesh, not resource.rs:75's (3k, 0, 0)
            apply_material(g, prim, m);
            for (size_t k = v0; k < vertices.size(); k++) {
                rt_vertex& v = vertices[k];
                const float px = v.pos[0], py = v.pos[1], pz = v.pos[2];
                v.pos[0] = a[0] * px + a[4] * py + a[8] * pz + a[12];
                v.pos[1] = a[1] * px + a[5] * py + a[9] * pz + a[13];
                v.pos[2] = a[2] * px + a[6] * py + a[10] * pz + a[14];
                if (det != 0.0f) {
                    const float nx = v.normal[0], ny = v.normal[1], nz = v.normal[2];
                    float qx = c00 * nx + c10 * ny + c20 * nz, qy = c01 * nx + c11 * ny + c21 * nz, qz = c02 * nx + c12 * ny + c22 * nz;
                    if (det < 0.0f) { qx = -qx; qy = -qy; qz = -qz; }
                    const float len = sqrtf(qx * qx + qy * qy + qz * qz);
                    if (len > 0.0f) { v.normal[0] = qx / len; v.normal[1] = qy / len; v.normal[2] = qz / len; }
                }
            }
            if (det < 0.0f)                                        // a mirroring transform flips the winding the one-sided test relies on
                for (size_t t = m.first; t + 2 < m.first + 3 * (size_t)m.triangles; t += 3) std::swap(indices[t + 1], indices[t + 2]);
        }
    }
    if (const Json* jc = node.get("children"))
        for (const Json& c : jc->arr) visit_node(g, buffers, c.index(), world, depth + 1, vertices, indices, meshes);
}
} // namespace

void load_scene_glb(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes)
{
    struct Guard { Guard() { g_keep_uv = true; } ~Guard() { g_keep_uv = false; } } guard;
    if (ends_with(path, ".obj")) { ObjScene sc; load_obj(path, vertices, indices, meshes, &sc); return; }   // OBJ + MTL material import
    Json g; std::vector<std::string> buffers;
    if (ends_with(path, ".gltf")) {                                // JSON file + external buffers or base64 data: URIs
        g = parse_json(read_file(path, false));
        if (const Json* jb = g.get("buffers"))
            for (const Json& b : jb->arr) { const Json* uri = b.get("uri"); buffers.push_back(uri ? resolve_uri(path, uri->str) : std::string()); }
    } else read_glb(path, g, buffers);
    const Json* scenes = g.get("scenes");
    if (!scenes || scenes->arr.empty()) {                          // no scene graph: every mesh once, untransformed
        if (const Json* jm = g.get("meshes"))
            for (const Json& mesh : jm->arr)
                for (const Json& prim : mesh.at("primitives").arr) append_primitive(g, buffers, prim, false, vertices, indices, meshes);
        return;
    }
    const Json* def = g.get("scene");
    const size_t si = def ? def->index() : 0;
    if (si >= scenes->arr.size()) throw std::runtime_error("GLB: bad default scene");
    if (const Json* roots = scenes->arr[si].get("nodes"))
        for (const Json& r : roots->arr) visit_node(g, buffers, r.index(), mat_identity(), 0, vertices, indices, meshes);
}

void load_model_gltf(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices, std::vector<rt_mesh>& meshes)
{
    Json g = parse_json(read_file(path, false));
    std::vector<std::string> buffers;
    if (const Json* jb = g.get("buffers"))
        for (const Json& b : jb->arr) {
            const Json* uri = b.get("uri");
            buffers.push_back(uri ? read_file(sibling(path, uri->str), true) : std::string());   // resource.rs:123-133
        }
    if (const Json* js = g.get("scenes"))
        for (const Json& sc : js->arr)
            for (const Json& ni : sc.at("nodes").arr) {
                const Json& node = g.at("nodes").arr.at(ni.index());
                const Json* jm = node.get("mesh");
                if (!jm) throw std::runtime_error("Failed to get mesh");   // resource.rs:142
                for (const Json& prim : g.at("meshes").arr.at(jm->index()).at("primitives").arr)
                    append_primitive(g, buffers, prim, true, vertices, indices, meshes);
            }
}

// ---- texture extension: UV-keeping load, PNG decode / encode, GLB-embedded images ---------
void load_model_uv(const std::string& path, std::vector<rt_vertex>& v, std::vector<uint32_t>& i, std::vector<rt_mesh>& m)
{
    struct Guard { Guard() { g_keep_uv = true; } ~Guard() { g_keep_uv = false; } } guard;
    load_model(path, v, i, m);
}

namespace {
uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
int paeth(int a, int b, int c) { int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c); return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); }
}

Image decode_png(const uint8_t* d, size_t size)
{
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (size < 8 || memcmp(d, sig, 8) != 0) throw std::runtime_error("PNG: bad signature");
    uint32_t W = 0, H = 0; int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    for (size_t off = 8; off + 12 <= size;) {
        uint32_t len = be32(d + off);
        if (off + 12 + (size_t)len > size) throw std::runtime_error("PNG: truncated chunk");
        const uint8_t* body = d + off + 8;
        if (!memcmp(d + off + 4, "IHDR", 4) && len >= 13) { W = be32(body); H = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12]; }
        else if (!memcmp(d + off + 4, "PLTE", 4)) plte.assign(body, body + len);
        else if (!memcmp(d + off + 4, "tRNS", 4)) trns.assign(body, body + len);
        else if (!memcmp(d + off + 4, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!memcmp(d + off + 4, "IEND", 4)) break;
        off += 12 + (size_t)len;
    }
    if (!W || !H || W > 32768 || H > 32768) throw std::runtime_error("PNG: missing or absurd IHDR");
    if (interlace > 1) throw std::runtime_error("PNG: unknown interlace method");
    int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    const bool small_ok = (ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4);
    if (!ch || !(depth == 8 || (depth == 16 && ctype != 3) || small_ok)) throw std::runtime_error("PNG: unsupported colour type / bit depth");
    const size_t bits = (size_t)ch * depth, bpp = bits >= 8 ? bits / 8 : 1;          // filter distance in bytes
    // one pass, or the seven passes of Adam7: (x0, y0, dx, dy)
    static const int adam7[7][4] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
    static const int whole[1][4] = {{0, 0, 1, 1}};
    const int (*pass)[4] = interlace ? adam7 : whole;
    const int n_pass = interlace ? 7 : 1;
    size_t raw_size = 0;
    for (int k = 0; k < n_pass; k++) {
        if ((uint32_t)pass[k][0] >= W || (uint32_t)pass[k][1] >= H) continue;
        const size_t pw = (W - pass[k][0] + pass[k][2] - 1) / pass[k][2], ph = (H - pass[k][1] + pass[k][3] - 1) / pass[k][3];
        raw_size += ph * ((pw * bits + 7) / 8 + 1);
    }
    std::vector<uint8_t> raw(raw_size);
    uLongf out_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size()) throw std::runtime_error("PNG: inflate failed");
    Image img; img.width = W; img.height = H; img.rgba.resize((size_t)W * H * 4);
    const size_t st = depth >= 8 ? depth / 8 : 0;                                    // 16-bit samples: keep the high byte
    const int grey_scale = depth < 8 ? 255 / ((1 << depth) - 1) : 1;
    size_t pos = 0;
    for (int k = 0; k < n_pass; k++) {
        if ((uint32_t)pass[k][0] >= W || (uint32_t)pass[k][1] >= H) continue;
        const size_t pw = (W - pass[k][0] + pass[k][2] - 1) / pass[k][2], ph = (H - pass[k][1] + pass[k][3] - 1) / pass[k][3];
        const size_t stride = (pw * bits + 7) / 8;
        std::vector<uint8_t> prev(stride, 0), cur(stride);
        for (size_t py = 0; py < ph; py++) {
            const uint8_t* row = &raw[pos]; pos += stride + 1;
            const int f = row[0];
            if (f > 4) throw std::runtime_error("PNG: bad filter type");
            for (size_t i = 0; i < stride; i++) {
                int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0, x = row[1 + i];
                cur[i] = (uint8_t)(f == 0 ? x : f == 1 ? x + a : f == 2 ? x + b : f == 3 ? x + ((a + b) >> 1) : x + paeth(a, b, c));
            }
            const size_t y = pass[k][1] + py * pass[k][3];
            for (size_t px = 0; px < pw; px++) {
                const size_t xx = pass[k][0] + px * pass[k][2];
                uint8_t* o = &img.rgba[(y * W + xx) * 4];
                uint8_t small = 0;
                const uint8_t* sp = &cur[0];
                if (depth < 8) { const size_t bit = px * depth; small = (uint8_t)((cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1)); }
                else sp = &cur[px * bpp];
                if (ctype == 0) { o[0] = o[1] = o[2] = depth < 8 ? (uint8_t)(small * grey_scale) : sp[0]; o[3] = 255; }
                else if (ctype == 2) { o[0] = sp[0]; o[1] = sp[st]; o[2] = sp[2 * st]; o[3] = 255; }
                else if (ctype == 3) { size_t q = depth < 8 ? small : sp[0]; if (3 * q + 2 >= plte.size()) throw std::runtime_error("PNG: palette index out of range");
                                       o[0] = plte[3 * q]; o[1] = plte[3 * q + 1]; o[2] = plte[3 * q + 2]; o[3] = q < trns.size() ? trns[q] : 255; }
                else if (ctype == 4) { o[0] = o[1] = o[2] = sp[0]; o[3] = sp[st]; }
                else { o[0] = sp[0]; o[1] = sp[st]; o[2] = sp[2 * st]; o[3] = sp[3 * st]; }
            }
            prev.swap(cur);
        }
    }
    return img;
}

// ---- JPEG (ITU T.81 DCT-based, Huffman, 8 bit): baseline / extended sequential (SOF0, SOF1) and progressive (SOF2:
// spectral selection + successive approximation, Annex G), grey or YCbCr with any sampling factors, interleaved and
// non-interleaved scans, restart intervals.  Arithmetic coding, lossless and 12-bit files are refused.  All scans are
// gathered into coefficient arrays first; chroma is upsampled bilinearly (the triangle filter libjpeg calls "fancy
// upsampling"), the inverse DCT is evaluated in float. --------------------------------------------------------------
namespace {
struct JpegHuff { uint16_t code_first[17]; int val_first[17]; int count[17]; uint8_t vals[256]; bool set = false; };
struct JpegBits {
    const uint8_t* p; const uint8_t* e; uint32_t acc = 0; int n = 0;
    int bit()
    {
        if (n == 0) {
            uint8_t b = p < e ? *p++ : 0;
            if (b == 0xFF) { uint8_t b2 = p < e ? *p : 0; if (b2 == 0) p++; else { p--; b = 0; } }   // stuffed zero, or a marker: feed zeros
            acc = b; n = 8;
        }
        return (acc >> --n) & 1;
    }
    int bits(int k) { int v = 0; while (k--) v = (v << 1) | bit(); return v; }
    void reset() { n = 0; }
};
int jpeg_decode(JpegBits& br, const JpegHuff& h)
{
    int code = 0;
    for (int len = 1; len <= 16; len++) {
        code = (code << 1) | br.bit();
        if (h.count[len] && code >= h.code_first[len] && code < h.code_first[len] + h.count[len]) return h.vals[h.val_first[len] + code - h.code_first[len]];
    }
    throw std::runtime_error("JPEG: bad Huffman code");
}
int jpeg_extend(int v, int t) { return (t && v < (1 << (t - 1))) ? v - (1 << t) + 1 : v; }
const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                             35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
void jpeg_idct(const float* in, uint8_t* out, size_t stride)
{
    static float c[8][8]; static bool init = false;
    if (!init) { for (int x = 0; x < 8; x++) for (int u = 0; u < 8; u++) c[x][u] = (u == 0 ? 0.35355339059327379f : 0.5f) * (float)cos((2 * x + 1) * u * 3.14159265358979323846 / 16.0); init = true; }
    float tmp[64];
    for (int v = 0; v < 8; v++) for (int x = 0; x < 8; x++) { float a = 0; for (int u = 0; u < 8; u++) a += c[x][u] * in[v * 8 + u]; tmp[v * 8 + x] = a; }
    for (int x = 0; x < 8; x++) for (int y = 0; y < 8; y++) {
        float a = 0; for (int v = 0; v < 8; v++) a += c[y][v] * tmp[v * 8 + x];
        float r = floorf(a + 128.5f);
        out[y * stride + x] = (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
    }
}
struct JpegComp {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0, pred = 0;
    size_t w = 0, hgt = 0;            // padded plane size (whole MCUs)
    size_t bw = 0, bh = 0;            // blocks that carry image samples (the block grid of a non-interleaved scan, T.81 A.2.2)
    size_t bstride = 0;               // blocks per padded row
    std::vector<int16_t> coef;        // 64 per block, zigzag order, as decoded (not dequantised)
    std::vector<uint8_t> px;
};
// one block of one scan into `blk` (zigzag order).  progressive == false: the sequential procedure of F.2.2.
void jpeg_block(JpegBits& br, JpegComp& c, const JpegHuff* hdc, const JpegHuff* hac, int16_t* blk, bool progressive,
                int ss, int se, int ah, int al, int& eobrun)
{
    if (!progressive) {
        const int t = jpeg_decode(br, hdc[c.td]);
        if (t > 11) throw std::runtime_error("JPEG: bad DC size");
        c.pred += jpeg_extend(br.bits(t), t);
        blk[0] = (int16_t)c.pred;
        for (int z = 1; z < 64;) {
            const int rs = jpeg_decode(br, hac[c.ta]), r = rs >> 4, sz = rs & 15;
            if (sz == 0) { if (r == 15) { z += 16; continue; } break; }
            z += r;
            if (z > 63) throw std::runtime_error("JPEG: bad AC run");
            blk[z] = (int16_t)jpeg_extend(br.bits(sz), sz);
            z++;
        }
        return;
    }
    if (ss == 0) {                                                   // DC scan (G.1.2.1)
        if (ah == 0) {
            const int t = jpeg_decode(br, hdc[c.td]);
            if (t > 11) throw std::runtime_error("JPEG: bad DC size");
            c.pred += jpeg_extend(br.bits(t), t);
            blk[0] = (int16_t)(c.pred * (1 << al));
        } else if (br.bit()) blk[0] = (int16_t)(blk[0] | (1 << al));
        return;
    }
    const JpegHuff& h = hac[c.ta];
    if (ah == 0) {                                                   // AC first pass (G.1.2.2)
        if (eobrun > 0) { eobrun--; return; }
        for (int k = ss; k <= se; k++) {
            const int rs = jpeg_decode(br, h), r = rs >> 4, sz = rs & 15;
            if (sz == 0) {
                if (r < 15) { eobrun = (1 << r) - 1; if (r) eobrun += br.bits(r); break; }
                k += 15;
            } else {
                k += r;
                if (k > se) throw std::runtime_error("JPEG: bad AC run");
                blk[k] = (int16_t)(jpeg_extend(br.bits(sz), sz) * (1 << al));
            }
        }
        return;
    }
    // AC refinement (G.1.2.3): correction bits for the coefficients that are already non-zero, new +-1 << al values between them
    const int p1 = 1 << al, m1 = -(1 << al);
    auto refine = [&](int16_t& v) { if (br.bit() && (v & p1) == 0) v = (int16_t)(v + (v >= 0 ? p1 : m1)); };
    int k = ss;
    if (eobrun == 0) {
        for (; k <= se; k++) {
            const int rs = jpeg_decode(br, h);
            int r = rs >> 4; const int sz = rs & 15;
            int value = 0;
            if (sz == 0) {
                if (r < 15) { eobrun = 1 << r; if (r) eobrun += br.bits(r); break; }
            } else {
                if (sz != 1) throw std::runtime_error("JPEG: bad refinement code");
                value = br.bit() ? p1 : m1;
            }
            for (; k <= se; k++) {
                if (blk[k] != 0) refine(blk[k]);
                else if (--r < 0) break;
            }
            if (value) { if (k > se) throw std::runtime_error("JPEG: bad refinement run"); blk[k] = (int16_t)value; }
        }
    }
    if (eobrun > 0) {
        for (; k <= se; k++) if (blk[k] != 0) refine(blk[k]);
        eobrun--;
    }
}
} // namespace

Image decode_jpeg(const uint8_t* d, size_t size)
{
    if (size < 4 || d[0] != 0xFF || d[1] != 0xD8) throw std::runtime_error("JPEG: bad signature");
    uint16_t qt[4][64] = {}; JpegHuff hdc[4], hac[4];
    JpegComp comp[3];
    int ncomp = 0, W = 0, H = 0, hmax = 1, vmax = 1, restart = 0;
    bool progressive = false, any_scan = false;
    size_t mcux = 0, mcuy = 0;
    size_t i = 2;
    for (;;) {
        if (i + 2 <= size && d[i] == 0xFF && d[i + 1] == 0xD9) break;                         // EOI
        if (i + 4 > size) { if (any_scan) break; throw std::runtime_error("JPEG: truncated"); }  // a missing EOI is tolerated after a scan
        if (d[i] != 0xFF) { if (any_scan) { i++; continue; } throw std::runtime_error("JPEG: marker expected"); }
        const int m = d[i + 1];
        if (m == 0xFF || m == 0x00 || (m >= 0xD0 && m <= 0xD7)) { i += (m == 0xFF) ? 1 : 2; continue; }   // fill bytes, leftovers of a scan
        const size_t len = ((size_t)d[i + 2] << 8) | d[i + 3];
        if (len < 2 || i + 2 + len > size) throw std::runtime_error("JPEG: truncated segment");
        const uint8_t* s = d + i + 4; const size_t n = len - 2;
        i += 2 + len;
        if (m == 0xDB) {
            for (size_t k = 0; k < n;) {
                const int pq = s[k] >> 4, tq = s[k] & 15; k++;
                if (tq > 3 || k + (pq ? 128u : 64u) > n) throw std::runtime_error("JPEG: bad DQT");
                for (int z = 0; z < 64; z++) { qt[tq][z] = pq ? (uint16_t)((s[k] << 8) | s[k + 1]) : s[k]; k += pq ? 2 : 1; }
            }
        } else if (m == 0xC0 || m == 0xC1 || m == 0xC2) {
            if (W) throw std::runtime_error("JPEG: more than one frame");
            if (n < 6 || s[0] != 8) throw std::runtime_error("JPEG: only 8-bit samples");
            progressive = m == 0xC2;
            H = (s[1] << 8) | s[2]; W = (s[3] << 8) | s[4]; ncomp = s[5];
            if ((ncomp != 1 && ncomp != 3) || n < 6 + 3 * (size_t)ncomp || W == 0 || H == 0) throw std::runtime_error("JPEG: unsupported frame");
            for (int k = 0; k < ncomp; k++) {
                comp[k].id = s[6 + 3 * k]; comp[k].h = s[7 + 3 * k] >> 4; comp[k].v = s[7 + 3 * k] & 15; comp[k].tq = s[8 + 3 * k] & 3;
                if (comp[k].h < 1 || comp[k].h > 4 || comp[k].v < 1 || comp[k].v > 4) throw std::runtime_error("JPEG: bad sampling factor");
                hmax = comp[k].h > hmax ? comp[k].h : hmax; vmax = comp[k].v > vmax ? comp[k].v : vmax;
            }
            mcux = ((size_t)W + 8 * hmax - 1) / (8 * hmax); mcuy = ((size_t)H + 8 * vmax - 1) / (8 * vmax);
            for (int k = 0; k < ncomp; k++) {
                JpegComp& c = comp[k];
                c.w = mcux * 8 * c.h; c.hgt = mcuy * 8 * c.v; c.bstride = mcux * c.h;
                c.bw = ((((size_t)W * c.h + hmax - 1) / hmax) + 7) / 8; c.bh = ((((size_t)H * c.v + vmax - 1) / vmax) + 7) / 8;
                c.coef.assign(c.bstride * mcuy * c.v * 64, 0);
            }
        } else if (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            throw std::runtime_error("JPEG: lossless / hierarchical / arithmetic coding is not supported");
        } else if (m == 0xC4) {
            for (size_t k = 0; k < n;) {
                const int tc = s[k] >> 4, th = s[k] & 15; k++;
                if (tc > 1 || th > 3 || k + 16 > n) throw std::runtime_error("JPEG: bad DHT");
                JpegHuff& h = tc ? hac[th] : hdc[th];
                int total = 0, code = 0;
                for (int l = 1; l <= 16; l++) { h.count[l] = s[k + l - 1]; h.val_first[l] = total; h.code_first[l] = (uint16_t)code; total += h.count[l]; code = (code + h.count[l]) << 1; }
                k += 16;
                if (total > 256 || k + (size_t)total > n) throw std::runtime_error("JPEG: bad DHT");
                memcpy(h.vals, s + k, (size_t)total); k += (size_t)total; h.set = true;
            }
        } else if (m == 0xDD) {
            if (n < 2) throw std::runtime_error("JPEG: bad DRI");
            restart = (s[0] << 8) | s[1];
        } else if (m == 0xDA) {
            if (!W || n < 1) throw std::runtime_error("JPEG: scan before frame");
            const int ns = s[0];
            if (ns < 1 || ns > ncomp || n < 4 + 2 * (size_t)ns) throw std::runtime_error("JPEG: unsupported scan");
            int order[3];
            for (int k = 0; k < ns; k++) {
                int c = -1; for (int q = 0; q < ncomp; q++) if (comp[q].id == s[1 + 2 * k]) c = q;
                if (c < 0) throw std::runtime_error("JPEG: scan component");
                comp[c].td = s[2 + 2 * k] >> 4; comp[c].ta = s[2 + 2 * k] & 15;
                if (comp[c].td > 3 || comp[c].ta > 3) throw std::runtime_error("JPEG: bad table selector");
                order[k] = c;
            }
            int ss = s[1 + 2 * ns], se = s[2 + 2 * ns], ah = s[3 + 2 * ns] >> 4, al = s[3 + 2 * ns] & 15;
            if (!progressive) { ss = 0; se = 63; ah = 0; al = 0; }
            else if (ss > se || se > 63 || (ss == 0 && se != 0) || (ss > 0 && ns != 1) || al > 13 || (ah != 0 && ah != al + 1))
                throw std::runtime_error("JPEG: bad progressive scan parameters");
            for (int k = 0; k < ns; k++) {
                const JpegComp& c = comp[order[k]];
                if ((ss == 0 && ah == 0 && !hdc[c.td].set) || ((se > 0) && !hac[c.ta].set)) throw std::runtime_error("JPEG: missing Huffman table");
            }
            JpegBits br{d + i, d + size};
            for (int k = 0; k < ncomp; k++) comp[k].pred = 0;
            int eobrun = 0, until_restart = restart;
            auto maybe_restart = [&]() {
                if (!restart) return;
                if (until_restart == 0) {
                    br.reset();
                    while (br.p + 1 < br.e && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) br.p++;
                    if (br.p + 1 < br.e) br.p += 2;
                    for (int k = 0; k < ncomp; k++) comp[k].pred = 0;
                    eobrun = 0; until_restart = restart;
                }
                until_restart--;
            };
            bool truncated = false;
            try {
            if (ns == 1) {                                           // non-interleaved: the component's own block raster
                JpegComp& c = comp[order[0]];
                for (size_t by = 0; by < c.bh; by++)
                    for (size_t bx = 0; bx < c.bw; bx++) {
                        maybe_restart();
                        jpeg_block(br, c, hdc, hac, &c.coef[(by * c.bstride + bx) * 64], progressive, ss, se, ah, al, eobrun);
                    }
            } else {
                for (size_t my = 0; my < mcuy; my++)
                    for (size_t mx = 0; mx < mcux; mx++) {
                        maybe_restart();
                        for (int k = 0; k < ns; k++) {
                            JpegComp& c = comp[order[k]];
                            for (int by = 0; by < c.v; by++)
                                for (int bx = 0; bx < c.h; bx++)
                                    jpeg_block(br, c, hdc, hac, &c.coef[((my * c.v + by) * c.bstride + mx * c.h + bx) * 64], progressive, ss, se, ah, al, eobrun);
                        }
                    }
            }
            } catch (const std::runtime_error&) {
                if (br.p + 2 < br.e) throw;                          // corrupt data in the middle of the file
                truncated = true;                                    // the file ends inside this scan: keep what was decoded (a coarser image)
            }
            any_scan = true;
            if (truncated) break;
            // skip the entropy-coded data up to the next marker that is not a stuffed zero or a restart marker
            i = (size_t)(br.p - d);
            while (i + 1 < size && !(d[i] == 0xFF && d[i + 1] != 0x00 && d[i + 1] != 0xFF && !(d[i + 1] >= 0xD0 && d[i + 1] <= 0xD7))) i++;
        }
    }
    if (!W || !any_scan) throw std::runtime_error("JPEG: no image data");
    // dequantise + inverse DCT
    for (int k = 0; k < ncomp; k++) {
        JpegComp& c = comp[k];
        c.px.assign(c.w * c.hgt, 0);
        const size_t rows = mcuy * c.v;
        for (size_t by = 0; by < rows; by++)
            for (size_t bx = 0; bx < c.bstride; bx++) {
                const int16_t* q = &c.coef[(by * c.bstride + bx) * 64];
                float blk[64];
                for (int z = 0; z < 64; z++) blk[kZigzag[z]] = (float)((int)q[z] * (int)qt[c.tq][z]);
                jpeg_idct(blk, c.px.data() + (by * 8) * c.w + bx * 8, c.w);
            }
    }
    typedef JpegComp Comp;
    Image img; img.width = (uint32_t)W; img.height = (uint32_t)H; img.rgba.resize((size_t)W * H * 4);
    auto sample = [&](const Comp& c, int x, int y) -> float {        // bilinear at the full-resolution pixel centre
        if (c.h == hmax && c.v == vmax) return (float)c.px[(size_t)y * c.w + x];
        const float fx = ((float)x + 0.5f) * (float)c.h / (float)hmax - 0.5f, fy = ((float)y + 0.5f) * (float)c.v / (float)vmax - 0.5f;
        const size_t cw = ((size_t)W * c.h + hmax - 1) / hmax, ch = ((size_t)H * c.v + vmax - 1) / vmax;   // valid samples
        int x0 = (int)floorf(fx), y0 = (int)floorf(fy); const float ax = fx - (float)x0, ay = fy - (float)y0;
        auto at = [&](int xx, int yy) { xx = xx < 0 ? 0 : (xx >= (int)cw ? (int)cw - 1 : xx); yy = yy < 0 ? 0 : (yy >= (int)ch ? (int)ch - 1 : yy); return (float)c.px[(size_t)yy * c.w + xx]; };
        return (at(x0, y0) * (1 - ax) + at(x0 + 1, y0) * ax) * (1 - ay) + (at(x0, y0 + 1) * (1 - ax) + at(x0 + 1, y0 + 1) * ax) * ay;
    };
    auto clamp8 = [](float v) { v = floorf(v + 0.5f); return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); };
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            uint8_t* o = &img.rgba[((size_t)y * W + x) * 4];
            const float Y = sample(comp[0], x, y);
            if (ncomp == 1) { o[0] = o[1] = o[2] = clamp8(Y); }
            else {
                const float cb = sample(comp[1], x, y) - 128.0f, cr = sample(comp[2], x, y) - 128.0f;
                o[0] = clamp8(Y + 1.402f * cr); o[1] = clamp8(Y - 0.344136f * cb - 0.714136f * cr); o[2] = clamp8(Y + 1.772f * cb);
            }
            o[3] = 255;
        }
    return img;
}

Image decode_image(const uint8_t* p, size_t size)
{
    if (size >= 2 && p[0] == 0xFF && p[1] == 0xD8) return decode_jpeg(p, size);
    return decode_png(p, size);
}

Image load_image(const std::string& path)
{
    std::string data = read_file(path, true);
    return decode_image((const uint8_t*)data.data(), data.size());
}

Image load_png(const std::string& path)
{
    std::string data = read_file(path, true);
    return decode_png((const uint8_t*)data.data(), data.size());
}

void write_png(const std::string& path, const uint8_t* rgba, uint32_t W, uint32_t H)
{
    std::vector<uint8_t> raw(((size_t)W * 4 + 1) * H);
    for (uint32_t y = 0; y < H; y++) { raw[((size_t)W * 4 + 1) * y] = 0; memcpy(&raw[((size_t)W * 4 + 1) * y + 1], rgba + (size_t)y * W * 4, (size_t)W * 4); }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) throw std::runtime_error("PNG: deflate failed");
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("PNG: cannot open " + path);
    auto put32 = [](uint8_t* p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; };
    auto chunk = [&](const char* tag, const uint8_t* body, uint32_t len) {
        std::vector<uint8_t> b(8 + (size_t)len + 4);
        put32(b.data(), len); memcpy(&b[4], tag, 4); if (len) memcpy(&b[8], body, len);
        put32(&b[8 + len], (uint32_t)crc32(0L, &b[4], 4 + len));
        f.write((const char*)b.data(), (std::streamsize)b.size());
    };
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    f.write((const char*)sig, 8);
    uint8_t ihdr[13]; put32(ihdr, W); put32(ihdr + 4, H); ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = ihdr[11] = ihdr[12] = 0;
    chunk("IHDR", ihdr, 13); chunk("IDAT", comp.data(), (uint32_t)clen); chunk("IEND", nullptr, 0);
    if (!f) throw std::runtime_error("PNG: write failed: " + path);
}

std::vector<Image> load_glb_images(const std::string& path)
{
    if (ends_with(path, ".obj")) {                                 // the map_Kd / map_Bump files of the .mtl, in handle order
        struct Guard { Guard() { g_keep_uv = true; } ~Guard() { g_keep_uv = false; } } guard;
        ObjScene sc; std::vector<rt_vertex> v; std::vector<uint32_t> i; std::vector<rt_mesh> m;
        load_obj(path, v, i, m, &sc);
        std::vector<Image> out;
        for (const std::string& f : sc.images) { Image img; try { img = load_image(sibling(path, f)); } catch (const std::exception&) { img = Image(); } out.push_back(std::move(img)); }
        return out;
    }
    Json g; std::vector<std::string> buffers;
    const bool text = ends_with(path, ".gltf");
    if (text) {
        g = parse_json(read_file(path, false));
        if (const Json* jb = g.get("buffers"))
            for (const Json& b : jb->arr) { const Json* uri = b.get("uri"); buffers.push_back(uri ? resolve_uri(path, uri->str) : std::string()); }
    } else read_glb(path, g, buffers);
    std::vector<Image> out;
    if (const Json* ji = g.get("images"))
        for (const Json& im : ji->arr) {
            Image img;
            std::string bytes;
            if (const Json* bv = im.get("bufferView")) {                       // embedded: a bufferView of one of the buffers
                const Json& v = g.at("bufferViews").arr.at(bv->index());
                const long long jb = v.integer("buffer", 0), jo = v.integer("byteOffset", 0), jl = v.integer("byteLength", 0);
                if (jb < 0 || jo < 0 || jl < 0 || (size_t)jb >= buffers.size() || (unsigned long long)jo > buffers[(size_t)jb].size() ||
                    (unsigned long long)jl > buffers[(size_t)jb].size() - (size_t)jo) throw std::runtime_error("GLB: image outside its buffer");
                const size_t bi = (size_t)jb, off = (size_t)jo, len = (size_t)jl;
                bytes.assign(buffers[bi], off, len);
            } else if (const Json* uri = im.get("uri")) {                      // .gltf: sibling file or data: URI; a missing file leaves the entry empty
                try { bytes = resolve_uri(path, uri->str); } catch (const std::exception&) { bytes.clear(); }
            }
            // by signature, not by mimeType (optional for uri images); unsupported encodings (arithmetic-coded JPEG, KTX, WebP) stay empty
            try { if (!bytes.empty()) img = decode_image((const uint8_t*)bytes.data(), bytes.size()); } catch (const std::exception&) { img = Image(); }
            out.push_back(std::move(img));
        }
    return out;
}

} // namespace rtb
