// obj_loader.cpp — Wavefront OBJ / MTL import behind include/rt_obj.h (replaces the reference's assimp path,
// triangle_mesh.h:51-143 and the host half of create_meshes, :208-352).  Host only.
#include "../../include/rt_obj.h"
#include "device_scene.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

using namespace rtb;

struct rt_obj {
    std::vector<float> vertices, uvs, normals;
    std::vector<uint32_t> indices;
    bool normals_generated = false;
    std::string diffuse_map;
    std::vector<std::string> mtllibs;
};

namespace {

int fail(int code, const std::string& msg) { set_error(msg); return code; }

struct Corner { int v, vt, vn; };
struct CornerKey {
    long long a, b;
    bool operator==(const CornerKey& o) const { return a == o.a && b == o.b; }
};
struct CornerHash {
    size_t operator()(const CornerKey& k) const { return (size_t)(k.a * 0x9E3779B97F4A7C15ull) ^ (size_t)(k.b * 0xC2B2AE3D27D4EB4Full); }
};

// one whitespace-delimited token of [p, end) on the current line; returns false at end of line
bool next_token(const char*& p, const char* end, const char*& tok, size_t& len) {
    while (p < end && (*p == ' ' || *p == '\t' || *p == '\r')) p++;
    if (p >= end || *p == '\n') return false;
    tok = p;
    while (p < end && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') p++;
    len = (size_t)(p - tok);
    return true;
}

float to_float(const char* tok, size_t len) {
    char buf[64];
    if (len >= sizeof(buf)) len = sizeof(buf) - 1;
    memcpy(buf, tok, len); buf[len] = 0;
    return strtof(buf, nullptr);
}

// "v", "v/vt", "v//vn", "v/vt/vn" -> 1-based (or negative, relative) indices; 0 = absent
bool parse_corner(const char* tok, size_t len, Corner& c) {
    int vals[3] = {0, 0, 0};
    int field = 0;
    size_t i = 0;
    while (i <= len && field < 3) {
        size_t j = i;
        while (j < len && tok[j] != '/') j++;
        if (j > i) {
            char buf[32];
            const size_t n = j - i < sizeof(buf) - 1 ? j - i : sizeof(buf) - 1;
            memcpy(buf, tok + i, n); buf[n] = 0;
            char* e = nullptr;
            const long v = strtol(buf, &e, 10);
            if (e == buf) return false;
            vals[field] = (int)v;
        }
        field++;
        i = j + 1;
    }
    c.v = vals[0]; c.vt = vals[1]; c.vn = vals[2];
    return c.v != 0;
}

// 1-based / negative-relative -> 0-based; -1 when absent or out of range
int resolve(int ix, size_t count) {
    if (ix > 0) return (size_t)ix <= count ? ix - 1 : -1;
    if (ix < 0) return (size_t)(-ix) <= count ? (int)count + ix : -1;
    return -1;
}

void parse_mtl(const char* text, size_t len, std::string& diffuse_map) {
    const char* p = text; const char* end = text + len;
    while (p < end) {
        const char* line_end = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!line_end) line_end = end;
        const char* q = p; const char* tok; size_t tl;
        if (next_token(q, line_end, tok, tl) && tl == 6 && !strncmp(tok, "map_Kd", 6) && diffuse_map.empty()) {
            // options (-s 1 1 1, -o ...) may precede the file name: the name is the last token
            std::string last;
            while (next_token(q, line_end, tok, tl)) last.assign(tok, tl);
            diffuse_map = last;
        }
        p = line_end < end ? line_end + 1 : end;
    }
}

int parse_obj(const char* text, size_t len, rt_obj& out) {
    std::vector<float> pos, tex, nrm;
    std::vector<Corner> corners;          // three per output triangle
    std::vector<Corner> face;
    const char* p = text; const char* end = text + len;
    size_t line_no = 0;
    while (p < end) {
        line_no++;
        const char* line_end = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!line_end) line_end = end;
        const char* q = p; const char* tok; size_t tl;
        if (next_token(q, line_end, tok, tl) && tok[0] != '#') {
            if (tl == 1 && tok[0] == 'v') {
                float f[3] = {0, 0, 0};
                for (int k = 0; k < 3; k++) { if (!next_token(q, line_end, tok, tl)) return fail(RT_ERR_INVALID, "obj: short v record at line " + std::to_string(line_no)); f[k] = to_float(tok, tl); }
                pos.insert(pos.end(), f, f + 3);
            } else if (tl == 2 && tok[0] == 'v' && tok[1] == 't') {
                float f[2] = {0, 0};
                for (int k = 0; k < 2; k++) { if (!next_token(q, line_end, tok, tl)) break; f[k] = to_float(tok, tl); }
                tex.insert(tex.end(), f, f + 2);
            } else if (tl == 2 && tok[0] == 'v' && tok[1] == 'n') {
                float f[3] = {0, 0, 0};
                for (int k = 0; k < 3; k++) { if (!next_token(q, line_end, tok, tl)) return fail(RT_ERR_INVALID, "obj: short vn record at line " + std::to_string(line_no)); f[k] = to_float(tok, tl); }
                nrm.insert(nrm.end(), f, f + 3);
            } else if (tl == 1 && tok[0] == 'f') {
                face.clear();
                while (next_token(q, line_end, tok, tl)) {
                    Corner c;
                    if (!parse_corner(tok, tl, c)) return fail(RT_ERR_INVALID, "obj: bad face corner at line " + std::to_string(line_no));
                    c.v = resolve(c.v, pos.size() / 3);
                    c.vt = resolve(c.vt, tex.size() / 2);
                    c.vn = resolve(c.vn, nrm.size() / 3);
                    if (c.v < 0) return fail(RT_ERR_INVALID, "obj: vertex index out of range at line " + std::to_string(line_no));
                    face.push_back(c);
                }
                if (face.size() < 3) return fail(RT_ERR_INVALID, "obj: face with fewer than 3 corners at line " + std::to_string(line_no));
                for (size_t i = 1; i + 1 < face.size(); i++) {        // fan triangulation
                    corners.push_back(face[0]); corners.push_back(face[i]); corners.push_back(face[i + 1]);
                }
            } else if (tl == 6 && !strncmp(tok, "mtllib", 6)) {
                while (next_token(q, line_end, tok, tl)) out.mtllibs.emplace_back(tok, tl);
            }
        }
        p = line_end < end ? line_end + 1 : end;
    }
    if (corners.empty()) return fail(RT_ERR_INVALID, "obj: no faces");

    // smooth normals per POSITION when the file has none: normalised sum of area-weighted face normals
    std::vector<float> gen;
    out.normals_generated = nrm.empty();
    if (out.normals_generated) {
        gen.assign(pos.size(), 0.f);
        for (size_t t = 0; t + 2 < corners.size(); t += 3) {
            const float* a = &pos[3 * (size_t)corners[t].v]; const float* b = &pos[3 * (size_t)corners[t + 1].v]; const float* c = &pos[3 * (size_t)corners[t + 2].v];
            const float e0[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e1[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
            const float n[3] = {e0[1] * e1[2] - e0[2] * e1[1], e0[2] * e1[0] - e0[0] * e1[2], e0[0] * e1[1] - e0[1] * e1[0]};
            for (int k = 0; k < 3; k++) for (int d = 0; d < 3; d++) gen[3 * (size_t)corners[t + k].v + d] += n[d];
        }
        for (size_t v = 0; v + 2 < gen.size(); v += 3) {
            const float l = std::sqrt(gen[v] * gen[v] + gen[v + 1] * gen[v + 1] + gen[v + 2] * gen[v + 2]);
            if (l > 0.f) { gen[v] /= l; gen[v + 1] /= l; gen[v + 2] /= l; } else { gen[v] = 0.f; gen[v + 1] = 1.f; gen[v + 2] = 0.f; }
        }
    }

    // unify (v, vt, vn) triples into output vertices, first use first
    std::unordered_map<CornerKey, uint32_t, CornerHash> seen;
    seen.reserve(corners.size());
    out.indices.reserve(corners.size());
    for (const Corner& c : corners) {
        const CornerKey key{(long long)c.v, ((long long)(c.vt + 1) << 32) | (uint32_t)(c.vn + 1)};
        auto it = seen.find(key);
        if (it == seen.end()) {
            const uint32_t ix = (uint32_t)(out.vertices.size() / 3);
            it = seen.emplace(key, ix).first;
            out.vertices.insert(out.vertices.end(), &pos[3 * (size_t)c.v], &pos[3 * (size_t)c.v] + 3);
            if (c.vt >= 0) out.uvs.insert(out.uvs.end(), &tex[2 * (size_t)c.vt], &tex[2 * (size_t)c.vt] + 2);
            else { out.uvs.push_back(0.f); out.uvs.push_back(0.f); }                   // triangle_mesh.h:66-69
            if (out.normals_generated) out.normals.insert(out.normals.end(), &gen[3 * (size_t)c.v], &gen[3 * (size_t)c.v] + 3);
            else if (c.vn >= 0) out.normals.insert(out.normals.end(), &nrm[3 * (size_t)c.vn], &nrm[3 * (size_t)c.vn] + 3);
            else { out.normals.push_back(0.f); out.normals.push_back(1.f); out.normals.push_back(0.f); }
        }
        out.indices.push_back(it->second);
    }
    return RT_OK;
}

bool read_file(const std::string& path, std::string& data) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) data.append(buf, n);
    fclose(f);
    return true;
}

} // namespace

extern "C" {

int rt_obj_parse(const char* obj_text, size_t obj_len, const char* mtl_text, size_t mtl_len, rt_obj** out) {
    if (!obj_text || !out) return fail(RT_ERR_INVALID, "obj: bad argument");
    rt_obj* o = new rt_obj;
    const int rc = parse_obj(obj_text, obj_len, *o);
    if (rc < 0) { delete o; return rc; }
    if (mtl_text) parse_mtl(mtl_text, mtl_len, o->diffuse_map);
    *out = o;
    return RT_OK;
}

int rt_obj_load(const char* path, rt_obj** out) {
    if (!path || !out) return fail(RT_ERR_INVALID, "obj: bad argument");
    std::string text;
    if (!read_file(path, text)) return fail(RT_ERR_INVALID, std::string("obj: cannot read ") + path);
    rt_obj* o = new rt_obj;
    const int rc = parse_obj(text.data(), text.size(), *o);
    if (rc < 0) { delete o; return rc; }
    std::string dir(path);
    const size_t slash = dir.find_last_of('/');
    dir = slash == std::string::npos ? std::string() : dir.substr(0, slash + 1);
    for (const std::string& lib : o->mtllibs) {           // a missing .mtl is not an error (triangle_mesh.h:275-281 just finds no texture)
        std::string mtl;
        if (read_file(dir + lib, mtl)) parse_mtl(mtl.data(), mtl.size(), o->diffuse_map);
    }
    *out = o;
    return RT_OK;
}

void rt_obj_free(rt_obj* o) { delete o; }
uint32_t rt_obj_num_vertices(const rt_obj* o) { return o ? (uint32_t)(o->vertices.size() / 3) : 0; }
uint32_t rt_obj_num_triangles(const rt_obj* o) { return o ? (uint32_t)(o->indices.size() / 3) : 0; }
const float* rt_obj_vertices(const rt_obj* o) { return o ? o->vertices.data() : nullptr; }
const float* rt_obj_uvs(const rt_obj* o) { return o ? o->uvs.data() : nullptr; }
const float* rt_obj_normals(const rt_obj* o) { return o ? o->normals.data() : nullptr; }
const uint32_t* rt_obj_indices(const rt_obj* o) { return o ? o->indices.data() : nullptr; }
int rt_obj_normals_generated(const rt_obj* o) { return o && o->normals_generated ? 1 : 0; }
const char* rt_obj_diffuse_map(const rt_obj* o) { return o ? o->diffuse_map.c_str() : ""; }

int rt_obj_add_to_scene(rt_scene* scene, const rt_obj* o, int mat) {
    if (!scene || !o) return fail(RT_ERR_INVALID, "obj: bad argument");
    return rt_mesh(scene, o->vertices.data(), (int)(o->vertices.size() / 3), o->indices.data(), (int)(o->indices.size() / 3),
                   o->uvs.data(), o->normals.data(), mat);
}

} // extern "C"
