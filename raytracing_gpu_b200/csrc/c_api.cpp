// c_api.cpp — extern "C" boundary (include/rt_b200.h): scene recording, program replay, flatten/export.
// The device-side entry points (commit / first_hit / render) live in device_api.cu.
#include "scene_host.h"
#include "device_scene.h"
#include "../../include/rt_catalog.h"
#include "../../include/rt_output.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

using namespace rtb;

struct rt_scene {
    HostScene host;
    DeviceScene* dev = nullptr;   // created by rt_scene_commit
    std::vector<DeviceScene*> replicas;   // further device copies made by rt_render_multi (one per extra device)
};

namespace rtb {
static thread_local std::string g_err;
void set_error(const std::string& e) { g_err = e; }
std::string last_error_copy() { return g_err; }
HostScene& host_of(rt_scene* s) { return s->host; }
DeviceScene*& device_of(rt_scene* s) { return s->dev; }
std::vector<DeviceScene*>& replicas_of(rt_scene* s) { return s->replicas; }
} // namespace rtb

static int fail(int code, const char* msg) { g_err = msg; return code; }

#define CHECK_SCENE(s) do { if (!(s)) return fail(RT_ERR_INVALID, "null scene"); } while (0)
#define CHECK_IX(ix, vec, what) do { if ((ix) < 0 || (size_t)(ix) >= (vec).size()) return fail(RT_ERR_INVALID, "bad " what " index"); } while (0)

extern "C" {

const char* rt_last_error(void) { return g_err.c_str(); }
const char* rt_version(void) { return "rt_b200 0.3 (sm_100a)"; }

size_t rt_struct_size(int id) {
    switch (id) {
    case RT_STRUCT_RENDER_SETTINGS: return sizeof(rt_render_settings);
    case RT_STRUCT_RENDER_STATS: return sizeof(rt_render_stats);
    case RT_STRUCT_BUILD_SETTINGS: return sizeof(rt_build_settings);
    case RT_STRUCT_BUILD_STATS: return sizeof(rt_build_stats);
    case RT_STRUCT_MULTI_SETTINGS: return sizeof(rt_multi_settings);
    case RT_STRUCT_MULTI_STATS: return sizeof(rt_multi_stats);
    case RT_STRUCT_DRAW_SETTINGS: return sizeof(rt_draw_settings);
    case RT_STRUCT_CHECKPOINT_INFO: return sizeof(rt_checkpoint_info);
    case RT_STRUCT_CATALOG_PARAMS: return sizeof(rt_catalog_params);
    default: return 0;
    }
}

int rt_scene_create(rt_scene** out) {
    if (!out) return fail(RT_ERR_INVALID, "null out pointer");
    *out = new (std::nothrow) rt_scene;
    return *out ? RT_OK : fail(RT_ERR_INVALID, "out of memory");
}

void rt_scene_destroy(rt_scene* s) {
    if (!s) return;
    if (s->dev) device_scene_destroy(s->dev);
    for (DeviceScene* r : s->replicas) device_scene_destroy(r);
    delete s;
}

/* ---- textures ---- */
static int add_tex(rt_scene* s, rt_flat_texture t) {
    s->host.flat.valid = false;
    s->host.textures.push_back(t);
    return (int)s->host.textures.size() - 1;
}

int rt_tex_solid(rt_scene* s, float r, float g, float b) {
    CHECK_SCENE(s);
    rt_flat_texture t{}; t.type = RT_TEX_SOLID; t.rgb[0] = r; t.rgb[1] = g; t.rgb[2] = b;
    return add_tex(s, t);
}

int rt_tex_checker(rt_scene* s, int even_tex, int odd_tex) {
    CHECK_SCENE(s); CHECK_IX(even_tex, s->host.textures, "texture"); CHECK_IX(odd_tex, s->host.textures, "texture");
    // the device resolves checker-of-checker chains iteratively, at most 8 levels (shade.cuh texture_value): refuse deeper
    // nesting here instead of shading it black.  Textures only refer to earlier ones, so the depth is a simple recursion.
    std::vector<int> depth(s->host.textures.size(), 0);
    for (size_t i = 0; i < s->host.textures.size(); i++) {
        const rt_flat_texture& x = s->host.textures[i];
        if (x.type == RT_TEX_CHECKER) depth[i] = 1 + std::max(depth[x.a], depth[x.b]);
    }
    if (1 + std::max(depth[(size_t)even_tex], depth[(size_t)odd_tex]) > 8) return fail(RT_ERR_UNSUPPORTED, "checker textures nested more than 8 deep");
    rt_flat_texture t{}; t.type = RT_TEX_CHECKER; t.a = (uint32_t)even_tex; t.b = (uint32_t)odd_tex;
    return add_tex(s, t);
}

int rt_perlin_tables(rt_scene* s, const float* ranvec768, const int32_t* perm768) {
    CHECK_SCENE(s);
    if (!ranvec768 || !perm768) return fail(RT_ERR_INVALID, "null perlin tables");
    for (int i = 0; i < 768; i++) if (perm768[i] < 0 || perm768[i] > 255) return fail(RT_ERR_INVALID, "perlin permutation entry outside 0..255");
    rt_flat_perlin p;
    std::memcpy(p.ranvec, ranvec768, sizeof(p.ranvec));
    std::memcpy(p.perm, perm768, sizeof(p.perm));
    s->host.perlins.push_back(p);
    return (int)s->host.perlins.size() - 1;
}

static int perlin_tex(rt_scene* s, uint32_t type, int perlin, float scale, int depth) {
    CHECK_SCENE(s); CHECK_IX(perlin, s->host.perlins, "perlin");
    if (depth < 0 || depth > 64) return fail(RT_ERR_INVALID, "turbulence depth outside 0..64 (perlin::turb loops `depth` times per shaded hit)");
    rt_flat_texture t{}; t.type = type; t.a = (uint32_t)perlin; t.scale = scale; t.depth = depth;
    return add_tex(s, t);
}
int rt_tex_noise(rt_scene* s, int perlin, float scale) { return perlin_tex(s, RT_TEX_NOISE, perlin, scale, 0); }
int rt_tex_turbulent(rt_scene* s, int perlin, float scale, int depth) { return perlin_tex(s, RT_TEX_TURB, perlin, scale, depth); }
int rt_tex_marble(rt_scene* s, int perlin, float scale) { return perlin_tex(s, RT_TEX_MARBLE, perlin, scale, 7); }

int rt_image(rt_scene* s, const uint8_t* pixels, int width, int height, int bpp) {
    CHECK_SCENE(s);
    if (!pixels || width <= 0 || height <= 0 || bpp < 3 || bpp > 4) return fail(RT_ERR_INVALID, "bad image (need >= 3 bytes per pixel)");
    const size_t n = (size_t)width * height * bpp;
    size_t off = (s->host.image_bytes.size() + 15) & ~size_t(15);
    if (off + n > 0xffffffffu) return fail(RT_ERR_UNSUPPORTED, "image pool exceeds 4 GiB");
    s->host.image_bytes.resize(off + n);
    std::memcpy(s->host.image_bytes.data() + off, pixels, n);
    s->host.images.push_back(rt_flat_image{(uint32_t)width, (uint32_t)height, (uint32_t)bpp, (uint32_t)off});
    return (int)s->host.images.size() - 1;
}

int rt_tex_image(rt_scene* s, int image) {
    CHECK_SCENE(s); CHECK_IX(image, s->host.images, "image");
    rt_flat_texture t{}; t.type = RT_TEX_IMAGE; t.a = (uint32_t)image;
    return add_tex(s, t);
}

/* ---- materials ---- */
static int add_mat(rt_scene* s, uint32_t type, int tex, float param, bool needs_tex) {
    CHECK_SCENE(s);
    if (needs_tex) CHECK_IX(tex, s->host.textures, "texture");
    s->host.flat.valid = false;
    s->host.materials.push_back(rt_flat_material{type, needs_tex ? (uint32_t)tex : 0u, param, 0});
    return (int)s->host.materials.size() - 1;
}
int rt_mat_lambertian(rt_scene* s, int tex) { return add_mat(s, RT_MAT_LAMBERTIAN, tex, 0, true); }
int rt_mat_metal(rt_scene* s, int tex, float fuzz) { return add_mat(s, RT_MAT_METAL, tex, fuzz, true); }
int rt_mat_dielectric(rt_scene* s, float ir) { return add_mat(s, RT_MAT_DIELECTRIC, 0, ir, false); }
int rt_mat_diffuse_light(rt_scene* s, int tex) { return add_mat(s, RT_MAT_DIFFUSE_LIGHT, tex, 0, true); }
int rt_mat_isotropic(rt_scene* s, int tex) { return add_mat(s, RT_MAT_ISOTROPIC, tex, 0, true); }

/* ---- hittables ---- */
static int add_node(rt_scene* s, Node&& n) {
    s->host.flat.valid = false;
    s->host.nodes.push_back(std::move(n));
    return (int)s->host.nodes.size() - 1;
}

static int add_leaf(rt_scene* s, NodeKind k, const float* f, int nf, int mat) {
    CHECK_SCENE(s); CHECK_IX(mat, s->host.materials, "material");
    Node n; n.kind = k; n.mat = mat;
    std::memcpy(n.f, f, sizeof(float) * nf);
    return add_node(s, std::move(n));
}

int rt_sphere(rt_scene* s, float cx, float cy, float cz, float radius, int mat) {
    const float f[4] = {cx, cy, cz, radius};
    return add_leaf(s, NodeKind::Sphere, f, 4, mat);
}

int rt_moving_sphere(rt_scene* s, const float c0[3], const float c1[3], float t0, float t1, float radius, int mat) {
    if (!c0 || !c1) return fail(RT_ERR_INVALID, "null center");
    const float f[9] = {c0[0], c0[1], c0[2], c1[0], c1[1], c1[2], t0, t1, radius};
    return add_leaf(s, NodeKind::MovingSphere, f, 9, mat);
}

static int rect(rt_scene* s, NodeKind k, float a0, float a1, float b0, float b1, float kk, int mat) {
    const float f[5] = {a0, a1, b0, b1, kk};
    return add_leaf(s, k, f, 5, mat);
}
int rt_xy_rect(rt_scene* s, float x0, float x1, float y0, float y1, float k, int mat) { return rect(s, NodeKind::XYRect, x0, x1, y0, y1, k, mat); }
int rt_xz_rect(rt_scene* s, float x0, float x1, float z0, float z1, float k, int mat) { return rect(s, NodeKind::XZRect, x0, x1, z0, z1, k, mat); }
int rt_yz_rect(rt_scene* s, float y0, float y1, float z0, float z1, float k, int mat) { return rect(s, NodeKind::YZRect, y0, y1, z0, z1, k, mat); }

int rt_box(rt_scene* s, const float p0[3], const float p1[3], int mat) {
    if (!p0 || !p1) return fail(RT_ERR_INVALID, "null corner");
    const float f[6] = {p0[0], p0[1], p0[2], p1[0], p1[1], p1[2]};
    return add_leaf(s, NodeKind::Box, f, 6, mat);
}

int rt_triangle(rt_scene* s, const float v[9], const float uv[6], const float* normals, int mat) {
    CHECK_SCENE(s); CHECK_IX(mat, s->host.materials, "material");
    if (!v || !uv) return fail(RT_ERR_INVALID, "null triangle data");
    Node n; n.kind = NodeKind::Triangle; n.mat = mat;
    rt_flat_tri_attr a;
    make_triangle(v, uv, normals, n.f, a);
    s->host.tri_attrs.push_back(a);
    n.tri_attr = (int)s->host.tri_attrs.size() - 1;
    return add_node(s, std::move(n));
}

int rt_mesh(rt_scene* s, const float* vertices, int n_vertices, const uint32_t* indices, int n_triangles,
            const float* uvs, const float* normals, int mat) {
    CHECK_SCENE(s); CHECK_IX(mat, s->host.materials, "material");
    if (!vertices || !indices || !uvs || !normals || n_vertices <= 0 || n_triangles <= 0) return fail(RT_ERR_INVALID, "bad mesh arrays");
    for (int k = 0; k < 3 * n_triangles; k++) if (indices[k] >= (uint32_t)n_vertices) return fail(RT_ERR_INVALID, "mesh index out of range");
    Node g; g.kind = NodeKind::Group; g.kids.reserve(n_triangles);
    s->host.nodes.reserve(s->host.nodes.size() + n_triangles + 1);
    s->host.tri_attrs.reserve(s->host.tri_attrs.size() + n_triangles);
    for (int k = 0; k < n_triangles; k++) {
        // triangle_mesh.h:158-192: vertex i of the triangle reads element indices[3k+i] of each array
        float v[9], nrm[9], uv[6];
        for (int i = 0; i < 3; i++) {
            const uint32_t ix = indices[3 * k + i];
            std::memcpy(v + 3 * i, vertices + 3 * (size_t)ix, 12);
            std::memcpy(nrm + 3 * i, normals + 3 * (size_t)ix, 12);
            std::memcpy(uv + 2 * i, uvs + 2 * (size_t)ix, 8);
        }
        const int t = rt_triangle(s, v, uv, nrm, mat);
        if (t < 0) return t;
        g.kids.push_back(t);
    }
    return add_node(s, std::move(g));
}

int rt_translate(rt_scene* s, int child, float dx, float dy, float dz) {
    CHECK_SCENE(s); CHECK_IX(child, s->host.nodes, "hittable");
    Node n; n.kind = NodeKind::Translate; n.child = child; n.f[0] = dx; n.f[1] = dy; n.f[2] = dz;
    return add_node(s, std::move(n));
}

int rt_rotate_y_sincos(rt_scene* s, int child, float sin_theta, float cos_theta) {
    CHECK_SCENE(s); CHECK_IX(child, s->host.nodes, "hittable");
    Node n; n.kind = NodeKind::RotateY; n.child = child; n.f[0] = sin_theta; n.f[1] = cos_theta;
    return add_node(s, std::move(n));
}

int rt_rotate_y(rt_scene* s, int child, float angle_degrees) {
    const float pi = 3.1415926535897932385f;
    const float radians = angle_degrees * pi / 180.0f;       // hittable.h:79, common.h:46
    return rt_rotate_y_sincos(s, child, sinf(radians), cosf(radians));
}

int rt_constant_medium_nid(rt_scene* s, int boundary, float neg_inv_density, int tex) {
    CHECK_SCENE(s); CHECK_IX(boundary, s->host.nodes, "hittable"); CHECK_IX(tex, s->host.textures, "texture");
    Node n; n.kind = NodeKind::Medium; n.child = boundary; n.mat = tex; n.f[0] = neg_inv_density;
    return add_node(s, std::move(n));
}

int rt_constant_medium(rt_scene* s, int boundary, float density, int tex) {
    return rt_constant_medium_nid(s, boundary, -1 / density, tex);   // constant_medium.h:16
}

static int group(rt_scene* s, const int* children, int n) {
    CHECK_SCENE(s);
    if (n < 0 || (n > 0 && !children)) return fail(RT_ERR_INVALID, "bad child list");
    Node g; g.kind = NodeKind::Group;
    for (int i = 0; i < n; i++) { CHECK_IX(children[i], s->host.nodes, "hittable"); g.kids.push_back(children[i]); }
    return add_node(s, std::move(g));
}
int rt_list(rt_scene* s, const int* children, int n) { return group(s, children, n); }
int rt_bvh(rt_scene* s, const int* children, int n) { return group(s, children, n); }

/* ---- world / camera ---- */
int rt_set_world(rt_scene* s, int h) {
    CHECK_SCENE(s); CHECK_IX(h, s->host.nodes, "hittable");
    s->host.world = h; s->host.flat.valid = false;
    return RT_OK;
}

int rt_scene_hash(rt_scene* s, uint64_t* out) {
    CHECK_SCENE(s);
    if (!out) return fail(RT_ERR_INVALID, "null out pointer");
    if (!s->host.flat.valid) { const int rc = rt_scene_flatten(s); if (rc < 0) return rc; }
    std::vector<uint8_t> blob(s->host.export_size());
    const int rc = rt_scene_export(s, blob.data(), blob.size());
    if (rc < 0) return rc;
    uint64_t h = 1469598103934665603ull;   // FNV-1a
    auto mix = [&](const uint8_t* p, size_t n) { for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ull; } };
    mix(blob.data(), blob.size());
    const uint32_t sem = s->host.semantics;
    mix(reinterpret_cast<const uint8_t*>(&sem), sizeof(sem));
    *out = h;
    return RT_OK;
}

int rt_set_semantics(rt_scene* s, uint32_t semantics) {
    CHECK_SCENE(s);
    if (semantics & ~(uint32_t)RT_SEM_FIXED) return fail(RT_ERR_INVALID, "set_semantics: unknown bits");
    s->host.semantics = semantics;
    return RT_OK;
}

int rt_set_background(rt_scene* s, float r, float g, float b) {
    CHECK_SCENE(s);
    s->host.background[0] = r; s->host.background[1] = g; s->host.background[2] = b; s->host.flat.valid = false;
    return RT_OK;
}

int rt_set_camera(rt_scene* s, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov,
                  float aspect, float aperture, float focus_dist, float t0, float t1) {
    CHECK_SCENE(s);
    if (!lookfrom || !lookat || !vup) return fail(RT_ERR_INVALID, "null camera vector");
    make_camera(s->host.camera, lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist, t0, t1);
    s->host.has_camera = true; s->host.flat.valid = false;
    return RT_OK;
}

int rt_set_camera_raw(rt_scene* s, const rt_flat_camera* cam) {
    CHECK_SCENE(s);
    if (!cam) return fail(RT_ERR_INVALID, "null camera");
    s->host.camera = *cam; s->host.has_camera = true; s->host.flat.valid = false;
    return RT_OK;
}

/* ---- flatten / export ---- */
int rt_scene_flatten(rt_scene* s) {
    CHECK_SCENE(s);
    if (s->host.flat.valid) return (int)s->host.flat.hdr.num_leaf_ids;   // unchanged since the last flattening (a committed device copy stays current)
    std::string e;
    const int r = s->host.flatten(e);
    if (r < 0) g_err = e;
    return r;
}

size_t rt_scene_export_size(rt_scene* s) { return s ? s->host.export_size() : 0; }

int rt_scene_export(rt_scene* s, void* buf, size_t cap) {
    CHECK_SCENE(s);
    if (!buf) return fail(RT_ERR_INVALID, "null buffer");
    std::string e;
    const int r = s->host.export_blob(buf, cap, e);
    if (r < 0) g_err = e;
    return r;
}

int rt_image_height(int width, float aspect) {
    const double a = aspect;                      // render_settings::aspect_ratio is double (render.h:22)
    return static_cast<int>((unsigned)width / a); // render.h:33
}

int rt_quantize_frame(const float* rgb, int width, int height, uint8_t* out) {
    if (!rgb || !out || width <= 0 || height <= 0) return fail(RT_ERR_INVALID, "bad frame");
    for (int j = height - 1, row = 0; j >= 0; j--, row++)      // color.h:23: top row first
        for (int i = 0; i < width; i++)
            for (int c = 0; c < 3; c++) {
                float v = rgb[((size_t)j * width + i) * 3 + c];
                v = std::sqrt(v);                                   // color.h:37-39 (gamma 2)
                v = v < 0.0f ? 0.0f : (v > 0.999f ? 0.999f : v);    // color.h:42, common.h:55
                out[((size_t)row * width + i) * 3 + c] = (uint8_t)static_cast<int>(256 * v);
            }
    return RT_OK;
}

} // extern "C"
