// scene_host.cpp — flattening of the recorded scene graph into include/rt_flat.h tables.
// Compiled with -ffp-contract=off: the few float expressions here (camera, triangle setup) follow the
// reference's expression order so the tables are bit-identical to what its constructors compute.
#include "scene_host.h"
#include <atomic>

#include <cmath>
#include <cstring>
#include <map>

namespace rtb {

namespace {

struct V3 { float x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 mul(float t, V3 v) { return {t * v.x, t * v.y, t * v.z}; }
inline V3 divf(V3 v, float t) { return mul(1 / t, v); }                       // vec3.h:114
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }      // vec3.h:118
inline V3 cross(V3 u, V3 v) { return {u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x}; } // vec3.h:122
inline float length(V3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }  // vec3.h:54-60
inline V3 unit(V3 v) { return divf(v, length(v)); }                             // vec3.h:126
inline void put(float* d, V3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

} // namespace

void make_camera(rt_flat_camera& c, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov,
                 float aspect, float aperture, float focus_dist, float t0, float t1) {
    const float pi = 3.1415926535897932385f;                 // common.h:43
    const float theta = vfov * pi / 180.0f;                  // common.h:46
    const float h = tanf(theta / 2.0f);                      // camera.h:27
    const float viewport_height = 2.0f * h;
    const float viewport_width = aspect * viewport_height;
    const V3 from{lookfrom[0], lookfrom[1], lookfrom[2]}, at{lookat[0], lookat[1], lookat[2]}, up{vup[0], vup[1], vup[2]};
    const V3 w = unit(sub(from, at));                        // camera.h:32-34
    const V3 u = unit(cross(up, w));
    const V3 v = cross(w, u);
    const V3 horizontal = mul(focus_dist * viewport_width, u);   // camera.h:38-39
    const V3 vertical = mul(focus_dist * viewport_height, v);
    const V3 llc = sub(sub(sub(from, divf(horizontal, 2.0f)), divf(vertical, 2.0f)), mul(focus_dist, w)); // camera.h:40
    put(c.origin, from); put(c.lower_left_corner, llc); put(c.horizontal, horizontal); put(c.vertical, vertical);
    put(c.u, u); put(c.v, v); put(c.w, w);
    c.lens_radius = aperture / 2;                            // camera.h:43
    c.time0 = t0; c.time1 = t1;
}

void make_triangle(const float v[9], const float uv[6], const float* normals, float prim_f[12], rt_flat_tri_attr& a) {
    const V3 p0{v[0], v[1], v[2]}, p1{v[3], v[4], v[5]}, p2{v[6], v[7], v[8]};
    const V3 e0 = sub(p1, p0), e1 = sub(p2, p0);             // triangle.h:25-26
    std::memset(prim_f, 0, sizeof(float) * 12);
    put(prim_f, p0); put(prim_f + 3, e0); put(prim_f + 6, e1);
    std::memset(&a, 0, sizeof(a));
    std::memcpy(a.uv, uv, sizeof(float) * 6);
    a.d00 = dot(e0, e0); a.d01 = dot(e0, e1); a.d11 = dot(e1, e1);     // triangle.h:28-30
    a.inv_denom = 1.0f / (a.d00 * a.d11 - a.d01 * a.d01);              // triangle.h:31
    if (normals) { std::memcpy(a.n, normals, sizeof(float) * 9); a.vertex_normals = 1; }
}

namespace {

struct Flattener {
    const HostScene& hs;
    FlatScene& out;
    std::string& err;
    std::vector<rt_flat_xop> stack;                 // current wrapper chain, outermost first
    std::map<std::vector<uint32_t>, uint32_t> chain_ix;
    uint32_t next_leaf = 0;
    bool in_medium = false;

    Flattener(const HostScene& h, FlatScene& o, std::string& e) : hs(h), out(o), err(e) {}

    uint32_t current_chain() {
        if (stack.empty()) return 0;
        std::vector<uint32_t> key(stack.size() * 4);
        std::memcpy(key.data(), stack.data(), stack.size() * sizeof(rt_flat_xop));
        auto it = chain_ix.find(key);
        if (it != chain_ix.end()) return it->second;
        rt_flat_chain c{(uint32_t)out.xops.size(), (uint32_t)stack.size(), 0, 0};
        out.xops.insert(out.xops.end(), stack.begin(), stack.end());
        out.chains.push_back(c);
        return chain_ix[key] = (uint32_t)out.chains.size() - 1;
    }

    void leaf(uint32_t type, const float* f, int nf, int mat) {
        rt_flat_prim p{};
        p.type = type; p.material = (uint32_t)mat; p.chain = current_chain(); p.leaf_id = next_leaf++;
        std::memcpy(p.f, f, sizeof(float) * nf);
        (in_medium ? out.bprims : out.prims).push_back(p);
    }

    bool walk(int ix, int depth) {
        if (ix < 0 || ix >= (int)hs.nodes.size()) { err = "flatten: dangling hittable index"; return false; }
        if (depth > 64) { err = "flatten: graph deeper than 64 levels (cycle?)"; return false; }
        const Node& n = hs.nodes[ix];
        switch (n.kind) {
        case NodeKind::Sphere: leaf(RT_PRIM_SPHERE, n.f, 4, n.mat); return true;
        case NodeKind::MovingSphere: leaf(RT_PRIM_MOVING_SPHERE, n.f, 9, n.mat); return true;
        case NodeKind::XYRect: leaf(RT_PRIM_XY_RECT, n.f, 5, n.mat); return true;
        case NodeKind::XZRect: leaf(RT_PRIM_XZ_RECT, n.f, 5, n.mat); return true;
        case NodeKind::YZRect: leaf(RT_PRIM_YZ_RECT, n.f, 5, n.mat); return true;
        case NodeKind::Triangle: {
            float f[12];
            std::memcpy(f, n.f, sizeof(f));
            const uint32_t a = (uint32_t)n.tri_attr;
            std::memcpy(&f[9], &a, 4);
            leaf(RT_PRIM_TRIANGLE, f, 12, n.mat);
            return true;
        }
        case NodeKind::Box: {
            // six sides in the order box.h:17-24 creates them
            const float *p0 = n.f, *p1 = n.f + 3;
            const float s0[5] = {p0[0], p1[0], p0[1], p1[1], p1[2]}, s1[5] = {p0[0], p1[0], p0[1], p1[1], p0[2]};
            const float s2[5] = {p0[0], p1[0], p0[2], p1[2], p1[1]}, s3[5] = {p0[0], p1[0], p0[2], p1[2], p0[1]};
            const float s4[5] = {p0[1], p1[1], p0[2], p1[2], p1[0]}, s5[5] = {p0[1], p1[1], p0[2], p1[2], p0[0]};
            leaf(RT_PRIM_XY_RECT, s0, 5, n.mat); leaf(RT_PRIM_XY_RECT, s1, 5, n.mat);
            leaf(RT_PRIM_XZ_RECT, s2, 5, n.mat); leaf(RT_PRIM_XZ_RECT, s3, 5, n.mat);
            leaf(RT_PRIM_YZ_RECT, s4, 5, n.mat); leaf(RT_PRIM_YZ_RECT, s5, 5, n.mat);
            return true;
        }
        case NodeKind::Translate: {
            stack.push_back(rt_flat_xop{RT_XOP_TRANSLATE, n.f[0], n.f[1], n.f[2]});
            const bool ok = walk(n.child, depth + 1);
            stack.pop_back();
            return ok;
        }
        case NodeKind::RotateY: {
            stack.push_back(rt_flat_xop{RT_XOP_ROTATE_Y, n.f[0], n.f[1], 0.0f});
            const bool ok = walk(n.child, depth + 1);
            stack.pop_back();
            return ok;
        }
        case NodeKind::Group:
            for (int k : n.kids) if (!walk(k, depth + 1)) return false;
            return true;
        case NodeKind::Medium: {
            if (in_medium) { err = "flatten: constant_medium nested inside a medium boundary is not supported"; return false; }
            rt_flat_medium m{};
            m.chain = current_chain();
            m.leaf_id = next_leaf++;
            m.neg_inv_density = n.f[0];
            // the phase function is an isotropic material owned by the medium (constant_medium.h:17, :22)
            rt_flat_material iso{RT_MAT_ISOTROPIC, (uint32_t)n.mat, 0.0f, 0};
            out.materials.push_back(iso);
            m.material = (uint32_t)out.materials.size() - 1;
            m.bprim_first = (uint32_t)out.bprims.size();
            std::vector<rt_flat_xop> saved;
            saved.swap(stack);          // boundary chains are relative to the medium's frame
            in_medium = true;
            const bool ok = walk(n.child, depth + 1);
            in_medium = false;
            stack.swap(saved);
            m.bprim_count = (uint32_t)out.bprims.size() - m.bprim_first;
            out.media.push_back(m);
            return ok;
        }
        }
        err = "flatten: unknown node kind";
        return false;
    }
};

} // namespace

int HostScene::flatten(std::string& err) {
    if (world < 0) { err = "flatten: no world set (rt_set_world)"; return RT_ERR_STATE; }
    if (!has_camera) { err = "flatten: no camera set (rt_set_camera)"; return RT_ERR_STATE; }
    flat = FlatScene{};
    flat.materials = materials;
    flat.textures = textures;
    flat.perlins = perlins;
    flat.images = images;
    flat.image_bytes = image_bytes;
    flat.tri_attrs = tri_attrs;
    flat.chains.push_back(rt_flat_chain{0, 0, 0, 0});    // chain 0 = identity
    Flattener fl(*this, flat, err);
    if (!fl.walk(world, 0)) return RT_ERR_UNSUPPORTED;
    rt_flat_header& h = flat.hdr;
    h.magic = RT_FLAT_MAGIC;
    h.num_leaf_ids = fl.next_leaf;
    h.n_prims = (uint32_t)flat.prims.size(); h.n_bprims = (uint32_t)flat.bprims.size();
    h.n_tri_attrs = (uint32_t)flat.tri_attrs.size(); h.n_xops = (uint32_t)flat.xops.size();
    h.n_chains = (uint32_t)flat.chains.size(); h.n_media = (uint32_t)flat.media.size();
    h.n_materials = (uint32_t)flat.materials.size(); h.n_textures = (uint32_t)flat.textures.size();
    h.n_perlins = (uint32_t)flat.perlins.size(); h.n_images = (uint32_t)flat.images.size();
    h.image_bytes = (uint32_t)flat.image_bytes.size();
    std::memcpy(h.background, background, sizeof(background));
    h.camera = camera;
    flat.valid = true;
    static std::atomic<uint64_t> next_generation{1};
    flat.generation = next_generation.fetch_add(1);
    return (int)fl.next_leaf;
}

namespace {
inline size_t pad16(size_t n) { return (n + 15) & ~size_t(15); }
template <class T> size_t arr_bytes(const std::vector<T>& v) { return pad16(v.size() * sizeof(T)); }
template <class T> void put_arr(uint8_t*& p, const std::vector<T>& v) {
    if (!v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(T));
    p += arr_bytes(v);
}
} // namespace

size_t HostScene::export_size() const {
    if (!flat.valid) return 0;
    return sizeof(rt_flat_header) + arr_bytes(flat.prims) + arr_bytes(flat.bprims) + arr_bytes(flat.tri_attrs) +
           arr_bytes(flat.xops) + arr_bytes(flat.chains) + arr_bytes(flat.media) + arr_bytes(flat.materials) +
           arr_bytes(flat.textures) + arr_bytes(flat.perlins) + arr_bytes(flat.images) + arr_bytes(flat.image_bytes);
}

int HostScene::export_blob(void* buf, size_t cap, std::string& err) const {
    const size_t need = export_size();
    if (!need) { err = "export: scene not flattened"; return RT_ERR_STATE; }
    if (cap < need) { err = "export: buffer too small"; return RT_ERR_INVALID; }
    std::memset(buf, 0, need);
    uint8_t* p = (uint8_t*)buf;
    std::memcpy(p, &flat.hdr, sizeof(flat.hdr)); p += sizeof(flat.hdr);
    put_arr(p, flat.prims); put_arr(p, flat.bprims); put_arr(p, flat.tri_attrs); put_arr(p, flat.xops);
    put_arr(p, flat.chains); put_arr(p, flat.media); put_arr(p, flat.materials); put_arr(p, flat.textures);
    put_arr(p, flat.perlins); put_arr(p, flat.images); put_arr(p, flat.image_bytes);
    return RT_OK;
}

} // namespace rtb
