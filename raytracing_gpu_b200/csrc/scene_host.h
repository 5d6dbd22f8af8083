// scene_host.h — host-side scene graph recording and flattening (no CUDA).
//
// The reference builds its world with device-side `new` inside <<<1,1>>> kernels (scenes.h:82-621).
// Here the same constructor calls are *recorded* as nodes on the host, then resolved by flatten() into
// the tables of include/rt_flat.h that the GPU consumes.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"

namespace rtb {

enum class NodeKind : uint8_t {
    Sphere, MovingSphere, XYRect, XZRect, YZRect, Box, Triangle, Translate, RotateY, Medium, Group
};

struct Node {
    NodeKind kind;
    int mat = -1;        // material (leaves), isotropic texture (Medium)
    int child = -1;      // Translate / RotateY / Medium
    float f[12] = {0};   // leaf parameters in rt_flat_prim order; Box: p0, p1; Translate: offset; RotateY: sin, cos;
                         // Medium: neg_inv_density
    int tri_attr = -1;   // Triangle
    std::vector<int> kids; // Group
};

struct FlatScene {
    rt_flat_header hdr{};
    std::vector<rt_flat_prim> prims, bprims;
    std::vector<rt_flat_tri_attr> tri_attrs;
    std::vector<rt_flat_xop> xops;
    std::vector<rt_flat_chain> chains;
    std::vector<rt_flat_medium> media;
    std::vector<rt_flat_material> materials;
    std::vector<rt_flat_texture> textures;
    std::vector<rt_flat_perlin> perlins;
    std::vector<rt_flat_image> images;
    std::vector<uint8_t> image_bytes;
    bool valid = false;
    uint64_t generation = 0;   // unique per successful flatten(): the device copy records which flattening it describes
};

struct HostScene {
    std::vector<rt_flat_texture> textures;
    std::vector<rt_flat_material> materials;
    std::vector<rt_flat_perlin> perlins;
    std::vector<rt_flat_image> images;
    std::vector<uint8_t> image_bytes;
    std::vector<rt_flat_tri_attr> tri_attrs;
    std::vector<Node> nodes;
    int world = -1;
    float background[3] = {0, 0, 0};
    rt_flat_camera camera{};
    bool has_camera = false;
    uint32_t semantics = 0;   // rt_semantics bits; device-side only (not part of the exported blob)
    FlatScene flat;

    // returns number of leaf ids, or a negative rt_status with `err` set
    int flatten(std::string& err);
    size_t export_size() const;
    int export_blob(void* buf, size_t cap, std::string& err) const;
};

// camera(lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist, t0, t1) — camera.h:18-47, same float
// expression order.
void make_camera(rt_flat_camera& c, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov,
                 float aspect, float aperture, float focus_dist, float t0, float t1);

// triangle constructor precomputation — triangle.h:19-41
void make_triangle(const float v[9], const float uv[6], const float* normals, float prim_f[12], rt_flat_tri_attr& attr);

} // namespace rtb
