// device_scene.cuh — scene data as it lives in HBM, shared by all CUDA translation units.
//
// Layout rules: every record is a multiple of 16 B and is fetched with 128-bit loads; primitives are stored in
// BVH leaf order so a leaf's primitives are contiguous.  Whole scenes are a few MB at most (<= ~70 MB with the
// largest image texture), i.e. L2-resident on B200 (126 MB); traversal traffic is L1/L2, not HBM.
#pragma once

#include <cuda_runtime.h>
#include <cstdint>
#include <string>

#include "../../include/rt_b200.h"

namespace rtb {

// 48-byte device primitive (3 x 128-bit).  Parameter order = rt_flat_prim::f (include/rt_flat.h).
struct __align__(16) DPrim {
    float4 a;            // f[0..3]
    float4 b;            // f[4..7]
    float c0;            // f[8]
    uint32_t attr;       // triangle: index into tri_attrs
    uint32_t type_chain; // rt_prim_type | chain << 4
    uint32_t material;
};
static_assert(sizeof(DPrim) == 48, "DPrim must be 48 bytes");

struct __align__(16) DMedium {   // 64 B
    uint32_t chain, bprim_first, bprim_count, material;
    uint32_t leaf_id;
    float neg_inv_density;
    uint32_t pad0, pad1;
    float4 lo, hi;   // world-space box of the boundary (padded), filled on the device at commit: a ray whose LINE misses it
                     // cannot cross the boundary at any t, so constant_medium::hit's first boundary query (-inf, +inf) fails
};

// Binary BVH node in the traversal layout: the PARENT holds both children's boxes, so one 64-byte fetch
// (4 x 128-bit) decides both descents.  child >= 0: internal node index; child < 0: ~child = primitive index.
struct __align__(16) BNode {
    float4 n0;   // c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y
    float4 n1;   // c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y
    float4 n2;   // c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z
    int child0, child1;
    int pad0, pad1;
};
static_assert(sizeof(BNode) == 64, "BNode must be 64 bytes");

struct __align__(16) DPerlin {
    float4 ranvec[256];   // xyz used
    uint8_t perm[3][256];
};

struct __align__(16) DImage {
    cudaTextureObject_t tex;  // RGBA8, point sampling, unnormalised coordinates
    uint32_t width, height;
};

// Everything a kernel needs to intersect and shade; passed by value (fits in the 4 KB parameter space).
struct SceneView {
    const DPrim* prims;            // surface primitives, BVH leaf order
    const uint32_t* prim_leaf_id;  // DFS leaf id per primitive (tie-breaks, first-hit output)
    const DPrim* bprims;           // medium boundary primitives
    const DMedium* media;
    const uint32_t* active_media;  // indices of the media that can report a hit at all (see commit_impl)
    const rt_flat_tri_attr* tri_attrs;
    const rt_flat_xop* xops;
    const rt_flat_chain* chains;
    const rt_flat_material* materials;
    const rt_flat_texture* textures;
    const DPerlin* perlins;
    const DImage* images;
    const BNode* nodes;            // nodes[0] = root (n_prims >= 1)
    uint32_t n_prims, n_media, n_nodes, n_active_media;
    rt_flat_camera camera;
    float background[3];
    uint32_t semantics;            // rt_semantics bits (rt_b200.h); 0 = the reference's behaviour
};

#define RT_STACK_DEPTH 96   // traversal stack entries (traverse.cuh); Karras trees are bounded by key + index bits, PLOC trees are checked at build

struct DeviceScene {
    int device = -1;
    SceneView view{};
    // owned allocations
    void* d_prims = nullptr; void* d_prim_leaf_id = nullptr; void* d_bprims = nullptr; void* d_media = nullptr;
    void* d_tri_attrs = nullptr; void* d_xops = nullptr; void* d_chains = nullptr; void* d_materials = nullptr;
    void* d_textures = nullptr; void* d_perlins = nullptr; void* d_images = nullptr; void* d_nodes = nullptr;
    void* d_active_media = nullptr;
    cudaArray_t* image_arrays = nullptr;
    cudaTextureObject_t* image_objs = nullptr;
    uint32_t n_images = 0;
    rt_build_stats build{};
    struct RenderPool* pool = nullptr;   // wavefront state, created lazily by the renderer
};

// error plumbing -----------------------------------------------------------------------------------
void set_error(const std::string& e);
bool cuda_fail(cudaError_t e, const char* what, const char* file, int line);   // true (and sets error) on failure
#define RT_CUDA(call) do { if (rtb::cuda_fail((call), #call, __FILE__, __LINE__)) return RT_ERR_CUDA; } while (0)

// bvh_build.cu: builds the LBVH over `prims` (device, flatten order) and returns them reordered.
int build_lbvh(DeviceScene& ds, const DPrim* d_prims_in, const uint32_t* d_leaf_ids_in, uint32_t n,
               float time_lo, float time_hi, cudaStream_t stream);

// bvh_build.cu: fills DMedium::lo / hi of every medium (in place, device)
int compute_medium_bounds(DeviceScene& ds, float time_lo, float time_hi, cudaStream_t stream);

// wavefront.cu
void render_pool_destroy(struct RenderPool* p);

} // namespace rtb
