// device_scene.cuh — scene data as it lives in HBM, shared by all CUDA translation units.
//
// Layout rules: every record is a multiple of 16 B and is fetched with 128-bit loads; primitives are stored in
// BVH leaf order so a leaf's primitives are contiguous.  Whole scenes are a few MB at most (<= ~70 MB with the
// largest image texture), i.e. L2-resident on B200 (126 MB); traversal traffic is L1/L2, not HBM.
#pragma once

#include <cuda_runtime.h>
#include <climits>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"

// Checked build (tools/build_variant.sh checked -DRT_CHECKED=1): device-side assertions on every index the kernels compute
// themselves — pool slots, stack depths, continuation records, queue slots.  compute-sanitizer is closed on this pool
// (profiles/r02_sanitizer.md), so this is the bounds check the round can run.
#if RT_CHECKED
#include <cstdio>
#define RT_DEV_ASSERT(cond) do { if (!(cond)) { printf("RT_DEV_ASSERT failed: %s (%s:%d) block %d thread %d\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, (int)threadIdx.x); __trap(); } } while (0)
#else
#define RT_DEV_ASSERT(cond) do { } while (0)
#endif

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

// wide_step's plane arithmetic and the builder's quantisation must agree (traverse.cuh, bvh_build.cu): 1 = folded planes + guard band
#ifndef RT_WIDE_FOLD
#define RT_WIDE_FOLD 1
#endif

// 4-wide BVH node with 8-bit child boxes (north_star (1): "wide-node SoA layout with quantised child AABBs and 128-bit
// vectorised node loads"): 64 bytes = four 128-bit loads decide FOUR descents (BNode: two).  A child's box is
// origin + q * 2^exp per axis, q in 0..255, rounded outwards; the planes are stored SoA (one 32-bit word holds the four
// children's bytes of one plane), so choosing near / far planes by the sign of the ray direction is one select per axis
// for all four children.  child >= 0: wide node index; < 0: ~primitive; RT_WIDE_EMPTY: unused slot (its box is inverted).
// Built by collapsing the binary tree (bvh_build.cu k_wide_level); used for scenes whose binary tree outgrows the L1.
struct __align__(16) WNode {
    float ox, oy, oz;              // origin = lo corner of the union of the children
    uint32_t exps;                 // biased float exponents of the three scales: ex | ey << 8 | ez << 16
    int child[4];
    uint32_t qlo_x, qlo_y, qlo_z, qhi_x;   // byte c of a word = child c
    uint32_t qhi_y, qhi_z, pad0, pad1;
};
static_assert(sizeof(WNode) == 64, "WNode must be 64 bytes");
#define RT_WIDE_EMPTY INT_MIN

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
    const WNode* wide;             // non-null: the traversal walks the 4-wide compressed tree instead (wide[0] = root)
    uint32_t n_prims, n_media, n_nodes, n_active_media;
    uint32_t n_wide;               // wide nodes in use (0 without the wide tree)
    rt_flat_camera camera;
    float background[3];
    uint32_t semantics;            // rt_semantics bits (rt_b200.h); 0 = the reference's behaviour
};

#define RT_STACK_DEPTH 96   // deepest binary tree the builders accept (a traversal stack never holds more entries than the tree has levels)

// Grow-only device / pinned-host buffers: a re-commit of a scene reuses every allocation of the previous one
// (cudaMalloc / cudaFree synchronise the device; round 1 paid ~40 of them per commit).
bool alloc_fault_due();   // device_api.cu: rt_debug_fail_allocation's countdown (tests inject an out-of-memory error)
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        const size_t want = bytes + bytes / 4 + 256;
        const cudaError_t e = alloc_fault_due() ? cudaErrorMemoryAllocation : cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct PinnedBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        const cudaError_t e = cudaHostAlloc(&p, bytes + 256, cudaHostAllocDefault);
        if (e == cudaSuccess) cap = bytes + 256;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct DeviceImage {
    cudaArray_t array = nullptr;
    cudaTextureObject_t tex = 0;
    uint32_t width = 0, height = 0;
};

struct DeviceScene {
    int device = -1;
    SceneView view{};
    // owned allocations, all grow-only and reused by the next commit on this device
    DevBuf tables;        // every small table of the scene + the primitives in flatten order: ONE H2D copy per commit
    DevBuf prims, prim_leaf_id, nodes, nodes_alt, wide;   // build outputs
    DevBuf scratch;       // build arena
    DevBuf query;         // rt_first_hit staging
    DevBuf accum;         // rt_render's accumulation buffer
    PinnedBuf staging;    // host image of `tables` (filled once per flattening)
    PinnedBuf rgba;       // RGBA8 expansion of every image texture (filled once per flattening)
    PinnedBuf readback;   // small results the host waits for (PLOC round totals, build control block)
    std::vector<DeviceImage> images;
    cudaEvent_t ev[10] = {};              // commit / build phase timers, created once
    cudaStream_t stream = nullptr;        // commit stream (non-blocking)
    uint64_t staged_generation = 0;       // FlatScene::generation the staging buffers were filled from
    uint64_t committed_generation = 0;    // ... and the one the device copy describes (0 = none)
    uint32_t committed_semantics = 0;
    rt_build_stats build{};
    struct { uint32_t sah_levels = 0; const DPrim* prims_in = nullptr; const uint32_t* leaf_in = nullptr; uint32_t n = 0; float t_lo = 0, t_hi = 0; } pending;   // build_bvh -> build_bvh_finish
    struct RenderPool* pool = nullptr;   // wavefront state, created lazily by the renderer
};

// error plumbing -----------------------------------------------------------------------------------
void set_error(const std::string& e);
bool cuda_fail(cudaError_t e, const char* what, const char* file, int line);   // true (and sets error) on failure
#define RT_CUDA(call) do { if (rtb::cuda_fail((call), #call, __FILE__, __LINE__)) return RT_ERR_CUDA; } while (0)

// bvh_build.cu: queues the build of the BVH over `prims` (device, flatten order) into ds.prims / ds.prim_leaf_id / ds.nodes.
int build_bvh(DeviceScene& ds, const rt_build_settings& bs, const DPrim* d_prims_in, const uint32_t* d_leaf_ids_in, uint32_t n,
              float time_lo, float time_hi, cudaStream_t stream);
// bvh_build.cu: after the commit's one synchronisation — reads the build's control block, fills the statistics, and rebuilds
// with PLOC in the one case the sweep gives up on (a degenerate chain deeper than RT_STACK_DEPTH)
int build_bvh_finish(DeviceScene& ds, const rt_build_settings& bs, cudaStream_t stream);
// bvh_build.cu: collapses the finished binary tree into ds.wide (4-wide, 8-bit boxes) where the settings / scene size ask
// for it; synchronises the stream once more.  Sets ds.view.wide and ds.build.n_wide_nodes.
int build_wide(DeviceScene& ds, const rt_build_settings& bs, cudaStream_t stream);

// bvh_build.cu: fills DMedium::lo / hi of every medium (in place, device)
int compute_medium_bounds(DeviceScene& ds, DMedium* media, float time_lo, float time_hi, cudaStream_t stream);

// wavefront.cu
void render_pool_destroy(struct RenderPool* p);

// device_api.cu / wavefront.cu: pieces rt_render_multi (multi_gpu.cu) drives per device
int commit_device_scene(rt_scene* s, DeviceScene& ds, int device, const rt_build_settings& bs);
int render_device_scene(DeviceScene& ds, const rt_render_settings& st, float* d_accum, cudaStream_t stream, rt_render_stats* stats);
DeviceScene* device_scene_new();
bool device_scene_describes(rt_scene* s, const DeviceScene& ds);   // committed from the scene's current flattening and semantics
std::string last_error_copy();

// device_api.cu: the scene's device copy, provided it describes the scene as it is now (RT_ERR_STATE otherwise)
int device_scene_current(rt_scene* s, DeviceScene** out, const char* who);

} // namespace rtb
