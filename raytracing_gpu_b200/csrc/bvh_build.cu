// bvh_build.cu — BVH construction (replaces the reference's one-thread bvh_node build, bvh.h:163-346).
//
// On the device, no CUB/thrust:
//   prim_bounds   world AABB per primitive (local box pushed through its translate/rotate_y chain) + centroid
//                 bounds of the scene (warp/block reduce, ordered-int atomics)
//   morton        63-bit Morton code of the centroid (21 bits per axis)
//   radix sort    LSD, 8 bits per pass, (u64 key, u32 value), 3 kernels per pass: per-tile histogram,
//                 single-block exclusive scan of the digit-major table, stable scatter (warp match ranking);
//                 passes whose digit is constant over all keys are skipped
// then ONE hierarchy over the sorted leaves (rt_build_settings::builder chooses, rt_build_stats::builder reports):
//   SAH sweep     scenes up to 65 536 primitives: full-sweep SAH, top-down, on the GPU (level-synchronous, block scans);
//                 11-16 % fewer node visits per ray than PLOC on every scene measured (profiles/r01_v8_sah.md).  Round 1
//                 ran this sweep on host threads; that builder is kept as the checker (same tree, bit for bit)
//   PLOC          everything larger, on the device: locally-ordered clustering, 1 M triangles in 4 ms; the top of the
//                 tree (the last 32 k clusters) goes to the SAH sweep
//   Karras        Morton-prefix hierarchy + atomic bottom-up refit (A/B only)
// and the primitives are permuted into leaf order.  Nodes are 64 bytes (the parent holds both child boxes), root = 0.
#include "device_scene.cuh"

#include <algorithm>
#include <cfloat>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace rtb {

namespace {

// ---------------------------------------------------------------------------------------------------
// bounds
// ---------------------------------------------------------------------------------------------------
struct Box { float3 lo, hi; };

__device__ __forceinline__ void grow(Box& b, float x, float y, float z) {
    b.lo.x = fminf(b.lo.x, x); b.lo.y = fminf(b.lo.y, y); b.lo.z = fminf(b.lo.z, z);
    b.hi.x = fmaxf(b.hi.x, x); b.hi.y = fmaxf(b.hi.y, y); b.hi.z = fmaxf(b.hi.z, z);
}

__device__ __forceinline__ Box empty_box() {
    Box b; b.lo = make_float3(FLT_MAX, FLT_MAX, FLT_MAX); b.hi = make_float3(-FLT_MAX, -FLT_MAX, -FLT_MAX); return b;
}

// object-space box of a primitive over the shutter interval [t_lo, t_hi]
__device__ Box local_box(const DPrim& p, float t_lo, float t_hi) {
    Box b = empty_box();
    const uint32_t type = p.type_chain & 15u;
    if (type == RT_PRIM_SPHERE) {
        const float r = fabsf(p.a.w);
        grow(b, p.a.x - r, p.a.y - r, p.a.z - r); grow(b, p.a.x + r, p.a.y + r, p.a.z + r);
    } else if (type == RT_PRIM_MOVING_SPHERE) {
        const float r = fabsf(p.c0);
        const float ts[2] = {t_lo, t_hi};
        for (int k = 0; k < 2; k++) {   // the centre moves linearly: the extremes are at the interval ends
            const float w = (ts[k] - p.b.z) / (p.b.w - p.b.z);
            const float cx = p.a.x + w * (p.a.w - p.a.x), cy = p.a.y + w * (p.b.x - p.a.y), cz = p.a.z + w * (p.b.y - p.a.z);
            grow(b, cx - r, cy - r, cz - r); grow(b, cx + r, cy + r, cz + r);
        }
    } else if (type == RT_PRIM_TRIANGLE) {
        grow(b, p.a.x, p.a.y, p.a.z);
        grow(b, p.a.x + p.a.w, p.a.y + p.b.x, p.a.z + p.b.y);
        grow(b, p.a.x + p.b.z, p.a.y + p.b.w, p.a.z + p.c0);
    } else {
        const float a0 = fminf(p.a.x, p.a.y), a1 = fmaxf(p.a.x, p.a.y), b0 = fminf(p.a.z, p.a.w), b1 = fmaxf(p.a.z, p.a.w), k = p.b.x;
        if (type == RT_PRIM_XY_RECT) { grow(b, a0, b0, k); grow(b, a1, b1, k); }
        else if (type == RT_PRIM_XZ_RECT) { grow(b, a0, k, b0); grow(b, a1, k, b1); }   // correct z extent (the reference's is degenerate)
        else { grow(b, k, a0, b0); grow(b, k, a1, b1); }
    }
    return b;
}

// object space -> world: undo the chain from the innermost wrapper outwards
__device__ Box world_box(const SceneView& sv, const DPrim& p, float t_lo, float t_hi) {
    Box b = local_box(p, t_lo, t_hi);
    const uint32_t chain = p.type_chain >> 4;
    const uint32_t type = p.type_chain & 15u;
    if (chain && type <= RT_PRIM_MOVING_SPHERE) {
        // a rotated sphere is a sphere: move the CENTRE(S) through the chain and keep the radius, instead of bounding the
        // rotated bounding cube (22 % wider in x and z at the 15 degrees of the final scene's 1000-sphere instance)
        const float r = fabsf(type == RT_PRIM_SPHERE ? p.a.w : p.c0);
        const rt_flat_chain c = sv.chains[chain];
        b = empty_box();
        const int n_centres = type == RT_PRIM_SPHERE ? 1 : 2;
        for (int k = 0; k < n_centres; k++) {
            float cx = p.a.x, cy = p.a.y, cz = p.a.z;
            if (type == RT_PRIM_MOVING_SPHERE) {
                const float w = ((k ? t_hi : t_lo) - p.b.z) / (p.b.w - p.b.z);
                cx = p.a.x + w * (p.a.w - p.a.x); cy = p.a.y + w * (p.b.x - p.a.y); cz = p.a.z + w * (p.b.y - p.a.z);
            }
            for (int j = (int)c.count - 1; j >= 0; j--) {
                const rt_flat_xop op = sv.xops[c.first + j];
                if (op.kind == RT_XOP_TRANSLATE) { cx += op.a; cy += op.b; cz += op.c; }
                else { const float x = op.b * cx + op.a * cz, z = -op.a * cx + op.b * cz; cx = x; cz = z; }   // hittable.h:131-132
            }
            grow(b, cx - r, cy - r, cz - r); grow(b, cx + r, cy + r, cz + r);
        }
    } else if (chain) {
        const rt_flat_chain c = sv.chains[chain];
        for (int k = (int)c.count - 1; k >= 0; k--) {
            const rt_flat_xop op = sv.xops[c.first + k];
            if (op.kind == RT_XOP_TRANSLATE) {
                b.lo.x += op.a; b.lo.y += op.b; b.lo.z += op.c; b.hi.x += op.a; b.hi.y += op.b; b.hi.z += op.c;
            } else {
                Box r = empty_box();
                for (int corner = 0; corner < 8; corner++) {
                    const float x = (corner & 1) ? b.hi.x : b.lo.x, y = (corner & 2) ? b.hi.y : b.lo.y, z = (corner & 4) ? b.hi.z : b.lo.z;
                    grow(r, op.b * x + op.a * z, y, -op.a * x + op.b * z);   // hittable.h:131-132
                }
                b = r;
            }
        }
    }
    // conservative padding: flat primitives get thickness, and every box absorbs the rounding of the
    // transforms above (absolute + relative)
    const float m = fmaxf(fmaxf(fmaxf(fabsf(b.lo.x), fabsf(b.hi.x)), fmaxf(fabsf(b.lo.y), fabsf(b.hi.y))), fmaxf(fabsf(b.lo.z), fabsf(b.hi.z)));
    const float pad = 1e-4f + 4e-6f * m;
    b.lo.x -= pad; b.lo.y -= pad; b.lo.z -= pad; b.hi.x += pad; b.hi.y += pad; b.hi.z += pad;
    return b;
}

// world box of one constant_medium: its boundary primitives through their own chains, then through the medium's
__device__ void apply_chain_to_box(const SceneView& sv, uint32_t chain, Box& b) {
    if (!chain) return;
    const rt_flat_chain c = sv.chains[chain];
    for (int k = (int)c.count - 1; k >= 0; k--) {
        const rt_flat_xop op = sv.xops[c.first + k];
        if (op.kind == RT_XOP_TRANSLATE) {
            b.lo.x += op.a; b.lo.y += op.b; b.lo.z += op.c; b.hi.x += op.a; b.hi.y += op.b; b.hi.z += op.c;
        } else {
            Box r = empty_box();
            for (int corner = 0; corner < 8; corner++) {
                const float x = (corner & 1) ? b.hi.x : b.lo.x, y = (corner & 2) ? b.hi.y : b.lo.y, z = (corner & 4) ? b.hi.z : b.lo.z;
                grow(r, op.b * x + op.a * z, y, -op.a * x + op.b * z);   // hittable.h:131-132
            }
            b = r;
        }
    }
}

__global__ void k_medium_bounds(SceneView sv, DMedium* media, uint32_t n_media, float t_lo, float t_hi) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_media) return;
    DMedium med = media[m];
    Box all = empty_box();
    for (uint32_t i = 0; i < med.bprim_count; i++) {
        const Box b = world_box(sv, sv.bprims[med.bprim_first + i], t_lo, t_hi);   // through the boundary primitive's own chain (padded)
        all.lo.x = fminf(all.lo.x, b.lo.x); all.lo.y = fminf(all.lo.y, b.lo.y); all.lo.z = fminf(all.lo.z, b.lo.z);
        all.hi.x = fmaxf(all.hi.x, b.hi.x); all.hi.y = fmaxf(all.hi.y, b.hi.y); all.hi.z = fmaxf(all.hi.z, b.hi.z);
    }
    apply_chain_to_box(sv, med.chain, all);
    const float mx = fmaxf(fmaxf(fmaxf(fabsf(all.lo.x), fabsf(all.hi.x)), fmaxf(fabsf(all.lo.y), fabsf(all.hi.y))), fmaxf(fabsf(all.lo.z), fabsf(all.hi.z)));
    const float pad = 1e-3f + 1.6e-5f * mx;      // two more transform levels of rounding than a primitive box
    media[m].lo = make_float4(all.lo.x - pad, all.lo.y - pad, all.lo.z - pad, 0.f);
    media[m].hi = make_float4(all.hi.x + pad, all.hi.y + pad, all.hi.z + pad, 0.f);
}

// order-preserving float <-> uint mapping for atomicMin / atomicMax
__device__ __forceinline__ uint32_t f2ord(float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ __forceinline__ float ord2f(uint32_t o) {
    const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

__global__ void k_prim_bounds(SceneView sv, const DPrim* prims, uint32_t n, float t_lo, float t_hi,
                              float4* box_lo, float4* box_hi, uint32_t* scene_ord /* [6]: min xyz, max xyz of centroids */) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float cx = FLT_MAX, cy = FLT_MAX, cz = FLT_MAX, dx = -FLT_MAX, dy = -FLT_MAX, dz = -FLT_MAX;
    if (i < n) {
        const Box b = world_box(sv, prims[i], t_lo, t_hi);
        box_lo[i] = make_float4(b.lo.x, b.lo.y, b.lo.z, 0.f);
        box_hi[i] = make_float4(b.hi.x, b.hi.y, b.hi.z, 0.f);
        cx = dx = 0.5f * (b.lo.x + b.hi.x); cy = dy = 0.5f * (b.lo.y + b.hi.y); cz = dz = 0.5f * (b.lo.z + b.hi.z);
    }
    for (int o = 16; o > 0; o >>= 1) {
        cx = fminf(cx, __shfl_xor_sync(0xffffffffu, cx, o)); cy = fminf(cy, __shfl_xor_sync(0xffffffffu, cy, o));
        cz = fminf(cz, __shfl_xor_sync(0xffffffffu, cz, o)); dx = fmaxf(dx, __shfl_xor_sync(0xffffffffu, dx, o));
        dy = fmaxf(dy, __shfl_xor_sync(0xffffffffu, dy, o)); dz = fmaxf(dz, __shfl_xor_sync(0xffffffffu, dz, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&scene_ord[0], f2ord(cx)); atomicMin(&scene_ord[1], f2ord(cy)); atomicMin(&scene_ord[2], f2ord(cz));
        atomicMax(&scene_ord[3], f2ord(dx)); atomicMax(&scene_ord[4], f2ord(dy)); atomicMax(&scene_ord[5], f2ord(dz));
    }
}

// ---------------------------------------------------------------------------------------------------
// Morton codes
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t spread21(uint32_t v) {   // 21 bits -> every third bit of 63
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const float4* box_lo, const float4* box_hi, uint32_t n, const uint32_t* scene_ord, uint64_t* keys, uint32_t* vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float lx = ord2f(scene_ord[0]), ly = ord2f(scene_ord[1]), lz = ord2f(scene_ord[2]);
    const float ex = fmaxf(ord2f(scene_ord[3]) - lx, 1e-30f), ey = fmaxf(ord2f(scene_ord[4]) - ly, 1e-30f), ez = fmaxf(ord2f(scene_ord[5]) - lz, 1e-30f);
    const float4 lo = box_lo[i], hi = box_hi[i];
    const float scale = 2097151.0f;   // 2^21 - 1
    const uint32_t qx = (uint32_t)fminf(fmaxf((0.5f * (lo.x + hi.x) - lx) / ex * scale, 0.f), scale);
    const uint32_t qy = (uint32_t)fminf(fmaxf((0.5f * (lo.y + hi.y) - ly) / ey * scale, 0.f), scale);
    const uint32_t qz = (uint32_t)fminf(fmaxf((0.5f * (lo.z + hi.z) - lz) / ez * scale, 0.f), scale);
    keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    vals[i] = i;
}

// ---------------------------------------------------------------------------------------------------
// radix sort: (u64 key, u32 value), LSD 8 bits per pass
// ---------------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_ROUNDS = 8;                         // keys per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_ROUNDS;  // keys per block

// blockIdx.y = batch: independent sorts of the same length laid out back to back (the three axis orders of the SAH sweep)
__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const uint64_t* keys, uint32_t n, int shift, uint32_t* counts /* [256][nblocks] */) {
    __shared__ uint32_t hist[256];
    keys += (size_t)blockIdx.y * n; counts += (size_t)blockIdx.y * 256 * gridDim.x;
    hist[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_TILE;
    for (int r = 0; r < SORT_ROUNDS; r++) {
        const uint32_t i = base + r * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[(uint32_t)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    counts[threadIdx.x * gridDim.x + blockIdx.x] = hist[threadIdx.x];
}

// exclusive scan of `len` counters by one block; also reports whether one digit holds all n keys
__global__ void __launch_bounds__(1024) k_sort_scan(uint32_t* counts, uint32_t len, uint32_t nblocks, uint32_t n, uint32_t* pass_is_trivial) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    __shared__ uint32_t trivial;
    counts += (size_t)blockIdx.y * len; pass_is_trivial += blockIdx.y;
    if (threadIdx.x == 0) { carry = 0; trivial = 0; }
    __syncthreads();
    // trivial-pass detection: digit d is constant iff sum_b counts[d][b] == n
    for (uint32_t d = threadIdx.x; d < 256; d += blockDim.x) {
        uint32_t s = 0;
        for (uint32_t b = 0; b < nblocks; b++) s += counts[d * nblocks + b];
        if (s == n) trivial = 1;
    }
    __syncthreads();
    for (uint32_t base = 0; base < len; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < len ? counts[i] : 0;
        uint32_t x = v;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((threadIdx.x & 31) >= o) x += y; }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x];
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if (threadIdx.x >= o) w += y; }
            warp_sums[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t warp_prefix = (threadIdx.x >> 5) ? warp_sums[(threadIdx.x >> 5) - 1] : 0;
        if (i < len) counts[i] = carry + warp_prefix + x - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry += warp_prefix + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) *pass_is_trivial = trivial;
}

__global__ void __launch_bounds__(SORT_THREADS) k_sort_scatter(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                                                               uint32_t n, int shift, const uint32_t* offsets /* scanned [256][nblocks] */,
                                                               const uint32_t* pass_is_trivial) {
    __shared__ uint32_t digit_base[256];                       // running global offset of this block per digit
    __shared__ uint32_t warp_count[SORT_THREADS / 32][256];    // per round
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = blockIdx.x * SORT_TILE;
    keys_in += (size_t)blockIdx.y * n; vals_in += (size_t)blockIdx.y * n; keys_out += (size_t)blockIdx.y * n; vals_out += (size_t)blockIdx.y * n;
    offsets += (size_t)blockIdx.y * 256 * gridDim.x; pass_is_trivial += blockIdx.y;
    if (*pass_is_trivial) {   // constant digit: order is unchanged, plain copy keeps the ping-pong consistent
        for (int r = 0; r < SORT_ROUNDS; r++) {
            const uint32_t i = base + r * SORT_THREADS + threadIdx.x;
            if (i < n) { keys_out[i] = keys_in[i]; vals_out[i] = vals_in[i]; }
        }
        return;
    }
    digit_base[threadIdx.x] = offsets[threadIdx.x * gridDim.x + blockIdx.x];
    for (int r = 0; r < SORT_ROUNDS; r++) {
        for (int w = 0; w < SORT_THREADS / 32; w++) warp_count[w][threadIdx.x] = 0;
        __syncthreads();
        const uint32_t i = base + r * SORT_THREADS + threadIdx.x;
        const bool valid = i < n;
        uint64_t key = 0; uint32_t val = 0; uint32_t digit = 256 + lane;   // invalid lanes never match anyone
        if (valid) { key = keys_in[i]; val = vals_in[i]; digit = (uint32_t)(key >> shift) & 255u; }
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        const uint32_t rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank_in_warp == 0) warp_count[warp][digit] = __popc(peers);
        __syncthreads();
        if (valid) {
            uint32_t before = 0;
            for (uint32_t w = 0; w < warp; w++) before += warp_count[w][digit];
            const uint32_t dst = digit_base[digit] + before + rank_in_warp;
            keys_out[dst] = key; vals_out[dst] = val;
        }
        __syncthreads();
        uint32_t total = 0;
        for (int w = 0; w < SORT_THREADS / 32; w++) total += warp_count[w][threadIdx.x];
        digit_base[threadIdx.x] += total;
        __syncthreads();
    }
}

} // namespace

namespace {

// ---------------------------------------------------------------------------------------------------
// Karras hierarchy over sorted keys
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ int delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);   // duplicates: fall back to the index
    return __clzll((long long)(a ^ b));
}

// children encoding: >= 0 internal node index, < 0 leaf (~leaf index in sorted order)
__global__ void k_karras(const uint64_t* keys, int n, int* left, int* right, int* parent_internal, int* parent_leaf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    if (lo == gamma) { left[i] = ~gamma; parent_leaf[gamma] = i; } else { left[i] = gamma; parent_internal[gamma] = i; }
    if (hi == gamma + 1) { right[i] = ~(gamma + 1); parent_leaf[gamma + 1] = i; } else { right[i] = gamma + 1; parent_internal[gamma + 1] = i; }
    if (i == 0) parent_internal[0] = -1;
}

// ---------------------------------------------------------------------------------------------------
// bottom-up refit
// ---------------------------------------------------------------------------------------------------
__global__ void k_refit(int n, const uint32_t* sorted_vals, const float4* prim_lo, const float4* prim_hi, const int* left, const int* right,
                        const int* parent_internal, const int* parent_leaf, float4* leaf_lo, float4* leaf_hi, float4* node_lo,
                        float4* node_hi, uint32_t* visit) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t p = sorted_vals[j];
    leaf_lo[j] = prim_lo[p];
    leaf_hi[j] = prim_hi[p];
    if (n == 1) return;
    __threadfence();
    int cur = parent_leaf[j];
    while (cur >= 0) {
        if (atomicAdd(&visit[cur], 1u) == 0) return;   // first arrival leaves; the second one owns the node
        __threadfence();
        const int l = left[cur], r = right[cur];
        const volatile float4* llo = l < 0 ? &leaf_lo[~l] : &node_lo[l];
        const volatile float4* lhi = l < 0 ? &leaf_hi[~l] : &node_hi[l];
        const volatile float4* rlo = r < 0 ? &leaf_lo[~r] : &node_lo[r];
        const volatile float4* rhi = r < 0 ? &leaf_hi[~r] : &node_hi[r];
        node_lo[cur] = make_float4(fminf(llo->x, rlo->x), fminf(llo->y, rlo->y), fminf(llo->z, rlo->z), 0.f);
        node_hi[cur] = make_float4(fmaxf(lhi->x, rhi->x), fmaxf(lhi->y, rhi->y), fmaxf(lhi->z, rhi->z), 0.f);
        __threadfence();
        cur = parent_internal[cur];
    }
}

// ---------------------------------------------------------------------------------------------------
// traversal layout + primitive permutation
// ---------------------------------------------------------------------------------------------------
__global__ void k_emit_nodes(int n, const int* left, const int* right, const float4* leaf_lo, const float4* leaf_hi,
                             const float4* node_lo, const float4* node_hi, BNode* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n == 1) {
        if (i == 0) {   // a single primitive: root with one real child and one empty slot
            BNode b;
            b.n0 = make_float4(leaf_lo[0].x, leaf_hi[0].x, leaf_lo[0].y, leaf_hi[0].y);
            b.n1 = make_float4(FLT_MAX, -FLT_MAX, FLT_MAX, -FLT_MAX);
            b.n2 = make_float4(leaf_lo[0].z, leaf_hi[0].z, FLT_MAX, -FLT_MAX);
            b.child0 = ~0; b.child1 = ~0; b.pad0 = b.pad1 = 0;
            out[0] = b;
        }
        return;
    }
    if (i >= n - 1) return;
    const int l = left[i], r = right[i];
    const float4 llo = l < 0 ? leaf_lo[~l] : node_lo[l], lhi = l < 0 ? leaf_hi[~l] : node_hi[l];
    const float4 rlo = r < 0 ? leaf_lo[~r] : node_lo[r], rhi = r < 0 ? leaf_hi[~r] : node_hi[r];
    BNode b;
    b.n0 = make_float4(llo.x, lhi.x, llo.y, lhi.y);
    b.n1 = make_float4(rlo.x, rhi.x, rlo.y, rhi.y);
    b.n2 = make_float4(llo.z, lhi.z, rlo.z, rhi.z);
    b.child0 = l; b.child1 = r; b.pad0 = b.pad1 = 0;
    out[i] = b;
}

__global__ void k_permute_prims(uint32_t n, const uint32_t* sorted_vals, const DPrim* in, const uint32_t* leaf_ids_in, DPrim* out, uint32_t* leaf_ids_out) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t p = sorted_vals[j];
    out[j] = in[p];
    leaf_ids_out[j] = leaf_ids_in[p];
}

// ---------------------------------------------------------------------------------------------------
// PLOC hierarchy (parallel locally-ordered clustering, after Meister & Bittner 2017) over the Morton-sorted
// primitives: every cluster looks RADIUS neighbours to each side for the partner that minimises the surface area
// of the merged box; mutual nearest neighbours merge; the survivors are compacted IN ORDER and the loop repeats
// until one cluster is left.  Bottom-up, so boxes are exact unions (no refit pass) and the tree is SAH-driven
// rather than dictated by the Morton prefix structure (Karras), which costs ~1.5-2x the node visits on meshes.
// Internal nodes are numbered from the top: the last merge is node 0 (the root), so the upper levels of the tree
// sit together at the start of the node array.
// ---------------------------------------------------------------------------------------------------
constexpr int PLOC_MAX_RADIUS = 64;
constexpr uint32_t PLOC_SWEEP_MAX = 131072;                     // scenes up to this many primitives are built with several radii
constexpr int PLOC_BLOCK = 256;
constexpr int SCAN_ITEMS = 8;                                   // per thread
constexpr int SCAN_TILE = PLOC_BLOCK * SCAN_ITEMS;

__device__ __forceinline__ float half_area_union(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y), dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_ploc_init(int n, const uint32_t* sorted_vals, const float4* prim_lo, const float4* prim_hi, int* cid, float4* lo, float4* hi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = sorted_vals[i];
    cid[i] = ~i;                       // leaf: ~(index in sorted order) = ~(primitive index after the permutation)
    float4 l = prim_lo[p];
    l.w = 0.f;                         // .w of lo carries the subtree depth
    lo[i] = l;
    hi[i] = prim_hi[p];
}

__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_nn(const float4* lo, const float4* hi, int n, int* nn, int radius) {
    __shared__ float4 slo[PLOC_BLOCK + 2 * PLOC_MAX_RADIUS], shi[PLOC_BLOCK + 2 * PLOC_MAX_RADIUS];
    const int first = blockIdx.x * PLOC_BLOCK - radius;
    for (int t = threadIdx.x; t < PLOC_BLOCK + 2 * radius; t += PLOC_BLOCK) {
        const int j = first + t;
        if (j >= 0 && j < n) { slo[t] = lo[j]; shi[t] = hi[j]; }
    }
    __syncthreads();
    const int i = blockIdx.x * PLOC_BLOCK + threadIdx.x;
    if (i >= n) return;
    const int c = threadIdx.x + radius;
    const float4 mlo = slo[c], mhi = shi[c];
    float best = INFINITY;
    int best_j = -1;
    #pragma unroll 4
    for (int d = -radius; d <= radius; d++) {      // ascending j and a strict '<': ties go to the lower index
        const int j = i + d;
        if (d == 0 || j < 0 || j >= n) continue;
        const float a = half_area_union(mlo, mhi, slo[c + d], shi[c + d]);
        if (a < best) { best = a; best_j = j; }
    }
    nn[i] = best_j;
}

// flags of cluster i packed for one scan: low word = stays in the list, high word = creates a node (the lower index
// of a mutual pair); the other member of the pair is absorbed
__device__ __forceinline__ unsigned long long ploc_flags(const int* nn, int n, int i) {
    if (i >= n) return 0ull;
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    const bool absorbed = mutual && i > j, leader = mutual && i < j;
    return (absorbed ? 0ull : 1ull) | (leader ? (1ull << 32) : 0ull);
}

// block-local exclusive scan of the packed flags; block totals go to sums[]
__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_scan_block(const int* nn, int n, unsigned long long* local, unsigned long long* sums) {
    __shared__ unsigned long long warp_tot[PLOC_BLOCK / 32];
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    unsigned long long v[SCAN_ITEMS], run = 0;
    #pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = run; run += ploc_flags(nn, n, base + k); }
    unsigned long long x = run;                               // inclusive scan of the per-thread totals
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_tot[warp] = x;
    __syncthreads();
    unsigned long long before = 0;
    for (int w = 0; w < warp; w++) before += warp_tot[w];
    const unsigned long long excl = before + x - run;
    #pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) local[base + k] = excl + v[k];
    if (threadIdx.x == PLOC_BLOCK - 1) sums[blockIdx.x] = before + x;
}

// exclusive scan of the block totals by one block; the grand total lands in sums[nblocks]
__global__ void __launch_bounds__(1024) k_ploc_scan_sums(unsigned long long* sums, int nblocks) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nblocks; base += 1024) {
        const int i = base + threadIdx.x;
        const unsigned long long v = i < nblocks ? sums[i] : 0ull;
        unsigned long long x = v;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        unsigned long long before = 0;
        for (int w = 0; w < warp; w++) before += warp_tot[w];
        if (i < nblocks) sums[i] = carry + before + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[nblocks] = carry;
}

// merge mutual pairs, emit their traversal node, compact the survivors in order
__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_merge(const int* nn, int n, const unsigned long long* local, const unsigned long long* sums,
                                                          const int* cid_in, const float4* lo_in, const float4* hi_in, int* cid_out, float4* lo_out,
                                                          float4* hi_out, BNode* nodes, int next_node /* index of the first node this round creates */, float* area_sum) {
    const int i = blockIdx.x * PLOC_BLOCK + threadIdx.x;
    if (i >= n) return;
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    if (mutual && i > j) return;                               // absorbed by its partner
    const unsigned long long pos = local[i] + sums[i / SCAN_TILE];
    const uint32_t p = (uint32_t)pos;
    float4 alo = lo_in[i], ahi = hi_in[i];
    int id = cid_in[i];
    if (mutual) {
        const float4 blo = lo_in[j], bhi = hi_in[j];
        const int node = next_node - (int)(pos >> 32);         // numbered downwards: the final merge is node 0
        BNode b;
        b.n0 = make_float4(alo.x, ahi.x, alo.y, ahi.y);
        b.n1 = make_float4(blo.x, bhi.x, blo.y, bhi.y);
        b.n2 = make_float4(alo.z, ahi.z, blo.z, bhi.z);
        b.child0 = id; b.child1 = cid_in[j]; b.pad0 = b.pad1 = 0;
        nodes[node] = b;
        atomicAdd(area_sum, half_area_union(alo, ahi, blo, bhi));          // SAH bookkeeping: sum of internal-node areas
        alo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), fmaxf(alo.w, blo.w) + 1.f);
        ahi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
        id = node;
    }
    cid_out[p] = id; lo_out[p] = alo; hi_out[p] = ahi;
}

// ---------------------------------------------------------------------------------------------------
// Full-sweep SAH hierarchy on the GPU (the default for scenes up to 65 536 primitives, and the top of PLOC trees)
//
// Top-down, one primitive per leaf: at every node every split position along every axis (primitives ordered by box
// centre) is scored with area(L) * |L| + area(R) * |R| and the best of the 3 x (n - 1) candidates wins — the sweep
// round 1 ran on host threads (SahBuilder below, kept as its checker), restated for the device:
//   * the three orders are sorted ONCE (radix sort of the centre keys, the three axes batched in one launch);
//     a split stably partitions the other two, so every subtree owns the same range of all three arrays;
//   * the tree is built level by level: the segments (= tree nodes) of one level sit in a queue, blocks of 256 threads
//     claim them one at a time; a segment's prefix / suffix box areas are chunked block scans (box union is exact and
//     associative, so the scan yields the very floats the sequential sweep computes), the best split a block-wide
//     lexicographic (cost, axis, position) minimum — the first strict minimum the host loop would keep;
//   * a subtree over k primitives owns k - 1 consecutive nodes in preorder (left child = me + 1, right child =
//     me + |L|), so node indices need no allocation, and both child boxes are known at split time: the parent writes
//     its finished 64-byte node immediately — no bottom-up pass;
//   * no host round trips: the level loop is a fixed number of launches (a level with an empty queue exits at once),
//     the host reads one control block at the end.
// The result equals the host sweep's tree bit for bit (tests/test_gpu.py::test_gpu_sah_sweep_equals_host_sweep).
// ---------------------------------------------------------------------------------------------------
constexpr int SAH_THREADS = 256, SAH_ITEMS = 4, SAH_CHUNK = SAH_THREADS * SAH_ITEMS, SAH_WARPS = SAH_THREADS / 32;
constexpr int SAH_MAX_LEVELS = RT_STACK_DEPTH;

struct B6 { float lx, ly, lz, hx, hy, hz; };
__device__ __forceinline__ B6 b6_empty() { return B6{INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY}; }
__device__ __forceinline__ B6 b6_unite(const B6& a, const B6& b) {
    return B6{fminf(a.lx, b.lx), fminf(a.ly, b.ly), fminf(a.lz, b.lz), fmaxf(a.hx, b.hx), fmaxf(a.hy, b.hy), fmaxf(a.hz, b.hz)};
}
__device__ __forceinline__ float b6_half_area(const B6& b) { const float x = b.hx - b.lx, y = b.hy - b.ly, z = b.hz - b.lz; return x * y + y * z + z * x; }
__device__ __forceinline__ B6 b6_shfl_up(const B6& b, int o) {
    return B6{__shfl_up_sync(0xffffffffu, b.lx, o), __shfl_up_sync(0xffffffffu, b.ly, o), __shfl_up_sync(0xffffffffu, b.lz, o),
              __shfl_up_sync(0xffffffffu, b.hx, o), __shfl_up_sync(0xffffffffu, b.hy, o), __shfl_up_sync(0xffffffffu, b.hz, o)};
}
__device__ __forceinline__ B6 b6_shfl_xor(const B6& b, int o) {
    return B6{__shfl_xor_sync(0xffffffffu, b.lx, o), __shfl_xor_sync(0xffffffffu, b.ly, o), __shfl_xor_sync(0xffffffffu, b.lz, o),
              __shfl_xor_sync(0xffffffffu, b.hx, o), __shfl_xor_sync(0xffffffffu, b.hy, o), __shfl_xor_sync(0xffffffffu, b.hz, o)};
}

struct SahSeg { uint32_t first, count; int me; uint32_t level; };
struct SahCtl {
    uint32_t nseg[SAH_MAX_LEVELS + 2];     // segments queued for level l
    uint32_t cursor[SAH_MAX_LEVELS + 2];   // ... and the next one to claim
    uint32_t depth, too_deep;
    double area_sum;
    float root_lo[3], root_hi[3];
};
struct SahArgs {
    const float4* lo; const float4* hi;    // leaf boxes; lo.w = depth of the subtree below the leaf when has_depth
    const int* leaf_code;                  // child code of leaf i (nullptr: ~i, primitive i)
    int has_depth;
    uint32_t* order[3];                    // the leaves in the three axis orders
    uint32_t* scratch;                     // [n] by position
    uint8_t* goes_right;                   // [n] by leaf
    float* right_area;                     // [n] by position
    SahSeg* queue[2];
    SahCtl* ctl;
    BNode* nodes;
    uint32_t n;
};

__global__ void k_sah_keys(const float4* lo, const float4* hi, uint32_t n, uint64_t* keys /* [3][n] */, uint32_t* vals /* [3][n] */) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 l = lo[i], h = hi[i];
    const float c[3] = {l.x + h.x, l.y + h.y, l.z + h.z};   // the host sweep's sort key: lo + hi, ties by index (the sort is stable)
    for (int a = 0; a < 3; a++) { keys[(size_t)a * n + i] = f2ord(c[a] + 0.0f); vals[(size_t)a * n + i] = i; }   // + 0: -0 and +0 tie, as they compare
}

__global__ void k_sah_init(SahArgs A) {
    SahCtl* c = A.ctl;
    for (int i = threadIdx.x; i < SAH_MAX_LEVELS + 2; i += blockDim.x) { c->nseg[i] = 0; c->cursor[i] = 0; }
    __syncthreads();
    if (threadIdx.x == 0) {
        c->depth = 0; c->too_deep = 0; c->area_sum = 0.0;
        c->nseg[0] = 1;
        A.queue[0][0] = SahSeg{0u, A.n, 0, 0u};
    }
}

// exclusive block scan of one box per thread (union); `total` = union over the block.  s_w: [SAH_WARPS] scratch.
__device__ __forceinline__ B6 sah_block_scan(const B6& v, B6& total, B6* s_w) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    B6 inc = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const B6 y = b6_shfl_up(inc, o); if (lane >= o) inc = b6_unite(inc, y); }
    if (lane == 31) s_w[warp] = inc;
    B6 ex = b6_shfl_up(inc, 1);
    if (lane == 0) ex = b6_empty();
    __syncthreads();
    B6 pre = b6_empty();
    total = b6_empty();
    #pragma unroll
    for (int w = 0; w < SAH_WARPS; w++) { const B6 t = s_w[w]; if (w < warp) pre = b6_unite(pre, t); total = b6_unite(total, t); }
    __syncthreads();
    return b6_unite(pre, ex);
}

__device__ __forceinline__ uint32_t sah_block_scan_u32(uint32_t v, uint32_t& total, uint32_t* s_w) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    uint32_t pre = 0;
    total = 0;
    #pragma unroll
    for (int w = 0; w < SAH_WARPS; w++) { const uint32_t t = s_w[w]; if (w < warp) pre += t; total += t; }
    __syncthreads();
    return pre + inc - v;
}

__device__ __forceinline__ B6 sah_leaf_box(const SahArgs& A, uint32_t leaf) {
    const float4 l = A.lo[leaf], h = A.hi[leaf];
    return B6{l.x, l.y, l.z, h.x, h.y, h.z};
}

// One sweep over the segment `ids[0..count)` in chunks of SAH_CHUNK; FORWARD: element e = 0, 1, ...; else e = count - 1,
// count - 2, ...  `visit(e, box)` sees the union of the elements swept so far, e included.
template <bool FORWARD, class Visit>
__device__ __forceinline__ void sah_sweep(const SahArgs& A, const uint32_t* ids, uint32_t count, B6* s_w, Visit&& visit) {
    B6 carry = b6_empty();
    for (uint32_t c0 = 0; c0 < count; c0 += SAH_CHUNK) {
        B6 loc[SAH_ITEMS];
        B6 run = b6_empty();
        #pragma unroll
        for (int k = 0; k < SAH_ITEMS; k++) {
            const uint32_t s = c0 + threadIdx.x * SAH_ITEMS + k;          // sweep step
            if (s < count) run = b6_unite(run, sah_leaf_box(A, ids[FORWARD ? s : count - 1 - s]));
            loc[k] = run;
        }
        B6 total;
        const B6 before = b6_unite(carry, sah_block_scan(run, total, s_w));
        #pragma unroll
        for (int k = 0; k < SAH_ITEMS; k++) {
            const uint32_t s = c0 + threadIdx.x * SAH_ITEMS + k;
            if (s < count) visit(FORWARD ? s : count - 1 - s, b6_unite(before, loc[k]));
        }
        carry = b6_unite(carry, total);
    }
}

__global__ void __launch_bounds__(SAH_THREADS) k_sah_level(SahArgs A, uint32_t level) {
    __shared__ B6 s_w[SAH_WARPS];
    __shared__ B6 s_lr[SAH_WARPS][2];
    __shared__ uint32_t s_u[SAH_WARPS];
    __shared__ float s_cost[SAH_WARPS];
    __shared__ uint32_t s_key[SAH_WARPS];
    __shared__ uint32_t s_claim, s_best;
    SahCtl* ctl = A.ctl;
    const uint32_t nseg = ctl->nseg[level];
    const SahSeg* in = A.queue[level & 1];
    SahSeg* out = A.queue[(level & 1) ^ 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (;;) {
        if (threadIdx.x == 0) s_claim = atomicAdd(&ctl->cursor[level], 1u);
        __syncthreads();
        const uint32_t si = s_claim;
        if (si >= nseg) return;
        const SahSeg seg = in[si];
        const uint32_t first = seg.first, count = seg.count;
        float* ra = A.right_area + first;

        // ---- score every split of every axis
        float best = INFINITY;
        uint32_t best_key = 0xffffffffu;                       // axis << 30 | split
        for (uint32_t axis = 0; axis < 3; axis++) {
            const uint32_t* ids = A.order[axis] + first;
            sah_sweep<false>(A, ids, count, s_w, [&](uint32_t e, const B6& b) { if (e) ra[e] = b6_half_area(b); });   // area of [e, count)
            __syncthreads();
            sah_sweep<true>(A, ids, count, s_w, [&](uint32_t e, const B6& b) {
                const uint32_t i = e + 1;                          // split after element e: |L| = i
                if (i < count) {
                    const float cost = b6_half_area(b) * (float)i + ra[i] * (float)(count - i);
                    if (cost < best) { best = cost; best_key = axis << 30 | i; }
                }
            });
            __syncthreads();
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float oc = __shfl_xor_sync(0xffffffffu, best, o);
            const uint32_t ok = __shfl_xor_sync(0xffffffffu, best_key, o);
            if (oc < best || (oc == best && ok < best_key)) { best = oc; best_key = ok; }
        }
        if (lane == 0) { s_cost[warp] = best; s_key[warp] = best_key; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float c = s_cost[0]; uint32_t k = s_key[0];
            for (int w = 1; w < SAH_WARPS; w++) if (s_cost[w] < c || (s_cost[w] == c && s_key[w] < k)) { c = s_cost[w]; k = s_key[w]; }
            if (k == 0xffffffffu) k = count / 2;               // no finite cost anywhere: the middle of the x order
            s_best = k;
        }
        __syncthreads();
        const uint32_t best_axis = s_best >> 30, split = s_best & 0x3fffffffu;

        // ---- partition the other two orders (stable), boxes of both sides
        const uint32_t* bid = A.order[best_axis] + first;
        B6 bl = b6_empty(), br = b6_empty();
        for (uint32_t i = threadIdx.x; i < count; i += SAH_THREADS) {
            const uint32_t leaf = bid[i];
            const bool right = i >= split;
            A.goes_right[leaf] = right;
            const B6 b = sah_leaf_box(A, leaf);
            if (right) br = b6_unite(br, b); else bl = b6_unite(bl, b);
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) { bl = b6_unite(bl, b6_shfl_xor(bl, o)); br = b6_unite(br, b6_shfl_xor(br, o)); }
        if (lane == 0) { s_lr[warp][0] = bl; s_lr[warp][1] = br; }
        __syncthreads();
        for (uint32_t axis = 0; axis < 3; axis++) {
            if (axis == best_axis) continue;
            uint32_t* ids = A.order[axis] + first;
            uint32_t* tmp = A.scratch + first;
            uint32_t lefts_before = 0;                          // block-uniform carry
            for (uint32_t c0 = 0; c0 < count; c0 += SAH_CHUNK) {
                uint32_t id[SAH_ITEMS], nl = 0;
                bool r[SAH_ITEMS];
                #pragma unroll
                for (int k = 0; k < SAH_ITEMS; k++) {
                    const uint32_t i = c0 + threadIdx.x * SAH_ITEMS + k;
                    id[k] = i < count ? ids[i] : 0u;
                    r[k] = i < count ? A.goes_right[id[k]] != 0 : true;
                    nl += r[k] ? 0u : 1u;
                }
                uint32_t total;
                uint32_t l = lefts_before + sah_block_scan_u32(nl, total, s_u);
                #pragma unroll
                for (int k = 0; k < SAH_ITEMS; k++) {
                    const uint32_t i = c0 + threadIdx.x * SAH_ITEMS + k;
                    if (i < count) { tmp[r[k] ? split + (i - l) : l] = id[k]; l += r[k] ? 0u : 1u; }
                }
                lefts_before += total;
            }
            __syncthreads();
            for (uint32_t i = threadIdx.x; i < count; i += SAH_THREADS) ids[i] = tmp[i];
            __syncthreads();
        }

        // ---- the node, and its children into the next level's queue
        if (threadIdx.x == 0) {
            B6 L = s_lr[0][0], R = s_lr[0][1];
            for (int w = 1; w < SAH_WARPS; w++) { L = b6_unite(L, s_lr[w][0]); R = b6_unite(R, s_lr[w][1]); }
            const uint32_t nl = split, nr = count - split;
            int c0, c1;
            uint32_t leaf_depth = 0;
            SahSeg kids[2];
            uint32_t nk = 0;
            if (nl == 1) {
                const uint32_t leaf = bid[0];
                c0 = A.leaf_code ? A.leaf_code[leaf] : ~(int)leaf;
                leaf_depth = max(leaf_depth, seg.level + 1 + (A.has_depth ? (uint32_t)A.lo[leaf].w : 0u));
            } else { c0 = seg.me + 1; kids[nk++] = SahSeg{first, nl, seg.me + 1, seg.level + 1}; }
            if (nr == 1) {
                const uint32_t leaf = bid[split];
                c1 = A.leaf_code ? A.leaf_code[leaf] : ~(int)leaf;
                leaf_depth = max(leaf_depth, seg.level + 1 + (A.has_depth ? (uint32_t)A.lo[leaf].w : 0u));
            } else { c1 = seg.me + (int)split; kids[nk++] = SahSeg{first + split, nr, seg.me + (int)split, seg.level + 1}; }
            BNode nd;
            nd.n0 = make_float4(L.lx, L.hx, L.ly, L.hy);
            nd.n1 = make_float4(R.lx, R.hx, R.ly, R.hy);
            nd.n2 = make_float4(L.lz, L.hz, R.lz, R.hz);
            nd.child0 = c0; nd.child1 = c1; nd.pad0 = 0; nd.pad1 = 0;
            A.nodes[seg.me] = nd;
            const B6 u = b6_unite(L, R);
            atomicAdd(&ctl->area_sum, (double)b6_half_area(u));
            if (leaf_depth) atomicMax(&ctl->depth, leaf_depth);
            if (seg.level == 0) { ctl->root_lo[0] = u.lx; ctl->root_lo[1] = u.ly; ctl->root_lo[2] = u.lz; ctl->root_hi[0] = u.hx; ctl->root_hi[1] = u.hy; ctl->root_hi[2] = u.hz; }
            if (nk) {
                if (seg.level + 3 > RT_STACK_DEPTH) ctl->too_deep = 1;   // a degenerate chain: the caller discards the tree
                else {
                    const uint32_t at = atomicAdd(&ctl->nseg[level + 1], nk);
                    for (uint32_t k = 0; k < nk; k++) out[at + k] = kids[k];
                }
            }
        }
        __syncthreads();
    }
}


// ---------------------------------------------------------------------------------------------------
// Collapse of the binary tree into 4-wide nodes with 8-bit child boxes (WNode, device_scene.cuh)
//
// Top-down, level by level, one thread per wide node: the node starts with the two children of its binary node and
// replaces, while it has fewer than four, its internal child of largest surface area by that child's two children (the
// SAH-greedy collapse: the biggest boxes are the ones most rays enter, so they are the ones worth skipping a level for).
// Internal children get consecutive wide-node indices from one atomic counter and go to the next level's queue.
// Quantisation: origin = lo corner of the union, scale 2^e per axis with e the smallest exponent whose 255 steps span
// the extent; lo planes are rounded down, hi planes up (directed-rounding subtraction), so a quantised box always
// contains the fp32 box — the closest hit is unchanged, only a few more boxes are entered.
// ---------------------------------------------------------------------------------------------------
struct WideItem { int bnode, wnode; };
struct WideCtl { uint32_t nq[RT_STACK_DEPTH + 2]; uint32_t n_wide, depth; };

__global__ void k_wide_init(WideCtl* ctl, WideItem* q0) {
    for (int i = threadIdx.x; i < RT_STACK_DEPTH + 2; i += blockDim.x) ctl->nq[i] = 0;
    __syncthreads();
    if (threadIdx.x == 0) { ctl->nq[0] = 1; ctl->n_wide = 1; ctl->depth = 0; q0[0] = WideItem{0, 0}; }
}

__device__ __forceinline__ void bnode_children(const BNode& b, B6 box[2], int code[2]) {
    box[0] = B6{b.n0.x, b.n0.z, b.n2.x, b.n0.y, b.n0.w, b.n2.y};
    box[1] = B6{b.n1.x, b.n1.z, b.n2.z, b.n1.y, b.n1.w, b.n2.w};
    code[0] = b.child0; code[1] = b.child1;
}

__global__ void __launch_bounds__(128) k_wide_level(const BNode* __restrict__ nodes, WNode* __restrict__ wide, const WideItem* __restrict__ q_in,
                                                    WideItem* __restrict__ q_out, WideCtl* ctl, uint32_t level) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nq = ctl->nq[level];
    for (uint32_t it = i; it < nq; it += gridDim.x * blockDim.x) {
        const WideItem item = q_in[it];
        B6 box[4];
        int code[4];
        int n = 2;
        bnode_children(nodes[item.bnode], box, code);
        if (box[1].lx > box[1].hx) n = 1;                       // the single-primitive root: its second slot is empty
        while (n < 4) {
            int k = -1;
            float best = -1.f;
            for (int c = 0; c < n; c++) if (code[c] >= 0) { const float a = b6_half_area(box[c]); if (a > best) { best = a; k = c; } }
            if (k < 0) break;
            B6 cb[2]; int cc[2];
            bnode_children(nodes[code[k]], cb, cc);
            box[k] = cb[0]; code[k] = cc[0];
            box[n] = cb[1]; code[n] = cc[1];
            n++;
        }
        B6 u = box[0];
        for (int c = 1; c < n; c++) u = b6_unite(u, box[c]);
        // scale 2^e per axis: the smallest power of two with 255 * 2^e >= extent (exponent bits straight from the float)
        uint32_t eb[3];
        float inv[3];
        const float ext[3] = {u.hx - u.lx, u.hy - u.ly, u.hz - u.lz};
        for (int a = 0; a < 3; a++) {
#if RT_WIDE_FOLD
            const float step = fmaxf(ext[a], 1e-30f) * (1.0f / 253.0f) * 1.0000002f;     // the grid starts one step below the box: planes at 1 .. 254
#else
            const float step = fmaxf(ext[a], 1e-30f) * (1.0f / 255.0f) * 1.0000002f;     // a hair more than extent / 255
#endif
            uint32_t e = (__float_as_uint(step) >> 23) & 0xffu;
            if (__float_as_uint(step) & 0x7fffffu) e++;                                   // round the step up to a power of two
            e = min(max(e, 1u), 254u);
            eb[a] = e;
            inv[a] = __uint_as_float((254u - e) << 23);                                   // 2^-(e - 127), exact
        }
#if RT_WIDE_FOLD
        const float org[3] = {__fsub_rd(u.lx, __uint_as_float(eb[0] << 23)), __fsub_rd(u.ly, __uint_as_float(eb[1] << 23)), __fsub_rd(u.lz, __uint_as_float(eb[2] << 23))};
        const float guard = 1.0f / 128.0f;     // of a step, on both sides of every plane: covers the folded plane arithmetic of wide_step
#else
        const float org[3] = {u.lx, u.ly, u.lz};
        const float guard = 0.f;
#endif
        uint32_t ql[3] = {0, 0, 0}, qh[3] = {0, 0, 0};
        int cnt_internal = 0;
        for (int c = 0; c < 4; c++) {
            uint32_t l[3] = {255u, 255u, 255u}, h[3] = {0u, 0u, 0u};                     // empty slot: inverted box
            if (c < n) {
                const float lo[3] = {box[c].lx, box[c].ly, box[c].lz}, hi[3] = {box[c].hx, box[c].hy, box[c].hz}, o[3] = {org[0], org[1], org[2]};
                for (int a = 0; a < 3; a++) {
                    const float fl = floorf(__fmul_rd(__fsub_rd(lo[a], o[a]), inv[a]) - guard), fh = ceilf(__fmul_ru(__fsub_ru(hi[a], o[a]), inv[a]) + guard);
                    l[a] = (uint32_t)fminf(fmaxf(fl, 0.f), 255.f);
                    h[a] = (uint32_t)fminf(fmaxf(fh, 0.f), 255.f);
                }
                if (code[c] >= 0) cnt_internal++;
            }
            for (int a = 0; a < 3; a++) { ql[a] |= l[a] << (8 * c); qh[a] |= h[a] << (8 * c); }
        }
        uint32_t base = 0, slot = 0;
        if (cnt_internal) { base = atomicAdd(&ctl->n_wide, (uint32_t)cnt_internal); slot = atomicAdd(&ctl->nq[level + 1], (uint32_t)cnt_internal); }
        RT_DEV_ASSERT(level + 1 < RT_STACK_DEPTH + 2);
        WNode w;
        w.ox = org[0]; w.oy = org[1]; w.oz = org[2];
        w.exps = eb[0] | eb[1] << 8 | eb[2] << 16;
        int k = 0;
        for (int c = 0; c < 4; c++) {
            if (c >= n) { w.child[c] = RT_WIDE_EMPTY; continue; }
            if (code[c] >= 0) { q_out[slot + k] = WideItem{code[c], (int)(base + k)}; w.child[c] = (int)(base + k); k++; }
            else w.child[c] = code[c];
        }
        w.qlo_x = ql[0]; w.qlo_y = ql[1]; w.qlo_z = ql[2]; w.qhi_x = qh[0]; w.qhi_y = qh[1]; w.qhi_z = qh[2]; w.pad0 = 0; w.pad1 = 0;
        wide[item.wnode] = w;
        if (it == 0) atomicMax(&ctl->depth, level + 1);
    }
}

// bump allocator over DeviceScene::scratch; pass 1 (base == nullptr) only measures
struct Arena {
    char* base; size_t off = 0;
    explicit Arena(void* b) : base((char*)b) {}
    template <class T> T* take(size_t count) {
        off = (off + 255) & ~size_t(255);
        T* p = base ? (T*)(base + off) : nullptr;
        off += (count ? count : 1) * sizeof(T);
        return p;
    }
};

} // namespace

int compute_medium_bounds(DeviceScene& ds, DMedium* media, float time_lo, float time_hi, cudaStream_t stream) {
    if (!ds.view.n_media) return RT_OK;
    k_medium_bounds<<<(ds.view.n_media + 63) / 64, 64, 0, stream>>>(ds.view, media, ds.view.n_media, time_lo, time_hi);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

// ---- host full-sweep SAH build: round 1's default, now the checker of the GPU sweep (RT_BUILD_HOST_SAH) ---------------
// The same algorithm, sequential per subtree; up to 16 host threads near the top.
// Input: boxes in leaf (sorted) order; output: nodes numbered in preorder, root = 0.
namespace {
struct HostBox { float lo[3], hi[3]; };
inline void grow(HostBox& b, const HostBox& o) { for (int a = 0; a < 3; a++) { b.lo[a] = std::min(b.lo[a], o.lo[a]); b.hi[a] = std::max(b.hi[a], o.hi[a]); } }
inline float half_area(const HostBox& b) { const float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2]; return x * y + y * z + z * x; }
const HostBox EMPTY_BOX = {{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};

struct SahBuilder {
    const std::vector<HostBox>& leaf;
    std::vector<uint32_t> order[3], scratch;   // the same primitive set in three orders; every subtree owns one range of all three
    std::vector<uint8_t> goes_right;
    std::vector<float> right_area;
    std::vector<BNode> nodes;                  // preorder: a subtree over k primitives owns k - 1 consecutive nodes, so subtrees
                                               // can be built by different threads without sharing anything but disjoint ranges
    struct Result { HostBox box; int code; double area_sum; uint32_t depth; };

    explicit SahBuilder(const std::vector<HostBox>& l)
        : leaf(l), scratch(l.size()), goes_right(l.size()), right_area(l.size()), nodes(l.size() > 1 ? l.size() - 1 : 1) {
        std::thread sorters[3];
        for (int axis = 0; axis < 3; axis++)
            sorters[axis] = std::thread([this, axis]() {
                order[axis].resize(leaf.size());
                for (uint32_t i = 0; i < leaf.size(); i++) order[axis][i] = i;
                std::sort(order[axis].begin(), order[axis].end(), [&](uint32_t a, uint32_t b) {
                    const float ca = leaf[a].lo[axis] + leaf[a].hi[axis], cb = leaf[b].lo[axis] + leaf[b].hi[axis];
                    return ca < cb || (ca == cb && a < b);
                });
            });
        for (auto& t : sorters) t.join();
    }

    // builds the subtree over range [first, first + count) into nodes[me ...]
    Result build(size_t first, size_t count, int me, uint32_t level) {
        if (count == 1 || level + 2 > RT_STACK_DEPTH) { // (too deep for the traversal stack: the caller discards the tree)
            const uint32_t i = order[0][first];
            return Result{leaf[i], ~(int)i, 0.0, level};
        }
        float best = INFINITY; int best_axis = 0; size_t best_split = count / 2;
        float* ra = right_area.data() + first;
        for (int axis = 0; axis < 3; axis++) {
            const uint32_t* ids = order[axis].data() + first;
            HostBox acc = EMPTY_BOX;
            for (size_t i = count - 1; i > 0; i--) { grow(acc, leaf[ids[i]]); ra[i] = half_area(acc); }
            acc = EMPTY_BOX;
            for (size_t i = 1; i < count; i++) {
                grow(acc, leaf[ids[i - 1]]);
                const float cost = half_area(acc) * (float)i + ra[i] * (float)(count - i);
                if (cost < best) { best = cost; best_axis = axis; best_split = i; }
            }
        }
        for (size_t i = 0; i < count; i++) goes_right[order[best_axis][first + i]] = i >= best_split;
        uint32_t* tmp = scratch.data() + first;
        for (int axis = 0; axis < 3; axis++) {          // stable partition of the other two orders
            if (axis == best_axis) continue;
            uint32_t* ids = order[axis].data() + first;
            size_t l = 0, r = 0;
            for (size_t i = 0; i < count; i++) { if (goes_right[ids[i]]) tmp[r++] = ids[i]; else ids[l++] = ids[i]; }
            std::copy(tmp, tmp + r, ids + l);
        }
        Result r0, r1;
        const int me0 = me + 1, me1 = me + (int)best_split;   // the left subtree owns best_split - 1 nodes
        if (count >= 8192 && level < 4) {                      // big subtrees near the top: one extra thread each (<= 16 in all)
            std::thread left([&]() { r0 = build(first, best_split, me0, level + 1); });
            r1 = build(first + best_split, count - best_split, me1, level + 1);
            left.join();
        } else {
            r0 = build(first, best_split, me0, level + 1);
            r1 = build(first + best_split, count - best_split, me1, level + 1);
        }
        BNode& n = nodes[me];
        const HostBox &b0 = r0.box, &b1 = r1.box;
        n.n0 = make_float4(b0.lo[0], b0.hi[0], b0.lo[1], b0.hi[1]);
        n.n1 = make_float4(b1.lo[0], b1.hi[0], b1.lo[1], b1.hi[1]);
        n.n2 = make_float4(b0.lo[2], b0.hi[2], b1.lo[2], b1.hi[2]);
        n.child0 = r0.code; n.child1 = r1.code; n.pad0 = 0; n.pad1 = 0;
        Result res{b0, me, 0.0, std::max(r0.depth, r1.depth)};
        grow(res.box, b1);
        res.area_sum = r0.area_sum + r1.area_sum + half_area(res.box);
        return res;
    }
};
constexpr uint32_t SAH_MAX_DEFAULT = 65536;
constexpr int PLOC_TOP_DEFAULT = 32768;       // clusters left when PLOC hands the top of the tree to the SAH sweep

struct SahScratch {      // carved from the build arena
    uint64_t *keys_a, *keys_b;
    uint32_t *vals_a, *vals_b, *counts, *trivial, *scratch;
    uint8_t* goes_right;
    float* right_area;
    SahSeg* queue[2];
    SahCtl* ctl;
};
SahScratch carve_sah(Arena& ar, uint32_t n) {
    SahScratch s;
    const uint32_t nblocks = (n + SORT_TILE - 1) / SORT_TILE;
    s.keys_a = ar.take<uint64_t>(3 * (size_t)n); s.keys_b = ar.take<uint64_t>(3 * (size_t)n);
    s.vals_a = ar.take<uint32_t>(3 * (size_t)n); s.vals_b = ar.take<uint32_t>(3 * (size_t)n);
    s.counts = ar.take<uint32_t>(3 * (size_t)256 * nblocks); s.trivial = ar.take<uint32_t>(4);
    s.scratch = ar.take<uint32_t>(n); s.goes_right = ar.take<uint8_t>(n); s.right_area = ar.take<float>(n);
    s.queue[0] = ar.take<SahSeg>(n / 2 + 2); s.queue[1] = ar.take<SahSeg>(n / 2 + 2);
    s.ctl = ar.take<SahCtl>(1);
    return s;
}

// Queues the whole sweep on `stream` (no synchronisation): leaves (lo, hi[, codes, depths in lo.w]) -> nodes[0 .. n - 2].
// The caller reads *ctl afterwards.  `levels_out`: how many level launches were queued.
int sah_sweep_gpu(const SahScratch& S, const float4* lo, const float4* hi, const int* codes, bool has_depth, uint32_t n, BNode* nodes,
                  int sms, cudaStream_t stream, uint32_t* levels_out) {
    const uint32_t nblocks = (n + SORT_TILE - 1) / SORT_TILE, gb = (n + 255) / 256;
    k_sah_keys<<<gb, 256, 0, stream>>>(lo, hi, n, S.keys_a, S.vals_a);
    uint64_t *kin = S.keys_a, *kout = S.keys_b;
    uint32_t *vin = S.vals_a, *vout = S.vals_b;
    for (int pass = 0; pass < 4; pass++) {   // 32-bit keys; the sort is stable, so equal centres keep index order
        k_sort_hist<<<dim3(nblocks, 3), SORT_THREADS, 0, stream>>>(kin, n, 8 * pass, S.counts);
        k_sort_scan<<<dim3(1, 3), 1024, 0, stream>>>(S.counts, 256 * nblocks, nblocks, n, S.trivial);
        k_sort_scatter<<<dim3(nblocks, 3), SORT_THREADS, 0, stream>>>(kin, vin, kout, vout, n, 8 * pass, S.counts, S.trivial);
        std::swap(kin, kout); std::swap(vin, vout);
    }
    SahArgs A;
    A.lo = lo; A.hi = hi; A.leaf_code = codes; A.has_depth = has_depth ? 1 : 0;
    for (int a = 0; a < 3; a++) A.order[a] = vin + (size_t)a * n;
    A.scratch = S.scratch; A.goes_right = S.goes_right; A.right_area = S.right_area;
    A.queue[0] = S.queue[0]; A.queue[1] = S.queue[1]; A.ctl = S.ctl; A.nodes = nodes; A.n = n;
    k_sah_init<<<1, 128, 0, stream>>>(A);
    // Level l holds at most min(2^l, n / 2) segments; SAH trees of real scenes are 1.5-2.5 x log2(n) deep.  Launches beyond the
    // last level find an empty queue and exit; a tree deeper than what is queued here is caught by the caller's check.
    uint32_t log2n = 0;
    while ((1u << log2n) < n) log2n++;
    const uint32_t levels = std::min<uint32_t>(SAH_MAX_LEVELS, 3 * log2n + 8);
    for (uint32_t l = 0; l < levels; l++) {
        const uint32_t cap = l < 31 ? std::min<uint32_t>(1u << l, n / 2 + 1) : n / 2 + 1;
        k_sah_level<<<std::min<uint32_t>(cap, (uint32_t)sms * 8u), SAH_THREADS, 0, stream>>>(A, l);
    }
    RT_CUDA(cudaGetLastError());
    *levels_out = levels;
    return RT_OK;
}

} // namespace

int build_bvh(DeviceScene& ds, const rt_build_settings& bs, const DPrim* d_prims_in, const uint32_t* d_leaf_ids_in, uint32_t n,
              float time_lo, float time_hi, cudaStream_t stream) {
    ds.pending.sah_levels = 0; ds.pending.prims_in = d_prims_in; ds.pending.leaf_in = d_leaf_ids_in; ds.pending.n = n;
    ds.pending.t_lo = time_lo; ds.pending.t_hi = time_hi;
    if (n == 0) { ds.view.n_nodes = 0; ds.build.n_nodes = 0; return RT_OK; }
    cudaEvent_t* ev = ds.ev + 2;   // [0], [1], [8] belong to commit_impl
    const uint32_t nblocks = (n + SORT_TILE - 1) / SORT_TILE;
    const uint32_t tb = 256, gb = (n + tb - 1) / tb;
    const uint32_t n_nodes = n > 1 ? n - 1 : 1;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ds.device);

    // builder choice
    const uint32_t sah_max = bs.sah_max_prims ? bs.sah_max_prims : SAH_MAX_DEFAULT;
    uint32_t choice = bs.builder;
    if (choice > RT_BUILD_GPU_SAH) { set_error("commit: unknown rt_build_settings::builder"); return RT_ERR_INVALID; }
    if (choice == RT_BUILD_AUTO) choice = n <= sah_max ? RT_BUILD_GPU_SAH : RT_BUILD_PLOC;
    if (n == 1) choice = RT_BUILD_KARRAS;

    RT_CUDA(ds.prims.reserve(sizeof(DPrim) * (size_t)n));
    RT_CUDA(ds.prim_leaf_id.reserve(sizeof(uint32_t) * (size_t)n));
    RT_CUDA(ds.nodes.reserve(sizeof(BNode) * (size_t)n_nodes));
    DPrim* prims_out = (DPrim*)ds.prims.p;
    uint32_t* leaf_ids_out = (uint32_t*)ds.prim_leaf_id.p;
    BNode* nodes = (BNode*)ds.nodes.p;

    // scratch: measured first, then carved from the scene's grow-only arena
    float4 *box_lo, *box_hi, *leaf_lo, *leaf_hi, *node_lo, *node_hi;
    uint64_t *keys_a, *keys_b;
    uint32_t *vals_a, *vals_b, *counts, *scene_ord, *trivial, *visit;
    int *left, *right, *parent_internal, *parent_leaf;
    unsigned long long *scan_local, *scan_sums;
    float* area_sum;
    SahScratch S{};
    const int max_sblk = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
    const uint32_t ploc_top = bs.ploc_top ? bs.ploc_top : (uint32_t)PLOC_TOP_DEFAULT;
    auto carve = [&](Arena& ar) {
        box_lo = ar.take<float4>(n); box_hi = ar.take<float4>(n); leaf_lo = ar.take<float4>(n); leaf_hi = ar.take<float4>(n);
        node_lo = ar.take<float4>(n); node_hi = ar.take<float4>(n);
        keys_a = ar.take<uint64_t>(n); keys_b = ar.take<uint64_t>(n); vals_a = ar.take<uint32_t>(n); vals_b = ar.take<uint32_t>(n);
        counts = ar.take<uint32_t>((size_t)256 * nblocks); scene_ord = ar.take<uint32_t>(8); trivial = ar.take<uint32_t>(4);
        visit = ar.take<uint32_t>(n);
        left = ar.take<int>(n); right = ar.take<int>(n); parent_internal = ar.take<int>(n); parent_leaf = ar.take<int>(n);
        scan_local = ar.take<unsigned long long>(n); scan_sums = ar.take<unsigned long long>((size_t)max_sblk + 1);
        area_sum = ar.take<float>(4);
        S = carve_sah(ar, choice == RT_BUILD_PLOC ? std::min<uint32_t>(n, ploc_top + 64u) : n);
    };
    { Arena measure(nullptr); carve(measure); RT_CUDA(ds.scratch.reserve(measure.off + 256)); }
    { Arena ar(ds.scratch.p); carve(ar); }
    RT_CUDA(ds.readback.reserve(sizeof(SahCtl) + 512));
    SahCtl* h_ctl = (SahCtl*)ds.readback.p;
    char* h_misc = (char*)ds.readback.p + ((sizeof(SahCtl) + 63) & ~size_t(63));
    unsigned long long* h_tot = (unsigned long long*)h_misc;              // PLOC round totals
    uint32_t* h_ord = (uint32_t*)(h_misc + 64);                           // initial centroid bounds
    float4* h_root = (float4*)(h_misc + 128);
    float* h_area = (float*)(h_misc + 192);

    const uint32_t ord_init[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
    memcpy(h_ord, ord_init, sizeof(ord_init));
    RT_CUDA(cudaMemcpyAsync(scene_ord, h_ord, sizeof(ord_init), cudaMemcpyHostToDevice, stream));

    RT_CUDA(cudaEventRecord(ev[0], stream));
    k_prim_bounds<<<gb, tb, 0, stream>>>(ds.view, d_prims_in, n, time_lo, time_hi, box_lo, box_hi, scene_ord);
    k_morton<<<gb, tb, 0, stream>>>(box_lo, box_hi, n, scene_ord, keys_a, vals_a);
    RT_CUDA(cudaEventRecord(ev[1], stream));
    uint64_t *kin = keys_a, *kout = keys_b;
    uint32_t *vin = vals_a, *vout = vals_b;
    for (int pass = 0; pass < 8; pass++) {   // 63-bit codes: 8 passes of 8 bits
        const int shift = 8 * pass;
        k_sort_hist<<<nblocks, SORT_THREADS, 0, stream>>>(kin, n, shift, counts);
        k_sort_scan<<<1, 1024, 0, stream>>>(counts, 256 * nblocks, nblocks, n, trivial);
        k_sort_scatter<<<nblocks, SORT_THREADS, 0, stream>>>(kin, vin, kout, vout, n, shift, counts, trivial);
        std::swap(kin, kout); std::swap(vin, vout);
    }
    RT_CUDA(cudaEventRecord(ev[2], stream));
    ds.build.build_rounds = 0; ds.build.ploc_radius = 0; ds.build.bvh_depth = 0; ds.build.sah_internal = 0.f;

    // Morton-prefix hierarchy: the A/B builder, and the last resort for inputs the SAH-driven builders make chains of
    // (thousands of identical boxes): equal keys split by index bits, so its depth is bounded by key + index bits
    auto build_karras = [&]() -> int {
        ds.build.builder = RT_BUILDER_KARRAS;
        ds.build.bvh_depth = 0; ds.build.sah_internal = 0.f; ds.build.build_rounds = 0; ds.build.ploc_radius = 0;
        RT_CUDA(cudaMemsetAsync(visit, 0, sizeof(uint32_t) * n, stream));
        if (n > 1) k_karras<<<gb, tb, 0, stream>>>(kin, (int)n, left, right, parent_internal, parent_leaf);
        RT_CUDA(cudaEventRecord(ev[3], stream));
        k_refit<<<gb, tb, 0, stream>>>((int)n, vin, box_lo, box_hi, left, right, parent_internal, parent_leaf, leaf_lo, leaf_hi, node_lo, node_hi, visit);
        RT_CUDA(cudaEventRecord(ev[4], stream));
        k_emit_nodes<<<gb, tb, 0, stream>>>((int)n, left, right, leaf_lo, leaf_hi, node_lo, node_hi, nodes);
        return RT_OK;
    };

    bool built = false;
    if (choice == RT_BUILD_GPU_SAH) {
        // leaves in Morton order (the order the primitives are stored in): leaf i = sorted position i
        k_ploc_init<<<gb, tb, 0, stream>>>((int)n, vin, box_lo, box_hi, left, leaf_lo, leaf_hi);
        uint32_t levels = 0;
        const int rc = sah_sweep_gpu(S, leaf_lo, leaf_hi, nullptr, false, n, nodes, sms, stream, &levels);
        if (rc < 0) return rc;
        RT_CUDA(cudaMemcpyAsync(h_ctl, S.ctl, sizeof(SahCtl), cudaMemcpyDeviceToHost, stream));
        RT_CUDA(cudaEventRecord(ev[3], stream));
        RT_CUDA(cudaEventRecord(ev[4], stream));
        ds.pending.sah_levels = levels;   // checked by build_bvh_finish after the commit's synchronisation
        ds.build.builder = RT_BUILDER_GPU_SAH;
        built = true;
    } else if (choice == RT_BUILD_HOST_SAH) {
        std::vector<float4> h_lo(n), h_hi(n);
        std::vector<uint32_t> h_ordv(n);
        RT_CUDA(cudaMemcpyAsync(h_lo.data(), box_lo, sizeof(float4) * n, cudaMemcpyDeviceToHost, stream));
        RT_CUDA(cudaMemcpyAsync(h_hi.data(), box_hi, sizeof(float4) * n, cudaMemcpyDeviceToHost, stream));
        RT_CUDA(cudaMemcpyAsync(h_ordv.data(), vin, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, stream));
        RT_CUDA(cudaStreamSynchronize(stream));
        std::vector<HostBox> leaf(n);
        for (uint32_t i = 0; i < n; i++) {
            const float4 l = h_lo[h_ordv[i]], h = h_hi[h_ordv[i]];
            leaf[i] = HostBox{{l.x, l.y, l.z}, {h.x, h.y, h.z}};
        }
        SahBuilder b(leaf);
        const SahBuilder::Result root = b.build(0, n, 0, 0);
        if (root.depth + 2 <= RT_STACK_DEPTH) {
            RT_CUDA(cudaMemcpyAsync(nodes, b.nodes.data(), sizeof(BNode) * n_nodes, cudaMemcpyHostToDevice, stream));
            RT_CUDA(cudaStreamSynchronize(stream));
            RT_CUDA(cudaEventRecord(ev[3], stream));
            RT_CUDA(cudaEventRecord(ev[4], stream));
            ds.build.bvh_depth = root.depth;
            ds.build.sah_internal = (float)(root.area_sum / std::max(half_area(root.box), 1e-30f));
            ds.build.builder = RT_BUILDER_HOST_SAH;
            built = true;
        } else choice = RT_BUILD_PLOC;
    }
    if (built) {
        // nodes are in place (GPU sweep: queued)
    } else if (choice == RT_BUILD_KARRAS) {
        const int rc = build_karras();
        if (rc < 0) return rc;
    } else {
        // PLOC: cluster lists ping-pong between (cid_a, leaf_*) and (cid_b, node_*); `left`/`right` double as cid arrays,
        // `parent_internal` as the nearest-neighbour array.  The search radius trades differently per scene (flat rect
        // soups like a wide window less than triangle meshes do), and a build is ~1 ms per 250 k primitives, so scenes up
        // to PLOC_SWEEP_MAX primitives are built with several radii and the tree with the smallest summed internal-node
        // area (the SAH traversal term) is kept; measured node visits per ray follow that sum (profiles/r01_v5_ploc.md).
        RT_CUDA(cudaEventRecord(ev[3], stream));
        std::vector<int> radii;
        if (bs.ploc_radius) radii.push_back(std::min<int>(PLOC_MAX_RADIUS, std::max<int>(1, (int)bs.ploc_radius)));
        else if (n <= PLOC_SWEEP_MAX) radii = {4, 8, 16, 32};
        else radii = {16};
        BNode* nodes_alt = nullptr;
        if (radii.size() > 1) { RT_CUDA(ds.nodes_alt.reserve(sizeof(BNode) * (size_t)n_nodes)); nodes_alt = (BNode*)ds.nodes_alt.p; }
        const int total_internal = (int)n - 1;
        float best_area = INFINITY;
        BNode* best_nodes = nullptr;
        for (const int radius : radii) {
            BNode* target = best_nodes == nodes ? nodes_alt : nodes;          // never overwrite the best tree so far
            int *cid_in = left, *cid_out = right, *nn = parent_internal;
            float4 *lo_in = leaf_lo, *hi_in = leaf_hi, *lo_out = node_lo, *hi_out = node_hi;
            RT_CUDA(cudaMemsetAsync(area_sum, 0, sizeof(float), stream));
            k_ploc_init<<<gb, tb, 0, stream>>>((int)n, vin, box_lo, box_hi, cid_in, lo_in, hi_in);
            int cur_n = (int)n, created = 0, rounds = 0;
            // Hybrid: agglomeration stops once at most `top_n` clusters are left and the TOP of the tree — the levels every
            // ray walks, and where bounded-radius clustering is weakest — is built over the cluster boxes by the SAH sweep
            // (on the GPU since round 2).  PLOC numbers its nodes from the high indices down, so the top gets [0, clusters - 1).
            // 1 M triangles: 26.7 -> 24.4 node visits per ray, extend -8 % (top of 32 k clusters; 25.0 with 2 k).
            // Scenes below 16 k primitives — PLOC only when asked for — stay pure PLOC for A/B.
            const int top_n = n >= 16384u ? std::min<int>((int)ploc_top, (int)(n / 8u)) : 1;
            while (cur_n > top_n) {
                const int nblk = (cur_n + PLOC_BLOCK - 1) / PLOC_BLOCK, sblk = (cur_n + SCAN_TILE - 1) / SCAN_TILE;
                k_ploc_nn<<<nblk, PLOC_BLOCK, 0, stream>>>(lo_in, hi_in, cur_n, nn, radius);
                k_ploc_scan_block<<<sblk, PLOC_BLOCK, 0, stream>>>(nn, cur_n, scan_local, scan_sums);
                k_ploc_scan_sums<<<1, 1024, 0, stream>>>(scan_sums, sblk);
                k_ploc_merge<<<nblk, PLOC_BLOCK, 0, stream>>>(nn, cur_n, scan_local, scan_sums, cid_in, lo_in, hi_in, cid_out, lo_out, hi_out, target,
                                                              total_internal - 1 - created, area_sum);
                RT_CUDA(cudaMemcpyAsync(h_tot, scan_sums + sblk, sizeof(*h_tot), cudaMemcpyDeviceToHost, stream));
                RT_CUDA(cudaStreamSynchronize(stream));
                const unsigned long long tot = *h_tot;
                const int survivors = (int)(uint32_t)tot, merges = (int)(tot >> 32);
                if (merges <= 0 || survivors != cur_n - merges) { set_error("bvh build: PLOC round made no progress (internal error)"); return RT_ERR_STATE; }
                cur_n = survivors; created += merges; rounds++;
                std::swap(cid_in, cid_out); std::swap(lo_in, lo_out); std::swap(hi_in, hi_out);
            }
            if (created + cur_n - 1 != total_internal) { set_error("bvh build: PLOC node count mismatch (internal error)"); return RT_ERR_STATE; }
            float root_lo[3], root_hi[3];
            float area = 0.f;
            uint32_t depth = 0;
            if (cur_n > 1) {   // top of the tree over the remaining clusters: SAH sweep, leaf codes = cluster ids, depths in lo.w
                uint32_t levels = 0;
                const int rc = sah_sweep_gpu(S, lo_in, hi_in, cid_in, true, (uint32_t)cur_n, target, sms, stream, &levels);
                if (rc < 0) return rc;
                RT_CUDA(cudaMemcpyAsync(h_ctl, S.ctl, sizeof(SahCtl), cudaMemcpyDeviceToHost, stream));
                RT_CUDA(cudaMemcpyAsync(h_area, area_sum, sizeof(float), cudaMemcpyDeviceToHost, stream));
                RT_CUDA(cudaStreamSynchronize(stream));
                if (h_ctl->too_deep || h_ctl->nseg[levels] != 0) continue;
                memcpy(root_lo, h_ctl->root_lo, 12); memcpy(root_hi, h_ctl->root_hi, 12);
                area = *h_area + (float)h_ctl->area_sum;
                depth = h_ctl->depth;
            } else {
                RT_CUDA(cudaMemcpyAsync(h_root, lo_in, sizeof(float4), cudaMemcpyDeviceToHost, stream));
                RT_CUDA(cudaMemcpyAsync(h_root + 1, hi_in, sizeof(float4), cudaMemcpyDeviceToHost, stream));
                RT_CUDA(cudaMemcpyAsync(h_area, area_sum, sizeof(float), cudaMemcpyDeviceToHost, stream));
                RT_CUDA(cudaStreamSynchronize(stream));
                root_lo[0] = h_root[0].x; root_lo[1] = h_root[0].y; root_lo[2] = h_root[0].z;
                root_hi[0] = h_root[1].x; root_hi[1] = h_root[1].y; root_hi[2] = h_root[1].z;
                area = *h_area;
                depth = (uint32_t)h_root[0].w;
            }
            if (depth + 2 > RT_STACK_DEPTH) continue;                         // deeper than the traversal supports: not a candidate
            if (area < best_area) {
                best_area = area; best_nodes = target;
                const float rdx = root_hi[0] - root_lo[0], rdy = root_hi[1] - root_lo[1], rdz = root_hi[2] - root_lo[2];
                ds.build.bvh_depth = depth;
                ds.build.sah_internal = area / fmaxf(rdx * rdy + rdy * rdz + rdz * rdx, 1e-30f);
                ds.build.build_rounds = (uint32_t)rounds;
                ds.build.ploc_radius = (uint32_t)radius;
                ds.build.builder = RT_BUILDER_PLOC;
            }
        }
        if (!best_nodes) {   // every candidate deeper than the traversal supports
            const int rc = build_karras();
            if (rc < 0) return rc;
        } else {
            if (best_nodes == nodes_alt) { std::swap(ds.nodes, ds.nodes_alt); nodes = (BNode*)ds.nodes.p; }
            RT_CUDA(cudaEventRecord(ev[4], stream));
        }
    }
    k_permute_prims<<<gb, tb, 0, stream>>>(n, vin, d_prims_in, d_leaf_ids_in, prims_out, leaf_ids_out);
    RT_CUDA(cudaEventRecord(ev[5], stream));
    RT_CUDA(cudaGetLastError());

    ds.view.prims = prims_out; ds.view.prim_leaf_id = leaf_ids_out; ds.view.nodes = nodes;
    ds.view.n_nodes = n_nodes;
    ds.build.n_nodes = n_nodes;
    return RT_OK;
}

int build_bvh_finish(DeviceScene& ds, const rt_build_settings& bs, cudaStream_t stream) {
    if (!ds.pending.sah_levels) return RT_OK;
    const SahCtl* h_ctl = (const SahCtl*)ds.readback.p;
    const uint32_t levels = ds.pending.sah_levels;
    ds.pending.sah_levels = 0;
    if (h_ctl->too_deep || h_ctl->nseg[levels] != 0) {
        // deeper than the launches queued (or than the traversal supports): thousands of identical boxes.  PLOC's balanced
        // merges take over — a second build, paid only by such scenes.
        rt_build_settings again = bs;
        again.builder = RT_BUILD_PLOC;
        const int rc = build_bvh(ds, again, ds.pending.prims_in, ds.pending.leaf_in, ds.pending.n, ds.pending.t_lo, ds.pending.t_hi, stream);
        if (rc < 0) return rc;
        RT_CUDA(cudaStreamSynchronize(stream));
        return RT_OK;
    }
    ds.build.bvh_depth = h_ctl->depth;
    const float rx = h_ctl->root_hi[0] - h_ctl->root_lo[0], ry = h_ctl->root_hi[1] - h_ctl->root_lo[1], rz = h_ctl->root_hi[2] - h_ctl->root_lo[2];
    ds.build.sah_internal = (float)(h_ctl->area_sum / std::max(rx * ry + ry * rz + rz * rx, 1e-30f));
    ds.build.build_rounds = h_ctl->depth;
    return RT_OK;
}

constexpr uint32_t WIDE_MIN_PRIMS = 65536;   // RT_WIDE_AUTO: scenes from this size on walk the 4-wide compressed tree

int build_wide(DeviceScene& ds, const rt_build_settings& bs, cudaStream_t stream) {
    ds.view.wide = nullptr; ds.view.n_wide = 0;
    ds.build.n_wide_nodes = 0;
    const uint32_t n = ds.view.n_prims;
    if (bs.wide_nodes > RT_WIDE_ON) { set_error("commit: unknown rt_build_settings::wide_nodes"); return RT_ERR_INVALID; }
    const bool want = bs.wide_nodes == RT_WIDE_ON || (bs.wide_nodes == RT_WIDE_AUTO && n >= WIDE_MIN_PRIMS);
    if (!want || n == 0) return RT_OK;
    const uint32_t n_nodes = ds.view.n_nodes;
    RT_CUDA(ds.wide.reserve(sizeof(WNode) * (size_t)n_nodes));
    // the collapse runs after the build's synchronisation: the build arena is free again
    Arena measure(nullptr);
    measure.take<WideItem>(n_nodes); measure.take<WideItem>(n_nodes); measure.take<WideCtl>(1);
    RT_CUDA(ds.scratch.reserve(measure.off + 256));
    Arena ar(ds.scratch.p);
    WideItem* q[2] = {ar.take<WideItem>(n_nodes), ar.take<WideItem>(n_nodes)};
    WideCtl* ctl = ar.take<WideCtl>(1);
    RT_CUDA(ds.readback.reserve(sizeof(WideCtl) + 64));
    WideCtl* h_ctl = (WideCtl*)ds.readback.p;
    RT_CUDA(cudaEventRecord(ds.ev[6], stream));
    k_wide_init<<<1, 128, 0, stream>>>(ctl, q[0]);
    // a wide level consumes one to two binary levels: at most depth + 1 of them (Karras trees report no depth: the bound)
    const uint32_t levels = ds.build.bvh_depth ? ds.build.bvh_depth + 1 : RT_STACK_DEPTH;
    for (uint32_t l = 0; l < levels; l++) {
        const uint32_t cap = l < 24 ? std::min<uint32_t>(1u << std::min<uint32_t>(2 * l, 30u), n_nodes) : n_nodes;   // a level holds at most 4^l nodes
        k_wide_level<<<std::max<uint32_t>(1u, std::min<uint32_t>((cap + 127) / 128, 148u * 16u)), 128, 0, stream>>>((const BNode*)ds.nodes.p, (WNode*)ds.wide.p, q[l & 1], q[(l & 1) ^ 1], ctl, l);
    }
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(h_ctl, ctl, sizeof(WideCtl), cudaMemcpyDeviceToHost, stream));
    RT_CUDA(cudaEventRecord(ds.ev[7], stream));
    RT_CUDA(cudaStreamSynchronize(stream));
    if (h_ctl->nq[levels] != 0) { set_error("bvh build: wide collapse did not finish (internal error)"); return RT_ERR_STATE; }
    ds.view.wide = (const WNode*)ds.wide.p;
    ds.build.n_wide_nodes = h_ctl->n_wide; ds.view.n_wide = h_ctl->n_wide;
    ds.build.wide_depth = h_ctl->depth;
    cudaEventElapsedTime(&ds.build.ms_wide, ds.ev[6], ds.ev[7]);
    return RT_OK;
}

} // namespace rtb
