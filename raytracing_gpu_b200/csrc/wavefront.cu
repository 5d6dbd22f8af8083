// wavefront.cu — streaming wavefront path tracer (replaces the render megakernel + color_f, render.h:55-113).
//
// Path state lives in HBM as SoA in TWO pools (ping-pong).  Every iteration reads pool `cur` and writes the
// survivors COMPACTED into pool `next`, so every global access of the hot kernels is a coalesced stream —
// there is no slot indirection and no index queue:
//
//   pool layout      bounced paths occupy positions [0, n_bounce), freshly generated camera paths occupy
//                    [capacity - n_fresh, capacity).  Keeping the two apart means extend warps are either all
//                    camera rays of neighbouring pixels (coherent) or all bounced rays.
//   extend           persistent warps claim 32 consecutive paths, traverse the BVH + media (render.h:63) and
//                    write an 8-byte hit record + a 1-byte material class at the same position.  Leaf tests are
//                    parked and run warp-wide (traverse.cuh).
//   shade            persistent WARPS claim tiles of 256 consecutive paths, rank them by material class with ballots
//                    (8 rounds of 32 paths, a round shades one class), evaluate hit attributes / textures / scatter,
//                    add thr*L of terminated paths to the accumulation buffer (float atomics), REGENERATE
//                    terminated paths from the global sample cursor, and append survivors / new camera paths to
//                    pool `next`.  Whether a path continues is known from its class byte alone (extend adds a
//                    last-bounce bit), so a warp allocates for its tile BEFORE it shades — three atomics per 256
//                    paths whose latency hides behind the ranking — and no barrier, block-wide sort or
//                    single-thread allocator is left (round 1: 37 % of k_shade's stall samples were barriers).
//   turnover         one thread: roll the counters and publish the live-path count to pinned host memory.
// The pools stay full until the samples run out; the host never synchronises inside a frame (it polls the
// live count a few iterations behind).
#include "device_scene.h"
#include "device_scene.cuh"
#include "shade.cuh"
#include "traverse.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace rtb {

// class byte written by extend: bits 0-2 = class, bit 3 = last bounce (the path cannot continue whatever it hits)
constexpr int Q_MISS = 0;             // 1 + rt_material_type = 1 .. 5
constexpr int Q_LIGHT = 1 + RT_MAT_DIFFUSE_LIGHT;
constexpr int Q_DEAD = 6;             // a slot reserved for a path that was then absorbed (metal scattering into the surface): ends here, regenerates
constexpr int Q_HOLE = 7;             // no path: the tile tail and (deferred extend) the slot of a suspended path
constexpr int Q_LAST = 8;
constexpr uint32_t DEAD_BOUNCE = 0xffu;   // bounce byte of a dead slot's ray word (live rays: bounce + 1 < max_depth <= 254)
constexpr int POLL_DEPTH = 4;         // host runs this many iterations ahead of the counter it polls
constexpr int POLL_SLOTS = 8;

// every hot counter on its own 128-byte line (its own L2 slice): same-line atomics serialise
struct __align__(128) Counters {
    uint32_t n_bounce[2];             uint32_t pad0[30];
    uint32_t n_fresh[2];              uint32_t pad1[30];
    uint32_t cur_extend;              uint32_t pad2[31];
    uint32_t cur_shade;               uint32_t pad3[31];
    unsigned long long next_sample;   unsigned long long pad4[15];   // global sample cursor (sample-major: g = sample * npix + pixel)
    unsigned long long total_rays, nodes, prim_tests, node_slots;
    unsigned long long iter_begin;    // min(next_sample, total) when the current iteration started (written by generate / turnover only)
    unsigned long long pad5[11];
    struct { uint32_t n, pad[31]; } cont[2];      // continuation records in list b (written by the extend of iterations with cur = b)
    struct { uint32_t n, pad[31]; } n_fin[2];     // resumed paths that finished their walk in this iteration: region [capacity, capacity + n) of pool b
    struct { uint32_t n, pad[31]; } n_dead[2];    // dead slots shade wrote into pool b (not path segments: subtracted from total_rays)
    struct { uint32_t n, pad[31]; } bin_n[8];     // class-binned shade: entries in each bin's list this iteration (k_bin)
    struct { uint32_t n, pad[31]; } shade_done;   // blocks of k_shade_bins that have finished: the last one runs the turnover
};

struct PoolView {
    float4* ray_o[2];    // origin.xyz, time
    float4* ray_d[2];    // direction.xyz, bits: sample | bounce << 24
    float4* thr[2];      // throughput rgb, bits: pixel
    uint2* hit;          // written by extend: bits of t, Closest::prim
    uint8_t* hit_class;  // written by extend: material class of the hit
    // continuation records of suspended walks (k_extend_defer), two lists: extend(cur) reads list cur ^ 1 and writes list cur.
    // A record owns its path: the pool slot it came from is a hole (class Q_HOLE) that shade skips.
    uint4* cont_walk[2];   // bits of the closest t so far, its primitive, node, parked leaf
    float4* cont_o[2];     // the path: ray origin + time,
    float4* cont_d[2];     //   direction + sample / bounce word,
    float4* cont_thr[2];   //   throughput + pixel
    uint8_t* cont_sp[2];   // stack entries
    int* cont_stack[2];    // [entry * capacity + record]
    uint32_t* bin_list;    // class-binned shade: NBINS lists of pool positions, bin b at [b * bin_stride, ...)
    uint32_t bin_stride;
    Counters* c;
    uint32_t capacity;
};

struct RenderPool {
    PoolView v{};
    int device = -1;
    uint32_t* h_live = nullptr;       // pinned: live-path count per iteration (ring of POLL_SLOTS)
    Counters* h_counters = nullptr;   // pinned copy for stats
    cudaEvent_t poll_ev[POLL_SLOTS]{};
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    int grid_extend = 0, grid_shade = 0, grid_shade_block = 0, grid_shade_bins = 0, grid_bin = 0;
    int extend_variant = -1;               // traversal variant the extend grid was sized for
    cudaGraphExec_t graph_exec = nullptr;   // GRAPH_ITERS wavefront iterations as one CUDA graph (render_impl)
    cudaStream_t capture_stream = nullptr;
    uint32_t suspend_below = 0;          // deferred extend (k_extend_defer): a full warp suspends once <= this many lanes are live; 0 = k_extend
    uint32_t requested = 0;              // the capacity asked for when the pool was made (the default halves until the pool fits the free memory)
    bool out_of_memory = false;          // pool_create failed on an allocation
    std::vector<cudaEvent_t> timer_ev;   // RT_RENDER_TIMERS: 3 events per iteration, reused across frames
};

void render_pool_destroy(RenderPool* p) {
    if (!p) return;
    if (p->graph_exec) cudaGraphExecDestroy(p->graph_exec);
    if (p->capture_stream) cudaStreamDestroy(p->capture_stream);
    for (int b = 0; b < 2; b++) { cudaFree(p->v.ray_o[b]); cudaFree(p->v.ray_d[b]); cudaFree(p->v.thr[b]); }
    cudaFree(p->v.hit); cudaFree(p->v.hit_class); cudaFree(p->v.bin_list); cudaFree(p->v.c);
    for (int b = 0; b < 2; b++) {
        cudaFree(p->v.cont_walk[b]); cudaFree(p->v.cont_o[b]); cudaFree(p->v.cont_d[b]); cudaFree(p->v.cont_thr[b]);
        cudaFree(p->v.cont_sp[b]); cudaFree(p->v.cont_stack[b]);
    }
    if (p->h_live) cudaFreeHost(p->h_live);
    if (p->h_counters) cudaFreeHost(p->h_counters);
    for (auto& e : p->poll_ev) if (e) cudaEventDestroy(e);
    if (p->ev_begin) cudaEventDestroy(p->ev_begin);
    if (p->ev_end) cudaEventDestroy(p->ev_end);
    for (auto& e : p->timer_ev) cudaEventDestroy(e);
    delete p;
}

namespace {

struct FrameParams {
    uint32_t width, height, max_depth, seed;
    uint32_t sample_first;
    uint32_t debug;                     // rt_debug_flags
    unsigned long long total_samples;   // samples this call renders = npix * sample_count
    unsigned long long npix, npix_magic; // film indices this call covers (the frame, or this tile-split share) and ceil(2^64 / npix)
    float* accum;                       // width*height*3 sums
    uint32_t tile_index, tile_count;    // tile-split: film blocks b with b % tile_count == tile_index (tile_count <= 1: all)
};

#ifndef RT_EXPERIMENTS
#define RT_EXPERIMENTS 0
#endif
constexpr int EXTEND_BLOCK = 128;
constexpr uint32_t SUSPEND_MIN_PRIMS = 65536, SUSPEND_BELOW_DEFAULT = 12;   // re-swept on the v18 pipeline: 4 / 8 / 12 / 16 / 20 = 120.5 / 115.5 / 114.4 / 116.7 / 122.8 ms
constexpr uint32_t DEFAULT_POOL = 1u << 25, MIN_DEFAULT_POOL = 1u << 20;   // paths (render_impl)
// blocks of k_extend per SM: 9 x 128 threads = 36 warps at 56 registers.  Same-box A/B on the final scene / 1 M-triangle mesh
// (profiles/r02_traversal_ab.md): 6 (77 registers, no spills) -8 % / -8 %, 7 -5 % / -3 %, 8 (64 registers) 0, 9 +3.0 % / +2.7 %,
// 10 (48 registers, 220 B of spill loads) -5 % / -4 %: occupancy pays until the spills reach the node loop
#ifndef EXTEND_MIN_BLOCKS
#define EXTEND_MIN_BLOCKS 9
#endif

// work index -> pool position: bounced paths first, fresh camera paths at the top end of the pool
__device__ __forceinline__ uint32_t pos_of(uint32_t idx, uint32_t n_bounce, uint32_t n, uint32_t capacity) {
    return idx < n_bounce ? idx : capacity - n + idx;
}

// persistent work fetch: a warp claims `n` consecutive paths
__device__ __forceinline__ uint32_t warp_claim(uint32_t* cursor, uint32_t n = 32u) {
    uint32_t base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, n);
    return __shfl_sync(0xffffffffu, base, 0);
}

// The same with the FIRST claim of every warp fixed by its position in the grid: the counter only hands out what lies behind the
// grid's first sweep.  A launch over a nearly empty pool (the drain of every frame, every iteration of a small one) then ends
// without a single atomic, instead of one per warp on one address (4736 of them: the 6 us floor of k_extend).
struct WarpClaim {
    uint32_t* cursor; uint32_t n, sweep; bool first;
    __device__ __forceinline__ WarpClaim(uint32_t* c, uint32_t n_) : cursor(c), n(n_), sweep(n_ * (gridDim.x * (blockDim.x >> 5))), first(true) {}
    __device__ __forceinline__ uint32_t next() {
        if (first) { first = false; return n * (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)); }
        return sweep + warp_claim(cursor, n);
    }
};

// g -> (pixel, sample): g = sample * npix + pixel.  Division by the frame size through a precomputed 64-bit
// reciprocal (one mul.hi + a fix-up) instead of a 64-bit divide per path.
__device__ __forceinline__ void split_sample_index(const FrameParams& fp, unsigned long long g, uint32_t& pixel, uint32_t& sample) {
    if (fp.npix == 1) { pixel = 0; sample = fp.sample_first + (uint32_t)g; return; }
    unsigned long long q = __umul64hi(g, fp.npix_magic);
    long long r = (long long)(g - q * fp.npix);
    if (r < 0) { q--; r += (long long)fp.npix; }
    else if (r >= (long long)fp.npix) { q++; r -= (long long)fp.npix; }
    pixel = (uint32_t)r;
    sample = fp.sample_first + (uint32_t)q;
}

// Film order: consecutive sample-cursor values walk the frame in 8x4-pixel tiles (one warp = one tile) inside bands
// of four rows, so the camera rays of a warp form a compact beam and the paths that descend from them stay
// neighbours in the pools.  Every pixel still receives every sample index exactly once; the Philox key is the
// (pixel, sample) pair, so the image does not depend on this order.
__device__ __forceinline__ void film_xy(const FrameParams& fp, uint32_t i, uint32_t& px, uint32_t& py) {
    const uint32_t band = i / (4u * fp.width);                 // bands of 4 rows (the last one may be shorter)
    const uint32_t y0 = 4u * band;
    const uint32_t hb = min(4u, fp.height - y0);
    const uint32_t r = i - band * 4u * fp.width;               // index inside the band
    const uint32_t full = fp.width >> 3;                       // whole 8-wide tiles
    const uint32_t tile_px = 8u * hb;
    if (r < full * tile_px) {
        const uint32_t t = r / tile_px, w = r - t * tile_px;
        px = 8u * t + (w & 7u);
        py = y0 + (w >> 3);
    } else {                                                    // the narrow last tile of the band
        const uint32_t j = r - full * tile_px, wr = fp.width & 7u;
        px = 8u * full + j % wr;
        py = y0 + j / wr;
    }
}

// camera ray generation into pool `buf` at `pos` (render.h:105-108)
__device__ __forceinline__ void start_path(const SceneView& sv, const PoolView& pv, const FrameParams& fp, int buf, uint32_t pos, unsigned long long g) {
    uint32_t index, sample, px, py;
    split_sample_index(fp, g, index, sample);
    if (fp.tile_count > 1u) {   // local film index -> global: the k-th owned block is global block k * tile_count + tile_index
        const uint32_t k = index / RT_TILE_BLOCK_PIXELS;
        index = (k * fp.tile_count + fp.tile_index) * RT_TILE_BLOCK_PIXELS + (index - k * RT_TILE_BLOCK_PIXELS);
    }
    if (fp.debug & RT_DEBUG_RASTER_FILM) { py = index / fp.width; px = index - py * fp.width; } else film_xy(fp, index, px, py);
    const uint32_t pixel = py * fp.width + px;
    const RngKey key{pixel, sample, 0u, fp.seed};
    const Ray r = camera_ray(sv.camera, key, (int)px, (int)py, (int)fp.width, (int)fp.height);
    pv.ray_o[buf][pos] = make_float4(r.o.x, r.o.y, r.o.z, r.tm);
    pv.ray_d[buf][pos] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(sample));   // bounce 0
    pv.thr[buf][pos] = make_float4(1.f, 1.f, 1.f, __uint_as_float(pixel));
}

// ---- generate: fills the fresh region of pool 0 at the start of a frame ---------------------------
__global__ void __launch_bounds__(256) k_generate(SceneView sv, PoolView pv, FrameParams fp, uint32_t n_fresh) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { pv.c->n_fresh[0] = n_fresh; pv.c->next_sample = n_fresh; pv.c->iter_begin = n_fresh; }
    if (i < n_fresh) start_path(sv, pv, fp, 0, pv.capacity - n_fresh + i, i);
}

// class byte of a finished walk
__device__ __forceinline__ uint8_t hit_class_of(const SceneView& sv, const Closest& c, uint32_t bounce, uint32_t max_depth) {
    int q = Q_MISS;
    if (c.prim != -1) {
        const uint32_t mat = c.prim >= 0 ? sv.prims[c.prim].material : sv.media[-2 - c.prim].material;
        q = 1 + (int)sv.materials[mat].type;
    }
    if (bounce + 1 >= max_depth) q |= Q_LAST;                   // render.h:60: the next segment would start at depth 0
    return (uint8_t)q;
}

constexpr int NBINS = 7, BIN_NONE = 7, BIN_CONT = 4;       // bins 0..3 continue (lambertian, metal, dielectric, isotropic), 4..6 end (light, miss, nothing to do)
#ifndef RT_BIN_THREADS
#define RT_BIN_THREADS 512
#endif
#ifndef RT_BIN_MIN_BLOCKS
#define RT_BIN_MIN_BLOCKS 3
#endif
constexpr int BIN_THREADS = RT_BIN_THREADS, BIN_WARPS = BIN_THREADS / 32, BIN_ROUNDS = 8, BIN_SPAN = BIN_THREADS * BIN_ROUNDS;
static_assert(BIN_WARPS >= 7 && BIN_ROUNDS <= 8, "one warp per bin scans; 4 bits per round in a 32-bit word");
#ifndef RT_BIN_CLAIM
#define RT_BIN_CLAIM 4
#endif
#ifndef RT_BINS_MIN_BLOCKS
#define RT_BINS_MIN_BLOCKS 4
#endif
constexpr int BIN_CLAIM = RT_BIN_CLAIM;                    // rounds of 32 list entries a shading warp claims at once

constexpr uint32_t bin_of_class(uint32_t c) {
    const uint32_t k = c & 7u;
    if (k == (uint32_t)Q_HOLE) return (uint32_t)BIN_NONE;
    if (k == (uint32_t)Q_MISS) return 5u;
    if (k == (uint32_t)Q_LIGHT) return 4u;
    if (k == (uint32_t)Q_DEAD || (c & (uint32_t)Q_LAST)) return 6u;       // absorbed earlier / depth exhausted: black, only regenerates
    return k == 1u + RT_MAT_ISOTROPIC ? 3u : k - 1u;                       // 1 + {lambertian 0, metal 1, dielectric 2}
}
constexpr uint32_t bin_table(uint32_t first) { uint32_t t = 0; for (uint32_t c = 0; c < 8; c++) t |= bin_of_class(first + c) << (4 * c); return t; }
constexpr uint32_t BIN_TABLE_LO = bin_table(0), BIN_TABLE_HI = bin_table(8);   // class byte -> bin, 4 bits each
__device__ __forceinline__ uint32_t bin_of(uint32_t c) { return (((c & 8u) ? BIN_TABLE_HI : BIN_TABLE_LO) >> (4u * (c & 7u))) & 7u; }

// ---- extend ---------------------------------------------------------------------------------------
// V: traversal variant (traverse.cuh: TV_TCULL | TV_SMEM | TV_MINMAX)
template <bool COUNT, int V>
__global__ void __launch_bounds__(EXTEND_BLOCK, EXTEND_MIN_BLOCKS) k_extend(SceneView sv, PoolView pv, FrameParams fp, int cur) {
    constexpr bool SM = (V & TV_SMEM) != 0, TC = (V & TV_TCULL) != 0;
    __shared__ int s_stack[SM ? RT_SHORT_STACK * (TC ? 2 : 1) * EXTEND_BLOCK : 1];
    int* column = s_stack + (SM ? threadIdx.x : 0);
    const uint32_t n_b = pv.c->n_bounce[cur], n = n_b + pv.c->n_fresh[cur];
    const uint32_t lane = threadIdx.x & 31;
    const float4* __restrict__ ray_o = pv.ray_o[cur];
    const float4* __restrict__ ray_d = pv.ray_d[cur];
    const float4* thr = pv.thr[cur];
    TraversalCounters tc{0, 0};
    unsigned long long slots = 0;
    WarpClaim claim(&pv.c->cur_extend, 32u);
    for (;;) {
        const uint32_t base = claim.next();
        if (base >= n) break;
        const uint32_t idx = base + lane;
        const bool in_range = idx < n;              // closest_hit is warp-collective
        const uint32_t nodes_before = tc.nodes;
        const uint32_t pos = pos_of(in_range ? idx : base, n_b, n, pv.capacity);
        const float4 o = ray_o[pos], d = ray_d[pos];
        const uint32_t sb = __float_as_uint(d.w);
        const bool dead = (sb >> 24) == DEAD_BOUNCE;
        Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
        // the free-flight draw of medium m is keyed (pixel, sample, bounce, m); the pixel is only fetched
        // when a ray really crosses a medium boundary
        const Closest c = closest_hit<COUNT, V, EXTEND_BLOCK>(sv, r, 0.001f, true, [&](uint32_t m) {                // render.h:63
            const RngKey key{__float_as_uint(thr[pos].w), sb & 0xffffffu, sb >> 24, fp.seed};
            return draw4(key, RT_STREAM_MEDIUM, m).x;
        }, &tc, in_range && !dead, column);
        if (COUNT) slots += 32ull * __reduce_max_sync(0xffffffffu, tc.nodes - nodes_before);   // node steps the warp paid for
        uint32_t cls = (uint32_t)Q_HOLE;
        if (in_range) {
            pv.hit[pos] = make_uint2(__float_as_uint(c.t), (uint32_t)c.prim);
            cls = dead ? (uint32_t)Q_DEAD : (uint32_t)hit_class_of(sv, c, sb >> 24, fp.max_depth);
            pv.hit_class[pos] = (uint8_t)cls;
        }
    }
    if (COUNT) {
        atomicAdd(&pv.c->nodes, (unsigned long long)tc.nodes);
        atomicAdd(&pv.c->prim_tests, (unsigned long long)tc.prim_tests);
        if (lane == 0) atomicAdd(&pv.c->node_slots, slots);
    }
}

// Deferred extend (walk_suspendable): a warp stops walking once no more than `suspend_below` of its 32 rays are still
// live and moves those paths — ray, throughput, traversal state — into a continuation list; their pool slots become holes
// (class Q_HOLE) that shade skips.  The NEXT iteration's extend takes the suspended walks up again, 32 stragglers to a
// warp, before it starts on the pool (longest work first); a resumed path that finishes its walk is written, with its hit
// record, to the region [capacity, capacity + n_fin) of the current pool, which shade treats like any other path.  A
// suspended path simply bounces an iteration later.  Nothing waits for stragglers: a second pass inside the same
// iteration raised the lane utilisation of the node loop just as much but paid for it with the latency of the longest
// walks at the end of every iteration (profiles/r01_v9_experiments.md).
#ifndef DEFER_MIN_BLOCKS
#define DEFER_MIN_BLOCKS EXTEND_MIN_BLOCKS
#endif
template <bool COUNT, bool WIDE>
__global__ void __launch_bounds__(EXTEND_BLOCK, DEFER_MIN_BLOCKS) k_extend_defer(SceneView sv, PoolView pv, FrameParams fp, int cur, uint32_t suspend_below) {
    const uint32_t n_b = pv.c->n_bounce[cur], n_pool = n_b + pv.c->n_fresh[cur];
    const int rd = cur ^ 1, wr = cur;
    const uint32_t n_c = pv.c->cont[rd].n, n_c32 = (n_c + 31u) & ~31u;   // work index space: continuation records (padded to whole warps), then the pool
    const uint32_t n = n_c32 + n_pool;
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    TraversalCounters tc{0, 0};
    uint32_t steps = 0;
    int stack[RT_STACK];
    WarpClaim claim(&pv.c->cur_extend, 32u);
    for (;;) {
        const uint32_t base = claim.next();
        if (base >= n) break;
        const bool resumed = base < n_c32;                       // warp-uniform
        const uint32_t idx = (resumed ? base : base - n_c32) + lane;
        const bool active = idx < (resumed ? n_c : n_pool);
        Walk w{RT_NODE_DONE, RT_NODE_DONE, 0};
        Closest c; c.t = INFINITY; c.prim = -1;
        uint32_t pos = 0;
        bool dead = false;
        float4 o, d, th = make_float4(0.f, 0.f, 0.f, 0.f);
        if (resumed) {
            const uint32_t rec = active ? idx : base;
            o = pv.cont_o[rd][rec]; d = pv.cont_d[rd][rec];
            if (active) {
                const uint4 ws = pv.cont_walk[rd][idx];
                th = pv.cont_thr[rd][idx];
                c.t = __uint_as_float(ws.x); c.prim = (int)ws.y; w.node = (int)ws.z; w.parked = (int)ws.w;
                w.sp = (int)pv.cont_sp[rd][idx];
                for (int k = 0; k < w.sp; k++) stack[k] = pv.cont_stack[rd][(size_t)k * pv.capacity + idx];
            }
        } else {
            pos = pos_of(active ? idx : idx - lane, n_b, n_pool, pv.capacity);
            o = pv.ray_o[cur][pos]; d = pv.ray_d[cur][pos];
            dead = (__float_as_uint(d.w) >> 24) == DEAD_BOUNCE;
            if (active && !dead && sv.n_prims) w.node = 0;
#if RT_STACK_SENTINEL
            stack[0] = RT_NODE_DONE; w.sp = 1;   // a fresh walk starts on the sentinel; a resumed one brings it along in its record
#endif
        }
        Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
        // only whole batches suspend: a few stragglers left over at the end of a list would otherwise advance a leaf phase per iteration
        const uint32_t below = __all_sync(0xffffffffu, active) ? suspend_below : 0u;
        walk_suspendable<COUNT, WIDE>(sv, r, 0.001f, c, &tc, w, stack, below, &steps);                                    // render.h:63
        const bool suspended = w.node != RT_NODE_DONE || w.parked != RT_NODE_DONE;
        const uint32_t ms = __ballot_sync(0xffffffffu, suspended);
        if (ms) {
            uint32_t slot = 0;
            if (lane == (uint32_t)(__ffs(ms) - 1)) slot = atomicAdd(&pv.c->cont[wr].n, (uint32_t)__popc(ms));
            slot = __shfl_sync(0xffffffffu, slot, __ffs(ms) - 1) + __popc(ms & lt);
            if (suspended) {
                RT_DEV_ASSERT(slot < pv.capacity && w.sp <= RT_SUSPEND_STACK);
                if (!resumed) { th = pv.thr[cur][pos]; pv.hit_class[pos] = (uint8_t)Q_HOLE; }
                pv.cont_o[wr][slot] = o; pv.cont_d[wr][slot] = d; pv.cont_thr[wr][slot] = th;
                pv.cont_walk[wr][slot] = make_uint4(__float_as_uint(c.t), (uint32_t)c.prim, (uint32_t)w.node, (uint32_t)w.parked);
                pv.cont_sp[wr][slot] = (uint8_t)w.sp;
                for (int k = 0; k < w.sp; k++) pv.cont_stack[wr][(size_t)k * pv.capacity + slot] = stack[k];
            }
        }
        const bool finished = active && !suspended;
        if (resumed) {   // finished resumed paths re-enter the pool behind its two regions
            const uint32_t mf = __ballot_sync(0xffffffffu, finished);
            uint32_t k = 0;
            if (mf && lane == (uint32_t)(__ffs(mf) - 1)) k = atomicAdd(&pv.c->n_fin[cur].n, (uint32_t)__popc(mf));
            if (mf) k = __shfl_sync(0xffffffffu, k, __ffs(mf) - 1) + __popc(mf & lt);
            if (finished) {
                RT_DEV_ASSERT(k < pv.capacity);
                pos = pv.capacity + k;
                pv.ray_o[cur][pos] = o; pv.ray_d[cur][pos] = d; pv.thr[cur][pos] = th;
            }
        }
        uint32_t cls = (uint32_t)Q_HOLE;
        if (finished) {
            const uint32_t sb = __float_as_uint(d.w);
            // the free-flight draw of medium m is keyed (pixel, sample, bounce, m); the pixel is only fetched when a ray really crosses a boundary
            if (sv.n_active_media && !dead) closest_medium(sv, r, 0.001f, c, [&](uint32_t m) {
                const RngKey key{__float_as_uint(resumed ? th.w : pv.thr[cur][pos].w), sb & 0xffffffu, sb >> 24, fp.seed};
                return draw4(key, RT_STREAM_MEDIUM, m).x;
            });
            pv.hit[pos] = make_uint2(__float_as_uint(c.t), (uint32_t)c.prim);
            cls = dead ? (uint32_t)Q_DEAD : (uint32_t)hit_class_of(sv, c, sb >> 24, fp.max_depth);
            pv.hit_class[pos] = (uint8_t)cls;
        }
    }
    if (COUNT) {
        atomicAdd(&pv.c->nodes, (unsigned long long)tc.nodes);
        atomicAdd(&pv.c->prim_tests, (unsigned long long)tc.prim_tests);
        if (lane == 0) atomicAdd(&pv.c->node_slots, 32ull * steps);
    }
}

// ---- shade ----------------------------------------------------------------------------------------
__device__ __forceinline__ void accumulate(float* accum, uint32_t pixel, float3 c) {
    if (c.x != 0.f) atomicAdd(&accum[3 * (size_t)pixel], c.x);
    if (c.y != 0.f) atomicAdd(&accum[3 * (size_t)pixel + 1], c.y);
    if (c.z != 0.f) atomicAdd(&accum[3 * (size_t)pixel + 2], c.z);
}

// what the class byte says about the path's future
__device__ __forceinline__ bool class_continues(uint32_t c) {   // lambertian, metal, dielectric, isotropic with bounces left
    const uint32_t k = c & 7u;
    return !(c & (uint32_t)Q_LAST) && k >= 1u && k <= 5u && k != (uint32_t)Q_LIGHT;
}
__device__ __forceinline__ bool class_terminates(uint32_t c) { return (c & 7u) != (uint32_t)Q_HOLE && !class_continues(c); }

// One path: hit attributes, emission / scatter.  Returns true when the path goes on (`out` filled); paths that end here
// add their radiance to the accumulation buffer.
struct ShadeOut { float4 o, d, thr; };
__device__ __forceinline__ bool shade_path(const SceneView& sv, const PoolView& pv, const FrameParams& fp, int cur, uint32_t pos, uint32_t k, bool last,
                                           ShadeOut& out) {
    if (k == (uint32_t)Q_DEAD || (last && k != (uint32_t)Q_LIGHT && k != (uint32_t)Q_MISS)) return false;   // absorbed earlier / depth exhausted (render.h:60): black
    const float4 th = pv.thr[cur][pos];
    const uint32_t pixel = __float_as_uint(th.w);
    float3 thr = make_float3(th.x, th.y, th.z);
    if (k == (uint32_t)Q_MISS) {
        accumulate(fp.accum, pixel, thr * f3(sv.background[0], sv.background[1], sv.background[2]));   // render.h:64
        return false;
    }
    const float4 d = pv.ray_d[cur][pos];
    const float4 o = pv.ray_o[cur][pos];
    const uint2 h = pv.hit[pos];
    const uint32_t sb = __float_as_uint(d.w), sample = sb & 0xffffffu, bounce = sb >> 24;
    Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
    const int prim = (int)h.y;
    const float t = __uint_as_float(h.x);
    HitAttr rec;
    uint32_t mat;
    if (prim >= 0) { const DPrim p = sv.prims[prim]; prim_attr(sv, p, r, t, rec); mat = p.material; }
    else { const DMedium m = sv.media[-2 - prim]; medium_attr(sv, m, r, t, rec); mat = m.material; }
    const rt_flat_material m = sv.materials[mat];
    if (m.type == RT_MAT_DIFFUSE_LIGHT) {
        accumulate(fp.accum, pixel, thr * texture_value(sv, m.texture, rec.u, rec.v, rec.p));           // render.h:68, :75
        return false;
    }
    const RngKey key{pixel, sample, bounce, fp.seed};
    float3 attenuation;
    Ray scattered;
    if (!material_scatter(sv, m, r, rec, key, attenuation, scattered)) return false;                    // render.h:70-73: absorbed, black
    thr = thr * attenuation;
    out.o = make_float4(scattered.o.x, scattered.o.y, scattered.o.z, scattered.tm);
    out.d = make_float4(scattered.d.x, scattered.d.y, scattered.d.z, __uint_as_float(sample | ((bounce + 1) << 24)));
    out.thr = make_float4(thr.x, thr.y, thr.z, th.w);
    return true;
}

#ifndef RT_WTILE
#define RT_WTILE 128
#endif
#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 4
#endif
#ifndef RT_SHADE_BLOCK
#define RT_SHADE_BLOCK 256
#endif
constexpr int SHADE_BLOCK = RT_SHADE_BLOCK, SHADE_WARPS = SHADE_BLOCK / 32;
constexpr int WTILE = RT_WTILE, WROUNDS = WTILE / 32;       // paths per warp tile / rounds of 32

// Warp-tile shade (RT_DEBUG_WARP_SHADE): round 2's attempt to remove the block-wide barriers of the block-tile kernel below
// (37 % of its stall samples).  A warp ranks a tile of RT_WTILE paths by class with ballots, allocates for the tile BEFORE it
// shades (possible because the class byte says whether a path can continue) and shades RT_WTILE / 32 rounds — no
// __syncthreads at all.  Measured same-box against the block-tile kernel (profiles/r02_shade_ab.md), best of seven
// tile-size / occupancy points: final scene 29.5 vs 27.1 ms per 128 spp, Cornell smoke 23.9 vs 23.3, big_scene1 1.00 vs 0.78,
// 1 M-triangle mesh 12.7 vs 12.3 — the barriers were not what the kernel waits for; a warp that owns its whole tile
// serialises RT_WTILE / 32 dependent-load chains that eight warps of a block overlap.  Kept as an A/B switch, not the default.
template <bool DEFER>
__device__ __forceinline__ void shade_warp_tiles(const SceneView& sv, const PoolView& pv, const FrameParams& fp, int cur) {
    __shared__ uint32_t s_perm[SHADE_WARPS][WTILE];   // sorted rank -> tile-local index | class byte << 8 | rank among the continuing (or ending) paths << 12
    const int next = cur ^ 1;
    // tile index space: bounced paths, fresh camera paths, (deferred extend) the resumed paths that finished at positions >= capacity
    const uint32_t n_b = pv.c->n_bounce[cur], n_pool = n_b + pv.c->n_fresh[cur];
    const uint32_t n = n_pool + (DEFER ? pv.c->n_fin[cur].n : 0u);
    const unsigned long long iter_begin = pv.c->iter_begin;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t* perm = s_perm[warp];
    auto pos_at = [&](uint32_t idx) { return DEFER && idx >= n_pool ? pv.capacity + (idx - n_pool) : pos_of(idx, n_b, n_pool, pv.capacity); };
    // mask of the lanes whose class has bit pattern `k`, from the three bit ballots
    auto class_mask = [](uint32_t b0, uint32_t b1, uint32_t b2, uint32_t k) {
        return ((k & 1u) ? b0 : ~b0) & ((k & 2u) ? b1 : ~b1) & ((k & 4u) ? b2 : ~b2);
    };

    uint32_t tile = warp_claim(&pv.c->cur_shade, (uint32_t)WTILE);
    while (tile < n) {
        // ---- class bytes of this lane's 8 paths (round r: path tile + 32 r + lane), 4 bits each
        uint32_t cls = 0;
        #pragma unroll
        for (int r = 0; r < WROUNDS; r++) {
            const uint32_t idx = tile + 32u * r + lane;
            const uint32_t c = idx < n ? (uint32_t)pv.hit_class[pos_at(idx)] & 15u : (uint32_t)Q_HOLE;
            cls |= c << (4 * r);
        }
        // ---- pass 1: class histogram (lane k < 8 counts class k), paths that continue / end
        uint32_t hist = 0, nc = 0, nt = 0;
        #pragma unroll
        for (int r = 0; r < WROUNDS; r++) {
            const uint32_t c = (cls >> (4 * r)) & 15u;
            const uint32_t b0 = __ballot_sync(0xffffffffu, c & 1u), b1 = __ballot_sync(0xffffffffu, c & 2u), b2 = __ballot_sync(0xffffffffu, c & 4u);
            hist += __popc(class_mask(b0, b1, b2, lane & 7u));
            nc += __popc(__ballot_sync(0xffffffffu, class_continues(c)));
            nt += __popc(__ballot_sync(0xffffffffu, class_terminates(c)));
        }
        // ---- allocation in pool `next` and the next tile: three INDEPENDENT atomics by lane 0, consumed after pass 2.
        //      The fresh region needs no counter of its own — samples [iter_begin, next_sample) are regenerated in this
        //      iteration, so a tile's offset in the region is g0 - iter_begin and k_turnover derives n_fresh from the cursor
        uint32_t base_b = 0, tile_next = 0;
        unsigned long long g0 = 0;
        if (lane == 0) {
            tile_next = atomicAdd(&pv.c->cur_shade, (uint32_t)WTILE);
            if (nc) base_b = atomicAdd(&pv.c->n_bounce[next], nc);
            if (nt) g0 = atomicAdd(&pv.c->next_sample, (unsigned long long)nt);
        }
        // ---- pass 2: sorted position of every path = class offset + paths of the class in earlier rounds + rank in the round
        uint32_t run = hist;                                    // lane k < 8: exclusive prefix over the classes
        #pragma unroll
        for (int o = 1; o < 8; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, run, o); if ((lane & 7u) >= (uint32_t)o) run += y; }
        run -= hist;
        uint32_t cbase = 0, tbase = 0;
        #pragma unroll
        for (int r = 0; r < WROUNDS; r++) {
            const uint32_t c = (cls >> (4 * r)) & 15u, k = c & 7u;
            const uint32_t b0 = __ballot_sync(0xffffffffu, c & 1u), b1 = __ballot_sync(0xffffffffu, c & 2u), b2 = __ballot_sync(0xffffffffu, c & 4u);
            const uint32_t peers = class_mask(b0, b1, b2, k);
            const uint32_t at = __shfl_sync(0xffffffffu, run, k) + __popc(peers & lt);
            run += __popc(class_mask(b0, b1, b2, lane & 7u));
            const bool cont = class_continues(c), term = class_terminates(c);
            const uint32_t mc = __ballot_sync(0xffffffffu, cont), mt = __ballot_sync(0xffffffffu, term);
            const uint32_t xr = cont ? cbase + __popc(mc & lt) : tbase + __popc(mt & lt);
            cbase += __popc(mc); tbase += __popc(mt);
            perm[at] = (32u * r + lane) | c << 8 | xr << 12;
        }
        __syncwarp();
        base_b = __shfl_sync(0xffffffffu, base_b, 0);
        tile_next = __shfl_sync(0xffffffffu, tile_next, 0);
        g0 = __shfl_sync(0xffffffffu, g0, 0);
        uint32_t n_regen = 0;
        if (nt && g0 < fp.total_samples) n_regen = (uint32_t)min((unsigned long long)nt, fp.total_samples - g0);
        const uint32_t base_f = n_regen ? (uint32_t)(g0 - iter_begin) : 0u;

        // ---- shade, a round = 32 paths of (mostly) one class; results go straight to their slots
        #pragma unroll 1
        for (int r = 0; r < WROUNDS; r++) {
            const uint32_t e = perm[32 * r + lane];
            const uint32_t c = (e >> 8) & 15u, k = c & 7u, xr = e >> 12;
            if (__all_sync(0xffffffffu, k == (uint32_t)Q_HOLE)) break;      // holes sort last
            if (k == (uint32_t)Q_HOLE) continue;
            const uint32_t pos = pos_at(tile + (e & 255u));
            ShadeOut out;
            const bool goes_on = shade_path(sv, pv, fp, cur, pos, k, (c & (uint32_t)Q_LAST) != 0, out);
            if (class_continues(c)) {
                const uint32_t dst = base_b + xr;                            // survivors keep tile order
                RT_DEV_ASSERT(dst < pv.capacity);
                if (goes_on) { pv.ray_o[next][dst] = out.o; pv.ray_d[next][dst] = out.d; pv.thr[next][dst] = out.thr; }
                else {   // absorbed after its slot was reserved: a dead slot, counted out of the ray total; its sample slot regenerates one iteration late
                    pv.ray_d[next][dst] = make_float4(0.f, 0.f, 1.f, __uint_as_float(DEAD_BOUNCE << 24));
                    pv.ray_o[next][dst] = make_float4(0.f, 0.f, 0.f, 0.f);
                    atomicAdd(&pv.c->n_dead[next].n, 1u);
                }
            } else if (xr < n_regen) start_path(sv, pv, fp, next, pv.capacity - (base_f + n_regen) + xr, g0 + xr);
        }
        __syncwarp();
        tile = tile_next;
    }
}

#if RT_EXPERIMENTS
__global__ void __launch_bounds__(SHADE_BLOCK, RT_SHADE_MIN_BLOCKS) k_shade(SceneView sv, PoolView pv, FrameParams fp, int cur) { shade_warp_tiles<false>(sv, pv, fp, cur); }
// shade of the deferred extend: holes in the tiles and the finished-resumed region behind the pool
__global__ void __launch_bounds__(SHADE_BLOCK, RT_SHADE_MIN_BLOCKS) k_shade_defer(SceneView sv, PoolView pv, FrameParams fp, int cur) { shade_warp_tiles<true>(sv, pv, fp, cur); }
#endif

// Round 1's shade (RT_DEBUG_BLOCK_SHADE; the default is the class-binned pipeline below): a BLOCK sorts a 256-path tile by class
// in shared memory (a warp shades one class, warps at class boundaries several), thread 0 allocates
// between two barriers — five __syncthreads per tile.
template <bool DEFER>
__device__ __forceinline__ void shade_block_tiles(const SceneView& sv, const PoolView& pv, const FrameParams& fp, int cur) {
    constexpr int NC = 8;
    __shared__ uint32_t s_tile[2];                   // this tile / the next one (claimed while this one is shaded)
    constexpr int CELLS = 2;                         // class-table cells per lane of the scanning warp
    __shared__ uint32_t s_cnt[32 * CELLS];           // [class][warp] counts -> exclusive offsets
    __shared__ uint16_t s_perm[SHADE_BLOCK];         // sorted rank -> tile-local index | class << 12
    __shared__ uint32_t s_wcont[SHADE_WARPS], s_wterm[SHADE_WARPS];
    __shared__ uint32_t s_base_b, s_base_f, s_nregen;
    __shared__ unsigned long long s_g0;
    static_assert(NC * SHADE_WARPS <= 32 * CELLS, "class table must fit one warp scan");
    const int next = cur ^ 1;
    const uint32_t n_b = pv.c->n_bounce[cur], n_pool = n_b + pv.c->n_fresh[cur];
    const uint32_t n = n_pool + (DEFER ? pv.c->n_fin[cur].n : 0u);
    const unsigned long long iter_begin = pv.c->iter_begin;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    auto pos_at = [&](uint32_t idx) { return DEFER && idx >= n_pool ? pv.capacity + (idx - n_pool) : pos_of(idx, n_b, n_pool, pv.capacity); };
    if (threadIdx.x < 32 * CELLS) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_tile[0] = atomicAdd(&pv.c->cur_shade, (uint32_t)SHADE_BLOCK);
    __syncthreads();
    auto class_of = [&](uint32_t tile) {
        const uint32_t idx = tile + threadIdx.x;
        return idx < n ? (int)(pv.hit_class[pos_at(idx)] & 15u) : (int)Q_HOLE;
    };
    uint32_t tile = s_tile[0];
    int cls = class_of(tile);
    for (int it = 0;; it ^= 1) {
        if (tile >= n) break;
        uint32_t rank = 0;
        #pragma unroll
        for (int k = 0; k < NC; k++) {
            const uint32_t m = __ballot_sync(0xffffffffu, (cls & 7) == k);
            if ((cls & 7) == k) rank = __popc(m & lt);
            if ((int)lane == k) s_cnt[k * SHADE_WARPS + warp] = __popc(m);
        }
        __syncthreads();
        if (warp == 0) {
            uint32_t v[CELLS], sum = 0;
            #pragma unroll
            for (int c = 0; c < CELLS; c++) { v[c] = s_cnt[CELLS * lane + c]; sum += v[c]; }
            uint32_t x = sum;
            #pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)lane >= o) x += y; }
            uint32_t ex = x - sum;
            #pragma unroll
            for (int c = 0; c < CELLS; c++) { s_cnt[CELLS * lane + c] = ex; ex += v[c]; }
        }
        __syncthreads();
        s_perm[s_cnt[(cls & 7) * SHADE_WARPS + warp] + rank] = (uint16_t)(threadIdx.x | (cls << 12));
        __syncthreads();
        const uint32_t e = s_perm[threadIdx.x];
        const uint32_t c = e >> 12, k = c & 7u;
        bool continues = false, terminated = false;
        ShadeOut out;
        if (k != (uint32_t)Q_HOLE) {
            continues = shade_path(sv, pv, fp, cur, pos_at(tile + (e & 0xfffu)), k, (c & (uint32_t)Q_LAST) != 0, out);
            terminated = !continues;
        }
        const uint32_t mc = __ballot_sync(0xffffffffu, continues), mt = __ballot_sync(0xffffffffu, terminated);
        if (lane == 0) { s_wcont[warp] = __popc(mc); s_wterm[warp] = __popc(mt); }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t nc = 0, nt = 0;
            #pragma unroll
            for (int w = 0; w < SHADE_WARPS; w++) {
                const uint32_t a = s_wcont[w], b = s_wterm[w];
                s_wcont[w] = nc; s_wterm[w] = nt;           // exclusive prefixes
                nc += a; nt += b;
            }
            const uint32_t tile_next = atomicAdd(&pv.c->cur_shade, (uint32_t)SHADE_BLOCK);
            const uint32_t base_b = nc ? atomicAdd(&pv.c->n_bounce[next], nc) : 0u;
            const unsigned long long g0 = nt ? atomicAdd(&pv.c->next_sample, (unsigned long long)nt) : 0ull;
            uint32_t n_regen = 0;
            if (nt && g0 < fp.total_samples) n_regen = (uint32_t)min((unsigned long long)nt, fp.total_samples - g0);
            s_base_b = base_b;
            s_base_f = n_regen ? (uint32_t)(g0 - iter_begin) : 0u;
            s_nregen = n_regen;
            s_g0 = g0;
            s_tile[it ^ 1] = tile_next;
        }
        __syncthreads();
        const uint32_t tile_next = s_tile[it ^ 1];
        const int cls_next = class_of(tile_next);    // in flight while this tile's results are stored
        if (continues) {
            const uint32_t dst = s_base_b + s_wcont[warp] + __popc(mc & lt);
            RT_DEV_ASSERT(dst < pv.capacity);
            pv.ray_o[next][dst] = out.o; pv.ray_d[next][dst] = out.d; pv.thr[next][dst] = out.thr;
        }
        if (terminated) {
            const uint32_t r = s_wterm[warp] + __popc(mt & lt);
            RT_DEV_ASSERT(r >= s_nregen || s_base_f + s_nregen <= pv.capacity);
            if (r < s_nregen) start_path(sv, pv, fp, next, pv.capacity - (s_base_f + s_nregen) + r, s_g0 + r);
        }
        tile = tile_next;
        cls = cls_next;
    }
}
__global__ void __launch_bounds__(SHADE_BLOCK) k_shade_block(SceneView sv, PoolView pv, FrameParams fp, int cur) { shade_block_tiles<false>(sv, pv, fp, cur); }
__global__ void __launch_bounds__(SHADE_BLOCK, 4) k_shade_block_defer(SceneView sv, PoolView pv, FrameParams fp, int cur) { shade_block_tiles<true>(sv, pv, fp, cur); }

// ---- turnover -------------------------------------------------------------------------------------
// End of an iteration: totals, the next iteration's region sizes, counters back to zero, the live count to the host.
// Run by one thread: k_turnover after the block-tile shade, the last block of k_shade_bins to finish (volatile: its L1 may
// hold lines other blocks' atomics have moved on).
__device__ __forceinline__ void turnover(volatile Counters* c, int cur, unsigned long long total_samples, uint32_t* live_out, bool binned) {
    // every load first (they are independent; a store in between would serialise them), then the stores
    uint32_t bins[NBINS];
    #pragma unroll
    for (int b = 0; b < NBINS; b++) bins[b] = binned ? c->bin_n[b].n : 0u;
    const uint32_t n_b = c->n_bounce[cur], n_f = c->n_fresh[cur], n_d = c->n_dead[cur].n, n_cont = c->cont[cur].n;
    uint32_t n_next = c->n_bounce[cur ^ 1];
    unsigned long long cursor = c->next_sample;
    const unsigned long long rays = c->total_rays, iter_begin = c->iter_begin;
    if (binned) {   // k_shade_bins allocates by list index: the totals are the bin sizes
        uint32_t goes_on = 0, ends = 0;
        #pragma unroll
        for (int b = 0; b < NBINS; b++) { (b < BIN_CONT ? goes_on : ends) += bins[b]; c->bin_n[b].n = 0; }
        n_next = goes_on;
        cursor += ends;
        c->n_bounce[cur ^ 1] = n_next;
        c->next_sample = cursor;
    }
    c->total_rays = rays + n_b + n_f - n_d;
    c->n_bounce[cur] = 0; c->n_fresh[cur] = 0; c->n_dead[cur].n = 0;
    const unsigned long long started = min(cursor, total_samples);   // samples regenerated by this iteration's shade
    const uint32_t fresh_next = (uint32_t)(started - iter_begin);
    c->n_fresh[cur ^ 1] = fresh_next;
    c->iter_begin = started;
    c->cur_extend = 0; c->cur_shade = 0;
    c->cont[cur ^ 1].n = 0;       // consumed by this iteration's extend; the next iteration's extend appends to it
    c->n_fin[cur].n = 0;
    *live_out = n_next + fresh_next + n_cont;   // + the suspended paths the next extend resumes
}
__global__ void k_turnover(PoolView pv, int cur, unsigned long long total_samples, uint32_t* live_out) { turnover(pv.c, cur, total_samples, live_out, false); }

// ---- class-binned shade (the default) -------------------------------------------------------------------
// VERDICT r1 #3, option A: no sort and no allocation inside shade.  k_bin lists the pool positions of every class ("bin")
// once per iteration; k_shade_bins walks the concatenation of the lists 32 entries at a time, so a warp shades paths of ONE
// class, and because the class byte says whether a path goes on, where its result lands is a pure function of its list
// index: the j-th entry of continuing bin b goes to slot (entries of the continuing bins before b) + j of the next pool,
// the r-th ending path regenerates sample next_sample + r.  No __syncthreads in the loop, no atomics but the work claim
// (and one per warp that meets an absorbed path: its reserved slot becomes a dead slot, as in the warp-tile kernel).
// lanes whose 3-bit value is `k`, from the three bit ballots
__device__ __forceinline__ uint32_t mask_of(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t k) {
    return ((k & 1u) ? b0 : ~b0) & ((k & 2u) ? b1 : ~b1) & ((k & 4u) ? b2 : ~b2);
}

template <bool DEFER>
__global__ void __launch_bounds__(BIN_THREADS, RT_BIN_MIN_BLOCKS) k_bin(PoolView pv, int cur) {
    __shared__ uint32_t s_cnt[8][BIN_WARPS];         // [bin][warp]: counts, then list offsets
    const uint32_t n_b = pv.c->n_bounce[cur], n_pool = n_b + pv.c->n_fresh[cur];
    const uint32_t n = n_pool + (DEFER ? pv.c->n_fin[cur].n : 0u);
    const uint32_t span = blockIdx.x * (uint32_t)BIN_SPAN;
    if (span >= n) return;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt = (1u << lane) - 1u;
    const uint32_t base = span + warp * 32u * BIN_ROUNDS;      // a warp lists 32 * BIN_ROUNDS consecutive paths
    auto pos_at = [&](uint32_t idx) { return DEFER && idx >= n_pool ? pv.capacity + (idx - n_pool) : pos_of(idx, n_b, n_pool, pv.capacity); };
    // lane k counts bin k & 7: the lanes of bin k are (b0 ^ x0) & (b1 ^ x1) & (b2 ^ x2) of the three bit ballots
    const uint32_t x0 = (lane & 1u) ? 0u : ~0u, x1 = (lane & 2u) ? 0u : ~0u, x2 = (lane & 4u) ? 0u : ~0u;
    uint32_t bins = 0, cnt = 0;                      // this lane's bin per round, 4 bits each
    uint32_t per_round[BIN_ROUNDS / 4] = {};         // lane k: entries of bin k in each round, 8 bits each
    #pragma unroll
    for (int r = 0; r < BIN_ROUNDS; r++) {
        const uint32_t idx = base + 32u * r + lane;
        const uint32_t b = idx < n ? bin_of(pv.hit_class[pos_at(idx)]) : (uint32_t)BIN_NONE;
        bins |= b << (4 * r);
        const uint32_t b0 = __ballot_sync(0xffffffffu, b & 1u), b1 = __ballot_sync(0xffffffffu, b & 2u), b2 = __ballot_sync(0xffffffffu, b & 4u);
        const uint32_t k = __popc((b0 ^ x0) & (b1 ^ x1) & (b2 ^ x2));
        cnt += k;
        per_round[r / 4] |= k << (8 * (r & 3));
    }
    if (lane < 8) s_cnt[lane][warp] = cnt;
    __syncthreads();
    if (warp < NBINS) {                              // warp b: exclusive scan of bin b over the block's warps, one atomic for the block
        const uint32_t v = lane < (uint32_t)BIN_WARPS ? s_cnt[warp][lane] : 0u;
        uint32_t x = v;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)lane >= o) x += y; }
        uint32_t at = 0;
        if (lane == 31 && x) at = atomicAdd(&pv.c->bin_n[warp].n, x);
        at = __shfl_sync(0xffffffffu, at, 31);
        if (lane < (uint32_t)BIN_WARPS) s_cnt[warp][lane] = at + x - v;
    }
    __syncthreads();
    uint32_t cursor = s_cnt[lane & 7u][warp];        // lane k: where this warp's next entry of bin k goes (row 7 is never used)
    #pragma unroll
    for (int r = 0; r < BIN_ROUNDS; r++) {
        const uint32_t b = (bins >> (4 * r)) & 7u;
        const uint32_t peers = __match_any_sync(0xffffffffu, b);
        const uint32_t at = __shfl_sync(0xffffffffu, cursor, b) + __popc(peers & lt);
        cursor += (per_round[r / 4] >> (8 * (r & 3))) & 0xffu;
        if (b != (uint32_t)BIN_NONE) {
            RT_DEV_ASSERT(at < pv.bin_stride);
            pv.bin_list[(size_t)b * pv.bin_stride + at] = pos_at(base + 32u * r + lane);
        }
    }
}

template <bool DEFER>
__global__ void __launch_bounds__(SHADE_BLOCK, RT_BINS_MIN_BLOCKS) k_shade_bins(SceneView sv, PoolView pv, FrameParams fp, int cur, uint32_t* live_out) {
    __shared__ uint32_t s_work[NBINS + 1];           // work index where bin b starts (each bin padded to whole warps)
    __shared__ uint32_t s_n[NBINS], s_rank[NBINS];   // entries of bin b; entries of the bins of its group (continuing / ending) before it
    const int next = cur ^ 1;
    const uint32_t lane = threadIdx.x & 31;
    if (threadIdx.x < 32) {   // lane b < NBINS owns bin b: one load latency for all counters, offsets by warp scans
        const uint32_t nb = lane < (uint32_t)NBINS ? pv.c->bin_n[lane].n : 0u;
        const uint32_t padded = (nb + 31u) & ~31u;
        uint32_t w = padded, rc = lane < (uint32_t)BIN_CONT ? nb : 0u, rt = lane < (uint32_t)BIN_CONT ? 0u : nb;
        #pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const uint32_t yw = __shfl_up_sync(0xffffffffu, w, o), yc = __shfl_up_sync(0xffffffffu, rc, o), yt = __shfl_up_sync(0xffffffffu, rt, o);
            if ((int)lane >= o) { w += yw; rc += yc; rt += yt; }
        }
        if (lane < (uint32_t)NBINS) { s_work[lane] = w - padded; s_n[lane] = nb; s_rank[lane] = (lane < (uint32_t)BIN_CONT ? rc : rt) - nb; }
        if (lane == (uint32_t)NBINS - 1u) s_work[NBINS] = w;
    }
    __syncthreads();                                  // the only barrier before the block's work: once per block
    const uint32_t n_work = s_work[NBINS];
    const uint32_t n_end = s_rank[NBINS - 1] + s_n[NBINS - 1];
    const unsigned long long g_first = pv.c->next_sample;     // no shade warp moves the cursor: the turnover adds n_end
    const uint32_t n_regen = g_first < fp.total_samples ? (uint32_t)min((unsigned long long)n_end, fp.total_samples - g_first) : 0u;
    // a warp claims BIN_CLAIM rounds at a time when there is enough work to keep every warp of the grid busy that way, else one
    // (mid-size iterations: 100 k paths are 3 k rounds for 4.7 k warps); the first claim is fixed by the warp's place in the grid
    const uint32_t grid_warps = gridDim.x * (SHADE_BLOCK / 32);
    const uint32_t rounds = n_work >= 32u * BIN_CLAIM * grid_warps ? (uint32_t)BIN_CLAIM : 1u;
    uint32_t w0 = 32u * rounds * (blockIdx.x * (SHADE_BLOCK / 32) + (threadIdx.x >> 5));
    while (w0 < n_work) {
        uint32_t w_next = 0;
        if (lane == 0) w_next = 32u * rounds * grid_warps + atomicAdd(&pv.c->cur_shade, 32u * rounds);   // in flight while this claim is shaded
        #pragma unroll 1
        for (uint32_t r = 0; r < rounds; r++) {
            const uint32_t w = w0 + 32u * r;
            if (w >= n_work) break;
            uint32_t b = 0;
            #pragma unroll
            for (int k = 1; k < NBINS; k++) b += w >= s_work[k];
            const uint32_t j = w - s_work[b] + lane;
            const bool valid = j < s_n[b];                        // false: padding of the bin's last warp
            const uint32_t part = __ballot_sync(0xffffffffu, valid);
            if (!valid) continue;
            const uint32_t pos = pv.bin_list[(size_t)b * pv.bin_stride + j];
            const uint32_t rank = s_rank[b] + j;
            if (b < (uint32_t)BIN_CONT) {
                ShadeOut out;
                const bool goes_on = shade_path(sv, pv, fp, cur, pos, 1u, false, out);   // the class only matters for paths that end
                RT_DEV_ASSERT(rank < pv.capacity);
                if (goes_on) { pv.ray_o[next][rank] = out.o; pv.ray_d[next][rank] = out.d; pv.thr[next][rank] = out.thr; }
                else {   // absorbed (metal scattering into the surface): a dead slot, regenerated one iteration late
                    pv.ray_d[next][rank] = make_float4(0.f, 0.f, 1.f, __uint_as_float(DEAD_BOUNCE << 24));
                    pv.ray_o[next][rank] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
                const uint32_t md = __ballot_sync(part, !goes_on);
                if (md && lane == (uint32_t)(__ffs(md) - 1)) atomicAdd(&pv.c->n_dead[next].n, (uint32_t)__popc(md));
            } else {
                ShadeOut out;
                if (b != 6u) shade_path(sv, pv, fp, cur, pos, b == 5u ? (uint32_t)Q_MISS : (uint32_t)Q_LIGHT, false, out);
                if (rank < n_regen) start_path(sv, pv, fp, next, pv.capacity - n_regen + rank, g_first + rank);
            }
        }
        w0 = __shfl_sync(0xffffffffu, w_next, 0);
    }
    // the last block to get here ends the iteration (k_turnover's job in the block-tile pipeline: one launch less)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&pv.c->shade_done.n, 1u) == gridDim.x - 1u) {
            __threadfence();
            pv.c->shade_done.n = 0;
            turnover(pv.c, cur, fp.total_samples, live_out, true);
        }
    }
}

__global__ void k_scale(const float* sums, float* out, size_t n, float scale) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = scale * sums[i];      // col / float(ns) = (1/ns) * col, render.h:112 + vec3.h:117
}

// FP32 peak probe (rt_measure_fp32_peak): 8 independent FMA chains per thread, operands the compiler cannot fold
__global__ void __launch_bounds__(256) k_fma_peak(float* out, float a, float b, int iters) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    #pragma unroll 1
    for (int i = 0; i < iters; i++) {
        #pragma unroll
        for (int k = 0; k < 16; k++) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

// Deferred extend: a whole warp suspends its stragglers once <= suspend_below lanes are live.  On by default for scenes
// whose walks are long and uneven — above SUSPEND_MIN_PRIMS primitives, i.e. triangle meshes: 1 M triangles +7.8 %
// (lane utilisation of the node loop 0.42 -> 0.61-0.68, extend -11 %) — and off where the node loop is already
// 0.64-0.87 full (every scene the reference ships: the holes suspended paths leave in the shade tiles cost more than
// extend gains; profiles/r01_v9_experiments.md).  rt_render_settings::extend_defer overrides.
uint32_t suspend_threshold(const DeviceScene& ds, const rt_render_settings& st) {
    if (st.extend_defer == RT_DEFER_OFF) return 0u;
    if (st.extend_defer >= 2u) return std::min(st.extend_defer - 1u, 31u);
    return ds.view.n_prims >= SUSPEND_MIN_PRIMS ? SUSPEND_BELOW_DEFAULT : 0u;
}

// traversal variant of k_extend: the caller's, else the default; the shared-memory stack only for trees it can hold
int extend_variant(const DeviceScene& ds, const rt_render_settings& st) {
    if (ds.view.wide)   // the committed scene carries the 4-wide compressed tree (A/B: entry-distance culling of its stack entries)
        return TV_WIDE | (((st.traversal & RT_TRAVERSAL_FORCE) && RT_EXPERIMENTS) ? (int)(st.traversal & TV_TCULL) : 0);
    int v = (st.traversal & RT_TRAVERSAL_FORCE) ? (int)(st.traversal & 23u) : (int)RT_TRAVERSAL_DEFAULT_BITS;
    if (v & TV_PACK) v = TV_PACK;
    if (ds.build.bvh_depth == 0 || ds.build.bvh_depth + 1 > RT_SHORT_STACK) v &= ~TV_SMEM;
    return v;
}

// The measured-and-rejected variants (traversal bits 1-7, the warp-tile shade) are compiled only with -DRT_EXPERIMENTS=1
// (tools/build_variant.sh exp -DRT_EXPERIMENTS=1): the product library carries the default kernels and the wide-tree ones.
template <bool COUNT> void launch_extend(int v, int grid, cudaStream_t s, const SceneView& sv, const PoolView& pv, const FrameParams& fp, int cur) {
    switch (v) {
    case 0: k_extend<COUNT, 0><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
#if RT_EXPERIMENTS
    case 1: k_extend<COUNT, 1><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 2: k_extend<COUNT, 2><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 3: k_extend<COUNT, 3><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 4: k_extend<COUNT, 4><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 5: k_extend<COUNT, 5><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 6: k_extend<COUNT, 6><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 7: k_extend<COUNT, 7><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 9: k_extend<COUNT, 9><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    case 16: k_extend<COUNT, 16><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
#endif
    default: k_extend<COUNT, TV_WIDE><<<grid, EXTEND_BLOCK, 0, s>>>(sv, pv, fp, cur); break;
    }
}
int extend_occupancy(int v, int* occ) {
    switch (v) {
    case 0: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 0>, EXTEND_BLOCK, 0)); break;
#if RT_EXPERIMENTS
    case 1: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 1>, EXTEND_BLOCK, 0)); break;
    case 2: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 2>, EXTEND_BLOCK, 0)); break;
    case 3: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 3>, EXTEND_BLOCK, 0)); break;
    case 4: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 4>, EXTEND_BLOCK, 0)); break;
    case 5: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 5>, EXTEND_BLOCK, 0)); break;
    case 6: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 6>, EXTEND_BLOCK, 0)); break;
    case 7: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 7>, EXTEND_BLOCK, 0)); break;
    case 9: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 9>, EXTEND_BLOCK, 0)); break;
    case 16: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, 16>, EXTEND_BLOCK, 0)); break;
#endif
    default: RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_extend<false, TV_WIDE>, EXTEND_BLOCK, 0)); break;
    }
    return RT_OK;
}

int pool_create(DeviceScene& ds, uint32_t capacity, uint32_t suspend_below) {
    RenderPool* p = new RenderPool;
    p->device = ds.device;
    p->v.capacity = capacity;
    ds.pool = p;                                  // owned by the scene from here on: a failure below is cleaned up with it
    p->suspend_below = suspend_below;
    auto pool_alloc = [p](auto** ptr, size_t bytes) {
        const cudaError_t e = alloc_fault_due() ? cudaErrorMemoryAllocation : cudaMalloc((void**)ptr, bytes);
        if (e == cudaErrorMemoryAllocation) p->out_of_memory = true;
        return e;
    };
    const size_t slots = p->suspend_below ? 2 * (size_t)capacity : capacity;   // + the finished-resumed region at positions >= capacity
    for (int b = 0; b < 2; b++) {
        RT_CUDA(pool_alloc(&p->v.ray_o[b], sizeof(float4) * slots));
        RT_CUDA(pool_alloc(&p->v.ray_d[b], sizeof(float4) * slots));
        RT_CUDA(pool_alloc(&p->v.thr[b], sizeof(float4) * slots));
    }
    RT_CUDA(pool_alloc(&p->v.hit, sizeof(uint2) * slots));
    RT_CUDA(pool_alloc(&p->v.hit_class, sizeof(uint8_t) * slots));
    RT_CUDA(pool_alloc(&p->v.bin_list, sizeof(uint32_t) * slots * NBINS));
    p->v.bin_stride = (uint32_t)slots;
    p->grid_bin = (int)((slots + BIN_SPAN - 1) / BIN_SPAN);
    RT_CUDA(pool_alloc(&p->v.c, sizeof(Counters)));
    RT_CUDA(cudaHostAlloc(&p->h_live, sizeof(uint32_t) * POLL_SLOTS, cudaHostAllocMapped));
    RT_CUDA(cudaHostAlloc(&p->h_counters, sizeof(Counters), cudaHostAllocDefault));
    for (auto& e : p->poll_ev) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    RT_CUDA(cudaEventCreate(&p->ev_begin)); RT_CUDA(cudaEventCreate(&p->ev_end));
    RT_CUDA(cudaStreamCreateWithFlags(&p->capture_stream, cudaStreamNonBlocking));
    int sms = 0, occ = 0;
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ds.device));
    if (p->suspend_below) {
        for (int b = 0; b < 2; b++) {
            RT_CUDA(pool_alloc(&p->v.cont_walk[b], sizeof(uint4) * capacity));
            RT_CUDA(pool_alloc(&p->v.cont_o[b], sizeof(float4) * capacity));
            RT_CUDA(pool_alloc(&p->v.cont_d[b], sizeof(float4) * capacity));
            RT_CUDA(pool_alloc(&p->v.cont_thr[b], sizeof(float4) * capacity));
            RT_CUDA(pool_alloc(&p->v.cont_sp[b], sizeof(uint8_t) * capacity));
            RT_CUDA(pool_alloc(&p->v.cont_stack[b], sizeof(int) * (size_t)capacity * RT_SUSPEND_STACK));
        }
        RT_CUDA((cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_extend_defer<false, false>, EXTEND_BLOCK, 0)));
        p->grid_extend = sms * std::max(occ, 1);
#if RT_EXPERIMENTS
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade_defer, SHADE_BLOCK, 0));
        p->grid_shade = sms * std::max(occ, 1);
#endif
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade_block_defer, SHADE_BLOCK, 0));
        p->grid_shade_block = sms * std::max(occ, 1);
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade_bins<true>, SHADE_BLOCK, 0));
        p->grid_shade_bins = sms * std::max(occ, 1);
    } else {
#if RT_EXPERIMENTS
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade, SHADE_BLOCK, 0));
        p->grid_shade = sms * std::max(occ, 1);
#endif
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade_block, SHADE_BLOCK, 0));
        p->grid_shade_block = sms * std::max(occ, 1);
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade_bins<false>, SHADE_BLOCK, 0));
        p->grid_shade_bins = sms * std::max(occ, 1);
    }
    return RT_OK;
}

int render_impl(DeviceScene& ds, const rt_render_settings& st, float* d_accum, cudaStream_t stream, rt_render_stats* stats) {
    if (st.width == 0 || st.height == 0 || st.samples_per_pixel == 0 || st.max_depth == 0 || st.max_depth > 254) {
        set_error("render: width/height/samples_per_pixel must be > 0 and 1 <= max_depth <= 254");
        return RT_ERR_INVALID;
    }
    const uint32_t s_first = st.sample_first;
    const uint32_t s_count = st.sample_count ? st.sample_count : st.samples_per_pixel - std::min(st.samples_per_pixel, s_first);
    if (s_count == 0) { set_error("render: empty sample range (sample_first >= samples_per_pixel)"); return RT_ERR_INVALID; }
    if ((unsigned long long)s_first + s_count > 0xffffffull) { set_error("render: sample index exceeds 2^24"); return RT_ERR_INVALID; }
    RT_CUDA(cudaSetDevice(ds.device));
    const unsigned long long frame_pix = (unsigned long long)st.width * st.height;
    if (frame_pix >= (1ull << 32)) { set_error("render: more than 2^32 pixels"); return RT_ERR_INVALID; }
    unsigned long long npix = frame_pix;
    const uint32_t tile_count = st.tile_count > 1 ? st.tile_count : 1u;
    if (st.tile_index >= tile_count) { set_error("render: tile_index >= tile_count"); return RT_ERR_INVALID; }
    if (tile_count > 1) {   // pixels in the film blocks this call owns; only the last block of the frame can be short
        const unsigned long long nb = (frame_pix + RT_TILE_BLOCK_PIXELS - 1) / RT_TILE_BLOCK_PIXELS;
        const unsigned long long owned = nb > st.tile_index ? (nb - st.tile_index + tile_count - 1) / tile_count : 0;
        npix = owned * RT_TILE_BLOCK_PIXELS;
        if (owned && (owned - 1) * tile_count + st.tile_index == nb - 1) npix -= nb * RT_TILE_BLOCK_PIXELS - frame_pix;
    }
    const unsigned long long total = npix * s_count;
    // Pool size: the caller's, or DEFAULT_POOL paths (half of it with the deferred extend, whose continuation records make a path
    // four times as big); never more than the frame has samples, a multiple of 32, at most 2^30.  The default is sized for the
    // B200's memory, not for a cache: every iteration ends in a tail in which the longest walks run on a nearly empty GPU, and
    // fewer, larger iterations have fewer tails (final scene 4 Mi -> 16 Mi -> 32 Mi paths: 3393 -> 3564 -> 3598 Msamples/s,
    // 1 M-triangle mesh 4 Mi -> 16 Mi: 1069 -> 1149; profiles/r02_pool_size.md).  133 B per path: 4.5 GB (deferred: ~590 B, 9.4 GB).
    const uint32_t susp = suspend_threshold(ds, st);
    const auto clip = [&](unsigned long long paths) {
        uint32_t c = (uint32_t)std::min<unsigned long long>(paths, 1ull << 30);
        c = (uint32_t)std::min<unsigned long long>(c, std::max<unsigned long long>(total, 32ull));
        return (c + 31u) & ~31u;
    };
    const uint32_t requested = clip(st.pool_paths ? st.pool_paths : (susp ? DEFAULT_POOL / 2 : DEFAULT_POOL));
    if (ds.pool && (ds.pool->requested != requested || ds.pool->suspend_below != susp)) { render_pool_destroy(ds.pool); ds.pool = nullptr; }
    for (uint32_t c = requested; !ds.pool; c = clip(c / 2)) {
        const int rc = pool_create(ds, c, susp);
        if (rc == RT_OK) { ds.pool->requested = requested; break; }
        const bool out_of_memory = ds.pool && ds.pool->out_of_memory;
        render_pool_destroy(ds.pool); ds.pool = nullptr;                          // a half-built pool is never kept
        if (!out_of_memory || st.pool_paths || c <= MIN_DEFAULT_POOL) return rc;   // the caller's size is the caller's choice;
        cudaGetLastError();                                                        // the default adapts: half, down to 1 Mi paths
    }
    const uint32_t capacity = ds.pool->v.capacity;
    RenderPool& P = *ds.pool;
    const int variant = extend_variant(ds, st);
#if !RT_EXPERIMENTS
    if ((variant != 0 && variant != TV_WIDE) || (st.debug_flags & RT_DEBUG_WARP_SHADE)) {   // (unreachable for variant here; the flag is)
        set_error("render: this variant (rt_render_settings::traversal bits / RT_DEBUG_WARP_SHADE) is compiled only into the experiments build (-DRT_EXPERIMENTS=1)");
        return RT_ERR_UNSUPPORTED;
    }
#endif
    if (!P.suspend_below && P.extend_variant != variant) {
        int occ = 0, sms = 0;
        RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ds.device));
        const int rc = extend_occupancy(variant, &occ);
        if (rc < 0) return rc;
        P.grid_extend = sms * std::max(occ, 1);
        P.extend_variant = variant;
    }
    const bool count = (st.flags & RT_RENDER_COUNTERS) != 0;
    const bool timers = (st.flags & RT_RENDER_TIMERS) != 0;
    const bool bin_shade = (st.debug_flags & (RT_DEBUG_BLOCK_SHADE | RT_DEBUG_WARP_SHADE)) == 0;   // the default pipeline
    const bool block_shade = (st.debug_flags & RT_DEBUG_WARP_SHADE) == 0;

    const unsigned long long magic = npix > 1 ? (~0ull / npix) + 1ull : 0ull;   // ceil(2^64 / npix) for npix >= 2 (exact unless npix is a power of two: the fix-up covers it)
    const FrameParams fp{st.width, st.height, st.max_depth, st.seed, s_first, st.debug_flags, total, npix, magic, d_accum, st.tile_index, tile_count};
    // hard bound: every iteration with live paths retires >= 1 segment of a path of <= max_depth segments (+ 1 for a dead slot)
    // (deferred extend: a suspended walk advances by at least one leaf phase per iteration, and only whole warps suspend)
    const uint64_t max_iterations = ((total / capacity + 2) * (uint64_t)(st.max_depth + 1) + POLL_DEPTH + 64) * (P.suspend_below ? 64u : 1u);
    uint32_t* d_live = nullptr;
    RT_CUDA(cudaHostGetDevicePointer(&d_live, P.h_live, 0));
    uint64_t it = 0, launches = 0;
    size_t n_timer = 0;
    bool done = total == 0;
    auto tick = [&]() -> int {   // records the next timer event
        if (!timers) return RT_OK;
        if (n_timer == P.timer_ev.size()) { cudaEvent_t e; RT_CUDA(cudaEventCreate(&e)); P.timer_ev.push_back(e); }
        RT_CUDA(cudaEventRecord(P.timer_ev[n_timer++], stream));
        return RT_OK;
    };

    RT_CUDA(cudaEventRecord(P.ev_begin, stream));
    RT_CUDA(cudaMemsetAsync(P.v.c, 0, sizeof(Counters), stream));
    if (!done) {   // frame start: the first min(capacity, total) samples fill pool 0
        const uint32_t n0 = (uint32_t)std::min<unsigned long long>(capacity, total);
        k_generate<<<(n0 + 255) / 256, 256, 0, stream>>>(ds.view, P.v, fp, n0);
        launches += 1;
    }
    // the three launches of one iteration (cur = parity of the iteration, slot = where turnover publishes the live count)
    auto launch_iteration = [&](int cur, uint32_t slot, cudaStream_t s, bool timed) -> int {
        if (timed && tick() < 0) return RT_ERR_CUDA;
        if (P.suspend_below) {
            if (ds.view.wide) {
                if (count) k_extend_defer<true, true><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, fp, cur, P.suspend_below);
                else k_extend_defer<false, true><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, fp, cur, P.suspend_below);
            } else if (count) k_extend_defer<true, false><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, fp, cur, P.suspend_below);
            else k_extend_defer<false, false><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, fp, cur, P.suspend_below);
        } else if (count) launch_extend<true>(variant, P.grid_extend, s, ds.view, P.v, fp, cur);
        else launch_extend<false>(variant, P.grid_extend, s, ds.view, P.v, fp, cur);
        if (timed && tick() < 0) return RT_ERR_CUDA;
        if (bin_shade) {
            if (P.suspend_below) { k_bin<true><<<P.grid_bin, BIN_THREADS, 0, s>>>(P.v, cur); k_shade_bins<true><<<P.grid_shade_bins, SHADE_BLOCK, 0, s>>>(ds.view, P.v, fp, cur, d_live + slot); }
            else { k_bin<false><<<P.grid_bin, BIN_THREADS, 0, s>>>(P.v, cur); k_shade_bins<false><<<P.grid_shade_bins, SHADE_BLOCK, 0, s>>>(ds.view, P.v, fp, cur, d_live + slot); }
        } else if (block_shade) {
            if (P.suspend_below) k_shade_block_defer<<<P.grid_shade_block, SHADE_BLOCK, 0, s>>>(ds.view, P.v, fp, cur);
            else k_shade_block<<<P.grid_shade_block, SHADE_BLOCK, 0, s>>>(ds.view, P.v, fp, cur);
        }
#if RT_EXPERIMENTS
        else if (P.suspend_below) k_shade_defer<<<P.grid_shade, SHADE_BLOCK, 0, s>>>(ds.view, P.v, fp, cur);
        else k_shade<<<P.grid_shade, SHADE_BLOCK, 0, s>>>(ds.view, P.v, fp, cur);
#endif
        if (timed && tick() < 0) return RT_ERR_CUDA;
        if (!bin_shade) k_turnover<<<1, 1, 0, s>>>(P.v, cur, fp.total_samples, d_live + slot);
        return RT_OK;
    };
    // No per-kernel timers: GRAPH_ITERS iterations go to the GPU as ONE CUDA graph.  Small frames and the drain of
    // every frame are launch-bound (3 launches of a few microseconds per iteration); the graph takes the host out of that
    // loop.  The graph is captured per call (the frame parameters are kernel arguments) and swapped into the pool's
    // instantiated graph with cudaGraphExecUpdate — same topology, no re-instantiation.  The host stays one graph ahead of
    // the live count it reads, so a finished frame costs at most one graph of empty iterations.
    const bool use_graph = !timers && !done && !(st.debug_flags & RT_DEBUG_NO_GRAPH);
    if (use_graph) {
        constexpr int GRAPH_ITERS = POLL_SLOTS;     // a whole period of (parity, live-count slot)
        cudaGraph_t g = nullptr;
        RT_CUDA(cudaStreamBeginCapture(P.capture_stream, cudaStreamCaptureModeThreadLocal));
        int rc = RT_OK;
        for (int k = 0; k < GRAPH_ITERS && rc == RT_OK; k++) rc = launch_iteration(k & 1, (uint32_t)k, P.capture_stream, false);
        const cudaError_t ce = cudaStreamEndCapture(P.capture_stream, &g);
        if (rc < 0 || ce != cudaSuccess || !g) { if (g) cudaGraphDestroy(g); cudaGetLastError(); set_error("render: graph capture failed"); return RT_ERR_CUDA; }
        if (P.graph_exec) {
            cudaGraphExecUpdateResultInfo info;
            if (cudaGraphExecUpdate(P.graph_exec, g, &info) != cudaSuccess) { cudaGetLastError(); cudaGraphExecDestroy(P.graph_exec); P.graph_exec = nullptr; }
        }
        if (!P.graph_exec && cuda_fail(cudaGraphInstantiate(&P.graph_exec, g, 0), "cudaGraphInstantiate", __FILE__, __LINE__)) { cudaGraphDestroy(g); return RT_ERR_CUDA; }
        cudaGraphDestroy(g);
        for (uint64_t gi = 0; !done; gi++) {
            RT_CUDA(cudaGraphLaunch(P.graph_exec, stream));
            RT_CUDA(cudaEventRecord(P.poll_ev[gi & 1], stream));
            launches += 3 * GRAPH_ITERS; it += GRAPH_ITERS;
            if (gi >= 1) {   // the graph before this one: its last iteration's live count (a later one is as good: zero stays zero)
                RT_CUDA(cudaEventSynchronize(P.poll_ev[(gi - 1) & 1]));
                if (P.h_live[GRAPH_ITERS - 1] == 0) done = true;
            }
            if (it > max_iterations + 2 * GRAPH_ITERS) { set_error("render: wavefront did not drain (internal error)"); return RT_ERR_STATE; }
        }
    }
    while (!done) {   // direct launches (per-kernel timers, RT_DEBUG_NO_GRAPH), polled POLL_DEPTH iterations behind
        const int rc = launch_iteration((int)(it & 1), (uint32_t)(it % POLL_SLOTS), stream, true);
        if (rc < 0) return rc;
        RT_CUDA(cudaEventRecord(P.poll_ev[it % POLL_SLOTS], stream));
        launches += 3;
        if (it >= POLL_DEPTH) {
            const uint64_t old = it - POLL_DEPTH;
            RT_CUDA(cudaEventSynchronize(P.poll_ev[old % POLL_SLOTS]));
            if (P.h_live[old % POLL_SLOTS] == 0) done = true;
        }
        if (it > max_iterations) { set_error("render: wavefront did not drain (internal error)"); return RT_ERR_STATE; }
        it++;
    }
    if (tick() < 0) return RT_ERR_CUDA;
    RT_CUDA(cudaMemcpyAsync(P.h_counters, P.v.c, sizeof(Counters), cudaMemcpyDeviceToHost, stream));
    RT_CUDA(cudaEventRecord(P.ev_end, stream));
    RT_CUDA(cudaStreamSynchronize(stream));
    RT_CUDA(cudaGetLastError());
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        const Counters& c = *P.h_counters;
        stats->samples = std::min<unsigned long long>(c.next_sample, total);
        stats->rays = c.total_rays;
        stats->iterations = it;
        stats->kernel_launches = launches;
        stats->nodes_visited = c.nodes;
        stats->prim_tests = c.prim_tests;
        stats->node_slots = c.node_slots;
        stats->pool_paths = capacity;
        if (timers) {   // per-kernel device time, summed over the iterations of this frame: 3 events per iteration + 1
            double ext = 0, shd = 0, fin = 0;
            const std::vector<cudaEvent_t>& ev = P.timer_ev;
            for (size_t k = 0; k + 3 < n_timer; k += 3) {
                float a = 0, b = 0, d = 0;
                cudaEventElapsedTime(&a, ev[k], ev[k + 1]);
                cudaEventElapsedTime(&b, ev[k + 1], ev[k + 2]);
                cudaEventElapsedTime(&d, ev[k + 2], ev[k + 3]);
                ext += a; shd += b; fin += d;
            }
            if (n_timer) cudaEventElapsedTime(&stats->ms_generate, P.ev_begin, ev[0]);
            stats->ms_extend = (float)ext; stats->ms_shade = (float)shd; stats->ms_finish = (float)fin;
        }
        cudaEventElapsedTime(&stats->ms_total, P.ev_begin, P.ev_end);
    }
    return RT_OK;
}

} // namespace
} // namespace rtb

namespace rtb {
int render_device_scene(DeviceScene& ds, const rt_render_settings& st, float* d_accum, cudaStream_t stream, rt_render_stats* stats) {
    return render_impl(ds, st, d_accum, stream, stats);
}
} // namespace rtb

using namespace rtb;

extern "C" int rt_render_device(rt_scene* s, const rt_render_settings* st, float* d_accum, void* cuda_stream, rt_render_stats* stats) {
    if (!s || !st || !d_accum) { set_error("render: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = nullptr;
    const int rc = device_scene_current(s, &ds, "render");
    if (rc < 0) return rc;
    return render_impl(*ds, *st, d_accum, (cudaStream_t)cuda_stream, stats);
}

extern "C" int rt_render(rt_scene* s, const rt_render_settings* st, float* out_rgb, rt_render_stats* stats) {
    if (!s || !st || !out_rgb) { set_error("render: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = nullptr;
    int rc = device_scene_current(s, &ds, "render");
    if (rc < 0) return rc;
    RT_CUDA(cudaSetDevice(ds->device));
    const size_t n = (size_t)st->width * st->height * 3;
    if (n == 0) { set_error("render: empty frame"); return RT_ERR_INVALID; }
    // the accumulation buffer lives with the scene (grow-only): no allocation per call, nothing to leak on an error path
    RT_CUDA(ds->accum.reserve(sizeof(float) * n));
    float* d_accum = (float*)ds->accum.p;
    cudaStream_t stream = ds->stream;
    const bool accumulate = (st->flags & RT_RENDER_ACCUMULATE) != 0;   // progressive: out_rgb carries the sums so far
    if (accumulate) RT_CUDA(cudaMemcpyAsync(d_accum, out_rgb, sizeof(float) * n, cudaMemcpyHostToDevice, stream));
    else RT_CUDA(cudaMemsetAsync(d_accum, 0, sizeof(float) * n, stream));
    rc = render_impl(*ds, *st, d_accum, stream, stats);
    if (rc < 0) return rc;
    if (!accumulate) {
        const uint32_t s_count = st->sample_count ? st->sample_count : st->samples_per_pixel - std::min(st->samples_per_pixel, st->sample_first);
        k_scale<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(d_accum, d_accum, n, 1 / (float)s_count);
    }
    RT_CUDA(cudaMemcpyAsync(out_rgb, d_accum, sizeof(float) * n, cudaMemcpyDeviceToHost, stream));
    RT_CUDA(cudaStreamSynchronize(stream));
    return RT_OK;
}

// Measured FP32 peak of `device` ("measure with an FMA microbenchmark on the box"): a grid of independent FMA
// chains, CUDA events, best of 5.  2 flop per FMA per lane.
extern "C" int rt_measure_fp32_peak(int device, double* tflops) {
    if (!tflops) { set_error("fp32_peak: null out pointer"); return RT_ERR_INVALID; }
    RT_CUDA(cudaSetDevice(device));
    int sms = 0;
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int blocks = sms * 8, threads = 256, iters = 4096;
    float* d = nullptr;
    RT_CUDA(cudaMalloc(&d, sizeof(float) * blocks * threads));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0);
        k_fma_peak<<<blocks, threads>>>(d, 1.0000001f, 1e-7f, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double tf = 2.0 * 8 * 16 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
        if (rep && tf > best) best = tf;      // the first launch warms the clocks up
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    const cudaError_t err = cudaGetLastError();
    cudaFree(d);
    if (cuda_fail(err, "k_fma_peak", __FILE__, __LINE__)) return RT_ERR_CUDA;
    *tflops = best;
    return RT_OK;
}
