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
//   shade            persistent blocks claim tiles of 256 consecutive paths, sort the tile by material class in
//                    shared memory (so a warp shades one class), evaluate hit attributes / textures / scatter,
//                    add thr*L of terminated paths to the accumulation buffer (float atomics), REGENERATE
//                    terminated paths from the global sample cursor, and append survivors / new camera paths to
//                    pool `next`.  All allocation is block-aggregated: three atomics per 256 paths.
//   turnover         one thread: roll the counters and publish the live-path count to pinned host memory.
// The pools stay full until the samples run out; the host never synchronises inside a frame (it polls the
// live count a few iterations behind).
#include "device_scene.h"
#include "device_scene.cuh"
#include "shade.cuh"
#include "traverse.cuh"

#include <cuda_pipeline.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace rtb {

constexpr int NQ = 6;                 // material classes: 0 = miss, 1 + rt_material_type
constexpr int Q_MISS = 0;
constexpr int POLL_DEPTH = 4;         // host runs this many iterations ahead of the counter it polls
constexpr int POLL_SLOTS = 8;
constexpr int Q_SUSP = NQ;            // deferred extend: the walk of this path is suspended and the path has moved into its continuation record

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
};

struct PoolView {
    float4* ray_o[2];    // origin.xyz, time
    float4* ray_d[2];    // direction.xyz, bits: sample | bounce << 24
    float4* thr[2];      // throughput rgb, bits: pixel
    uint2* hit;          // written by extend: bits of t, Closest::prim
    uint8_t* hit_class;  // written by extend: material class of the hit
    // continuation records of suspended walks (k_extend_defer), two lists: extend(cur) reads list cur ^ 1 and writes list cur.
    // A record owns its path: the pool slot it came from is a hole (class Q_SUSP) that shade skips.
    uint4* cont_walk[2];   // bits of the closest t so far, its primitive, node, parked leaf
    float4* cont_o[2];     // the path: ray origin + time,
    float4* cont_d[2];     //   direction + sample / bounce word,
    float4* cont_thr[2];   //   throughput + pixel
    uint8_t* cont_sp[2];   // stack entries
    int* cont_stack[2];    // [entry * capacity + record]
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
    int grid_extend = 0, grid_shade = 0, shade_block = 256;
    cudaGraphExec_t graph_exec = nullptr;   // GRAPH_ITERS wavefront iterations as one CUDA graph (render_impl)
    cudaStream_t capture_stream = nullptr;
    RenderPool* second = nullptr;        // two-lane rendering: the second lane's pools (owned)
    cudaStream_t own_stream = nullptr;   //   and its stream
    cudaEvent_t ev_extend = nullptr;     //   recorded after each extend launch of this lane
    int lanes = 1;
    uint32_t suspend_below = 0;          // deferred extend (k_extend_defer): a full warp suspends once <= this many lanes are live; 0 = k_extend
    std::vector<cudaEvent_t> timer_ev;   // RT_RENDER_TIMERS: 3 events per iteration, reused across frames
};

void render_pool_destroy(RenderPool* p) {
    if (!p) return;
    render_pool_destroy(p->second);
    if (p->graph_exec) cudaGraphExecDestroy(p->graph_exec);
    if (p->capture_stream) cudaStreamDestroy(p->capture_stream);
    if (p->own_stream) cudaStreamDestroy(p->own_stream);
    if (p->ev_extend) cudaEventDestroy(p->ev_extend);
    for (int b = 0; b < 2; b++) { cudaFree(p->v.ray_o[b]); cudaFree(p->v.ray_d[b]); cudaFree(p->v.thr[b]); }
    cudaFree(p->v.hit); cudaFree(p->v.hit_class); cudaFree(p->v.c);
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
    uint32_t pad;
    unsigned long long total_samples;   // samples this call renders = npix * sample_count
    unsigned long long npix, npix_magic; // film indices this call covers (the frame, or this tile-split share) and ceil(2^64 / npix)
    float* accum;                       // width*height*3 sums
    uint32_t tile_index, tile_count;    // tile-split: film blocks b with b % tile_count == tile_index (tile_count <= 1: all)
};

constexpr int EXTEND_BLOCK = 128;
constexpr uint32_t SUSPEND_MIN_PRIMS = 65536, SUSPEND_BELOW_DEFAULT = 8;
#ifndef EXTEND_MIN_BLOCKS
#define EXTEND_MIN_BLOCKS 8
#endif

// work index -> pool position: bounced paths first, fresh camera paths at the top end of the pool
__device__ __forceinline__ uint32_t pos_of(uint32_t idx, uint32_t n_bounce, uint32_t n, uint32_t capacity) {
    return idx < n_bounce ? idx : capacity - n + idx;
}

// persistent work fetch: a warp claims 32 consecutive paths
__device__ __forceinline__ uint32_t warp_claim(uint32_t* cursor) {
    uint32_t base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32u);
    return __shfl_sync(0xffffffffu, base, 0);
}

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
    if (fp.pad & 1u) { py = index / fp.width; px = index - py * fp.width; } else film_xy(fp, index, px, py);
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

// ---- extend ---------------------------------------------------------------------------------------
template <bool COUNT, bool PARKED = false>
__global__ void __launch_bounds__(EXTEND_BLOCK, EXTEND_MIN_BLOCKS) k_extend(SceneView sv, PoolView pv, FrameParams fp, int cur) {
    const uint32_t n_b = pv.c->n_bounce[cur], n = n_b + pv.c->n_fresh[cur];
    const uint32_t lane = threadIdx.x & 31;
    const float4* __restrict__ ray_o = pv.ray_o[cur];
    const float4* __restrict__ ray_d = pv.ray_d[cur];
    const float4* thr = pv.thr[cur];
    TraversalCounters tc{0, 0};
    unsigned long long slots = 0;
    for (;;) {
        const uint32_t base = warp_claim(&pv.c->cur_extend);
        if (base >= n) break;
        const uint32_t idx = base + lane;
        const bool active = idx < n;                // closest_hit is warp-collective
        const uint32_t nodes_before = tc.nodes;
        const uint32_t pos = pos_of(active ? idx : base, n_b, n, pv.capacity);
        const float4 o = ray_o[pos], d = ray_d[pos];
        Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
        // the free-flight draw of medium m is keyed (pixel, sample, bounce, m); the pixel is only fetched
        // when a ray really crosses a medium boundary
        const Closest c = closest_hit<COUNT, PARKED>(sv, r, 0.001f, true, [&](uint32_t m) {                      // render.h:63
            const uint32_t sb = __float_as_uint(d.w);
            const RngKey key{__float_as_uint(thr[pos].w), sb & 0xffffffu, sb >> 24, fp.seed};
            return draw4(key, RT_STREAM_MEDIUM, m).x;
        }, &tc, active);
        if (COUNT) slots += 32ull * __reduce_max_sync(0xffffffffu, tc.nodes - nodes_before);   // node steps the warp paid for
        if (active) {
            int q = Q_MISS;
            if (c.prim != -1) {
                const uint32_t mat = c.prim >= 0 ? sv.prims[c.prim].material : sv.media[-2 - c.prim].material;
                q = 1 + (int)sv.materials[mat].type;
            }
            pv.hit[pos] = make_uint2(__float_as_uint(c.t), (uint32_t)c.prim);
            pv.hit_class[pos] = (uint8_t)q;
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
// (class Q_SUSP) that shade skips.  The NEXT iteration's extend takes the suspended walks up again, 32 stragglers to a
// warp, before it starts on the pool (longest work first); a resumed path that finishes its walk is written, with its hit
// record, to the region [capacity, capacity + n_fin) of the current pool, which shade treats like any other path.  A
// suspended path simply bounces an iteration later.  Nothing waits for stragglers: a second pass inside the same
// iteration raised the lane utilisation of the node loop just as much but paid for it with the latency of the longest
// walks at the end of every iteration (profiles/r01_v9_experiments.md).
template <bool COUNT>
#ifndef DEFER_MIN_BLOCKS
#define DEFER_MIN_BLOCKS EXTEND_MIN_BLOCKS
#endif
__global__ void __launch_bounds__(EXTEND_BLOCK, DEFER_MIN_BLOCKS) k_extend_defer(SceneView sv, PoolView pv, FrameParams fp, int cur, uint32_t suspend_below) {
    const uint32_t n_b = pv.c->n_bounce[cur], n_pool = n_b + pv.c->n_fresh[cur];
    const int rd = cur ^ 1, wr = cur;
    const uint32_t n_c = pv.c->cont[rd].n, n_c32 = (n_c + 31u) & ~31u;   // work index space: continuation records (padded to whole warps), then the pool
    const uint32_t n = n_c32 + n_pool;
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    TraversalCounters tc{0, 0};
    uint32_t steps = 0;
    int stack[RT_STACK];
    for (;;) {
        const uint32_t base = warp_claim(&pv.c->cur_extend);
        if (base >= n) break;
        const bool resumed = base < n_c32;                       // warp-uniform
        const uint32_t idx = (resumed ? base : base - n_c32) + lane;
        const bool active = idx < (resumed ? n_c : n_pool);
        Walk w{RT_NODE_DONE, RT_NODE_DONE, 0};
        Closest c; c.t = INFINITY; c.prim = -1;
        uint32_t pos = 0;
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
            if (active && sv.n_prims) w.node = 0;
        }
        Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
        // only whole batches suspend: a few stragglers left over at the end of a list would otherwise advance a leaf phase per iteration
        const uint32_t below = __all_sync(0xffffffffu, active) ? suspend_below : 0u;
        walk_suspendable<COUNT>(sv, r, 0.001f, c, &tc, w, stack, below, &steps);                                    // render.h:63
        const bool suspended = w.node != RT_NODE_DONE || w.parked != RT_NODE_DONE;
        const uint32_t ms = __ballot_sync(0xffffffffu, suspended);
        if (ms) {
            uint32_t slot = 0;
            if (lane == (uint32_t)(__ffs(ms) - 1)) slot = atomicAdd(&pv.c->cont[wr].n, (uint32_t)__popc(ms));
            slot = __shfl_sync(0xffffffffu, slot, __ffs(ms) - 1) + __popc(ms & lt);
            if (suspended) {
                if (!resumed) { th = pv.thr[cur][pos]; pv.hit_class[pos] = (uint8_t)Q_SUSP; }
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
                pos = pv.capacity + k;
                pv.ray_o[cur][pos] = o; pv.ray_d[cur][pos] = d; pv.thr[cur][pos] = th;
            }
        }
        if (finished) {
            // the free-flight draw of medium m is keyed (pixel, sample, bounce, m); the pixel is only fetched when a ray really crosses a boundary
            if (sv.n_active_media) closest_medium(sv, r, 0.001f, c, [&](uint32_t m) {
                const uint32_t sb = __float_as_uint(d.w);
                const RngKey key{__float_as_uint(resumed ? th.w : pv.thr[cur][pos].w), sb & 0xffffffu, sb >> 24, fp.seed};
                return draw4(key, RT_STREAM_MEDIUM, m).x;
            });
            int q = Q_MISS;
            if (c.prim != -1) {
                const uint32_t mat = c.prim >= 0 ? sv.prims[c.prim].material : sv.media[-2 - c.prim].material;
                q = 1 + (int)sv.materials[mat].type;
            }
            pv.hit[pos] = make_uint2(__float_as_uint(c.t), (uint32_t)c.prim);
            pv.hit_class[pos] = (uint8_t)q;
        }
    }
    if (COUNT) {
        atomicAdd(&pv.c->nodes, (unsigned long long)tc.nodes);
        atomicAdd(&pv.c->prim_tests, (unsigned long long)tc.prim_tests);
        if (lane == 0) atomicAdd(&pv.c->node_slots, 32ull * steps);
    }
}

// Streaming variant (closest_surface_stream): finished lanes are refilled from the queue while the rest of the warp
// keeps walking.  Same results ray by ray; only which lane traces which ray changes.
// A refill must not cost a round trip to the queue cursor plus one to the pool (both DRAM latencies, paid with most of
// the warp waiting): every warp keeps a ring of 64 staged rays in shared memory, filled 32 at a time by cp.async from
// a chunk whose queue position was claimed one refill earlier, so the atomic and the loads are in flight while the
// warp traverses and a refill reads shared memory only.
#ifndef RT_REFILL_BELOW
#define RT_REFILL_BELOW 24
#endif
constexpr uint32_t RING = 64, RING_INVALID = 0xffffffffu;
struct StreamShared {
    float4 ring_o[EXTEND_BLOCK / 32][RING], ring_d[EXTEND_BLOCK / 32][RING];   // staged rays, per warp
    float4 lane_o[EXTEND_BLOCK], lane_d[EXTEND_BLOCK];                          // the ray each lane traces (LaneSlots)
    float4 obj_o[EXTEND_BLOCK], obj_d[EXTEND_BLOCK];
    float4 inv_d[EXTEND_BLOCK], inv_o[EXTEND_BLOCK];
    uint32_t ring_pos[EXTEND_BLOCK / 32][RING];
    uint32_t lane_pos[EXTEND_BLOCK];                                            // its pool slot
    uint32_t q[EXTEND_BLOCK / 32][4];                                           // per warp: handed out, staged, -, staged all
};
template <bool COUNT>
__global__ void __launch_bounds__(EXTEND_BLOCK, EXTEND_MIN_BLOCKS) k_extend_stream(SceneView sv, PoolView pv, FrameParams fp, int cur) {
    __shared__ StreamShared sh;
    const uint32_t n_b = pv.c->n_bounce[cur], n = n_b + pv.c->n_fresh[cur];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    TraversalCounters tc{0, 0};
    unsigned long long slots = 0;
    uint32_t next_base = 0;              // lane 0: queue position of the next chunk to stage, claimed one refill ahead
    if (lane == 0) {                     // (a register, so that nobody waits for the atomic before the value is needed)
        sh.q[warp][0] = 0; sh.q[warp][1] = 0; sh.q[warp][3] = 0;
        next_base = atomicAdd(&pv.c->cur_extend, 32u);
    }
    __syncwarp();
    const LaneSlots ls{&sh.lane_o[threadIdx.x], &sh.lane_d[threadIdx.x], &sh.obj_o[threadIdx.x], &sh.obj_d[threadIdx.x],
                       &sh.inv_d[threadIdx.x], &sh.inv_o[threadIdx.x]};
    closest_surface_stream<COUNT, RT_REFILL_BELOW>(sv, 0.001f, ls,                                              // render.h:63
        [&](bool fin, uint32_t fm, bool& valid, Closest& c) {
            const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt = (1u << lane) - 1u;
            float4* ring_o = sh.ring_o[warp];
            float4* ring_d = sh.ring_d[warp];
            uint32_t* ring_pos = sh.ring_pos[warp];
            volatile uint32_t* q = sh.q[warp];
            if (fin && valid) {
                const uint32_t pos = sh.lane_pos[threadIdx.x];
                if (sv.n_active_media) {
                    const Ray r = lane_ray(ls);
                    const uint32_t sb = __float_as_uint(lds4(ls.ray_d).w);
                    closest_medium(sv, r, 0.001f, c, [&](uint32_t m) {   // free-flight draw keyed (pixel, sample, bounce, m)
                        const RngKey key{__float_as_uint(pv.thr[cur][pos].w), sb & 0xffffffu, sb >> 24, fp.seed};
                        return draw4(key, RT_STREAM_MEDIUM, m).x;
                    });
                }
                int qc = Q_MISS;
                if (c.prim != -1) {
                    const uint32_t mat = c.prim >= 0 ? sv.prims[c.prim].material : sv.media[-2 - c.prim].material;
                    qc = 1 + (int)sv.materials[mat].type;
                }
                pv.hit[pos] = make_uint2(__float_as_uint(c.t), (uint32_t)c.prim);
                pv.hit_class[pos] = (uint8_t)qc;
                valid = false;
            }
            // hand out staged rays
            __pipeline_wait_prior(0);
            __syncwarp();
            uint32_t head = q[0], tail = q[1];
            const uint32_t avail = tail - head, want = (uint32_t)__popc(fm), rank = (uint32_t)__popc(fm & lt);
            if (fin && rank < avail) {
                const uint32_t slot = (head + rank) & (RING - 1);
                const uint32_t p = ring_pos[slot];
                if (p != RING_INVALID) {
                    sts4(ls.ray_o, lds4(&ring_o[slot]));
                    sts4(ls.ray_d, lds4(&ring_d[slot]));
                    sh.lane_pos[threadIdx.x] = p;
                    valid = true;
                }
            }
            head += min(want, avail);
            __syncwarp();
            // stage the chunk claimed earlier, claim the one after it
            bool staged_all = q[3] != 0;
            while (!staged_all && tail - head <= RING - 32) {
                const uint32_t base = __shfl_sync(0xffffffffu, next_base, 0);
                if (base >= n) { staged_all = true; break; }
                const uint32_t idx = base + lane, slot = (tail + lane) & (RING - 1);
                if (idx < n) {
                    const uint32_t p = pos_of(idx, n_b, n, pv.capacity);
                    ring_pos[slot] = p;
                    __pipeline_memcpy_async(&ring_o[slot], &pv.ray_o[cur][p], 16);
                    __pipeline_memcpy_async(&ring_d[slot], &pv.ray_d[cur][p], 16);
                } else ring_pos[slot] = RING_INVALID;
                __pipeline_commit();
                tail += 32;
                if (lane == 0) next_base = atomicAdd(&pv.c->cur_extend, 32u);
            }
            if (lane == 0) { q[0] = head; q[1] = tail; q[3] = staged_all ? 1u : 0u; }
            __syncwarp();
            return staged_all && head == tail;
        }, &tc, &slots);
    if (COUNT) {
        atomicAdd(&pv.c->nodes, (unsigned long long)tc.nodes);
        atomicAdd(&pv.c->prim_tests, (unsigned long long)tc.prim_tests);
        if (lane == 0) atomicAdd(&pv.c->node_slots, slots);
    }
}

// ---- shade ----------------------------------------------------------------------------------------
__device__ __forceinline__ void accumulate(float* accum, uint32_t pixel, float3 c) {
    if (c.x != 0.f) atomicAdd(&accum[3 * (size_t)pixel], c.x);
    if (c.y != 0.f) atomicAdd(&accum[3 * (size_t)pixel + 1], c.y);
    if (c.z != 0.f) atomicAdd(&accum[3 * (size_t)pixel + 2], c.z);
}

template <int SHADE_BLOCK, bool DEFER>
__device__ __forceinline__ void shade_tiles(const SceneView& sv, const PoolView& pv, const FrameParams& fp, int cur) {
    constexpr int SHADE_WARPS = SHADE_BLOCK / 32;
    constexpr int NP = NQ;                           // "no path": the tile tail and (deferred extend) the holes suspended paths leave
    constexpr int NC = NP + 1;
    __shared__ uint32_t s_tile[2];                   // this tile / the next one (claimed while this one is shaded)
    constexpr int CELLS = NC * SHADE_WARPS <= 64 ? 2 : 4;   // class-table cells per lane of the scanning warp
    __shared__ uint32_t s_cnt[32 * CELLS];           // [class][warp] counts -> exclusive offsets
    __shared__ uint16_t s_perm[SHADE_BLOCK];         // sorted rank -> tile-local index | class << 12
    __shared__ uint32_t s_wcont[SHADE_WARPS], s_wterm[SHADE_WARPS];
    __shared__ uint32_t s_base_b, s_base_f, s_nregen;
    __shared__ unsigned long long s_g0;
    static_assert(NC * SHADE_WARPS <= 32 * CELLS, "class table must fit one warp scan");

    const int next = cur ^ 1;
    // tile index space: bounced paths, fresh camera paths, (deferred extend) the resumed paths that finished at positions >= capacity
    const uint32_t n_b = pv.c->n_bounce[cur], n_pool = n_b + pv.c->n_fresh[cur];
    const uint32_t n = n_pool + (DEFER ? pv.c->n_fin[cur].n : 0u);
    const unsigned long long iter_begin = pv.c->iter_begin;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    auto pos_at = [&](uint32_t idx) { return DEFER && idx >= n_pool ? pv.capacity + (idx - n_pool) : pos_of(idx, n_b, n_pool, pv.capacity); };
    if (threadIdx.x < 32 * CELLS) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_tile[0] = atomicAdd(&pv.c->cur_shade, (uint32_t)SHADE_BLOCK);
    __syncthreads();
    auto class_of = [&](uint32_t tile) {             // material class of this thread's path in `tile` (NQ past the end)
        const uint32_t idx = tile + threadIdx.x;
        return idx < n ? min((int)pv.hit_class[pos_at(idx)], NP) : (int)NP;      // Q_SUSP == NP
    };
    uint32_t tile = s_tile[0];
    int cls = class_of(tile);
    for (int it = 0;; it ^= 1) {
        if (tile >= n) break;

        // -- sort the tile by material class (counting sort: ballots -> [class][warp] table -> one warp scan)
        uint32_t e = threadIdx.x | ((uint32_t)cls << 12);
        if (!(fp.pad & 32u)) {   // RT_EXP=32: unsorted tiles (A/B of the sort: profiles/r01_v9_experiments.md)
            uint32_t rank = 0;
            #pragma unroll
            for (int k = 0; k < NC; k++) {
                const uint32_t m = __ballot_sync(0xffffffffu, cls == k);
                if (cls == k) rank = __popc(m & lt);
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
            s_perm[s_cnt[cls * SHADE_WARPS + warp] + rank] = (uint16_t)(threadIdx.x | (cls << 12));
            __syncthreads();
            e = s_perm[threadIdx.x];
        }
        const int k = (int)(e >> 12);

        // -- shade one path
        bool continues = false, terminated = false;
        float4 new_o, new_d, new_thr;
        if (k < NP) {
            const uint32_t pos = pos_at(tile + (e & 0xfffu));
            const float4 th = pv.thr[cur][pos];
            const float4 d = pv.ray_d[cur][pos];
            const uint32_t pixel = __float_as_uint(th.w), sb = __float_as_uint(d.w);
            const uint32_t sample = sb & 0xffffffu, bounce = sb >> 24;
            float3 thr = make_float3(th.x, th.y, th.z);
            terminated = true;
            if (k == Q_MISS) {
                accumulate(fp.accum, pixel, thr * f3(sv.background[0], sv.background[1], sv.background[2]));   // render.h:64
            } else {
                const float4 o = pv.ray_o[cur][pos];
                const uint2 h = pv.hit[pos];
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
                } else {
                    const RngKey key{pixel, sample, bounce, fp.seed};
                    float3 attenuation;
                    Ray scattered;
                    if (material_scatter(sv, m, r, rec, key, attenuation, scattered) && bounce + 1 < fp.max_depth) {   // render.h:70-73, :80
                        thr = thr * attenuation;
                        new_o = make_float4(scattered.o.x, scattered.o.y, scattered.o.z, scattered.tm);
                        new_d = make_float4(scattered.d.x, scattered.d.y, scattered.d.z, __uint_as_float(sample | ((bounce + 1) << 24)));
                        new_thr = make_float4(thr.x, thr.y, thr.z, th.w);
                        continues = true;
                        terminated = false;
                    }
                    // absorbed (metal scattering into the surface) or depth exhausted: contributes black
                }
            }
        }

        // -- block-aggregated allocation in pool `next`: survivors at the bottom (tile order kept), regenerated
        //    camera paths (the next samples of the global cursor) at the top; the next tile is claimed here too
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
            // three INDEPENDENT atomics (one round trip, not three): the fresh region needs no counter of its own —
            // samples [iter_begin, next_sample) are regenerated in this iteration, so a block's offset in the region
            // is g0 - iter_begin and k_turnover derives n_fresh from the cursor
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
            pv.ray_o[next][dst] = new_o;
            pv.ray_d[next][dst] = new_d;
            pv.thr[next][dst] = new_thr;
        }
        if (terminated) {
            const uint32_t r = s_wterm[warp] + __popc(mt & lt);
            if (r < s_nregen) start_path(sv, pv, fp, next, pv.capacity - (s_base_f + s_nregen) + r, s_g0 + r);
        }
        tile = tile_next;
        cls = cls_next;
    }
}

template <int SHADE_BLOCK>
__global__ void __launch_bounds__(SHADE_BLOCK) k_shade(SceneView sv, PoolView pv, FrameParams fp, int cur) { shade_tiles<SHADE_BLOCK, false>(sv, pv, fp, cur); }
// shade of the deferred extend: holes (Q_SUSP) in the tiles and the finished-resumed region behind the pool; held to the 64 registers the plain kernel gets
__global__ void __launch_bounds__(256, 4) k_shade_defer(SceneView sv, PoolView pv, FrameParams fp, int cur) { shade_tiles<256, true>(sv, pv, fp, cur); }

// ---- turnover -------------------------------------------------------------------------------------
__global__ void k_turnover(PoolView pv, int cur, unsigned long long total_samples, uint32_t* live_out) {
    Counters* c = pv.c;
    c->total_rays += c->n_bounce[cur] + c->n_fresh[cur];
    c->n_bounce[cur] = 0; c->n_fresh[cur] = 0;
    const unsigned long long started = min(c->next_sample, total_samples);   // samples regenerated by this iteration's shade
    c->n_fresh[cur ^ 1] = (uint32_t)(started - c->iter_begin);
    c->iter_begin = started;
    c->cur_extend = 0; c->cur_shade = 0;
    c->cont[cur ^ 1].n = 0;       // consumed by this iteration's extend; the next iteration's extend appends to it
    c->n_fin[cur].n = 0;
    *live_out = c->n_bounce[cur ^ 1] + c->n_fresh[cur ^ 1] + c->cont[cur].n;   // + the suspended paths the next extend resumes
}

__global__ void k_scale(const float* sums, float* out, size_t n, float scale) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = scale * sums[i];      // col / float(ns) = (1/ns) * col, render.h:112 + vec3.h:117
}

uint32_t suspend_threshold(const DeviceScene& ds) {
    if (const char* e = getenv("RT_SUSPEND")) return (uint32_t)std::min(std::max(atoi(e), 0), 31);
    return ds.view.n_prims >= SUSPEND_MIN_PRIMS ? SUSPEND_BELOW_DEFAULT : 0u;
}

// `owner`: where the pool is kept (DeviceScene::pool, or RenderPool::second for the second lane); set before the first
// allocation so that a failure is cleaned up with the scene.
int pool_create(DeviceScene& ds, RenderPool** owner, uint32_t capacity, int lanes) {
    RenderPool* p = new RenderPool;
    p->device = ds.device;
    p->v.capacity = capacity;
    p->lanes = lanes;
    *owner = p;
    // Deferred extend: a whole warp suspends its stragglers once <= suspend_below lanes are live.  On by default for scenes
    // whose walks are long and uneven — above SUSPEND_MIN_PRIMS primitives, i.e. triangle meshes: 1 M triangles +7.8 %
    // (lane utilisation of the node loop 0.42 -> 0.61-0.68, extend -11 %) — and off where the node loop is already
    // 0.64-0.87 full (every scene the reference ships: the holes suspended paths leave in the shade tiles cost more than
    // extend gains; profiles/r01_v9_experiments.md).  RT_SUSPEND=n forces a threshold, 0 = off.
    p->suspend_below = suspend_threshold(ds);
    const size_t slots = p->suspend_below ? 2 * (size_t)capacity : capacity;   // + the suspended region at positions >= capacity
    for (int b = 0; b < 2; b++) {
        RT_CUDA(cudaMalloc(&p->v.ray_o[b], sizeof(float4) * slots));
        RT_CUDA(cudaMalloc(&p->v.ray_d[b], sizeof(float4) * slots));
        RT_CUDA(cudaMalloc(&p->v.thr[b], sizeof(float4) * slots));
    }
    RT_CUDA(cudaMalloc(&p->v.hit, sizeof(uint2) * slots));
    RT_CUDA(cudaMalloc(&p->v.hit_class, sizeof(uint8_t) * slots));
    RT_CUDA(cudaMalloc(&p->v.c, sizeof(Counters)));
    RT_CUDA(cudaHostAlloc(&p->h_live, sizeof(uint32_t) * POLL_SLOTS, cudaHostAllocMapped));
    RT_CUDA(cudaHostAlloc(&p->h_counters, sizeof(Counters), cudaHostAllocDefault));
    for (auto& e : p->poll_ev) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    RT_CUDA(cudaEventCreate(&p->ev_begin)); RT_CUDA(cudaEventCreate(&p->ev_end));
    RT_CUDA(cudaEventCreateWithFlags(&p->ev_extend, cudaEventDisableTiming));
    RT_CUDA(cudaStreamCreateWithFlags(&p->capture_stream, cudaStreamNonBlocking));
    int sms = 0;
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ds.device));
    int occ = 0;
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_extend<false, true>, EXTEND_BLOCK, 0));
    p->grid_extend = sms * std::max(occ, 1);
    p->shade_block = (getenv("RT_EXP") && (atoi(getenv("RT_EXP")) & 8)) ? 128 : (getenv("RT_EXP") && (atoi(getenv("RT_EXP")) & 64)) ? 512 : 256;
    if (p->shade_block == 128) RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade<128>, 128, 0));
    else if (p->shade_block == 512) RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade<512>, 512, 0));
    else RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade<256>, 256, 0));
    p->grid_shade = sms * std::max(occ, 1);
    if (p->suspend_below) {
        for (int b = 0; b < 2; b++) {
            RT_CUDA(cudaMalloc(&p->v.cont_walk[b], sizeof(uint4) * capacity));
            RT_CUDA(cudaMalloc(&p->v.cont_o[b], sizeof(float4) * capacity));
            RT_CUDA(cudaMalloc(&p->v.cont_d[b], sizeof(float4) * capacity));
            RT_CUDA(cudaMalloc(&p->v.cont_thr[b], sizeof(float4) * capacity));
            RT_CUDA(cudaMalloc(&p->v.cont_sp[b], sizeof(uint8_t) * capacity));
            RT_CUDA(cudaMalloc(&p->v.cont_stack[b], sizeof(int) * (size_t)capacity * RT_SUSPEND_STACK));
        }
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_extend_defer<false>, EXTEND_BLOCK, 0));
        p->grid_extend = sms * std::max(occ, 1);
        RT_CUDA((cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade_defer, 256, 0)));
        p->grid_shade = sms * std::max(occ, 1);
        p->shade_block = 256;
    }
    if (lanes > 1) {
        // two lanes share the SMs: the persistent grids are sized so that one lane's extend and the other lane's shade
        // are resident together (RT_LANE_GRID="extend blocks per SM,shade blocks per SM")
        int eb = 6, sb = 1;
        if (const char* e = getenv("RT_LANE_GRID")) sscanf(e, "%d,%d", &eb, &sb);
        p->grid_extend = sms * std::max(1, std::min(eb, p->grid_extend / sms));
        p->grid_shade = sms * std::max(1, std::min(sb, p->grid_shade / sms));
    }
    if (const char* e = getenv("RT_GRID_DIV")) {   // experiment: several renders sharing one GPU (tools/overlap_probe.py)
        const int div = std::max(1, atoi(e));
        p->grid_extend = std::max(sms, p->grid_extend / div);
        p->grid_shade = std::max(sms, p->grid_shade / div);
    }
    return RT_OK;
}

// One wavefront pipeline of a frame: its pools, its share of the samples, its stream.
struct LaneRun {
    RenderPool* P = nullptr;
    FrameParams fp{};
    cudaStream_t stream = nullptr;
    uint32_t* d_live = nullptr;
    uint64_t it = 0, launches = 0, iterations = 0, max_iterations = 0;
    size_t n_timer = 0;
    bool done = false;
};

int render_impl(DeviceScene& ds, const rt_render_settings& st, float* d_accum, cudaStream_t stream, rt_render_stats* stats) {
    if (st.width == 0 || st.height == 0 || st.samples_per_pixel == 0 || st.max_depth == 0 || st.max_depth > 255) {
        set_error("render: width/height/samples_per_pixel must be > 0 and 1 <= max_depth <= 255");
        return RT_ERR_INVALID;
    }
    const uint32_t s_first = st.sample_first;
    const uint32_t s_count = st.sample_count ? st.sample_count : st.samples_per_pixel - std::min(st.samples_per_pixel, s_first);
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
    // Two lanes (RT_LANES=2): the sample range is cut in two and each half runs as its own wavefront pipeline — pools,
    // counters, stream — into the same accumulation buffer.  The lanes' extend kernels are chained by events so that they
    // alternate, which leaves one lane's latency-bound shade running beside the other lane's extend.  Which lane traces a
    // sample does not change it (Philox keys), so the image is the same up to the order of the float atomics.
    int lanes = getenv("RT_LANES") ? std::min(2, std::max(1, atoi(getenv("RT_LANES")))) : 1;
    if (s_count < 2 || total < (1ull << 16)) lanes = 1;
    uint32_t capacity = st.pool_paths ? st.pool_paths : (1u << 22);
    capacity = (uint32_t)std::min<unsigned long long>(capacity, std::max<unsigned long long>((total + lanes - 1) / lanes, 32ull));
    capacity = (capacity + 31u) & ~31u;
    if (ds.pool && (ds.pool->v.capacity != capacity || ds.pool->lanes != lanes || ds.pool->suspend_below != suspend_threshold(ds))) { render_pool_destroy(ds.pool); ds.pool = nullptr; }
    if (!ds.pool) {
        int rc = pool_create(ds, &ds.pool, capacity, lanes);
        if (rc < 0) return rc;
        if (lanes > 1) {
            rc = pool_create(ds, &ds.pool->second, capacity, lanes);
            if (rc < 0) return rc;
            RT_CUDA(cudaStreamCreateWithFlags(&ds.pool->second->own_stream, cudaStreamNonBlocking));
        }
    }
    const bool count = (st.flags & RT_RENDER_COUNTERS) != 0;
    const bool timers = (st.flags & RT_RENDER_TIMERS) != 0;
    const uint32_t exp_bits = (uint32_t)(getenv("RT_EXP") ? atoi(getenv("RT_EXP")) : 0);

    LaneRun run[2];
    uint32_t first = s_first;
    for (int l = 0; l < lanes; l++) {
        LaneRun& L = run[l];
        L.P = l ? ds.pool->second : ds.pool;
        L.stream = l ? L.P->own_stream : stream;
        const uint32_t share = lanes == 1 ? s_count : (l == 0 ? (s_count + 1) / 2 : s_count / 2);
        const unsigned long long lane_total = npix * share;
        const unsigned long long magic = npix > 1 ? (~0ull / npix) + 1ull : 0ull;   // ceil(2^64 / npix) for npix >= 2 (exact unless npix is a power of two: the fix-up covers it)
        L.fp = FrameParams{st.width, st.height, st.max_depth, st.seed, first, exp_bits, lane_total, npix, magic, d_accum, st.tile_index, tile_count};
        first += share;
        // hard bound: every iteration with live paths retires >= 1 segment of a path of <= max_depth segments
        // (deferred extend: a suspended walk advances by at least one leaf phase per iteration, and only whole warps suspend)
        L.max_iterations = ((lane_total / capacity + 2) * (uint64_t)st.max_depth + POLL_DEPTH + 64) * (L.P->suspend_below ? 64u : 1u);
        RT_CUDA(cudaHostGetDevicePointer(&L.d_live, L.P->h_live, 0));
        L.done = lane_total == 0;
    }
    auto tick = [&](LaneRun& L) -> int {   // records the lane's next timer event on its stream
        if (!timers) return RT_OK;
        RenderPool& P = *L.P;
        if (L.n_timer == P.timer_ev.size()) { cudaEvent_t e; RT_CUDA(cudaEventCreate(&e)); P.timer_ev.push_back(e); }
        RT_CUDA(cudaEventRecord(P.timer_ev[L.n_timer++], L.stream));
        return RT_OK;
    };

    RenderPool& P0 = *ds.pool;
    RT_CUDA(cudaEventRecord(P0.ev_begin, stream));
    for (int l = 0; l < lanes; l++) {
        LaneRun& L = run[l];
        if (l) RT_CUDA(cudaStreamWaitEvent(L.stream, P0.ev_begin, 0));   // the caller's stream orders the buffer before this call
        RT_CUDA(cudaMemsetAsync(L.P->v.c, 0, sizeof(Counters), L.stream));
        if (L.done) continue;
        // frame start: the first min(capacity, total) samples fill pool 0
        const uint32_t n0 = (uint32_t)std::min<unsigned long long>(capacity, L.fp.total_samples);
        k_generate<<<(n0 + 255) / 256, 256, 0, L.stream>>>(ds.view, L.P->v, L.fp, n0);
        L.launches += 1;
    }
    // the three launches of one iteration (cur = parity of the iteration, slot = where turnover publishes the live count)
    auto launch_iteration = [&](LaneRun& L, int cur, uint32_t slot, cudaStream_t s, bool timed) -> int {
        RenderPool& P = *L.P;
        if (timed && tick(L) < 0) return RT_ERR_CUDA;
        if (P.suspend_below) {
            if (count) k_extend_defer<true><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur, P.suspend_below);
            else k_extend_defer<false><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur, P.suspend_below);
        } else if (exp_bits & 16u) {   // RT_EXP=16: streaming refill
            if (count) k_extend_stream<true><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur);
            else k_extend_stream<false><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur);
        } else if (count) k_extend<true, true><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur);
        else if (exp_bits & 4u) k_extend<false, false><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur);   // RT_EXP=4: fused loop
        else k_extend<false, true><<<P.grid_extend, EXTEND_BLOCK, 0, s>>>(ds.view, P.v, L.fp, cur);
        if (lanes > 1) RT_CUDA(cudaEventRecord(P.ev_extend, s));
        if (timed && tick(L) < 0) return RT_ERR_CUDA;
        if (P.suspend_below) k_shade_defer<<<P.grid_shade, 256, 0, s>>>(ds.view, P.v, L.fp, cur);
        else if (P.shade_block == 128) k_shade<128><<<P.grid_shade, 128, 0, s>>>(ds.view, P.v, L.fp, cur);
        else if (P.shade_block == 512) k_shade<512><<<P.grid_shade, 512, 0, s>>>(ds.view, P.v, L.fp, cur);   // RT_EXP=64
        else k_shade<256><<<P.grid_shade, 256, 0, s>>>(ds.view, P.v, L.fp, cur);
        if (timed && tick(L) < 0) return RT_ERR_CUDA;
        k_turnover<<<1, 1, 0, s>>>(P.v, cur, L.fp.total_samples, L.d_live + slot);
        return RT_OK;
    };
    // One lane, no per-kernel timers: GRAPH_ITERS iterations go to the GPU as ONE CUDA graph.  Small frames and the drain of
    // every frame are launch-bound (3 launches of a few microseconds per iteration); the graph takes the host out of that
    // loop.  The graph is captured per call (the frame parameters are kernel arguments) and swapped into the pool's
    // instantiated graph with cudaGraphExecUpdate — same topology, no re-instantiation.  The host stays one graph ahead of
    // the live count it reads, so a finished frame costs at most one graph of empty iterations.
    const bool use_graph = lanes == 1 && !timers && !run[0].done && !getenv("RT_NO_GRAPH");
    if (use_graph) {
        LaneRun& L = run[0];
        RenderPool& P = *L.P;
        constexpr int GRAPH_ITERS = POLL_SLOTS;     // a whole period of (parity, live-count slot)
        cudaGraph_t g = nullptr;
        RT_CUDA(cudaStreamBeginCapture(P.capture_stream, cudaStreamCaptureModeThreadLocal));
        int rc = RT_OK;
        for (int k = 0; k < GRAPH_ITERS && rc == RT_OK; k++) rc = launch_iteration(L, k & 1, (uint32_t)k, P.capture_stream, false);
        const cudaError_t ce = cudaStreamEndCapture(P.capture_stream, &g);
        if (rc < 0 || ce != cudaSuccess || !g) { if (g) cudaGraphDestroy(g); cudaGetLastError(); set_error("render: graph capture failed"); return RT_ERR_CUDA; }
        if (P.graph_exec) {
            cudaGraphExecUpdateResultInfo info;
            if (cudaGraphExecUpdate(P.graph_exec, g, &info) != cudaSuccess) { cudaGetLastError(); cudaGraphExecDestroy(P.graph_exec); P.graph_exec = nullptr; }
        }
        if (!P.graph_exec && cuda_fail(cudaGraphInstantiate(&P.graph_exec, g, 0), "cudaGraphInstantiate", __FILE__, __LINE__)) { cudaGraphDestroy(g); return RT_ERR_CUDA; }
        cudaGraphDestroy(g);
        for (uint64_t gi = 0; !L.done; gi++) {
            RT_CUDA(cudaGraphLaunch(P.graph_exec, stream));
            RT_CUDA(cudaEventRecord(P.poll_ev[gi & 1], stream));
            L.launches += 3 * GRAPH_ITERS; L.iterations += GRAPH_ITERS; L.it += GRAPH_ITERS;
            if (gi >= 1) {   // the graph before this one: its last iteration's live count (a later one is as good: zero stays zero)
                RT_CUDA(cudaEventSynchronize(P.poll_ev[(gi - 1) & 1]));
                if (P.h_live[GRAPH_ITERS - 1] == 0) L.done = true;
            }
            if (L.it > L.max_iterations + 2 * GRAPH_ITERS) { set_error("render: wavefront did not drain (internal error)"); return RT_ERR_STATE; }
        }
    }
    for (bool all_done = use_graph; !all_done;) {
        all_done = true;
        for (int l = 0; l < lanes; l++) {
            LaneRun& L = run[l];
            if (L.done) continue;
            all_done = false;
            RenderPool& P = *L.P;
            // two lanes: this extend starts when the other lane's latest extend has finished, so the two alternate
            if (lanes > 1 && run[l ^ 1].it > 0) RT_CUDA(cudaStreamWaitEvent(L.stream, run[l ^ 1].P->ev_extend, 0));
            const int rc = launch_iteration(L, (int)(L.it & 1), (uint32_t)(L.it % POLL_SLOTS), L.stream, true);
            if (rc < 0) return rc;
            RT_CUDA(cudaEventRecord(P.poll_ev[L.it % POLL_SLOTS], L.stream));
            L.launches += 3;
            L.iterations++;
            if (L.it >= POLL_DEPTH) {
                const uint64_t old = L.it - POLL_DEPTH;
                RT_CUDA(cudaEventSynchronize(P.poll_ev[old % POLL_SLOTS]));
                if (P.h_live[old % POLL_SLOTS] == 0) L.done = true;
            }
            if (L.it > L.max_iterations) { set_error("render: wavefront did not drain (internal error)"); return RT_ERR_STATE; }
            L.it++;
        }
    }
    for (int l = 0; l < lanes; l++) {
        LaneRun& L = run[l];
        if (tick(L) < 0) return RT_ERR_CUDA;
        RT_CUDA(cudaMemcpyAsync(L.P->h_counters, L.P->v.c, sizeof(Counters), cudaMemcpyDeviceToHost, L.stream));
        if (l) {   // the caller's stream continues after the second lane
            RT_CUDA(cudaEventRecord(L.P->ev_end, L.stream));
            RT_CUDA(cudaStreamWaitEvent(stream, L.P->ev_end, 0));
        }
    }
    RT_CUDA(cudaEventRecord(P0.ev_end, stream));
    RT_CUDA(cudaStreamSynchronize(stream));
    RT_CUDA(cudaGetLastError());
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        double ext = 0, shd = 0, fin = 0;
        float ms_gen = 0.f;
        for (int l = 0; l < lanes; l++) {
            LaneRun& L = run[l];
            const Counters& c = *L.P->h_counters;
            stats->samples += std::min<unsigned long long>(c.next_sample, L.fp.total_samples);
            stats->rays += c.total_rays;
            stats->iterations += L.iterations;
            stats->kernel_launches += L.launches;
            stats->nodes_visited += c.nodes;
            stats->prim_tests += c.prim_tests;
            stats->node_slots += c.node_slots;
            if (timers) {   // per-kernel device time, summed over the iterations of this frame: 3 events per iteration + 1
                            // (two lanes: the intervals of the lanes overlap in time; the sums say what each kernel occupied)
                const std::vector<cudaEvent_t>& ev = L.P->timer_ev;
                for (size_t k = 0; k + 3 < L.n_timer; k += 3) {
                    float a = 0, b = 0, d = 0;
                    cudaEventElapsedTime(&a, ev[k], ev[k + 1]);
                    cudaEventElapsedTime(&b, ev[k + 1], ev[k + 2]);
                    cudaEventElapsedTime(&d, ev[k + 2], ev[k + 3]);
                    ext += a; shd += b; fin += d;
                }
                if (l == 0 && L.n_timer) cudaEventElapsedTime(&ms_gen, P0.ev_begin, ev[0]);
            }
        }
        cudaEventElapsedTime(&stats->ms_total, P0.ev_begin, P0.ev_end);
        stats->ms_extend = (float)ext; stats->ms_shade = (float)shd; stats->ms_generate = ms_gen; stats->ms_finish = (float)fin;
    }
    return RT_OK;
}

} // namespace
} // namespace rtb

using namespace rtb;

extern "C" int rt_render_device(rt_scene* s, const rt_render_settings* st, float* d_accum, void* cuda_stream, rt_render_stats* stats) {
    if (!s || !st || !d_accum) { set_error("render: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = device_of(s);
    if (!ds) { set_error("render: scene not committed (rt_scene_commit)"); return RT_ERR_STATE; }
    return render_impl(*ds, *st, d_accum, (cudaStream_t)cuda_stream, stats);
}

extern "C" int rt_render(rt_scene* s, const rt_render_settings* st, float* out_rgb, rt_render_stats* stats) {
    if (!s || !st || !out_rgb) { set_error("render: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = device_of(s);
    if (!ds) { set_error("render: scene not committed (rt_scene_commit)"); return RT_ERR_STATE; }
    RT_CUDA(cudaSetDevice(ds->device));
    const size_t n = (size_t)st->width * st->height * 3;
    float* d_accum = nullptr;
    RT_CUDA(cudaMalloc(&d_accum, sizeof(float) * n));
    const bool accumulate = (st->flags & RT_RENDER_ACCUMULATE) != 0;   // progressive: out_rgb carries the sums so far
    if (accumulate) RT_CUDA(cudaMemcpyAsync(d_accum, out_rgb, sizeof(float) * n, cudaMemcpyHostToDevice, 0));
    else RT_CUDA(cudaMemsetAsync(d_accum, 0, sizeof(float) * n, 0));
    const int rc = render_impl(*ds, *st, d_accum, 0, stats);
    if (rc < 0) { cudaFree(d_accum); return rc; }
    if (!accumulate) {
        const uint32_t s_count = st->sample_count ? st->sample_count : st->samples_per_pixel - std::min(st->samples_per_pixel, st->sample_first);
        k_scale<<<(unsigned)((n + 255) / 256), 256>>>(d_accum, d_accum, n, 1 / (float)s_count);
    }
    RT_CUDA(cudaMemcpy(out_rgb, d_accum, sizeof(float) * n, cudaMemcpyDeviceToHost));
    cudaFree(d_accum);
    return RT_OK;
}
