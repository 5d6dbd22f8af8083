// wavefront.cu — streaming wavefront path tracer (replaces the render megakernel + color_f, render.h:55-113).
//
// A fixed pool of path slots lives in HBM as SoA.  One iteration =
//   extend    persistent warps pull 32 queued paths at a time, traverse the LBVH + media, write the hit, and append
//             the path to the shade queue of its material class (warp-ballot aggregated appends)
//   shade     one launch over all material queues; each warp works inside ONE queue, so a warp shades one
//             material class; terminated paths add thr*L to the accumulation buffer (float atomics) and are
//             REGENERATED in place: the slot fetches the next (pixel, sample) from a global cursor (one atomic
//             per warp) and starts a new camera path; survivors and regenerated paths are appended to the next
//             extend queue (warp-ballot compaction)
//   turnover  a one-thread kernel that rolls queue counters over and publishes the live-path count
// (generate, the camera ray-generation kernel, fills the pool once at the start of a frame.)
// The pool stays full until the samples run out, so kernels stay large; no host synchronisation inside the loop
// (the host polls a pinned counter a few iterations behind).
#include "device_scene.h"
#include "device_scene.cuh"
#include "shade.cuh"
#include "traverse.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

namespace rtb {

constexpr int NQ = 6;                 // shade queues: 0 = miss, 1 + rt_material_type
constexpr int Q_MISS = 0;
constexpr int POLL_DEPTH = 4;         // host runs this many iterations ahead of the counter it polls
constexpr int POLL_SLOTS = 8;

struct Counters {
    uint32_t n_extend[2];
    uint32_t n_shade[NQ];
    uint32_t n_new;
    uint32_t cur_extend, cur_shade, cur_new;
    unsigned long long next_sample;   // global sample cursor (sample-major: g = sample * npix + pixel)
    unsigned long long total_rays, started, nodes, prim_tests;
};

struct PoolView {
    float4* ray_o;     // origin.xyz, time
    float4* ray_d;     // direction.xyz, hit t (written by extend)
    int* hit_prim;     // Closest::prim
    float4* thr;       // throughput rgb
    uint2* pid;        // pixel, sample | bounce << 24
    uint32_t* q_extend[2];
    uint32_t* q_shade; // NQ * capacity
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
    int grid_extend = 0, grid_shade = 0, grid_gen = 0;
    std::vector<cudaEvent_t> timer_ev;   // RT_RENDER_TIMERS: 4 events per iteration, reused across frames
};

void render_pool_destroy(RenderPool* p) {
    if (!p) return;
    cudaFree(p->v.ray_o); cudaFree(p->v.ray_d); cudaFree(p->v.hit_prim); cudaFree(p->v.thr); cudaFree(p->v.pid);
    cudaFree(p->v.q_extend[0]); cudaFree(p->v.q_extend[1]); cudaFree(p->v.q_shade); cudaFree(p->v.c);
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
    unsigned long long npix, npix_magic; // frame size and ceil(2^64 / npix)
    float* accum;                       // width*height*3 sums
};

constexpr int BLOCK = 128;

// warp-aggregated append: all lanes call; lanes with `pred` get consecutive slots in the queue
__device__ __forceinline__ void warp_append(bool pred, uint32_t* queue, uint32_t* counter, uint32_t value) {
    const uint32_t mask = __ballot_sync(0xffffffffu, pred);
    if (!mask) return;
    const uint32_t lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == (uint32_t)(__ffs(mask) - 1)) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
    if (pred) queue[base + __popc(mask & ((1u << lane) - 1u))] = value;
}

// persistent work fetch: a warp claims 32 consecutive queue entries
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

// camera ray generation for one slot (render.h:105-108)
__device__ __forceinline__ void start_path(const SceneView& sv, const PoolView& pv, const FrameParams& fp, uint32_t slot, unsigned long long g) {
    uint32_t pixel, sample;
    split_sample_index(fp, g, pixel, sample);
    const RngKey key{pixel, sample, 0u, fp.seed};
    const uint32_t py = pixel / fp.width, px = pixel - py * fp.width;
    const Ray r = camera_ray(sv.camera, key, (int)px, (int)py, (int)fp.width, (int)fp.height);
    pv.ray_o[slot] = make_float4(r.o.x, r.o.y, r.o.z, r.tm);
    pv.ray_d[slot] = make_float4(r.d.x, r.d.y, r.d.z, 0.f);
    pv.thr[slot] = make_float4(1.f, 1.f, 1.f, 0.f);
    pv.pid[slot] = make_uint2(pixel, sample);   // bounce 0
}

// lanes with `want` claim consecutive global sample indices: one 64-bit atomic per warp
__device__ __forceinline__ unsigned long long claim_samples(bool want, unsigned long long* cursor) {
    const uint32_t mask = __ballot_sync(0xffffffffu, want);
    if (!mask) return ~0ull;
    const uint32_t lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    unsigned long long g0 = 0;
    if ((int)lane == leader) g0 = atomicAdd(cursor, (unsigned long long)__popc(mask));
    g0 = __shfl_sync(0xffffffffu, g0, leader);
    return g0 + __popc(mask & ((1u << lane) - 1u));
}

// ---- generate -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_generate(SceneView sv, PoolView pv, FrameParams fp, int next) {
    // start of a frame: every slot [0, capacity) is free
    const uint32_t n = pv.capacity;
    const uint32_t lane = threadIdx.x & 31;
    for (;;) {
        const uint32_t base = warp_claim(&pv.c->cur_new);
        if (base >= n) break;
        const uint32_t slot = base + lane;
        const bool have_slot = slot < n;
        const unsigned long long g = claim_samples(have_slot, &pv.c->next_sample);
        const bool live = have_slot && g < fp.total_samples;
        if (live) start_path(sv, pv, fp, slot, g);
        warp_append(live, pv.q_extend[next], &pv.c->n_extend[next], slot);
    }
}

// ---- extend ---------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(BLOCK) k_extend(SceneView sv, PoolView pv, FrameParams fp, int cur) {
    const uint32_t n = pv.c->n_extend[cur];
    const uint32_t lane = threadIdx.x & 31;
    TraversalCounters tc{0, 0};
    for (;;) {
        const uint32_t base = warp_claim(&pv.c->cur_extend);
        if (base >= n) break;
        const uint32_t idx = base + lane;
        const bool valid = idx < n;
        uint32_t slot = 0;
        int q = -1;
        if (valid) {
            slot = pv.q_extend[cur][idx];
            const float4 o = pv.ray_o[slot], d = pv.ray_d[slot];
            const uint2 id = pv.pid[slot];
            Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
            const RngKey key{id.x, id.y & 0xffffffu, id.y >> 24, fp.seed};
            const Closest c = closest_hit<COUNT>(sv, r, 0.001f, &key, nullptr, &tc);      // render.h:63
            pv.ray_d[slot].w = c.t;
            pv.hit_prim[slot] = c.prim;
            if (c.prim == -1) q = Q_MISS;
            else {
                const uint32_t mat = c.prim >= 0 ? sv.prims[c.prim].material : sv.media[-2 - c.prim].material;
                q = 1 + (int)sv.materials[mat].type;
            }
        }
        #pragma unroll
        for (int k = 0; k < NQ; k++) warp_append(q == k, pv.q_shade + (size_t)k * pv.capacity, &pv.c->n_shade[k], slot);
    }
    if (COUNT) {
        atomicAdd(&pv.c->nodes, (unsigned long long)tc.nodes);
        atomicAdd(&pv.c->prim_tests, (unsigned long long)tc.prim_tests);
    }
}

// ---- shade ----------------------------------------------------------------------------------------
__device__ __forceinline__ void accumulate(float* accum, uint32_t pixel, float3 c) {
    if (c.x != 0.f) atomicAdd(&accum[3 * (size_t)pixel], c.x);
    if (c.y != 0.f) atomicAdd(&accum[3 * (size_t)pixel + 1], c.y);
    if (c.z != 0.f) atomicAdd(&accum[3 * (size_t)pixel + 2], c.z);
}

__global__ void __launch_bounds__(BLOCK) k_shade(SceneView sv, PoolView pv, FrameParams fp, int next) {
    // chunk table: queue k owns warp-chunks [first[k], first[k+1])
    uint32_t first[NQ + 1];
    first[0] = 0;
    #pragma unroll
    for (int k = 0; k < NQ; k++) first[k + 1] = first[k] + (pv.c->n_shade[k] + 31u) / 32u;
    const uint32_t lane = threadIdx.x & 31;
    for (;;) {
        uint32_t chunk = 0;
        if (lane == 0) chunk = atomicAdd(&pv.c->cur_shade, 1u);
        chunk = __shfl_sync(0xffffffffu, chunk, 0);
        if (chunk >= first[NQ]) break;
        int k = 0;
        #pragma unroll
        for (int t = 1; t < NQ; t++) if (chunk >= first[t]) k = t;
        const uint32_t idx = (chunk - first[k]) * 32u + lane;
        const bool valid = idx < pv.c->n_shade[k];
        bool continues = false, terminated = false;
        uint32_t slot = 0;
        if (valid) {
            slot = pv.q_shade[(size_t)k * pv.capacity + idx];
            const float4 th = pv.thr[slot];
            const uint2 id = pv.pid[slot];
            float3 thr = make_float3(th.x, th.y, th.z);
            const uint32_t bounce = id.y >> 24;
            terminated = true;
            if (k == Q_MISS) {
                accumulate(fp.accum, id.x, thr * f3(sv.background[0], sv.background[1], sv.background[2]));   // render.h:64
            } else {
                const float4 o = pv.ray_o[slot], d = pv.ray_d[slot];
                Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
                const int prim = pv.hit_prim[slot];
                HitAttr rec;
                uint32_t mat;
                if (prim >= 0) { const DPrim p = sv.prims[prim]; prim_attr(sv, p, r, d.w, rec); mat = p.material; }
                else { const DMedium m = sv.media[-2 - prim]; medium_attr(sv, m, r, d.w, rec); mat = m.material; }
                const rt_flat_material m = sv.materials[mat];
                if (m.type == RT_MAT_DIFFUSE_LIGHT) {
                    accumulate(fp.accum, id.x, thr * texture_value(sv, m.texture, rec.u, rec.v, rec.p));           // render.h:68, :75
                } else {
                    const RngKey key{id.x, id.y & 0xffffffu, bounce, fp.seed};
                    float3 attenuation;
                    Ray scattered;
                    if (material_scatter(sv, m, r, rec, key, attenuation, scattered) && bounce + 1 < fp.max_depth) {   // render.h:70-73, :80
                        thr = thr * attenuation;
                        pv.ray_o[slot] = make_float4(scattered.o.x, scattered.o.y, scattered.o.z, scattered.tm);
                        pv.ray_d[slot] = make_float4(scattered.d.x, scattered.d.y, scattered.d.z, 0.f);
                        pv.thr[slot] = make_float4(thr.x, thr.y, thr.z, 0.f);
                        pv.pid[slot] = make_uint2(id.x, (id.y & 0xffffffu) | ((bounce + 1) << 24));
                        continues = true;
                        terminated = false;
                    }
                    // absorbed (metal scattering into the surface) or depth exhausted: contributes black
                }
            }
        }
        // regeneration: a terminated path's slot immediately starts the next camera sample (if any are left)
        const unsigned long long g = claim_samples(terminated, &pv.c->next_sample);
        if (terminated && g < fp.total_samples) { start_path(sv, pv, fp, slot, g); continues = true; }
        warp_append(continues, pv.q_extend[next], &pv.c->n_extend[next], slot);
    }
}

// ---- turnover -------------------------------------------------------------------------------------
__global__ void k_turnover(PoolView pv, int cur, int next, uint32_t* live_out) {
    Counters* c = pv.c;
    c->total_rays += c->n_extend[cur];
    c->n_extend[cur] = 0;
    for (int k = 0; k < NQ; k++) c->n_shade[k] = 0;
    c->n_new = 0;
    c->cur_extend = 0; c->cur_shade = 0; c->cur_new = 0;
    *live_out = c->n_extend[next];
}

__global__ void k_scale(const float* sums, float* out, size_t n, float scale) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = scale * sums[i];      // col / float(ns) = (1/ns) * col, render.h:112 + vec3.h:117
}

int pool_create(DeviceScene& ds, uint32_t capacity) {
    RenderPool* p = new RenderPool;
    p->device = ds.device;
    p->v.capacity = capacity;
    RT_CUDA(cudaMalloc(&p->v.ray_o, sizeof(float4) * capacity));
    RT_CUDA(cudaMalloc(&p->v.ray_d, sizeof(float4) * capacity));
    RT_CUDA(cudaMalloc(&p->v.hit_prim, sizeof(int) * capacity));
    RT_CUDA(cudaMalloc(&p->v.thr, sizeof(float4) * capacity));
    RT_CUDA(cudaMalloc(&p->v.pid, sizeof(uint2) * capacity));
    RT_CUDA(cudaMalloc(&p->v.q_extend[0], sizeof(uint32_t) * capacity));
    RT_CUDA(cudaMalloc(&p->v.q_extend[1], sizeof(uint32_t) * capacity));
    RT_CUDA(cudaMalloc(&p->v.q_shade, sizeof(uint32_t) * (size_t)capacity * NQ));
    RT_CUDA(cudaMalloc(&p->v.c, sizeof(Counters)));
    RT_CUDA(cudaHostAlloc(&p->h_live, sizeof(uint32_t) * POLL_SLOTS, cudaHostAllocMapped));
    RT_CUDA(cudaHostAlloc(&p->h_counters, sizeof(Counters), cudaHostAllocDefault));
    for (auto& e : p->poll_ev) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    RT_CUDA(cudaEventCreate(&p->ev_begin)); RT_CUDA(cudaEventCreate(&p->ev_end));
    int sms = 0;
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ds.device));
    int occ = 0;
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_extend<false>, BLOCK, 0));
    p->grid_extend = sms * std::max(occ, 1);
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade, BLOCK, 0));
    p->grid_shade = sms * std::max(occ, 1);
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_generate, BLOCK, 0));
    p->grid_gen = sms * std::max(occ, 1);
    ds.pool = p;
    return RT_OK;
}

int render_impl(DeviceScene& ds, const rt_render_settings& st, float* d_accum, cudaStream_t stream, rt_render_stats* stats) {
    if (st.width == 0 || st.height == 0 || st.samples_per_pixel == 0 || st.max_depth == 0 || st.max_depth > 255) {
        set_error("render: width/height/samples_per_pixel must be > 0 and 1 <= max_depth <= 255");
        return RT_ERR_INVALID;
    }
    const uint32_t s_first = st.sample_first;
    const uint32_t s_count = st.sample_count ? st.sample_count : st.samples_per_pixel - std::min(st.samples_per_pixel, s_first);
    if ((unsigned long long)s_first + s_count > 0xffffffull) { set_error("render: sample index exceeds 2^24"); return RT_ERR_INVALID; }
    RT_CUDA(cudaSetDevice(ds.device));
    const unsigned long long npix = (unsigned long long)st.width * st.height;
    const unsigned long long total = npix * s_count;
    uint32_t capacity = st.pool_paths ? st.pool_paths : (1u << 21);
    capacity = (uint32_t)std::min<unsigned long long>(capacity, std::max<unsigned long long>(total, 32ull));
    capacity = (capacity + 31u) & ~31u;
    if (ds.pool && ds.pool->v.capacity != capacity) { render_pool_destroy(ds.pool); ds.pool = nullptr; }
    if (!ds.pool) { const int rc = pool_create(ds, capacity); if (rc < 0) return rc; }
    RenderPool& P = *ds.pool;
    const bool count = (st.flags & RT_RENDER_COUNTERS) != 0;
    const bool timers = (st.flags & RT_RENDER_TIMERS) != 0;
    size_t n_timer = 0;
    auto tick = [&]() -> int {   // records the next timer event on the stream
        if (!timers) return RT_OK;
        if (n_timer == P.timer_ev.size()) { cudaEvent_t e; RT_CUDA(cudaEventCreate(&e)); P.timer_ev.push_back(e); }
        RT_CUDA(cudaEventRecord(P.timer_ev[n_timer++], stream));
        return RT_OK;
    };

    const unsigned long long magic = npix > 1 ? (~0ull / npix) + 1ull : 0ull;   // ceil(2^64 / npix) for npix >= 2 (exact unless npix is a power of two: the fix-up covers it)
    FrameParams fp{st.width, st.height, st.max_depth, st.seed, s_first, 0u, total, npix, magic, d_accum};
    uint64_t launches = 0, iterations = 0;
    RT_CUDA(cudaMemsetAsync(P.v.c, 0, sizeof(Counters), stream));
    RT_CUDA(cudaEventRecord(P.ev_begin, stream));
    if (total > 0) {
        // frame start: fill the pool
        k_generate<<<P.grid_gen, BLOCK, 0, stream>>>(ds.view, P.v, fp, 0);
        uint32_t* d_live = nullptr;
        RT_CUDA(cudaHostGetDevicePointer(&d_live, P.h_live, 0));
        k_turnover<<<1, 1, 0, stream>>>(P.v, 1, 0, d_live);     // clears cursors; "cur" = the empty queue 1
        launches += 2;
        bool done = false;
        // hard bound: every iteration with live paths retires >= 1 segment of a path of <= max_depth segments
        const uint64_t max_iterations = (total / capacity + 2) * (uint64_t)st.max_depth + POLL_DEPTH + 64;
        for (uint64_t it = 0; !done; it++) {
            const int cur = (int)(it & 1), next = cur ^ 1;
            if (tick() < 0) return RT_ERR_CUDA;
            if (count) k_extend<true><<<P.grid_extend, BLOCK, 0, stream>>>(ds.view, P.v, fp, cur);
            else k_extend<false><<<P.grid_extend, BLOCK, 0, stream>>>(ds.view, P.v, fp, cur);
            if (tick() < 0) return RT_ERR_CUDA;
            k_shade<<<P.grid_shade, BLOCK, 0, stream>>>(ds.view, P.v, fp, next);
            if (tick() < 0) return RT_ERR_CUDA;
            if (tick() < 0) return RT_ERR_CUDA;
            k_turnover<<<1, 1, 0, stream>>>(P.v, cur, next, d_live + (it % POLL_SLOTS));
            RT_CUDA(cudaEventRecord(P.poll_ev[it % POLL_SLOTS], stream));
            launches += 3;
            iterations++;
            if (it >= POLL_DEPTH) {
                const uint64_t old = it - POLL_DEPTH;
                RT_CUDA(cudaEventSynchronize(P.poll_ev[old % POLL_SLOTS]));
                if (P.h_live[old % POLL_SLOTS] == 0) done = true;
            }
            if (it > max_iterations) { set_error("render: wavefront did not drain (internal error)"); return RT_ERR_STATE; }
        }
    }
    RT_CUDA(cudaEventRecord(P.ev_end, stream));
    RT_CUDA(cudaMemcpyAsync(P.h_counters, P.v.c, sizeof(Counters), cudaMemcpyDeviceToHost, stream));
    RT_CUDA(cudaStreamSynchronize(stream));
    RT_CUDA(cudaGetLastError());
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->samples = std::min<unsigned long long>(P.h_counters->next_sample, total);
        stats->rays = P.h_counters->total_rays;
        stats->iterations = iterations;
        stats->kernel_launches = launches;
        stats->nodes_visited = P.h_counters->nodes;
        stats->prim_tests = P.h_counters->prim_tests;
        cudaEventElapsedTime(&stats->ms_total, P.ev_begin, P.ev_end);
        if (timers) {   // per-kernel device time, summed over the iterations of this frame
            double ext = 0, shd = 0, gen = 0, fin = 0;
            for (size_t k = 0; k + 3 < n_timer; k += 4) {
                float a = 0, b = 0, c = 0;
                cudaEventElapsedTime(&a, P.timer_ev[k], P.timer_ev[k + 1]);
                cudaEventElapsedTime(&b, P.timer_ev[k + 1], P.timer_ev[k + 2]);
                cudaEventElapsedTime(&c, P.timer_ev[k + 2], P.timer_ev[k + 3]);
                ext += a; shd += b; gen += c;
                if (k + 4 < n_timer) { float d = 0; cudaEventElapsedTime(&d, P.timer_ev[k + 3], P.timer_ev[k + 4]); fin += d; }
            }
            stats->ms_extend = (float)ext; stats->ms_shade = (float)shd; stats->ms_generate = (float)gen; stats->ms_finish = (float)fin;
        }
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
    RT_CUDA(cudaMemsetAsync(d_accum, 0, sizeof(float) * n, 0));
    const int rc = render_impl(*ds, *st, d_accum, 0, stats);
    if (rc < 0) { cudaFree(d_accum); return rc; }
    const uint32_t s_count = st->sample_count ? st->sample_count : st->samples_per_pixel - std::min(st->samples_per_pixel, st->sample_first);
    k_scale<<<(unsigned)((n + 255) / 256), 256>>>(d_accum, d_accum, n, 1 / (float)s_count);
    RT_CUDA(cudaMemcpy(out_rgb, d_accum, sizeof(float) * n, cudaMemcpyDeviceToHost));
    cudaFree(d_accum);
    return RT_OK;
}
