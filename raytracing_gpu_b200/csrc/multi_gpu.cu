// multi_gpu.cu — rt_render_multi: one frame on several GPUs of one node behind the C-ABI (the reference is
// single-GPU: draw() launches one render<<<>>> per batch, render.h:118-162).
//
// One host thread per device commits the scene there (if that device's copy is not current) and renders the device's share
// into its own zero-initialised accumulation buffer; the root then runs ONE reduce kernel that reads every other device's
// buffer over NVLink through peer mappings, sums in device order, scales by 1 / samples and leaves the frame on the root.
// No NCCL: the exchange is one read of (n - 1) x W*H*3 floats by one GPU (99.5 MB per peer at 3840 x 2159), which a plain
// peer-load kernel moves at link speed; bench.py's torchrun path keeps the NCCL reduce for the one-process-per-GPU layout.
#include "device_scene.h"
#include "device_scene.cuh"

#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

namespace rtb {
namespace {

struct PeerBuffers { const float* p[RT_MAX_DEVICES]; int n; };

// out[i] = scale * (((p0[i] + p1[i]) + p2[i]) + ...): fixed device order, so a given device count reproduces bit for bit.
// 128-bit loads; every block streams one contiguous span of every buffer (peer reads are plain coalesced global loads).
__global__ void __launch_bounds__(256) k_reduce_peers(PeerBuffers in, float* out, size_t n4, size_t n, float scale) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 a = reinterpret_cast<const float4*>(in.p[0])[i];
        for (int k = 1; k < in.n; k++) {
            const float4 b = reinterpret_cast<const float4*>(in.p[k])[i];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        reinterpret_cast<float4*>(out)[i] = make_float4(scale * a.x, scale * a.y, scale * a.z, scale * a.w);
    }
    if (blockIdx.x == 0)   // the tail (n is W*H*3: not always a multiple of 4)
        for (size_t i = 4 * n4 + threadIdx.x; i < n; i += blockDim.x) {
            float a = in.p[0][i];
            for (int k = 1; k < in.n; k++) a += in.p[k][i];
            out[i] = scale * a;
        }
}

struct Worker {
    DeviceScene* ds = nullptr;
    int device = -1;
    int rc = RT_OK;
    std::string err;
    rt_render_stats stats{};
    float ms_commit = 0.f;
};

} // namespace
} // namespace rtb

using namespace rtb;

extern "C" int rt_render_multi(rt_scene* s, const rt_render_settings* st_in, const rt_multi_settings* ms, float* out_rgb, rt_multi_stats* stats) {
    if (!s || !st_in || !out_rgb) { set_error("render_multi: bad argument"); return RT_ERR_INVALID; }
    if (st_in->tile_count > 1 || st_in->tile_index) { set_error("render_multi: tile_index / tile_count must be 0 (the split is this call's)"); return RT_ERR_INVALID; }
    if (st_in->flags & RT_RENDER_ACCUMULATE) { set_error("render_multi: RT_RENDER_ACCUMULATE is not supported"); return RT_ERR_INVALID; }
    const auto t0 = std::chrono::steady_clock::now();
    int visible = 0;
    RT_CUDA(cudaGetDeviceCount(&visible));
    if (visible <= 0) { set_error("render_multi: no CUDA device"); return RT_ERR_CUDA; }
    std::vector<int> devs;
    if (!ms || ms->n_devices == 0) for (int d = 0; d < visible && d < RT_MAX_DEVICES; d++) devs.push_back(d);
    else {
        if (ms->n_devices > RT_MAX_DEVICES) { set_error("render_multi: more than RT_MAX_DEVICES devices"); return RT_ERR_INVALID; }
        for (uint32_t k = 0; k < ms->n_devices; k++) {
            const int d = ms->device_ids[k];
            if (d < 0 || d >= visible) { set_error("render_multi: no such CUDA device"); return RT_ERR_INVALID; }
            for (int e : devs) if (e == d) { set_error("render_multi: a device is listed twice"); return RT_ERR_INVALID; }
            devs.push_back(d);
        }
    }
    const uint32_t split = ms ? ms->split_mode : (uint32_t)RT_SPLIT_SAMPLE;
    if (split > RT_SPLIT_TILE) { set_error("render_multi: unknown split_mode"); return RT_ERR_INVALID; }
    const int n = (int)devs.size();
    // the sample range this call renders (rt_render's convention), split over the devices below
    const uint32_t s_first = st_in->sample_first;
    const uint32_t spp = st_in->sample_count ? st_in->sample_count : st_in->samples_per_pixel - std::min(st_in->samples_per_pixel, s_first);
    const size_t nf = (size_t)st_in->width * st_in->height * 3;
    if (nf == 0 || spp == 0) { set_error("render_multi: empty frame"); return RT_ERR_INVALID; }

    // host side of the scene: flatten once, before the threads read it
    HostScene& hs = host_of(s);
    if (!hs.flat.valid) {
        std::string e;
        const int r = hs.flatten(e);
        if (r < 0) { set_error(e); return r; }
    }
    // a device copy per device: the scene's own (rt_scene_commit) where the device matches, replicas elsewhere
    std::vector<Worker> W(n);
    for (int k = 0; k < n; k++) {
        W[k].device = devs[k];
        DeviceScene*& own = device_of(s);
        if (own && own->device == devs[k]) { W[k].ds = own; continue; }
        for (DeviceScene* r : replicas_of(s)) if (r->device == devs[k]) W[k].ds = r;
        if (!W[k].ds) {
            if (!own) { own = device_scene_new(); own->device = devs[k]; W[k].ds = own; }
            else { DeviceScene* r = device_scene_new(); r->device = devs[k]; replicas_of(s).push_back(r); W[k].ds = r; }
        }
    }

    auto work = [&](int k) {
        Worker& w = W[k];
        DeviceScene& ds = *w.ds;
        auto fail = [&](int rc) { w.rc = rc; w.err = last_error_copy(); };
        if (cuda_fail(cudaSetDevice(w.device), "cudaSetDevice", __FILE__, __LINE__)) return fail(RT_ERR_CUDA);
        if (!device_scene_describes(s, ds)) {
            const rt_build_settings bs{};
            const int rc = commit_device_scene(s, ds, w.device, bs);
            if (rc < 0) return fail(rc);
            w.ms_commit = ds.build.ms_host;
        }
        if (cuda_fail(ds.accum.reserve(sizeof(float) * nf), "accum", __FILE__, __LINE__)) return fail(RT_ERR_CUDA);
        if (cuda_fail(cudaMemsetAsync(ds.accum.p, 0, sizeof(float) * nf, ds.stream), "memset", __FILE__, __LINE__)) return fail(RT_ERR_CUDA);
        rt_render_settings st = *st_in;
        bool idle = false;
        if (split == RT_SPLIT_TILE) { st.tile_index = (uint32_t)k; st.tile_count = (uint32_t)n; st.sample_first = s_first; st.sample_count = spp; }
        else {   // contiguous sample ranges, the remainder to the first devices (raytracing_gpu_b200/parallel.py::sample_range)
            const uint32_t base = spp / (uint32_t)n, rem = spp % (uint32_t)n;
            st.sample_count = base + ((uint32_t)k < rem ? 1u : 0u);
            st.sample_first = s_first + (uint32_t)k * base + std::min<uint32_t>((uint32_t)k, rem);
            idle = st.sample_count == 0;
        }
        if (!idle) {
            const int rc = render_device_scene(ds, st, (float*)ds.accum.p, ds.stream, &w.stats);
            if (rc < 0) return fail(rc);
        } else if (cuda_fail(cudaStreamSynchronize(ds.stream), "sync", __FILE__, __LINE__)) return fail(RT_ERR_CUDA);
    };
    if (n == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int k = 0; k < n; k++) th.emplace_back(work, k);
        for (auto& t : th) t.join();
    }
    for (int k = 0; k < n; k++) if (W[k].rc < 0) { set_error("render_multi: device " + std::to_string(W[k].device) + ": " + W[k].err); return W[k].rc; }

    // ---- the one reduce, on the root
    DeviceScene& root = *W[0].ds;
    RT_CUDA(cudaSetDevice(root.device));
    bool peer = true;
    for (int k = 1; k < n && peer; k++) {
        int can = 0;
        RT_CUDA(cudaDeviceCanAccessPeer(&can, root.device, W[k].device));
        if (!can) { peer = false; break; }
        const cudaError_t e = cudaDeviceEnablePeerAccess(W[k].device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) peer = false;
        cudaGetLastError();
    }
    PeerBuffers pb{};
    pb.n = n;
    pb.p[0] = (const float*)root.accum.p;
    if (peer) for (int k = 1; k < n; k++) pb.p[k] = (const float*)W[k].ds->accum.p;
    else if (n > 1) {   // staged: the peers' buffers are copied next to the root's first
        RT_CUDA(root.query.reserve(sizeof(float) * nf * (size_t)(n - 1)));
        for (int k = 1; k < n; k++) {
            float* dst = (float*)root.query.p + nf * (size_t)(k - 1);
            RT_CUDA(cudaMemcpyPeerAsync(dst, root.device, W[k].ds->accum.p, W[k].device, sizeof(float) * nf, root.stream));
            pb.p[k] = dst;
        }
    }
    RT_CUDA(cudaEventRecord(root.ev[6], root.stream));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, root.device);
    k_reduce_peers<<<sms * 8, 256, 0, root.stream>>>(pb, (float*)root.accum.p, nf / 4, nf, 1 / (float)spp);   // col / float(ns), render.h:112
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaEventRecord(root.ev[7], root.stream));
    RT_CUDA(cudaMemcpyAsync(out_rgb, root.accum.p, sizeof(float) * nf, cudaMemcpyDeviceToHost, root.stream));
    RT_CUDA(cudaStreamSynchronize(root.stream));
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->n_devices = (uint32_t)n;
        stats->peer_reduce = peer ? 1u : 0u;
        for (int k = 0; k < n; k++) {
            stats->ms_render[k] = W[k].stats.ms_total;
            stats->ms_commit_max = std::max(stats->ms_commit_max, W[k].ms_commit);
            stats->samples += W[k].stats.samples; stats->rays += W[k].stats.rays;
        }
        cudaEventElapsedTime(&stats->ms_reduce, root.ev[6], root.ev[7]);
        stats->ms_wall = (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return RT_OK;
}
