// device_api.cu — rt_scene_commit (upload + BVH build) and rt_first_hit.
//
// A commit costs the host next to nothing (round 1: ~40 cudaMalloc/cudaFree, a pageable 20 MB image copy, two stream
// synchronisations around a host-thread SAH build — 13 % of the headline e2e on a 16-core box):
//   * every device allocation belongs to the DeviceScene and only ever grows; a re-commit on the same device reuses all;
//   * the small tables and the primitives go up as ONE copy from a pinned staging image that is filled once per
//     flattening; image textures are expanded to RGBA8 once, into pinned memory, and copied into CUDA arrays whose
//     texture objects survive re-commits of images of the same size;
//   * the hierarchy is built on the device (bvh_build.cu); the host synchronises once, at the end.
#include "device_scene.h"
#include "device_scene.cuh"
#include "traverse.cuh"

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace rtb {

bool cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    if (e == cudaSuccess) return false;
    char buf[512];
    snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
    set_error(buf);
    cudaGetLastError();   // clear the sticky-less error state
    return true;
}

// fault injection for the error-path tests: the n-th growing DevBuf::reserve from now fails with cudaErrorMemoryAllocation
static std::atomic<int> g_alloc_fault{-1};
bool alloc_fault_due() {
    int v = g_alloc_fault.load();
    while (v >= 0) {
        if (g_alloc_fault.compare_exchange_weak(v, v - 1)) return v == 0;
    }
    return false;
}

void device_scene_destroy(DeviceScene* d) {
    if (!d) return;
    if (d->device >= 0) cudaSetDevice(d->device);
    if (d->pool) render_pool_destroy(d->pool);
    for (DevBuf* b : {&d->tables, &d->prims, &d->prim_leaf_id, &d->nodes, &d->nodes_alt, &d->wide, &d->scratch, &d->query, &d->accum}) b->release();
    for (PinnedBuf* b : {&d->staging, &d->rgba, &d->readback}) b->release();
    for (DeviceImage& im : d->images) {
        if (im.tex) cudaDestroyTextureObject(im.tex);
        if (im.array) cudaFreeArray(im.array);
    }
    for (cudaEvent_t& e : d->ev) if (e) cudaEventDestroy(e);
    if (d->stream) cudaStreamDestroy(d->stream);
    delete d;
}

namespace {

DPrim to_dprim(const rt_flat_prim& p) {
    DPrim d;
    d.a = make_float4(p.f[0], p.f[1], p.f[2], p.f[3]);
    d.b = make_float4(p.f[4], p.f[5], p.f[6], p.f[7]);
    d.c0 = p.f[8];
    std::memcpy(&d.attr, &p.f[9], 4);
    d.type_chain = p.type | (p.chain << 4);
    d.material = p.material;
    return d;
}

// layout of DeviceScene::tables (and of its pinned staging image): sections aligned to 256 bytes
struct TableLayout {
    size_t prims_in, leaf_in, bprims, media, active_media, tri_attrs, xops, chains, materials, textures, perlins, images, total;
};
template <class T> size_t section(size_t& off, size_t count) {
    off = (off + 255) & ~size_t(255);
    const size_t at = off;
    off += (count ? count : 1) * sizeof(T);
    return at;
}

int commit_impl(HostScene& hs, DeviceScene& ds, int device, const rt_build_settings& bs) {
    const auto t_begin = std::chrono::steady_clock::now();
    const FlatScene& fs = hs.flat;
    for (const auto& c : fs.chains)
        if (c.count > RT_MAX_CHAIN) { set_error("commit: more than 8 nested translate/rotate_y wrappers"); return RT_ERR_UNSUPPORTED; }
    int ndev = 0;
    RT_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { set_error("commit: no such CUDA device"); return RT_ERR_CUDA; }
    RT_CUDA(cudaSetDevice(device));
    ds.device = device;
    ds.committed_generation = 0;          // nothing valid until this commit succeeds
    if (!ds.stream) RT_CUDA(cudaStreamCreateWithFlags(&ds.stream, cudaStreamNonBlocking));
    for (cudaEvent_t& e : ds.ev) if (!e) RT_CUDA(cudaEventCreate(&e));
    cudaStream_t st = ds.stream;
    const bool trace = (bs.flags & RT_BUILD_TRACE) != 0;
    auto t_last = t_begin;
    auto lap = [&](const char* what) {
        if (!trace) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[commit] %-34s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };

    // A constant_medium whose boundary is ONE sphere can never report a hit under the reference's semantics
    // (SURVEY Q1/Q2): its second boundary query (constant_medium.h:39, t_min = t1 + 0.0001) meets the same near
    // root t1 again, which is either rejected (t1 < t_min) or accepted as t2 == t1 and then dropped by
    // `t1 >= t2` (:46).  Such media are left out of the per-ray loop; indices (RNG blocks, hit encoding) keep
    // counting all media, so results are unchanged.
    std::vector<uint32_t> active_media;
    for (size_t i = 0; i < fs.media.size(); i++) {
        const rt_flat_medium& m = fs.media[i];
        const bool inert = m.bprim_count == 1 && (fs.bprims[m.bprim_first].type == RT_PRIM_SPHERE || fs.bprims[m.bprim_first].type == RT_PRIM_MOVING_SPHERE);
        if (!inert || (hs.semantics & RT_SEM_SPHERE_FAR_ROOT)) active_media.push_back((uint32_t)i);   // with the far root they do scatter
    }

    const size_t n_prims = fs.prims.size();
    TableLayout L{};
    {
        size_t off = 0;
        L.prims_in = section<DPrim>(off, n_prims); L.leaf_in = section<uint32_t>(off, n_prims);
        L.bprims = section<DPrim>(off, fs.bprims.size()); L.media = section<DMedium>(off, fs.media.size());
        L.active_media = section<uint32_t>(off, fs.media.size()); L.tri_attrs = section<rt_flat_tri_attr>(off, fs.tri_attrs.size());
        L.xops = section<rt_flat_xop>(off, fs.xops.size()); L.chains = section<rt_flat_chain>(off, fs.chains.size());
        L.materials = section<rt_flat_material>(off, fs.materials.size()); L.textures = section<rt_flat_texture>(off, fs.textures.size());
        L.perlins = section<DPerlin>(off, fs.perlins.size()); L.images = section<DImage>(off, fs.images.size());
        L.total = (off + 255) & ~size_t(255);
    }
    RT_CUDA(ds.tables.reserve(L.total));
    RT_CUDA(ds.staging.reserve(L.total));
    char* stg = (char*)ds.staging.p;

    // image textures: RGBA8 CUDA arrays behind texture objects (point sampling, unnormalised coordinates)
    size_t rgba_total = 0;
    std::vector<size_t> rgba_off(fs.images.size());
    for (size_t i = 0; i < fs.images.size(); i++) { rgba_off[i] = rgba_total; rgba_total += (((size_t)fs.images[i].width * fs.images[i].height * 4) + 255) & ~size_t(255); }
    RT_CUDA(ds.rgba.reserve(rgba_total));
    const bool restage = ds.staged_generation != fs.generation;
    ds.staged_generation = 0;
    while (ds.images.size() > fs.images.size()) {
        DeviceImage& im = ds.images.back();
        if (im.tex) cudaDestroyTextureObject(im.tex);
        if (im.array) cudaFreeArray(im.array);
        ds.images.pop_back();
    }
    ds.images.resize(fs.images.size());
    for (size_t i = 0; i < fs.images.size(); i++) {
        const rt_flat_image& im = fs.images[i];
        DeviceImage& di = ds.images[i];
        if (di.width != im.width || di.height != im.height || !di.tex) {
            if (di.tex) { cudaDestroyTextureObject(di.tex); di.tex = 0; }
            if (di.array) { cudaFreeArray(di.array); di.array = nullptr; }
            di.width = di.height = 0;
            cudaChannelFormatDesc desc = cudaCreateChannelDesc<uchar4>();
            RT_CUDA(cudaMallocArray(&di.array, &desc, im.width, im.height));
            cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = di.array;
            cudaTextureDesc td{};
            td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
            td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
            RT_CUDA(cudaCreateTextureObject(&di.tex, &rd, &td, nullptr));
            di.width = im.width; di.height = im.height;
        }
        uint8_t* rgba = (uint8_t*)ds.rgba.p + rgba_off[i];
        if (restage) {
            const uint8_t* src = fs.image_bytes.data() + im.offset;
            const size_t npx = (size_t)im.width * im.height, bpp = im.bytes_per_pixel;
            for (size_t k = 0; k < npx; k++) { rgba[4 * k] = src[k * bpp]; rgba[4 * k + 1] = src[k * bpp + 1]; rgba[4 * k + 2] = src[k * bpp + 2]; rgba[4 * k + 3] = 255; }
        }
        RT_CUDA(cudaMemcpy2DToArrayAsync(di.array, 0, 0, rgba, (size_t)im.width * 4, (size_t)im.width * 4, im.height, cudaMemcpyHostToDevice, st));
    }
    lap("image arrays (+ RGBA staging)");

    if (restage) {   // first commit of this flattening: fill the pinned image of the tables
        DPrim* dp = (DPrim*)(stg + L.prims_in);
        uint32_t* dl = (uint32_t*)(stg + L.leaf_in);
        for (size_t i = 0; i < n_prims; i++) { dp[i] = to_dprim(fs.prims[i]); dl[i] = fs.prims[i].leaf_id; }
        DPrim* db = (DPrim*)(stg + L.bprims);
        for (size_t i = 0; i < fs.bprims.size(); i++) db[i] = to_dprim(fs.bprims[i]);
        DMedium* dm = (DMedium*)(stg + L.media);
        for (size_t i = 0; i < fs.media.size(); i++) {
            const rt_flat_medium& m = fs.media[i];
            dm[i] = DMedium{m.chain, m.bprim_first, m.bprim_count, m.material, m.leaf_id, m.neg_inv_density, 0, 0,
                            make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
        }
        auto put = [&](size_t at, const void* src, size_t bytes) { if (bytes) std::memcpy(stg + at, src, bytes); };
        put(L.tri_attrs, fs.tri_attrs.data(), fs.tri_attrs.size() * sizeof(rt_flat_tri_attr));
        put(L.xops, fs.xops.data(), fs.xops.size() * sizeof(rt_flat_xop));
        put(L.chains, fs.chains.data(), fs.chains.size() * sizeof(rt_flat_chain));
        put(L.materials, fs.materials.data(), fs.materials.size() * sizeof(rt_flat_material));
        put(L.textures, fs.textures.data(), fs.textures.size() * sizeof(rt_flat_texture));
        DPerlin* pp = (DPerlin*)(stg + L.perlins);
        for (size_t i = 0; i < fs.perlins.size(); i++) {
            for (int k = 0; k < 256; k++) pp[i].ranvec[k] = make_float4(fs.perlins[i].ranvec[3 * k], fs.perlins[i].ranvec[3 * k + 1], fs.perlins[i].ranvec[3 * k + 2], 0.f);
            for (int a = 0; a < 3; a++) for (int k = 0; k < 256; k++) pp[i].perm[a][k] = (uint8_t)fs.perlins[i].perm[a][k];
        }
    }
    // per-commit words of the staging image: the media that can scatter (depends on the semantics), the texture objects
    if (!active_media.empty()) std::memcpy(stg + L.active_media, active_media.data(), active_media.size() * sizeof(uint32_t));
    DImage* dimg = (DImage*)(stg + L.images);
    for (size_t i = 0; i < fs.images.size(); i++) dimg[i] = DImage{ds.images[i].tex, fs.images[i].width, fs.images[i].height};
    lap(restage ? "tables -> pinned staging (first)" : "tables (staged earlier)");

    RT_CUDA(cudaEventRecord(ds.ev[0], st));
    RT_CUDA(cudaMemcpyAsync(ds.tables.p, stg, L.total, cudaMemcpyHostToDevice, st));
    char* tb = (char*)ds.tables.p;
    SceneView& v = ds.view;
    v.bprims = (const DPrim*)(tb + L.bprims); v.media = (const DMedium*)(tb + L.media);
    v.tri_attrs = (const rt_flat_tri_attr*)(tb + L.tri_attrs); v.xops = (const rt_flat_xop*)(tb + L.xops);
    v.chains = (const rt_flat_chain*)(tb + L.chains); v.materials = (const rt_flat_material*)(tb + L.materials);
    v.textures = (const rt_flat_texture*)(tb + L.textures); v.perlins = (const DPerlin*)(tb + L.perlins);
    v.images = (const DImage*)(tb + L.images);
    v.n_prims = (uint32_t)n_prims; v.n_media = (uint32_t)fs.media.size();
    v.active_media = (const uint32_t*)(tb + L.active_media); v.n_active_media = (uint32_t)active_media.size();
    v.camera = fs.hdr.camera;
    std::memcpy(v.background, fs.hdr.background, sizeof(v.background));
    v.semantics = hs.semantics;
    RT_CUDA(cudaEventRecord(ds.ev[1], st));

    // shutter interval the boxes of moving spheres must cover: the camera's, widened to the reference's
    // bvh_node(…, 0, 1) convention (scenes.h:192)
    const float t_lo = fminf(0.0f, v.camera.time0), t_hi = fmaxf(1.0f, v.camera.time1);
    ds.build = rt_build_stats{};
    int rc = build_bvh(ds, bs, (const DPrim*)(tb + L.prims_in), (const uint32_t*)(tb + L.leaf_in), v.n_prims, t_lo, t_hi, st);
    if (rc < 0) return rc;
    lap("build_bvh (queued)");
    if ((rc = compute_medium_bounds(ds, (DMedium*)(tb + L.media), t_lo, t_hi, st)) < 0) return rc;
    RT_CUDA(cudaEventRecord(ds.ev[8], st));
    RT_CUDA(cudaStreamSynchronize(st));
    lap("synchronise");
    if ((rc = build_bvh_finish(ds, bs, st)) < 0) return rc;
    if ((rc = build_wide(ds, bs, st)) < 0) return rc;
    if (v.n_prims) {
        cudaEvent_t* ev = ds.ev + 2;
        cudaEventElapsedTime(&ds.build.ms_morton, ev[0], ev[1]);
        cudaEventElapsedTime(&ds.build.ms_sort, ev[1], ev[2]);
        cudaEventElapsedTime(&ds.build.ms_hierarchy, ev[2], ev[3]);
        cudaEventElapsedTime(&ds.build.ms_refit, ev[3], ev[4]);
        cudaEventElapsedTime(&ds.build.ms_collapse, ev[4], ev[5]);
    }
    cudaEventElapsedTime(&ds.build.ms_upload, ds.ev[0], ds.ev[1]);
    cudaEventElapsedTime(&ds.build.ms_total, ds.ev[0], ds.ev[8]);
    cudaGetLastError();
    ds.build.n_prims = v.n_prims;
    ds.build.ms_host = (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    ds.staged_generation = fs.generation;
    ds.committed_generation = fs.generation;
    ds.committed_semantics = hs.semantics;
    lap("stats");
    return RT_OK;
}

template <int V>
__global__ void k_first_hit(SceneView sv, const float* rays, int n, float t_min, const float* xi, int32_t* out_id, float* out_t) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < n;                     // closest_hit is warp-collective: idle lanes take part in its votes
    Ray r; r.o = make_float3(0.f, 0.f, 0.f); r.d = make_float3(0.f, 0.f, 1.f); r.tm = 0.f;
    float x = 1.0f;
    if (active) {
        const float* q = rays + 7 * (size_t)i;
        r.o = make_float3(q[0], q[1], q[2]); r.d = make_float3(q[3], q[4], q[5]); r.tm = q[6];
        if (xi) x = xi[i];
    }
    const Closest c = closest_hit<false, V, 1>(sv, r, t_min, xi != nullptr, [x](uint32_t) { return x; }, nullptr, active, nullptr);
    if (!active) return;
    int32_t id = -1;
    if (c.prim >= 0) id = (int32_t)sv.prim_leaf_id[c.prim];
    else if (c.prim <= -2) id = (int32_t)sv.media[-2 - c.prim].leaf_id;
    out_id[i] = id;
    out_t[i] = c.t;
}

// ---- rt_first_hit_static: the first-hit query with its device staging in a file-scope array (no allocation per call) ----
// Layout of the array: rays [7 n] | xi [n] | t [n] | id [n], each section at a 256-byte boundary for n = RT_FIRST_HIT_STATIC_MAX.
constexpr size_t FHS_MAX = RT_FIRST_HIT_STATIC_MAX;
constexpr size_t FHS_OFF_XI = sizeof(float) * 7 * FHS_MAX, FHS_OFF_T = FHS_OFF_XI + sizeof(float) * FHS_MAX,
                 FHS_OFF_ID = FHS_OFF_T + sizeof(float) * FHS_MAX, FHS_BYTES = FHS_OFF_ID + sizeof(int32_t) * FHS_MAX;
static_assert(FHS_OFF_XI % 256 == 0 && FHS_OFF_T % 256 == 0 && FHS_OFF_ID % 256 == 0, "sections stay 256-byte aligned");
static __device__ __align__(256) unsigned char g_first_hit_scratch[FHS_BYTES];   // 2.5 MiB on every GPU that loads the library, until exit

// the same query as k_first_hit; device code names the array directly
template <int V>
__global__ void k_first_hit_static(SceneView sv, int n, float t_min, int with_xi) {
    const float* rays = reinterpret_cast<const float*>(g_first_hit_scratch);
    const float* xi = with_xi ? reinterpret_cast<const float*>(g_first_hit_scratch + FHS_OFF_XI) : nullptr;
    float* out_t = reinterpret_cast<float*>(g_first_hit_scratch + FHS_OFF_T);
    int32_t* out_id = reinterpret_cast<int32_t*>(g_first_hit_scratch + FHS_OFF_ID);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < n;
    Ray r; r.o = make_float3(0.f, 0.f, 0.f); r.d = make_float3(0.f, 0.f, 1.f); r.tm = 0.f;
    float x = 1.0f;
    if (active) {
        const float* q = rays + 7 * (size_t)i;
        r.o = make_float3(q[0], q[1], q[2]); r.d = make_float3(q[3], q[4], q[5]); r.tm = q[6];
        if (xi) x = xi[i];
    }
    const Closest c = closest_hit<false, V, 1>(sv, r, t_min, xi != nullptr, [x](uint32_t) { return x; }, nullptr, active, nullptr);
    if (!active) return;
    int32_t id = -1;
    if (c.prim >= 0) id = (int32_t)sv.prim_leaf_id[c.prim];
    else if (c.prim <= -2) id = (int32_t)sv.media[-2 - c.prim].leaf_id;
    out_id[i] = id;
    out_t[i] = c.t;
}

// HOST address of the array on each device: in host code the array's name is a stand-in, not a device address — it is
// looked up once per device, with that device current (cudaGetSymbolAddress), and cached
constexpr int FHS_MAX_DEVICES = 64;
unsigned char* g_fhs_addr[FHS_MAX_DEVICES] = {};

} // namespace

// The device copy must describe the scene as it is NOW: an edit (or rt_set_semantics) after the commit would otherwise
// render — and checkpoint under the new rt_scene_hash — the old scene without a word.
bool device_scene_describes(rt_scene* s, const DeviceScene& ds) {
    const HostScene& hs = host_of(s);
    return ds.committed_generation && hs.flat.valid && hs.flat.generation == ds.committed_generation && hs.semantics == ds.committed_semantics;
}
bool device_scene_is_current(rt_scene* s) {
    const DeviceScene* ds = device_of(s);
    return ds && device_scene_describes(s, *ds);
}
DeviceScene* device_scene_new() { return new DeviceScene; }
int commit_device_scene(rt_scene* s, DeviceScene& ds, int device, const rt_build_settings& bs) { return commit_impl(host_of(s), ds, device, bs); }

int device_scene_current(rt_scene* s, DeviceScene** out, const char* who) {
    DeviceScene* ds = device_of(s);
    if (!ds || !ds->committed_generation) { set_error(std::string(who) + ": scene not committed (rt_scene_commit)"); return RT_ERR_STATE; }
    if (!device_scene_is_current(s)) { set_error(std::string(who) + ": scene changed since rt_scene_commit (commit again)"); return RT_ERR_STATE; }
    *out = ds;
    return RT_OK;
}

} // namespace rtb

using namespace rtb;

extern "C" int rt_scene_commit_ex(rt_scene* s, int device, const rt_build_settings* settings, rt_build_stats* stats) {
    if (!s) { set_error("null scene"); return RT_ERR_INVALID; }
    HostScene& hs = host_of(s);
    if (!hs.flat.valid) {
        std::string e;
        const int r = hs.flatten(e);
        if (r < 0) { set_error(e); return r; }
    }
    rt_build_settings bs{};
    if (settings) bs = *settings;
    DeviceScene*& slot = device_of(s);
    if (slot && slot->device != device) { device_scene_destroy(slot); slot = nullptr; }   // another device: nothing to reuse
    if (!slot) slot = new DeviceScene;
    const int rc = commit_impl(hs, *slot, device, bs);
    if (rc < 0) return rc;                // the allocations stay with the scene (freed by rt_scene_destroy or reused by a retry)
    if (stats) *stats = slot->build;
    return RT_OK;
}

extern "C" void rt_debug_fail_allocation(int nth) { g_alloc_fault.store(nth); }

extern "C" int rt_scene_commit(rt_scene* s, int device, rt_build_stats* stats) { return rt_scene_commit_ex(s, device, nullptr, stats); }

extern "C" int rt_scene_bvh_nodes(rt_scene* s, void* buf, size_t cap) {
    if (!s || !buf) { set_error("bvh_nodes: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = nullptr;
    const int rc = device_scene_current(s, &ds, "bvh_nodes");
    if (rc < 0) return rc;
    const size_t bytes = sizeof(BNode) * (size_t)ds->view.n_nodes;
    if (cap < bytes) { set_error("bvh_nodes: buffer too small"); return RT_ERR_INVALID; }
    RT_CUDA(cudaSetDevice(ds->device));
    if (bytes) RT_CUDA(cudaMemcpy(buf, ds->view.nodes, bytes, cudaMemcpyDeviceToHost));
    return (int)ds->view.n_nodes;
}

extern "C" int rt_first_hit_static(rt_scene* s, const float* rays, int n, float t_min, const float* xi, int32_t* prim_id, float* t) {
    if (!s || !rays || !prim_id || !t || n < 0) { set_error("first_hit_static: bad argument"); return RT_ERR_INVALID; }
    if ((size_t)n > FHS_MAX) {
        set_error("first_hit_static: n exceeds RT_FIRST_HIT_STATIC_MAX (65536 rays per call): split the ray set or use rt_first_hit");
        return RT_ERR_INVALID;
    }
    DeviceScene* ds = nullptr;
    const int rc = device_scene_current(s, &ds, "first_hit_static");
    if (rc < 0) return rc;
    if (n == 0) return RT_OK;
    if (ds->device < 0 || ds->device >= FHS_MAX_DEVICES) { set_error("first_hit_static: device index out of range"); return RT_ERR_INVALID; }
    RT_CUDA(cudaSetDevice(ds->device));
    unsigned char*& base = g_fhs_addr[ds->device];
    if (!base) {   // once per device, with that device current
        void* p = nullptr;
        RT_CUDA(cudaGetSymbolAddress(&p, g_first_hit_scratch));
        base = (unsigned char*)p;
    }
    const size_t nn = (size_t)n;
    cudaStream_t st = ds->stream;
    RT_CUDA(cudaMemcpyAsync(base, rays, sizeof(float) * 7 * nn, cudaMemcpyHostToDevice, st));
    if (xi) RT_CUDA(cudaMemcpyAsync(base + FHS_OFF_XI, xi, sizeof(float) * nn, cudaMemcpyHostToDevice, st));
    // nothing in the array has to start at zero: every word the kernel reads was just written, every output word below n is overwritten
    if (ds->view.wide) k_first_hit_static<TV_WIDE><<<(n + 127) / 128, 128, 0, st>>>(ds->view, n, t_min, xi ? 1 : 0);
    else k_first_hit_static<TV_PLAIN><<<(n + 127) / 128, 128, 0, st>>>(ds->view, n, t_min, xi ? 1 : 0);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(prim_id, base + FHS_OFF_ID, sizeof(int32_t) * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(t, base + FHS_OFF_T, sizeof(float) * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}

extern "C" int rt_first_hit(rt_scene* s, const float* rays, int n, float t_min, const float* xi, int32_t* prim_id, float* t) {
    if (!s || !rays || !prim_id || !t || n < 0) { set_error("first_hit: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = nullptr;
    const int rc = device_scene_current(s, &ds, "first_hit");
    if (rc < 0) return rc;
    if (n == 0) return RT_OK;
    RT_CUDA(cudaSetDevice(ds->device));
    // one grow-only staging buffer owned by the scene: rays | xi | t | id (nothing to leak on an error path)
    const size_t nn = (size_t)n, off_xi = (sizeof(float) * 7 * nn + 255) & ~size_t(255), off_t = off_xi + ((sizeof(float) * nn + 255) & ~size_t(255)),
                 off_id = off_t + ((sizeof(float) * nn + 255) & ~size_t(255)), total = off_id + sizeof(int32_t) * nn;
    RT_CUDA(ds->query.reserve(total));
    char* q = (char*)ds->query.p;
    float *d_rays = (float*)q, *d_xi = xi ? (float*)(q + off_xi) : nullptr, *d_t = (float*)(q + off_t);
    int32_t* d_id = (int32_t*)(q + off_id);
    cudaStream_t st = ds->stream;
    RT_CUDA(cudaMemcpyAsync(d_rays, rays, sizeof(float) * 7 * nn, cudaMemcpyHostToDevice, st));
    if (xi) RT_CUDA(cudaMemcpyAsync(d_xi, xi, sizeof(float) * nn, cudaMemcpyHostToDevice, st));
    if (ds->view.wide) k_first_hit<TV_WIDE><<<(n + 127) / 128, 128, 0, st>>>(ds->view, d_rays, n, t_min, d_xi, d_id, d_t);
    else k_first_hit<TV_PLAIN><<<(n + 127) / 128, 128, 0, st>>>(ds->view, d_rays, n, t_min, d_xi, d_id, d_t);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(prim_id, d_id, sizeof(int32_t) * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(t, d_t, sizeof(float) * nn, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    return RT_OK;
}
