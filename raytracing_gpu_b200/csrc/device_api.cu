// device_api.cu — rt_scene_commit (upload + LBVH build) and rt_first_hit.
#include "device_scene.h"
#include "device_scene.cuh"
#include "traverse.cuh"

#include <cstdio>
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

void device_scene_destroy(DeviceScene* d) {
    if (!d) return;
    if (d->device >= 0) cudaSetDevice(d->device);
    if (d->pool) render_pool_destroy(d->pool);
    void* ptrs[] = {d->d_prims, d->d_prim_leaf_id, d->d_bprims, d->d_media, d->d_tri_attrs, d->d_xops, d->d_chains,
                    d->d_materials, d->d_textures, d->d_perlins, d->d_images, d->d_nodes, d->d_active_media};
    for (void* p : ptrs) if (p) cudaFree(p);
    for (uint32_t i = 0; i < d->n_images; i++) {
        if (d->image_objs && d->image_objs[i]) cudaDestroyTextureObject(d->image_objs[i]);
        if (d->image_arrays && d->image_arrays[i]) cudaFreeArray(d->image_arrays[i]);
    }
    delete[] d->image_objs;
    delete[] d->image_arrays;
    delete d;
}

namespace {

template <class T> int upload(void** dst, const std::vector<T>& v, cudaStream_t st) {
    RT_CUDA(cudaMalloc(dst, (v.empty() ? 1 : v.size()) * sizeof(T)));
    if (!v.empty()) RT_CUDA(cudaMemcpyAsync(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, st));
    return RT_OK;
}

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

int commit_impl(HostScene& hs, DeviceScene& ds, int device) {
    const FlatScene& fs = hs.flat;
    for (const auto& c : fs.chains)
        if (c.count > RT_MAX_CHAIN) { set_error("commit: more than 8 nested translate/rotate_y wrappers"); return RT_ERR_UNSUPPORTED; }
    int ndev = 0;
    RT_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { set_error("commit: no such CUDA device"); return RT_ERR_CUDA; }
    RT_CUDA(cudaSetDevice(device));
    ds.device = device;
    cudaStream_t st = 0;
    cudaEvent_t e0, e1;
    RT_CUDA(cudaEventCreate(&e0)); RT_CUDA(cudaEventCreate(&e1));
    RT_CUDA(cudaEventRecord(e0, st));
    // RT_COMMIT_TRACE=1: host wall clock of the commit phases on stderr (tools/commit_probe.py)
    const bool trace = getenv("RT_COMMIT_TRACE") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!trace) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[commit] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };

    if (fs.commit_leaf_ids.size() != fs.prims.size() || fs.commit_bprims.size() != fs.bprims.size() * sizeof(DPrim)) {   // first commit of this flattening
        fs.commit_prims.resize(fs.prims.size() * sizeof(DPrim));
        fs.commit_bprims.resize(fs.bprims.size() * sizeof(DPrim));
        fs.commit_leaf_ids.resize(fs.prims.size());
        DPrim* dp = reinterpret_cast<DPrim*>(fs.commit_prims.data());
        DPrim* db = reinterpret_cast<DPrim*>(fs.commit_bprims.data());
        for (size_t i = 0; i < fs.prims.size(); i++) { dp[i] = to_dprim(fs.prims[i]); fs.commit_leaf_ids[i] = fs.prims[i].leaf_id; }
        for (size_t i = 0; i < fs.bprims.size(); i++) db[i] = to_dprim(fs.bprims[i]);
    }
    const std::vector<uint8_t>&prims = fs.commit_prims, &bprims = fs.commit_bprims;
    const std::vector<uint32_t>& leaf_ids = fs.commit_leaf_ids;
    const size_t n_prims = fs.prims.size();
    std::vector<DMedium> media(fs.media.size());
    for (size_t i = 0; i < fs.media.size(); i++) {
        const rt_flat_medium& m = fs.media[i];
        media[i] = DMedium{m.chain, m.bprim_first, m.bprim_count, m.material, m.leaf_id, m.neg_inv_density, 0, 0,
                           make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
    }
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
    lap("prims -> device records");
    std::vector<DPerlin> perlins(fs.perlins.size());
    for (size_t i = 0; i < fs.perlins.size(); i++) {
        for (int k = 0; k < 256; k++) perlins[i].ranvec[k] = make_float4(fs.perlins[i].ranvec[3 * k], fs.perlins[i].ranvec[3 * k + 1], fs.perlins[i].ranvec[3 * k + 2], 0.f);
        for (int a = 0; a < 3; a++) for (int k = 0; k < 256; k++) perlins[i].perm[a][k] = (uint8_t)fs.perlins[i].perm[a][k];
    }

    // image textures -> RGBA8 CUDA arrays behind texture objects (point sampling, unnormalised coordinates)
    ds.n_images = (uint32_t)fs.images.size();
    ds.image_arrays = new cudaArray_t[ds.n_images ? ds.n_images : 1]();
    ds.image_objs = new cudaTextureObject_t[ds.n_images ? ds.n_images : 1]();
    std::vector<DImage> images(fs.images.size());
    for (size_t i = 0; i < fs.images.size(); i++) {
        const rt_flat_image& im = fs.images[i];
        if (fs.commit_rgba.size() != fs.images.size()) fs.commit_rgba.assign(fs.images.size(), std::vector<uint8_t>());
        std::vector<uint8_t>& rgba = fs.commit_rgba[i];
        if (rgba.size() != (size_t)im.width * im.height * 4) {   // first commit of this flattening
            rgba.resize((size_t)im.width * im.height * 4);
            const uint8_t* src = fs.image_bytes.data() + im.offset;
            for (size_t k = 0; k < (size_t)im.width * im.height; k++) {
                rgba[4 * k] = src[k * im.bytes_per_pixel]; rgba[4 * k + 1] = src[k * im.bytes_per_pixel + 1];
                rgba[4 * k + 2] = src[k * im.bytes_per_pixel + 2]; rgba[4 * k + 3] = 255;
            }
        }
        cudaChannelFormatDesc desc = cudaCreateChannelDesc<uchar4>();
        RT_CUDA(cudaMallocArray(&ds.image_arrays[i], &desc, im.width, im.height));
        RT_CUDA(cudaMemcpy2DToArray(ds.image_arrays[i], 0, 0, rgba.data(), (size_t)im.width * 4, (size_t)im.width * 4, im.height, cudaMemcpyHostToDevice));
        cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = ds.image_arrays[i];
        cudaTextureDesc td{};
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
        RT_CUDA(cudaCreateTextureObject(&ds.image_objs[i], &rd, &td, nullptr));
        images[i] = DImage{ds.image_objs[i], im.width, im.height};
    }

    lap("perlin + image textures");
    void *d_prims_in = nullptr, *d_leaf_in = nullptr;
    int rc;
    if ((rc = upload(&d_prims_in, prims, st)) < 0) return rc;
    if ((rc = upload(&d_leaf_in, leaf_ids, st)) < 0) return rc;
    if ((rc = upload(&ds.d_bprims, bprims, st)) < 0) return rc;
    if ((rc = upload(&ds.d_media, media, st)) < 0) return rc;
    if ((rc = upload(&ds.d_active_media, active_media, st)) < 0) return rc;
    if ((rc = upload(&ds.d_tri_attrs, fs.tri_attrs, st)) < 0) return rc;
    if ((rc = upload(&ds.d_xops, fs.xops, st)) < 0) return rc;
    if ((rc = upload(&ds.d_chains, fs.chains, st)) < 0) return rc;
    if ((rc = upload(&ds.d_materials, fs.materials, st)) < 0) return rc;
    if ((rc = upload(&ds.d_textures, fs.textures, st)) < 0) return rc;
    if ((rc = upload(&ds.d_perlins, perlins, st)) < 0) return rc;
    if ((rc = upload(&ds.d_images, images, st)) < 0) return rc;

    lap("cudaMalloc + H2D (queued)");
    SceneView& v = ds.view;
    v.bprims = (const DPrim*)ds.d_bprims; v.media = (const DMedium*)ds.d_media;
    v.tri_attrs = (const rt_flat_tri_attr*)ds.d_tri_attrs; v.xops = (const rt_flat_xop*)ds.d_xops;
    v.chains = (const rt_flat_chain*)ds.d_chains; v.materials = (const rt_flat_material*)ds.d_materials;
    v.textures = (const rt_flat_texture*)ds.d_textures; v.perlins = (const DPerlin*)ds.d_perlins;
    v.images = (const DImage*)ds.d_images;
    v.n_prims = (uint32_t)n_prims; v.n_media = (uint32_t)media.size();
    v.active_media = (const uint32_t*)ds.d_active_media; v.n_active_media = (uint32_t)active_media.size();
    v.camera = fs.hdr.camera;
    std::memcpy(v.background, fs.hdr.background, sizeof(v.background));
    v.semantics = hs.semantics;
    RT_CUDA(cudaEventRecord(e1, st));

    // shutter interval the boxes of moving spheres must cover: the camera's, widened to the reference's
    // bvh_node(…, 0, 1) convention (scenes.h:192)
    const float t_lo = fminf(0.0f, v.camera.time0), t_hi = fmaxf(1.0f, v.camera.time1);
    cudaEvent_t b0, b1;
    RT_CUDA(cudaEventCreate(&b0)); RT_CUDA(cudaEventCreate(&b1));
    RT_CUDA(cudaEventRecord(b0, st));
    rc = build_lbvh(ds, (const DPrim*)d_prims_in, (const uint32_t*)d_leaf_in, v.n_prims, t_lo, t_hi, st);
    if (rc < 0) return rc;
    lap("build_lbvh");
    if ((rc = compute_medium_bounds(ds, t_lo, t_hi, st)) < 0) return rc;
    RT_CUDA(cudaEventRecord(b1, st));
    RT_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_prims_in); cudaFree(d_leaf_in);
    lap("medium bounds + sync + free");
    cudaEventElapsedTime(&ds.build.ms_upload, e0, e1);
    cudaEventElapsedTime(&ds.build.ms_total, e0, b1);
    ds.build.n_prims = v.n_prims;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(b0); cudaEventDestroy(b1);
    return RT_OK;
}

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
    const Closest c = closest_hit<false>(sv, r, t_min, xi != nullptr, [x](uint32_t) { return x; }, nullptr, active);
    if (!active) return;
    int32_t id = -1;
    if (c.prim >= 0) id = (int32_t)sv.prim_leaf_id[c.prim];
    else if (c.prim <= -2) id = (int32_t)sv.media[-2 - c.prim].leaf_id;
    out_id[i] = id;
    out_t[i] = c.t;
}

} // namespace
} // namespace rtb

using namespace rtb;

extern "C" int rt_scene_commit(rt_scene* s, int device, rt_build_stats* stats) {
    if (!s) { set_error("null scene"); return RT_ERR_INVALID; }
    HostScene& hs = host_of(s);
    if (!hs.flat.valid) {
        std::string e;
        const int r = hs.flatten(e);
        if (r < 0) { set_error(e); return r; }
    }
    DeviceScene*& slot = device_of(s);
    RenderPool* kept_pool = nullptr;      // the wavefront pools (hundreds of MB) survive a re-commit on the same device
    int kept_device = -1;
    if (slot) { kept_pool = slot->pool; kept_device = slot->device; slot->pool = nullptr; device_scene_destroy(slot); slot = nullptr; }
    DeviceScene* ds = new DeviceScene;
    const int rc = commit_impl(hs, *ds, device);
    if (rc < 0 || kept_device != device) {
        if (kept_pool) { if (kept_device >= 0) cudaSetDevice(kept_device); render_pool_destroy(kept_pool); kept_pool = nullptr; }
    }
    if (rc < 0) { device_scene_destroy(ds); return rc; }
    ds->pool = kept_pool;
    slot = ds;
    if (stats) *stats = ds->build;
    return RT_OK;
}

extern "C" int rt_first_hit(rt_scene* s, const float* rays, int n, float t_min, const float* xi, int32_t* prim_id, float* t) {
    if (!s || !rays || !prim_id || !t || n < 0) { set_error("first_hit: bad argument"); return RT_ERR_INVALID; }
    DeviceScene* ds = device_of(s);
    if (!ds) { set_error("first_hit: scene not committed (rt_scene_commit)"); return RT_ERR_STATE; }
    if (n == 0) return RT_OK;
    RT_CUDA(cudaSetDevice(ds->device));
    float *d_rays = nullptr, *d_xi = nullptr, *d_t = nullptr;
    int32_t* d_id = nullptr;
    RT_CUDA(cudaMalloc(&d_rays, sizeof(float) * 7 * (size_t)n));
    RT_CUDA(cudaMalloc(&d_t, sizeof(float) * (size_t)n));
    RT_CUDA(cudaMalloc(&d_id, sizeof(int32_t) * (size_t)n));
    RT_CUDA(cudaMemcpy(d_rays, rays, sizeof(float) * 7 * (size_t)n, cudaMemcpyHostToDevice));
    if (xi) {
        RT_CUDA(cudaMalloc(&d_xi, sizeof(float) * (size_t)n));
        RT_CUDA(cudaMemcpy(d_xi, xi, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    }
    k_first_hit<<<(n + 127) / 128, 128>>>(ds->view, d_rays, n, t_min, d_xi, d_id, d_t);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpy(prim_id, d_id, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost));
    RT_CUDA(cudaMemcpy(t, d_t, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(d_rays); cudaFree(d_xi); cudaFree(d_t); cudaFree(d_id);
    return RT_OK;
}
