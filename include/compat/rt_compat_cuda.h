// rt_compat_cuda.h — the CUDA-isms the reference's scene files use around its constructors (scenes.h,
// triangle_mesh.h:208-352, texture.h:166-203), emulated on the host so those files compile against rt_compat.h:
// cudaMalloc'd pointer arrays, checkCudaErrors, thrust::device_ptr, launch built-ins, make_image, create_meshes.
#pragma once
#include <functional>
#include <iostream>
#include "../rt_compat.h"
#include "../rt_obj.h"

// device allocations of the scene files are small host arrays here (hittable** lists, curandState, ...)
static inline cudaError_t rtc_host_alloc(void** p, size_t n) { *p = calloc(1, n ? n : 1); return cudaSuccess; }
static inline cudaError_t rtc_host_ok() { return cudaSuccess; }
#define cudaMalloc(p, n) rtc_host_alloc((void**)(p), (n))
#define cudaFree(p) rtc_host_ok()          /* scene objects stay alive until the rt_scene is committed */
#define cudaGetLastError() rtc_host_ok()
#define cudaDeviceSynchronize() rtc_host_ok()
#define cudaDeviceReset() rtc_host_ok()
#define checkCudaErrors(val) do { if ((val) != cudaSuccess) throw std::runtime_error("rt_compat: emulated CUDA call failed"); } while (0)

struct rtc_idx3 { unsigned x = 0, y = 0, z = 0; };
static thread_local rtc_idx3 rtc_thread_idx, rtc_block_idx;
#define threadIdx rtc_thread_idx
#define blockIdx rtc_block_idx

namespace thrust {
template <typename T> struct device_ptr {
    T* p = nullptr;
    device_ptr() {}
    explicit device_ptr(T* q) : p(q) {}
    T& operator[](size_t i) const { return p[i]; }
};
template <typename T> device_ptr<T> device_malloc(size_t n) { return device_ptr<T>((T*)calloc(n ? n : 1, sizeof(T))); }
template <typename T> void device_free(device_ptr<T>) {}
template <typename T> T* raw_pointer_cast(const device_ptr<T>& d) { return d.p; }
} // namespace thrust

template <typename T> void upload_to_device(thrust::device_ptr<T>& d_ptr, std::vector<T>& h) {   // common.h:92
    d_ptr = thrust::device_malloc<T>(h.size());
    for (size_t i = 0; i < h.size(); i++) d_ptr.p[i] = h[i];
}

inline int num_bvh_nodes(int n) { return (int)pow(2, ceil(log2(n))) - 1 + n; }   // bvh.h:18
inline int size_of_bvh(int n) { return 64 * num_bvh_nodes(n) + 64; }            // only sizes a pointer array here

namespace rtc {
// Image decoding is the host application's job (out of scope).  Binary PPM (P6) is built in; anything
// else goes through this hook: bool(path, bytes, width, height, channels).
using image_decoder_fn = std::function<bool(const char*, std::vector<unsigned char>&, int&, int&, int&)>;
inline image_decoder_fn& image_decoder() { static image_decoder_fn f; return f; }

inline bool read_ppm(const char* path, std::vector<unsigned char>& out, int& w, int& h, int& ch) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    char magic[3] = {0};
    int maxv = 0;
    const bool ok = fscanf(f, "%2s %d %d %d", magic, &w, &h, &maxv) == 4 && strcmp(magic, "P6") == 0 && maxv == 255;
    if (ok) { fgetc(f); ch = 3; out.resize((size_t)w * h * 3); if (fread(out.data(), 1, out.size(), f) != out.size()) { fclose(f); return false; } }
    fclose(f);
    return ok;
}

// Mesh loading hook: bool(path, vertices, indices, uvs, texture_paths, normals) — same outputs as the
// reference's load_model (triangle_mesh.h:129).  When no hook is installed the library's own OBJ loader is used.
using mesh_loader_fn = std::function<bool(const std::string&, std::vector<float>&, std::vector<unsigned>&, std::vector<float>&,
                                          std::vector<std::string>&, std::vector<float>&)>;
inline mesh_loader_fn& mesh_loader() { static mesh_loader_fn f; return f; }

// Default mesh loader: the library's own OBJ / MTL reader (include/rt_obj.h) in place of assimp.
inline bool load_obj_builtin(const std::string& path, std::vector<float>& vertices, std::vector<unsigned>& indices, std::vector<float>& uvs,
                             std::vector<std::string>& tex_paths, std::vector<float>& norms) {
    rt_obj* o = nullptr;
    if (rt_obj_load(path.c_str(), &o) < 0) { std::cerr << "OBJ loader error:\n\t" << rt_last_error() << std::endl; return false; }   // triangle_mesh.h:137-140
    const uint32_t nv = rt_obj_num_vertices(o), nt = rt_obj_num_triangles(o);
    vertices.assign(rt_obj_vertices(o), rt_obj_vertices(o) + 3 * (size_t)nv);
    uvs.assign(rt_obj_uvs(o), rt_obj_uvs(o) + 2 * (size_t)nv);
    norms.assign(rt_obj_normals(o), rt_obj_normals(o) + 3 * (size_t)nv);
    indices.assign(rt_obj_indices(o), rt_obj_indices(o) + 3 * (size_t)nt);
    if (rt_obj_diffuse_map(o)[0]) tex_paths.push_back(rt_obj_diffuse_map(o));
    rt_obj_free(o);
    return true;
}
} // namespace rtc

// make_image(paths, data, widths, heights, channels) — texture.h:188-203
inline void make_image(std::vector<const char*> impaths, thrust::device_ptr<unsigned char>& imdata, thrust::device_ptr<int>& imwidths,
                       thrust::device_ptr<int>& imhs, thrust::device_ptr<int>& imch) {
    std::vector<int> ws, hs, chs;
    std::vector<unsigned char> all;
    for (const char* path : impaths) {
        std::vector<unsigned char> px;
        int w = 0, h = 0, ch = 0;
        const bool ok = (rtc::image_decoder() && rtc::image_decoder()(path, px, w, h, ch)) || rtc::read_ppm(path, px, w, h, ch);
        if (!ok) { std::cerr << "Failed to load texture " << path << std::endl; return; }   // texture.h:175
        ws.push_back(w); hs.push_back(h); chs.push_back(ch);
        all.insert(all.end(), px.begin(), px.end());
    }
    upload_to_device(imdata, all); upload_to_device(imwidths, ws); upload_to_device(imhs, hs); upload_to_device(imch, chs);
}

// create_meshes_d — triangle_mesh.h:147-204 (same indexing, including its offset arithmetic for object 0)
inline void create_meshes_d(hittable** hits, unsigned* num_tris, unsigned num_objs, unsigned char* imdata, int* widths, int* heights,
                            int* bytes_per_pixels, float* vertices, unsigned* indices, float* uvs, float* norms) {
    unsigned hit_counter = 0, ind_counter = 0;
    for (unsigned i = 0; i < num_objs; i++) {
        if (i > 0) throw std::runtime_error("rt_compat: create_meshes_d with more than one object is undefined in the reference (SURVEY Q11)");
        material* mat = new lambertian(new image_texture(imdata, widths, heights, bytes_per_pixels, (int)i));
        for (unsigned j = 0; j < num_tris[i]; j++) {
            vec3 v[3], n[3]; float uv[6];
            for (int k = 0; k < 3; k++) {
                const unsigned ix = indices[ind_counter + k];
                v[k] = vec3(vertices[3 * ix], vertices[3 * ix + 1], vertices[3 * ix + 2]);
                n[k] = vec3(norms[3 * ix], norms[3 * ix + 1], norms[3 * ix + 2]);
                uv[2 * k] = uvs[2 * ix]; uv[2 * k + 1] = uvs[2 * ix + 1];
            }
            hits[hit_counter++] = new triangle(v[0], v[1], v[2], n[0], n[1], n[2], uv[0], uv[1], uv[2], uv[3], uv[4], uv[5], mat);
            ind_counter += 3;
        }
    }
}

// create_meshes(objs, hits, num_data, size) — triangle_mesh.h:208-352
inline void create_meshes(std::vector<std::string> objs, hittable**& hits, thrust::device_ptr<unsigned>& num_data, int& size) {
    if (objs.size() != 1) throw std::runtime_error("rt_compat: create_meshes supports one .obj per call (SURVEY Q11)");
    std::vector<float> vertices, uvs, norms; std::vector<unsigned> indices; std::vector<std::string> tex_paths;
    const bool loaded = rtc::mesh_loader() ? rtc::mesh_loader()(objs[0], vertices, indices, uvs, tex_paths, norms)
                                           : rtc::load_obj_builtin(objs[0], vertices, indices, uvs, tex_paths, norms);
    if (!loaded) throw std::runtime_error("rt_compat: cannot load " + objs[0]);
    if (tex_paths.size() != 1) { std::cerr << "Meshes need exactly 1 texture" << std::endl; return; }   // triangle_mesh.h:275-281
    const std::string dir = objs[0].substr(0, objs[0].find_last_of('/') + 1);
    const std::string tex = dir + tex_paths[0];
    thrust::device_ptr<unsigned char> imdata; thrust::device_ptr<int> iw, ih, ic;
    make_image({tex.c_str()}, imdata, iw, ih, ic);
    std::vector<unsigned> num_t{(unsigned)(indices.size() / 3)};
    upload_to_device(num_data, num_t);
    hits = (hittable**)calloc(num_t[0] ? num_t[0] : 1, sizeof(hittable*));
    create_meshes_d(hits, num_data.p, 1, imdata.p, iw.p, ih.p, ic.p, vertices.data(), indices.data(), uvs.data(), norms.data());
    size = size_of_bvh((int)num_t[0]);
}
