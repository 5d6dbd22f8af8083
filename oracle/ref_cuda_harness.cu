/* oracle/ref_cuda_harness.cu — TEST / BASELINE INFRASTRUCTURE, not product code.
 *
 * "The reference's own CUDA build on one B200" (BASELINE.json north_star; SURVEY §8c route B): the UNMODIFIED
 * reference headers are compiled by nvcc straight from /root/reference (oracle/Makefile, target _ref/ref_cuda;
 * assimp / png++ are satisfied by the stub headers in oracle/shim/), and this file supplies what main.cu
 * hard-codes: scene choice, image size, sample count, CUDA-event timing around the reference's own
 * render_init<<<>>> / render<<<>>> launches (render.h:84, :94) with the reference's launch shape (8x8 blocks,
 * render.h:129-130), and the device heap / stack limits the reference never sets.
 * The Next-Week final scene does not exist in the reference: it is composed here from the
 * reference's device constructors inside a <<<1,1>>> builder kernel, exactly like the reference's own
 * create_*_world kernels and in the same order as oracle/ref_harness.cpp::build_final.
 *
 * usage: ref_cuda <big1|cornell_smoke|cornell|final> <width> <spp> <max_depth> [earth.rgb8 w h] [image_out.f32]
 *        ref_cuda <scene> probe <rays.f32 (n x 7)> <out (n float t, n x 3 float normals, n int32 leaf ids)> [earth.rgb8 w h]
 * prints one JSON line.  Only bench.py's reference legs and tests/ may execute it.
 * Everything in this file is harness code written for this repo; no reference code is copied.
 */
#include "scenes.h"
#include "render.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

__global__ void create_final_world(hittable** top, hittable** d_world, camera** d_camera, curandState* rs, hittable** boxes1,
                                   hittable** boxes2, unsigned char* img, int* ws, int* hs, int* bs, float aspect, int* n_top) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int k = 0;
    material* ground = new lambertian(color(0.48f, 0.83f, 0.53f));
    const int boxes_per_side = 20;
    int nb = 0;
    for (int i = 0; i < boxes_per_side; i++)
        for (int j = 0; j < boxes_per_side; j++) {
            const float bw = 100.0f;
            const float x0 = -1000.0f + i * bw, z0 = -1000.0f + j * bw, y0 = 0.0f;
            const float x1 = x0 + bw, z1 = z0 + bw;
            const float y1 = random_float(rs, 1, 101);
            boxes1[nb++] = new box(point3(x0, y0, z0), point3(x1, y1, z1), ground);
        }
    const int ns = 1000;
    material* white = new lambertian(color(0.73f, 0.73f, 0.73f));
    for (int j = 0; j < ns; j++) {
        const float x = random_float(rs, 0, 165);
        const float y = random_float(rs, 0, 165);
        const float z = random_float(rs, 0, 165);
        boxes2[j] = new sphere(point3(x, y, z), 10, white);
    }
    texturez* pertext = new noise_texture(rs, 0.1f);
    top[k++] = new bvh_node(boxes1, nb, 0, 1, rs);
    top[k++] = new xz_rect(123, 423, 147, 412, 554, new diffuse_light(color(7, 7, 7)));
    const point3 center1(400, 400, 200);
    const point3 center2 = center1 + vec3(30, 0, 0);
    top[k++] = new moving_sphere(center1, center2, 0, 1, 50, new lambertian(color(0.7f, 0.3f, 0.1f)));
    top[k++] = new sphere(point3(260, 150, 45), 50, new dielectric(1.5f));
    top[k++] = new sphere(point3(0, 150, 145), 50, new metal(color(0.8f, 0.8f, 0.9f), 1.0f));
    hittable* boundary = new sphere(point3(360, 150, 145), 70, new dielectric(1.5f));
    top[k++] = boundary;
    top[k++] = new constant_medium(boundary, 0.2f, color(0.2f, 0.4f, 0.9f));
    hittable* fog = new sphere(point3(0, 0, 0), 5000, new dielectric(1.5f));
    top[k++] = new constant_medium(fog, 0.0001f, color(1, 1, 1));
    top[k++] = new sphere(point3(400, 200, 400), 100, new lambertian(new image_texture(img, ws, hs, bs, 0)));
    top[k++] = new sphere(point3(220, 280, 300), 80, new lambertian(pertext));
    hittable* cluster = new bvh_node(boxes2, ns, 0, 1, rs);
    top[k++] = new translate(new rotate_y(cluster, 15), vec3(-100, 270, 395));
    *d_world = new hittable_list(top, k);
    *d_camera = new camera(vec3(478, 278, -600), vec3(278, 278, 0), vec3(0, 1, 0), 40, aspect, 0.0f, 10.0f, 0, 1);
    *n_top = k;
}

struct final_scene : public scene {
    final_scene(float aspect, const std::vector<unsigned char>& img, int w, int h) : scene(aspect, 16, background_color::black) {
        curandState* rand_state;
        checkCudaErrors(cudaMalloc((void**)&rand_state, sizeof(curandState)));
        world_init<<<1, 1>>>(rand_state);
        checkCudaErrors(cudaGetLastError());
        hittable **boxes1, **boxes2;
        checkCudaErrors(cudaMalloc((void**)&boxes1, 400 * sizeof(hittable*)));
        checkCudaErrors(cudaMalloc((void**)&boxes2, 1000 * sizeof(hittable*)));
        unsigned char* d_img;
        int *d_w, *d_h, *d_b, *d_n;
        const int bpp = 3;
        checkCudaErrors(cudaMalloc((void**)&d_img, img.size()));
        checkCudaErrors(cudaMemcpy(d_img, img.data(), img.size(), cudaMemcpyHostToDevice));
        checkCudaErrors(cudaMalloc((void**)&d_w, sizeof(int))); checkCudaErrors(cudaMalloc((void**)&d_h, sizeof(int)));
        checkCudaErrors(cudaMalloc((void**)&d_b, sizeof(int))); checkCudaErrors(cudaMalloc((void**)&d_n, sizeof(int)));
        checkCudaErrors(cudaMemcpy(d_w, &w, sizeof(int), cudaMemcpyHostToDevice));
        checkCudaErrors(cudaMemcpy(d_h, &h, sizeof(int), cudaMemcpyHostToDevice));
        checkCudaErrors(cudaMemcpy(d_b, &bpp, sizeof(int), cudaMemcpyHostToDevice));
        create_final_world<<<1, 1>>>(d_list, d_world, d_camera, rand_state, boxes1, boxes2, d_img, d_w, d_h, d_b, aspect, d_n);
        checkCudaErrors(cudaGetLastError());
        checkCudaErrors(cudaDeviceSynchronize());
        checkCudaErrors(cudaMemcpy(&no_hittables, d_n, sizeof(int), cudaMemcpyDeviceToHost));
    }
};

/* ---- leaf ids on the device (same DFS order as the product's flattener and as oracle/ref_harness.cpp's tagger) ------------
 * Device code has no RTTI; the dynamic type of an object is read from its vtable pointer, compared with the vtable pointer
 * of one exemplar per class constructed in the same kernel.  Every leaf's material is wrapped in a tag_material that
 * carries the leaf id and forwards scatter / emitted, so rendering is unchanged; the probe reads the id back from
 * rec.mat_ptr.  Boundary leaves of a constant_medium consume ids but never report (the medium's phase function does). */
struct tag_material : public material {
    material* real;
    int leaf_id;
    __device__ tag_material(material* r, int id) : real(r), leaf_id(id) {}
    __device__ bool scatter(const ray& ray_in, const hit_record& rec, color& attenuation, ray& scattered, curandState* s) const override {
        return real->scatter(ray_in, rec, attenuation, scattered, s);
    }
    __device__ color emitted(const float u, const float v, const point3& p) const override { return real->emitted(u, v, p); }
};
struct vtables { void *sphere_, *moving_sphere_, *xy_, *xz_, *yz_, *box_, *translate_, *rotate_y_, *medium_, *list_, *bvh_, *tag_; };
__device__ inline void* vptr_of(const void* obj) { return *(void* const*)obj; }

__device__ material* tag_wrap(material* m, bool visible, int& next, const vtables& vt) {
    const int id = next++;
    if (!visible || vptr_of(m) == vt.tag_) return m;
    return new tag_material(m, id);
}
__device__ void tag_walk(hittable* h, bool visible, int& next, const vtables& vt, int* unknown) {
    void* v = vptr_of(h);
    if (v == vt.sphere_) { sphere* p = (sphere*)h; p->mat_ptr = tag_wrap(p->mat_ptr, visible, next, vt); return; }
    if (v == vt.moving_sphere_) { moving_sphere* p = (moving_sphere*)h; p->mat_ptr = tag_wrap(p->mat_ptr, visible, next, vt); return; }
    if (v == vt.xy_) { xy_rect* p = (xy_rect*)h; p->mp = tag_wrap(p->mp, visible, next, vt); return; }
    if (v == vt.xz_) { xz_rect* p = (xz_rect*)h; p->mp = tag_wrap(p->mp, visible, next, vt); return; }
    if (v == vt.yz_) { yz_rect* p = (yz_rect*)h; p->mp = tag_wrap(p->mp, visible, next, vt); return; }
    if (v == vt.box_) { box* p = (box*)h; for (int i = 0; i < 6; i++) tag_walk(p->temp_hittable[i], visible, next, vt, unknown); return; }
    if (v == vt.translate_) { tag_walk(((translate*)h)->ptr, visible, next, vt, unknown); return; }
    if (v == vt.rotate_y_) { tag_walk(((rotate_y*)h)->ptr, visible, next, vt, unknown); return; }
    if (v == vt.medium_) {
        constant_medium* p = (constant_medium*)h;
        p->phase_function = tag_wrap(p->phase_function, visible, next, vt);
        tag_walk(p->boundary, false, next, vt, unknown);
        return;
    }
    if (v == vt.list_) { hittable_list* p = (hittable_list*)h; for (int i = 0; i < p->list_size; i++) tag_walk(p->objects[i], visible, next, vt, unknown); return; }
    if (v == vt.bvh_) { bvh_node* p = (bvh_node*)h; for (int i = 0; i < p->n; i++) tag_walk(p->objs[i], visible, next, vt, unknown); return; }
    atomicAdd(unknown, 1);
}
__global__ void tag_world(hittable** world, vtables* out_vt, int* out_counts /* leaves, unknown */) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    curandState rs;
    curand_init(1, 0, 0, &rs);
    material* m = new lambertian(color(0, 0, 0));
    hittable** two = new hittable*[2];
    two[0] = new sphere(point3(0, 0, 0), 1, m); two[1] = new sphere(point3(3, 0, 0), 1, m);
    vtables vt;
    vt.sphere_ = vptr_of(two[0]);
    vt.moving_sphere_ = vptr_of(new moving_sphere(point3(0, 0, 0), point3(1, 0, 0), 0, 1, 1, m));
    vt.xy_ = vptr_of(new xy_rect(0, 1, 0, 1, 0, m)); vt.xz_ = vptr_of(new xz_rect(0, 1, 0, 1, 0, m)); vt.yz_ = vptr_of(new yz_rect(0, 1, 0, 1, 0, m));
    vt.box_ = vptr_of(new box(point3(0, 0, 0), point3(1, 1, 1), m));
    vt.translate_ = vptr_of(new translate(two[0], vec3(0, 0, 0)));
    vt.rotate_y_ = vptr_of(new rotate_y(two[0], 0));
    vt.medium_ = vptr_of(new constant_medium(two[0], 1.0f, color(1, 1, 1)));
    vt.list_ = vptr_of(new hittable_list(two, 2));
    vt.bvh_ = vptr_of(new bvh_node());   // (the default constructor: only the vtable pointer is read)
    vt.tag_ = vptr_of(new tag_material(m, -1));
    int next = 0, unknown = 0;
    tag_walk(*world, true, next, vt, &unknown);
    *out_vt = vt;
    out_counts[0] = next; out_counts[1] = unknown;
}

/* Route-B probe: the reference's own device intersection code asked for the closest hit of caller-supplied
 * rays, exactly as color_f does at render.h:63.  One thread per ray; t = +inf on a miss.  constant_medium draws from the
 * per-thread cuRAND state, so media hits are stochastic: compare those scenes on surface hits only. */
__global__ void probe_first_hit(hittable** world, const float* rays, int n, float* out_t, float* out_normal, int* out_id, const vtables* vt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    curandState rs;
    curand_init(1984, i, 0, &rs);
    const float* q = rays + 7 * (size_t)i;
    ray r(point3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), q[6]);
    hit_record rec;
    const bool hit = (*world)->hit(r, 0.001f, infinity, rec, &rs);
    out_t[i] = hit ? rec.t : infinity;
    out_id[i] = (hit && vt && vptr_of(rec.mat_ptr) == vt->tag_) ? ((const tag_material*)rec.mat_ptr)->leaf_id : -1;
    out_normal[3 * i] = hit ? rec.normal.x() : 0.f; out_normal[3 * i + 1] = hit ? rec.normal.y() : 0.f; out_normal[3 * i + 2] = hit ? rec.normal.z() : 0.f;
}

static int run_probe(scene* sc, const char* rays_path, const char* out_path) {
    FILE* f = fopen(rays_path, "rb");
    if (!f) { fprintf(stderr, "cannot read %s\n", rays_path); return 2; }
    fseek(f, 0, SEEK_END); const long bytes = ftell(f); fseek(f, 0, SEEK_SET);
    const int n = (int)(bytes / (7 * sizeof(float)));
    std::vector<float> rays((size_t)n * 7), t(n), nrm((size_t)n * 3);
    std::vector<int> ids(n);
    if (fread(rays.data(), sizeof(float), rays.size(), f) != rays.size()) { fclose(f); return 2; }
    fclose(f);
    float *d_rays, *d_t, *d_n;
    checkCudaErrors(cudaMalloc((void**)&d_rays, rays.size() * sizeof(float)));
    checkCudaErrors(cudaMalloc((void**)&d_t, n * sizeof(float)));
    checkCudaErrors(cudaMalloc((void**)&d_n, 3 * (size_t)n * sizeof(float)));
    checkCudaErrors(cudaMemcpy(d_rays, rays.data(), rays.size() * sizeof(float), cudaMemcpyHostToDevice));
    vtables* d_vt; int *d_counts, *d_id, counts[2] = {0, 0};
    checkCudaErrors(cudaMalloc((void**)&d_vt, sizeof(vtables)));
    checkCudaErrors(cudaMalloc((void**)&d_counts, 2 * sizeof(int)));
    checkCudaErrors(cudaMalloc((void**)&d_id, n * sizeof(int)));
    tag_world<<<1, 1>>>(sc->d_world, d_vt, d_counts);
    checkCudaErrors(cudaGetLastError());
    checkCudaErrors(cudaDeviceSynchronize());
    checkCudaErrors(cudaMemcpy(counts, d_counts, sizeof(counts), cudaMemcpyDeviceToHost));
    probe_first_hit<<<(n + 63) / 64, 64>>>(sc->d_world, d_rays, n, d_t, d_n, d_id, d_vt);
    checkCudaErrors(cudaGetLastError());
    checkCudaErrors(cudaDeviceSynchronize());
    checkCudaErrors(cudaMemcpy(t.data(), d_t, n * sizeof(float), cudaMemcpyDeviceToHost));
    checkCudaErrors(cudaMemcpy(nrm.data(), d_n, 3 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
    f = fopen(out_path, "wb");
    if (!f) return 2;
    fwrite(t.data(), sizeof(float), t.size(), f);
    fwrite(nrm.data(), sizeof(float), nrm.size(), f);
    checkCudaErrors(cudaMemcpy(ids.data(), d_id, n * sizeof(int), cudaMemcpyDeviceToHost));
    fwrite(ids.data(), sizeof(int), ids.size(), f);
    fclose(f);
    printf("{\"impl\": \"reference_cuda\", \"mode\": \"probe\", \"rays\": %d, \"leaves\": %d, \"unknown_types\": %d}\n", n, counts[0], counts[1]);
    return 0;
}

static double wall_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

int main(int argc, char** argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s <big1|cornell_smoke|cornell|final> <width> <spp> <max_depth> [earth.rgb8 w h]\n", argv[0]); return 2; }
    const std::string name = argv[1];
    const bool probe = std::string(argv[2]) == "probe";
    const int width = probe ? 16 : atoi(argv[2]), spp = probe ? 1 : atoi(argv[3]), max_depth = probe ? 1 : atoi(argv[4]);
    // the reference never raises these: bvh_node::hit allocates per ray, merge() leaks during the build
    checkCudaErrors(cudaDeviceSetLimit(cudaLimitMallocHeapSize, (size_t)4 << 30));
    checkCudaErrors(cudaDeviceSetLimit(cudaLimitStackSize, 8192));

    const double t_build0 = wall_ms();
    scene* sc = nullptr;
    if (name == "big1") sc = new big_scene1();
    else if (name == "cornell_smoke") sc = new cornell_smoke_box_scene();
    else if (name == "cornell") sc = new cornell_box_scene();
    else if (name == "final") {
        if (argc < 8) { fprintf(stderr, "final needs earth.rgb8 w h\n"); return 2; }
        const int w = atoi(argv[6]), h = atoi(argv[7]);
        std::vector<unsigned char> img((size_t)w * h * 3);
        FILE* f = fopen(argv[5], "rb");
        if (!f || fread(img.data(), 1, img.size(), f) != img.size()) { fprintf(stderr, "cannot read %s\n", argv[5]); return 2; }
        fclose(f);
        sc = new final_scene(16.0f / 9.0f, img, w, h);
    } else { fprintf(stderr, "unknown scene %s\n", name.c_str()); return 2; }
    checkCudaErrors(cudaDeviceSynchronize());
    const double ms_build = wall_ms() - t_build0;
    if (probe) return run_probe(sc, argv[3], argv[4]);

    render_settings settings;                      // render.h:21-50
    settings.aspect_ratio = sc->aspect;
    settings.image_width = width;
    settings.samples_per_pixel_per_fb = spp;
    settings.no_fb = 1;
    settings.max_depth = max_depth;
    settings.calc_all();

    vec3* fb;
    checkCudaErrors(cudaMalloc((void**)&fb, settings.num_pixels * sizeof(vec3)));
    dim3 blocks(settings.image_width / settings.threads_x + 1, settings.image_height / settings.threads_y + 1);   // render.h:129
    dim3 threads(settings.threads_x, settings.threads_y);
    curandState* d_rand_state;
    checkCudaErrors(cudaMalloc((void**)&d_rand_state, settings.num_pixels * sizeof(curandState)));
    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    cudaEventRecord(e0);
    render_init<<<blocks, threads>>>(settings.image_width, settings.image_height, d_rand_state);
    cudaEventRecord(e1);
    render<<<blocks, threads>>>(fb, settings.image_width, settings.image_height, settings.samples_per_pixel_per_fb, sc->d_camera,
                                d_rand_state, sc->d_world, settings.max_depth, 0, sc->background);
    cudaEventRecord(e2);
    checkCudaErrors(cudaGetLastError());
    checkCudaErrors(cudaDeviceSynchronize());
    float ms_init = 0, ms_render = 0;
    cudaEventElapsedTime(&ms_init, e0, e1);
    cudaEventElapsedTime(&ms_render, e1, e2);

    // image mean (sanity: comparable with the product's image mean)
    std::vector<vec3> host(settings.num_pixels);
    checkCudaErrors(cudaMemcpy(host.data(), fb, settings.num_pixels * sizeof(vec3), cudaMemcpyDeviceToHost));
    double mean[3] = {0, 0, 0};
    size_t bad = 0;
    for (const vec3& c : host) {
        if (!(c.x() == c.x()) || !(c.y() == c.y()) || !(c.z() == c.z()) || fabsf(c.x()) > 1e30f || fabsf(c.y()) > 1e30f || fabsf(c.z()) > 1e30f) { bad++; continue; }   // NaN / inf pixels are counted, not averaged
        mean[0] += c.x(); mean[1] += c.y(); mean[2] += c.z();
    }
    for (double& m : mean) m /= (double)settings.num_pixels;
    if (argc >= 9 || (name != "final" && argc >= 6)) {   // optional raw float dump
        const char* out = name == "final" ? argv[8] : argv[5];
        FILE* f = fopen(out, "wb");
        if (f) { fwrite(host.data(), sizeof(vec3), host.size(), f); fclose(f); }
    }
    const double samples = (double)settings.num_pixels * spp;
    for (double& m : mean) if (!(m == m) || m > 1e300 || m < -1e300) m = -1.0;      // keep the line valid JSON (mean over the non-NaN pixels; -1 = not finite)
    printf("{\"impl\": \"reference_cuda\", \"scene\": \"%s\", \"width\": %u, \"height\": %u, \"spp\": %d, \"max_depth\": %d, "
           "\"ms_build\": %.3f, \"ms_render_init\": %.3f, \"ms_render\": %.3f, \"msamples_per_s\": %.6f, "
           "\"image_mean\": [%.6f, %.6f, %.6f], \"nan_pixels\": %zu}\n",
           name.c_str(), settings.image_width, settings.image_height, spp, max_depth, ms_build, ms_init, ms_render,
           samples / (ms_render * 1e-3) / 1e6, mean[0], mean[1], mean[2], bad);
    return 0;
}
