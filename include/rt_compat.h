// rt_compat.h — header-only C++ host mirror of the reference's scene-description API.
//
// The reference constructs its world with device-side `new` inside <<<1,1>>> kernels (scenes.h:82-621).  Here the
// same class names and constructor signatures exist as HOST classes that record the scene through the C-ABI
// (include/rt_b200.h).  A statement list such as
//     d_list[0] = new sphere(point3(0,0,-1), 0.5, new lambertian(color(0,1,0)));
//     *d_world  = new hittable_list(d_list, 1);
//     *d_camera = new camera(vec3(0,0,-3), vec3(0,0,0), vec3(0,1,0), 40, 16.0f/9.0f, 0.0f, 10.0f, 0, 1);
// compiles unchanged; the objects hold handles, not geometry.  Random-driven scenes reproduce because
// curandState/curand_init/curand_uniform come from the CUDA toolkit's own host-compilable curand_kernel.h.
//
// Usage: rtc::scope sc(my_rt_scene);  ...constructor calls...  rtc::finish(world_ptr, camera_ptr, background);
#ifndef RT_COMPAT_H
#define RT_COMPAT_H

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include <curand_kernel.h>   // host-compilable XORWOW: the scene-building stream of scenes.h:28-32

#include "rt_b200.h"

#ifndef __device__
#define __device__
#endif
#ifndef __host__
#define __host__
#endif
#ifndef __global__
#define __global__
#endif

namespace rtc {

inline rt_scene*& current() { static thread_local rt_scene* s = nullptr; return s; }

struct scope {   // binds constructor calls on this thread to one rt_scene
    rt_scene* prev;
    explicit scope(rt_scene* s) : prev(current()) { current() = s; }
    ~scope() { current() = prev; }
};

inline int ok(int rc) {
    if (rc < 0) throw std::runtime_error(std::string("rt_b200: ") + rt_last_error());
    return rc;
}
inline rt_scene* sc() {
    if (!current()) throw std::runtime_error("rt_compat: no rtc::scope active");
    return current();
}

} // namespace rtc

// ---- vec3.h / ray.h / common.h --------------------------------------------------------------------
inline float random_float(curandState* s) { return curand_uniform(s); }                                    // vec3.h:7
inline float random_float(curandState* s, const float min, const float max) { return min + (max - min) * random_float(s); } // vec3.h:11

struct vec3 {
    float e[3];
    vec3() : e{0, 0, 0} {}
    vec3(float e0, float e1, float e2) : e{e0, e1, e2} {}
    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
    vec3 operator-() const { return vec3(-e[0], -e[1], -e[2]); }
    float operator[](int i) const { return e[i]; }
    float& operator[](int i) { return e[i]; }
    vec3& operator+=(const vec3& v) { e[0] += v.e[0]; e[1] += v.e[1]; e[2] += v.e[2]; return *this; }
    vec3& operator*=(const float t) { e[0] *= t; e[1] *= t; e[2] *= t; return *this; }
    vec3& operator*=(const vec3& v) { e[0] *= v.e[0]; e[1] *= v.e[1]; e[2] *= v.e[2]; return *this; }
    vec3& operator/=(const float t) { return *this *= 1 / t; }
    float length_squared() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    float length() const { return sqrtf(length_squared()); }
    bool near_zero() const { const float s = 1e-6f; return (fabsf(e[0]) < s) && (fabsf(e[1]) < s) && (fabsf(e[2]) < s); }
    // vec3.h:62-69 passes three draws as constructor arguments, whose evaluation order C++ leaves to the compiler
    // (SURVEY §3.4): nvcc's device code — what the reference actually runs — evaluates left to right (x first), g++
    // right to left (z first).  rtc::device_argument_order() selects nvcc's order explicitly; otherwise the expression
    // is the reference's own and the order is the host compiler's.
    static inline vec3 random(curandState* s);
    static inline vec3 random(curandState* s, const float min, const float max);
};
using point3 = vec3;
using color = vec3;

namespace rtc {
// true: draws that the reference passes as function arguments are taken in nvcc's (left-to-right) order, so
// RNG-built scenes equal what the reference's CUDA build constructs on the device (checked against that build on the
// B200: tests/test_gpu.py::test_reference_cuda_build_agrees_on_first_hits).  false: the host compiler's order.
inline bool& device_argument_order() { static thread_local bool f = false; return f; }
}
inline vec3 vec3::random(curandState* s) {
    if (rtc::device_argument_order()) { const float x = random_float(s); const float y = random_float(s); const float z = random_float(s); return vec3(x, y, z); }
    return vec3(random_float(s), random_float(s), random_float(s));
}
inline vec3 vec3::random(curandState* s, const float min, const float max) {
    if (rtc::device_argument_order()) { const float x = random_float(s, min, max); const float y = random_float(s, min, max); const float z = random_float(s, min, max); return vec3(x, y, z); }
    return vec3(random_float(s, min, max), random_float(s, min, max), random_float(s, min, max));
}

inline vec3 operator+(const vec3& u, const vec3& v) { return vec3(u.e[0] + v.e[0], u.e[1] + v.e[1], u.e[2] + v.e[2]); }
inline vec3 operator-(const vec3& u, const vec3& v) { return vec3(u.e[0] - v.e[0], u.e[1] - v.e[1], u.e[2] - v.e[2]); }
inline vec3 operator*(const vec3& u, const vec3& v) { return vec3(u.e[0] * v.e[0], u.e[1] * v.e[1], u.e[2] * v.e[2]); }
inline vec3 operator*(const float t, const vec3& v) { return vec3(t * v.e[0], t * v.e[1], t * v.e[2]); }
inline vec3 operator*(const vec3& v, const float t) { return t * v; }
inline vec3 operator/(const vec3& v, const float t) { return (1 / t) * v; }
inline float dot(const vec3& u, const vec3& v) { return u.e[0] * v.e[0] + u.e[1] * v.e[1] + u.e[2] * v.e[2]; }
inline vec3 cross(const vec3& u, const vec3& v) {
    return vec3(u.e[1] * v.e[2] - u.e[2] * v.e[1], u.e[2] * v.e[0] - u.e[0] * v.e[2], u.e[0] * v.e[1] - u.e[1] * v.e[0]);
}
inline vec3 unit_vector(const vec3 v) { return v / v.length(); }
inline vec3 random_in_unit_sphere(curandState* s) { for (;;) { const vec3 p = vec3::random(s, -1, 1); if (p.length_squared() < 1) return p; } }
inline vec3 random_in_unit_disk(curandState* s) {
    for (;;) { const vec3 p = vec3(random_float(s, -1, 1), random_float(s, -1, 1), 0); if (p.length_squared() < 1) return p; }
}
inline vec3 random_unit_vector(curandState* s) { return unit_vector(random_in_unit_sphere(s)); }
inline vec3 reflect(const vec3& v, const vec3& n) { return v - 2 * dot(v, n) * n; }

struct ray {
    point3 orig; vec3 dir; float tm;
    ray() {}
    ray(point3 origin, vec3 direction, const float time = 0.0f) : orig(origin), dir(direction), tm(time) {}
    point3 origin() const { return orig; }
    vec3 direction() const { return dir; }
    float time() const { return tm; }
    point3 at(const float t) const { return orig + t * dir; }
};

constexpr float infinity = std::numeric_limits<float>::infinity();
constexpr float pi = 3.1415926535897932385f;
inline float degrees_to_radians(float degrees) { return degrees * pi / 180.0f; }
inline float random_int(curandState* s, const int min, const int max) { return static_cast<int>(random_float(s, min, max + 1)); } // common.h:49
inline float clamp(const float x, const float min, const float max) { return x < min ? min : (x > max ? max : x); }
inline float clamp_d(const float x, const float min, const float max) { return x < min ? min : (x > max ? max : x); }

// ---- texture.h / perlin.h -------------------------------------------------------------------------
struct texturez { int h = -1; virtual ~texturez() {} };

struct solid_color : public texturez {
    solid_color() {}
    solid_color(const color c) { h = rtc::ok(rt_tex_solid(rtc::sc(), c.x(), c.y(), c.z())); }
    solid_color(const float r, const float g, const float b) : solid_color(color(r, g, b)) {}
};

struct checker_texture : public texturez {
    texturez *odd = nullptr, *even = nullptr;
    checker_texture() {}
    checker_texture(texturez* _even, texturez* _odd) : odd(_odd), even(_even) { h = rtc::ok(rt_tex_checker(rtc::sc(), even->h, odd->h)); }
    checker_texture(color c1, color c2) {
        even = new solid_color(c1);
        odd = new solid_color(c2);
        h = rtc::ok(rt_tex_checker(rtc::sc(), even->h, odd->h));
    }
};

// perlin(curandState*) — perlin.h:63-76, :18-35: 256 unit gradients, then three Fisher-Yates permutations
struct perlin {
    int tables = -1;
    perlin() {}
    explicit perlin(curandState* s) {
        float ranvec[768];
        int32_t perm[768];
        for (int i = 0; i < 256; i++) {
            const vec3 v = unit_vector(vec3::random(s, -1, 1));
            ranvec[3 * i] = v.x(); ranvec[3 * i + 1] = v.y(); ranvec[3 * i + 2] = v.z();
        }
        for (int axis = 0; axis < 3; axis++) {
            int32_t* p = perm + 256 * axis;
            for (int i = 0; i < 256; i++) p[i] = i;
            for (int i = 255; i > 0; i--) {
                const int target = random_int(s, 0, i);
                const int tmp = p[i]; p[i] = p[target]; p[target] = tmp;
            }
        }
        tables = rtc::ok(rt_perlin_tables(rtc::sc(), ranvec, perm));
    }
};

struct noise_texture : public texturez {
    perlin noise; float scale;
    noise_texture(curandState* s, const float sc = 1.0f) : scale(sc) { noise = perlin(s); h = rtc::ok(rt_tex_noise(rtc::sc(), noise.tables, scale)); }
};
struct turbulent_texture : public texturez {
    perlin noise; float scale; int depth;
    turbulent_texture(curandState* s, const float sc = 1.0f, const int dpt = 7) : scale(sc), depth(dpt) {
        noise = perlin(s); h = rtc::ok(rt_tex_turbulent(rtc::sc(), noise.tables, scale, depth));
    }
};
struct marble_texture : public texturez {
    perlin noise; float scale;
    marble_texture(curandState* s, const float sc = 1.0f) : scale(sc) { noise = perlin(s); h = rtc::ok(rt_tex_marble(rtc::sc(), noise.tables, scale)); }
};

// image_texture(data, widths, heights, bytes_per_pixel, index) — texture.h:125-141 (the library copies the bytes)
struct image_texture : public texturez {
    image_texture(unsigned char* d, int* ws, int* hs, int* bpps, int ind) {
        size_t start = 0;
        for (int i = 0; i < ind; i++) start += (size_t)ws[i] * hs[i] * bpps[i];
        const int img = rtc::ok(rt_image(rtc::sc(), d + start, ws[ind], hs[ind], bpps[ind]));
        h = rtc::ok(rt_tex_image(rtc::sc(), img));
    }
};

// ---- material.h -----------------------------------------------------------------------------------
struct material { int h = -1; virtual ~material() {} };

struct lambertian : public material {
    texturez* albedo;
    lambertian(const color& a) { albedo = new solid_color(a); h = rtc::ok(rt_mat_lambertian(rtc::sc(), albedo->h)); }
    lambertian(texturez* a) : albedo(a) { h = rtc::ok(rt_mat_lambertian(rtc::sc(), albedo->h)); }
};
struct metal : public material {
    texturez* albedo; float fuzz;
    metal(const color& a, const float f = 0) : fuzz(f) { albedo = new solid_color(a); h = rtc::ok(rt_mat_metal(rtc::sc(), albedo->h, fuzz)); }
    metal(texturez* a) : albedo(a), fuzz(0) { h = rtc::ok(rt_mat_metal(rtc::sc(), albedo->h, fuzz)); }   // fuzz is uninitialised in the reference
};
struct dielectric : public material {
    float ir;
    dielectric(const float index_of_refraction) : ir(index_of_refraction) { h = rtc::ok(rt_mat_dielectric(rtc::sc(), ir)); }
};
struct diffuse_light : public material {
    texturez* emit;
    diffuse_light(texturez* a) : emit(a) { h = rtc::ok(rt_mat_diffuse_light(rtc::sc(), emit->h)); }
    diffuse_light(const color c) { emit = new solid_color(c); h = rtc::ok(rt_mat_diffuse_light(rtc::sc(), emit->h)); }
};
struct isotropic : public material {
    texturez* albedo;
    isotropic(const color c) { albedo = new solid_color(c); h = rtc::ok(rt_mat_isotropic(rtc::sc(), albedo->h)); }
    isotropic(texturez* a) : albedo(a) { h = rtc::ok(rt_mat_isotropic(rtc::sc(), albedo->h)); }
};

// ---- hittables ------------------------------------------------------------------------------------
struct hittable { int h = -1; virtual ~hittable() {} };

struct sphere : public hittable {
    sphere() {}
    sphere(const point3 cen, const float r, material* m) { h = rtc::ok(rt_sphere(rtc::sc(), cen.x(), cen.y(), cen.z(), r, m->h)); }
};
struct moving_sphere : public hittable {
    moving_sphere(const point3 cen0, const point3 cen1, const float t0, const float t1, const float r, material* m) {
        h = rtc::ok(rt_moving_sphere(rtc::sc(), cen0.e, cen1.e, t0, t1, r, m->h));
    }
};
struct xy_rect : public hittable {
    xy_rect() {}
    xy_rect(const float x0, const float x1, const float y0, const float y1, const float k, material* m) { h = rtc::ok(rt_xy_rect(rtc::sc(), x0, x1, y0, y1, k, m->h)); }
};
struct xz_rect : public hittable {
    xz_rect() {}
    xz_rect(const float x0, const float x1, const float z0, const float z1, const float k, material* m) { h = rtc::ok(rt_xz_rect(rtc::sc(), x0, x1, z0, z1, k, m->h)); }
};
struct yz_rect : public hittable {
    yz_rect() {}
    yz_rect(const float y0, const float y1, const float z0, const float z1, const float k, material* m) { h = rtc::ok(rt_yz_rect(rtc::sc(), y0, y1, z0, z1, k, m->h)); }
};
struct box : public hittable {
    box(point3 p0, point3 p1, material* m) { h = rtc::ok(rt_box(rtc::sc(), p0.e, p1.e, m->h)); }
};
struct triangle : public hittable {
    triangle() {}
    triangle(const vec3 v0, const vec3 v1, const vec3 v2, const float u0, const float vv0, const float u1, const float vv1, const float u2,
             const float vv2, material* m) {
        const float v[9] = {v0.x(), v0.y(), v0.z(), v1.x(), v1.y(), v1.z(), v2.x(), v2.y(), v2.z()};
        const float uv[6] = {u0, vv0, u1, vv1, u2, vv2};
        h = rtc::ok(rt_triangle(rtc::sc(), v, uv, nullptr, m->h));
    }
    triangle(const vec3 v0, const vec3 v1, const vec3 v2, const vec3 n0, const vec3 n1, const vec3 n2, const float u0, const float vv0,
             const float u1, const float vv1, const float u2, const float vv2, material* m) {
        const float v[9] = {v0.x(), v0.y(), v0.z(), v1.x(), v1.y(), v1.z(), v2.x(), v2.y(), v2.z()};
        const float n[9] = {n0.x(), n0.y(), n0.z(), n1.x(), n1.y(), n1.z(), n2.x(), n2.y(), n2.z()};
        const float uv[6] = {u0, vv0, u1, vv1, u2, vv2};
        h = rtc::ok(rt_triangle(rtc::sc(), v, uv, n, m->h));
    }
};
struct translate : public hittable {
    translate(hittable* p, const vec3& d) { h = rtc::ok(rt_translate(rtc::sc(), p->h, d.x(), d.y(), d.z())); }
};
struct rotate_y : public hittable {
    rotate_y(hittable* p, const float angle) { h = rtc::ok(rt_rotate_y(rtc::sc(), p->h, angle)); }
};
struct constant_medium : public hittable {
    constant_medium(hittable* b, const float d, texturez* a) { h = rtc::ok(rt_constant_medium(rtc::sc(), b->h, d, a->h)); }
    constant_medium(hittable* b, const float d, color c) {
        texturez* a = new solid_color(c);
        h = rtc::ok(rt_constant_medium(rtc::sc(), b->h, d, a->h));
    }
};

namespace rtc {
inline std::vector<int> handles(hittable** l, int n, int offset = 0) {
    std::vector<int> v((size_t)n);
    for (int i = 0; i < n; i++) v[(size_t)i] = l[i + offset]->h;
    return v;
}
} // namespace rtc

struct hittable_list : public hittable {
    hittable** objects = nullptr; int list_size = 0;
    hittable_list() {}
    hittable_list(hittable** l, int n) : objects(l), list_size(n) {
        const std::vector<int> v = rtc::handles(l, n);
        h = rtc::ok(rt_list(rtc::sc(), v.data(), n));
    }
};

// bvh_node(hits, n, time0, time1, rng, offset) — bvh.h:270.  Grouping hint only (everything lands in the one
// LBVH).  The reference draws one random axis per internal node (bvh.h:294); the same number of values is
// consumed here so that anything drawn AFTER a bvh_node sees the same stream position.
struct bvh_node : public hittable {
    int n = 0;
    bvh_node() {}
    bvh_node(hittable**& hits, int num_obj, const float, const float, curandState* s, int offset = 0) : n(num_obj) {
        const std::vector<int> v = rtc::handles(hits, num_obj, offset);
        h = rtc::ok(rt_bvh(rtc::sc(), v.data(), num_obj));
        const int internal = static_cast<int>(powf(2, ceilf(log2f((float)num_obj)))) - 1;   // num_nodes() - n, bvh.h:143
        for (int i = 0; i < internal; i++) (void)random_int(s, 0, 2);
    }
};

struct triangle_mesh : public hittable {
    bvh_node* tris = nullptr; int n = 0;
    triangle_mesh() {}
    triangle_mesh(hittable**& triangles, int num, const float t0, const float t1, curandState* s) : n(num) {
        tris = new bvh_node(triangles, num, t0, t1, s);
        h = tris->h;
    }
    triangle_mesh(hittable**& triangles, unsigned* num, const float t0, const float t1, curandState* s, int id) : n((int)num[id]) {
        int offset = 0;
        for (int i = 0; i < id; i++) offset += (int)num[i];
        tris = new bvh_node(triangles, (int)num[id], t0, t1, s, offset);
        h = tris->h;
    }
};

// ---- camera.h -------------------------------------------------------------------------------------
struct camera {
    vec3 lookfrom, lookat, vup; float vfov, aspect_ratio, aperture, focus_dist, time0, time1;
    camera() {}
    camera(const point3 from, const vec3 at, const vec3 up, const float fov, const float aspect, const float ap, const float focus,
           const float t0, const float t1)
        : lookfrom(from), lookat(at), vup(up), vfov(fov), aspect_ratio(aspect), aperture(ap), focus_dist(focus), time0(t0), time1(t1) {}
};

namespace rtc {
// hand the finished world to the scene: *d_world / *d_camera / scene::background of the reference (scenes.h:36-44)
inline void finish(hittable* world, const camera* cam, const color& background) {
    ok(rt_set_world(sc(), world->h));
    ok(rt_set_background(sc(), background.x(), background.y(), background.z()));
    ok(rt_set_camera(sc(), cam->lookfrom.e, cam->lookat.e, cam->vup.e, cam->vfov, cam->aspect_ratio, cam->aperture, cam->focus_dist,
                     cam->time0, cam->time1));
}
} // namespace rtc

#endif // RT_COMPAT_H
