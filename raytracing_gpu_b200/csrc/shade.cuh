// shade.cuh — device textures, materials and camera sampling as tagged unions (replaces the virtual
// texturez::value / material::scatter / material::emitted / camera::get_ray calls of the reference).
// Operation order follows the reference; compiled with -fmad=false (see intersect.cuh).
#pragma once

#include "intersect.cuh"
#include "rng.cuh"

namespace rtb {

#define RT_MAX_REJECT 64   // rejection loops are bounded; P(overflow) < 1e-20

// ---- perlin.h ------------------------------------------------------------------------------------
__device__ __forceinline__ float perlin_noise(const DPerlin& pn, float3 p) {   // perlin.h:85-112 + perlin_interp :37-57
    const float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    const float u = p.x - fx, v = p.y - fy, w = p.z - fz;
    const int i = (int)fx, j = (int)fy, k = (int)fz;
    const float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
    float accum = 0.0f;
    #pragma unroll
    for (int a = 0; a < 2; a++)
        #pragma unroll
        for (int b = 0; b < 2; b++)
            #pragma unroll
            for (int c = 0; c < 2; c++) {
                const int ix = pn.perm[0][(i + a) & 255] ^ pn.perm[1][(j + b) & 255] ^ pn.perm[2][(k + c) & 255];
                const float4 g = __ldg(&pn.ranvec[ix]);
                const float3 weight_v = f3(u - a, v - b, w - c);
                accum += (a * uu + (1 - a) * (1 - uu)) * (b * vv + (1 - b) * (1 - vv)) * (c * ww + (1 - c) * (1 - ww)) *
                         dot3(f3(g.x, g.y, g.z), weight_v);
            }
    return accum;
}

__device__ __forceinline__ float perlin_turb(const DPerlin& pn, float3 p, int depth) {   // perlin.h:114-126 (double accumulator)
    double accum = 0.0, weight = 1.0;
    float3 temp_p = p;
    for (int i = 0; i < depth; i++) {
        accum += weight * (double)perlin_noise(pn, temp_p);
        weight *= 0.5;
        temp_p = 2.0f * temp_p;
    }
    return (float)fabs(accum);
}

// ---- texture.h -----------------------------------------------------------------------------------
__device__ __forceinline__ float clamp_d(float x, float mn, float mx) { return x < mn ? mn : (x > mx ? mx : x); }   // common.h:62

__device__ __forceinline__ float3 texture_value(const SceneView& sv, uint32_t ti, float u, float v, float3 p) {
    rt_flat_texture t = sv.textures[ti];
    for (int guard = 0; t.type == RT_TEX_CHECKER && guard < 8; guard++) {          // texture.h:37-45
        const float sines = sinf(10 * p.x) * sinf(10 * p.y) * sinf(10 * p.z);
        t = sv.textures[sines < 0 ? t.b : t.a];
    }
    switch (t.type) {
    case RT_TEX_SOLID: return f3(t.rgb[0], t.rgb[1], t.rgb[2]);                     // texture.h:17
    case RT_TEX_NOISE: {                                                            // texture.h:57-59
        const float g = 0.5f * (1.0f + perlin_noise(sv.perlins[t.a], t.scale * p));
        return f3(g, g, g);
    }
    case RT_TEX_TURB: {                                                             // texture.h:74-76
        const float g = perlin_turb(sv.perlins[t.a], t.scale * p, t.depth);
        return f3(g, g, g);
    }
    case RT_TEX_MARBLE: {                                                           // texture.h:88-90
        const float g = 0.5f * (1 + sinf(t.scale * p.z + 10 * perlin_turb(sv.perlins[t.a], t.scale * p, 7)));
        return f3(g, g, g);
    }
    case RT_TEX_IMAGE: {                                                            // image_texture::value, texture.h:143-163
        const DImage im = sv.images[t.a];
        const float uu = clamp_d(u, 0.0f, 1.0f);
        const double vv = 1.0 - (double)clamp_d(v, 0.0f, 1.0f);
        int i = (int)(uu * im.width), j = (int)(vv * im.height);
        if (i >= (int)im.width) i = (int)im.width - 1;
        if (j >= (int)im.height) j = (int)im.height - 1;
        const uchar4 px = tex2D<uchar4>(im.tex, i + 0.5f, j + 0.5f);
        const float color_scale = 1.0f / 255.0f;
        return f3(color_scale * px.x, color_scale * px.y, color_scale * px.z);
    }
    default: return f3(0, 0, 0);
    }
}

// ---- vec3.h samplers -----------------------------------------------------------------------------
__device__ __forceinline__ float3 random_in_unit_sphere(const RngKey& g) {         // vec3.h:129-134; iteration k = block k
    float3 p = f3(0, 0, 0);
    for (uint32_t k = 0; k < RT_MAX_REJECT; k++) {
        const float4 u = draw4(g, RT_STREAM_SCATTER, k);
        p = f3(rf(u.x, -1, 1), rf(u.y, -1, 1), rf(u.z, -1, 1));
        if (len2(p) < 1) break;
    }
    return p;
}

__device__ __forceinline__ bool near_zero(float3 v) { const float s = 1e-6f; return fabsf(v.x) < s && fabsf(v.y) < s && fabsf(v.z) < s; }   // vec3.h:72
__device__ __forceinline__ float3 reflect3(float3 v, float3 n) { return v - (2 * dot3(v, n)) * n; }                                         // vec3.h:147
__device__ __forceinline__ float3 refract3(float3 uv, float3 n, float etai_over_etat) {                                                     // vec3.h:152
    const float cos_theta = fminf(dot3(-uv, n), 1.0f);
    const float3 r_out_perp = etai_over_etat * (uv + cos_theta * n);
    const float3 r_out_parallel = (-sqrtf(fabsf(1.0f - len2(r_out_perp)))) * n;
    return r_out_perp + r_out_parallel;
}

// ---- material.h ----------------------------------------------------------------------------------
// Returns true when the path continues; `scattered` inherits the incoming ray's time.
__device__ __forceinline__ bool material_scatter(const SceneView& sv, const rt_flat_material& m, const Ray& ray_in, const HitAttr& rec,
                                                 const RngKey& g, float3& attenuation, Ray& scattered) {
    scattered.o = rec.p;
    scattered.tm = ray_in.tm;
    switch (m.type) {
    case RT_MAT_LAMBERTIAN: {                                                       // material.h:25-35
        float3 scatter_direction = rec.normal + unit3(random_in_unit_sphere(g));
        if (near_zero(scatter_direction)) scatter_direction = rec.normal;
        scattered.d = scatter_direction;
        attenuation = texture_value(sv, m.texture, rec.u, rec.v, rec.p);
        return true;
    }
    case RT_MAT_METAL: {                                                            // material.h:50-55
        const float3 reflected = reflect3(unit3(ray_in.d), rec.normal);
        scattered.d = reflected + m.param * random_in_unit_sphere(g);
        attenuation = texture_value(sv, m.texture, rec.u, rec.v, rec.p);
        return dot3(scattered.d, rec.normal) > 0;
    }
    case RT_MAT_DIELECTRIC: {                                                       // material.h:64-95
        const float ir = m.param;
        attenuation = f3(1.0f, 1.0f, 1.0f);
        const float refraction_ratio = rec.front_face ? (1.0f / ir) : ir;
        const float3 unit_direction = unit3(ray_in.d);
        const float cos_theta = fminf(dot3(-unit_direction, rec.normal), 1.0f);
        const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
        bool reflect_it = refraction_ratio * sin_theta > 1.0f;
        if (!reflect_it) {                                                          // reflectance, material.h:98-103
            const float sqrt_r0 = (1.0f - refraction_ratio) / (1.0f + refraction_ratio);
            const float r0 = sqrt_r0 * sqrt_r0;
            const float x = 1.0f - cos_theta;
            const float refl = r0 + (1.0f - r0) * (x * x * x * x * x);
            reflect_it = refl > draw4(g, RT_STREAM_SCATTER, 0).x;
        }
        scattered.d = reflect_it ? reflect3(unit_direction, rec.normal) : refract3(unit_direction, rec.normal, refraction_ratio);
        return true;
    }
    case RT_MAT_ISOTROPIC: {                                                        // material.h:133-137
        scattered.d = random_in_unit_sphere(g);
        attenuation = texture_value(sv, m.texture, rec.u, rec.v, rec.p);
        return true;
    }
    default: return false;                                                          // diffuse_light, material.h:115-117
    }
}

// ---- camera::get_ray — camera.h:49-58, pixel jitter render.h:105-106 -----------------------------
// draws: camera block 0 = (jitter u, jitter v, time, -); lens iteration k = words (2(k&1), 2(k&1)+1) of block 1 + k/2
__device__ __forceinline__ Ray camera_ray(const rt_flat_camera& c, const RngKey& g, int i, int j, int max_x, int max_y) {
    const float4 b0 = draw4(g, RT_STREAM_CAMERA, 0);
    const float su = (float)(i + b0.x) / max_x;
    const float sv_ = (float)(j + b0.y) / max_y;
    float3 pd = f3(0, 0, 0);
    if (c.lens_radius != 0.0f) {          // with a zero radius the reference multiplies the sample by 0: skipped, same ray
        for (uint32_t k = 0; k < RT_MAX_REJECT; k++) {                              // random_in_unit_disk, vec3.h:136-141
            const float4 u = draw4(g, RT_STREAM_CAMERA, 1 + k / 2);
            pd = (k & 1) ? f3(rf(u.z, -1, 1), rf(u.w, -1, 1), 0) : f3(rf(u.x, -1, 1), rf(u.y, -1, 1), 0);
            if (len2(pd) < 1) break;
        }
    }
    const float3 rd = c.lens_radius * pd;
    const float3 cu = f3(c.u[0], c.u[1], c.u[2]), cv = f3(c.v[0], c.v[1], c.v[2]);
    const float3 offset = rd.x * cu + rd.y * cv;
    const float3 origin = f3(c.origin[0], c.origin[1], c.origin[2]);
    const float3 llc = f3(c.lower_left_corner[0], c.lower_left_corner[1], c.lower_left_corner[2]);
    const float3 hor = f3(c.horizontal[0], c.horizontal[1], c.horizontal[2]), ver = f3(c.vertical[0], c.vertical[1], c.vertical[2]);
    Ray r;
    r.o = origin + offset;
    r.d = llc + su * hor + sv_ * ver - origin - offset;
    r.tm = rf(b0.z, c.time0, c.time1);
    return r;
}

} // namespace rtb
