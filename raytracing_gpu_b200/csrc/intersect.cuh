// intersect.cuh — device intersection + hit-attribute routines over tagged-union primitives.
//
// These replace the reference's virtual hittable::hit family.  The arithmetic that decides (primitive, t) keeps
// the reference's operation order and is compiled with -fmad=false, so t is bit-identical to the reference's
// code compiled without FMA contraction (oracle/).  FMAs appear only where written as fmaf() (BVH slab tests).
#pragma once

#include "device_scene.cuh"

namespace rtb {

#define RT_MAX_CHAIN 8
#define RT_PI_F 3.1415926535897932385f   // common.h:43

struct Ray { float3 o, d; float tm; };

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(float t, float3 v) { return f3(t * v.x, t * v.y, t * v.z); }   // vec3.h:109
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 vdiv(float3 v, float t) { return (1 / t) * v; }                           // vec3.h:117
__device__ __forceinline__ float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }     // vec3.h:121
__device__ __forceinline__ float3 cross3(float3 u, float3 v) {                                                // vec3.h:125
    return f3(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
__device__ __forceinline__ float len2(float3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }               // vec3.h:58
__device__ __forceinline__ float len3(float3 v) { return sqrtf(len2(v)); }
__device__ __forceinline__ float3 unit3(float3 v) { return vdiv(v, len3(v)); }                               // vec3.h:129
__device__ __forceinline__ float3 ray_at(const Ray& r, float t) { return r.o + t * r.d; }                   // ray.h:17

__device__ __forceinline__ uint32_t prim_type(const DPrim& p) { return p.type_chain & 15u; }
__device__ __forceinline__ uint32_t prim_chain(const DPrim& p) { return p.type_chain >> 4; }

// ---- ray -> object space through a wrapper chain (translate::hit hittable.h:38, rotate_y::hit :113-123) ----
__device__ __forceinline__ void apply_xop(const rt_flat_xop& op, Ray& r) {
    if (op.kind == RT_XOP_TRANSLATE) {
        r.o = r.o - f3(op.a, op.b, op.c);
    } else {
        const float s = op.a, c = op.b;
        const float ox = c * r.o.x - s * r.o.z, oz = s * r.o.x + c * r.o.z;
        const float dx = c * r.d.x - s * r.d.z, dz = s * r.d.x + c * r.d.z;
        r.o.x = ox; r.o.z = oz; r.d.x = dx; r.d.z = dz;
    }
}

__device__ __forceinline__ Ray to_object_space(const SceneView& sv, uint32_t chain, Ray r) {
    if (chain) {
        const rt_flat_chain c = sv.chains[chain];
        #pragma unroll 1   // chains are 1-2 ops long; unrolled copies only cost instruction-cache space
        for (uint32_t k = 0; k < c.count; k++) apply_xop(sv.xops[c.first + k], r);
    }
    return r;
}

// ---- t-only intersection in object space: returns true and t when the reference's ::hit would accept ----
__device__ __forceinline__ bool sphere_t(float3 center, float radius, const Ray& r, float t_min, float t_max, bool far_root, float& t) {
    // sphere.h:38-55 — near root only (the reference's "second root" repeats the first);
    // far_root (RT_SEM_SPHERE_FAR_ROOT): the book's second root when the near one is out of range
    const float3 oc = r.o - center;
    const float a = len2(r.d);
    const float half_b = dot3(oc, r.d);
    const float c = len2(oc) - radius * radius;
    const float discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return false;
    const float sqrtd = sqrtf(discriminant);
    float root = (-half_b - sqrtd) / a;
    if (root < t_min || t_max < root) {
        if (!far_root) return false;
        root = (-half_b + sqrtd) / a;
        if (root < t_min || t_max < root) return false;
    }
    t = root;
    return true;
}

__device__ __forceinline__ float3 moving_center(const DPrim& p, float time) {   // moving_sphere.h:20-22
    const float3 c0 = f3(p.a.x, p.a.y, p.a.z), c1 = f3(p.a.w, p.b.x, p.b.y);
    return c0 + ((time - p.b.z) / (p.b.w - p.b.z)) * (c1 - c0);
}

__device__ __forceinline__ bool rect_t(uint32_t type, const DPrim& p, const Ray& r, float t_min, float t_max, float& t) {
    // aarect.h:63-75, :94-105, :124-135.  The three orientations differ only in which components play (k, a, b): the
    // components are selected first so that there is ONE copy of the arithmetic (and of the division) in the kernel —
    // less code, and rects of different orientation in one warp stay converged
    const bool xy = type == RT_PRIM_XY_RECT, xz = type == RT_PRIM_XZ_RECT;
    const float ok = xy ? r.o.z : (xz ? r.o.y : r.o.x), dk = xy ? r.d.z : (xz ? r.d.y : r.d.x);
    const float oa = (xy | xz) ? r.o.x : r.o.y, da = (xy | xz) ? r.d.x : r.d.y;
    const float ob = xy ? r.o.y : r.o.z, db = xy ? r.d.y : r.d.z;
    const float tt = (p.b.x - ok) / dk;
    const float a = oa + tt * da, b = ob + tt * db;
    if (tt < t_min || tt > t_max) return false;
    if (a < p.a.x || a > p.a.y || b < p.a.z || b > p.a.w) return false;
    t = tt;
    return true;
}

__device__ __forceinline__ bool triangle_t(const DPrim& p, const Ray& r, float t_min, float t_max, float& t) {
    // triangle.h:124-152 (Möller–Trumbore, epsilon 1e-7)
    const float epsilon = 0.0000001f;
    const float3 vertex0 = f3(p.a.x, p.a.y, p.a.z), e0 = f3(p.a.w, p.b.x, p.b.y), e1 = f3(p.b.z, p.b.w, p.c0);
    const float3 h = cross3(r.d, e1);
    const float a = dot3(e0, h);
    if (a > -epsilon && a < epsilon) return false;
    const float f = 1.0f / a;
    const float3 s = r.o - vertex0;
    const float u = f * dot3(s, h);
    if (u < 0.0f || u > 1.0f) return false;
    const float3 q = cross3(s, e0);
    const float v = f * dot3(r.d, q);
    if ((v < 0.0f) | (u + v > 1.0f)) return false;
    const float tt = f * dot3(e1, q);
    if (tt < t_min || tt > t_max || tt < epsilon) return false;
    t = tt;
    return true;
}

// object-space test of one primitive (ray already in the primitive's frame)
__device__ __forceinline__ bool prim_t_local(const DPrim& p, const Ray& r, float t_min, float t_max, uint32_t semantics, float& t) {
    const uint32_t type = prim_type(p);
    if (type <= RT_PRIM_MOVING_SPHERE) {   // one copy of the sphere test for both kinds: only centre and radius differ
        static_assert(RT_PRIM_SPHERE == 0 && RT_PRIM_MOVING_SPHERE == 1, "sphere kinds first");
        float3 center = f3(p.a.x, p.a.y, p.a.z);
        float radius = p.a.w;
        if (type == RT_PRIM_MOVING_SPHERE) { center = moving_center(p, r.tm); radius = p.c0; }
        return sphere_t(center, radius, r, t_min, t_max, (semantics & RT_SEM_SPHERE_FAR_ROOT) != 0, t);
    }
    if (type == RT_PRIM_TRIANGLE) return triangle_t(p, r, t_min, t_max, t);
    return rect_t(type, p, r, t_min, t_max, t);
}

// world-space test (applies the primitive's wrapper chain)
__device__ __forceinline__ bool prim_t(const SceneView& sv, const DPrim& p, const Ray& world, float t_min, float t_max, float& t) {
    const Ray r = to_object_space(sv, prim_chain(p), world);
    return prim_t_local(p, r, t_min, t_max, sv.semantics, t);
}

// ---- hit attributes of the accepted hit: what the reference leaves in hit_record ---------------------
struct HitAttr {
    float3 p, normal;
    float u, v;
    bool front_face;
};

__device__ __forceinline__ void set_face_normal(HitAttr& h, float3 ray_dir, float3 outward_normal) {   // hittable.h:17-22
    h.front_face = dot3(ray_dir, outward_normal) < 0;
    h.normal = h.front_face ? outward_normal : -outward_normal;
}

// object-space attributes at distance t (sphere.h:57-70, moving_sphere.h:48-57, aarect.h:78-89 etc., triangle.h:154-175)
__device__ __forceinline__ void prim_attr_local(const SceneView& sv, const DPrim& p, const Ray& r, float t, HitAttr& h) {
    const uint32_t type = prim_type(p);
    h.p = ray_at(r, t);
    if (type == RT_PRIM_SPHERE) {
        const float3 outward_normal = vdiv(h.p - f3(p.a.x, p.a.y, p.a.z), p.a.w);
        set_face_normal(h, r.d, outward_normal);
        const float theta = acosf(-outward_normal.y);                      // get_sphere_uv, sphere.h:19-32
        const float phi = atan2f(-outward_normal.z, outward_normal.x) + RT_PI_F;
        h.u = phi / (2 * RT_PI_F);
        h.v = theta / RT_PI_F;
    } else if (type == RT_PRIM_MOVING_SPHERE) {
        const float3 outward_normal = vdiv(h.p - moving_center(p, r.tm), p.c0);
        set_face_normal(h, r.d, outward_normal);
        h.u = 0.0f; h.v = 0.0f;                                            // not set by the reference
    } else if (type == RT_PRIM_TRIANGLE) {
        const rt_flat_tri_attr& at = sv.tri_attrs[p.attr];
        const float3 vertex0 = f3(p.a.x, p.a.y, p.a.z), e0 = f3(p.a.w, p.b.x, p.b.y), e1 = f3(p.b.z, p.b.w, p.c0);
        const float3 v2 = h.p - vertex0;                                   // barycentric_coords, triangle.h:101-111
        const float d20 = dot3(v2, e0), d21 = dot3(v2, e1);
        const float B0 = (at.d11 * d20 - at.d01 * d21) * at.inv_denom;
        const float B1 = (at.d00 * d21 - at.d01 * d20) * at.inv_denom;
        const float B2 = 1.0f - B0 - B1;
        h.u = B2 * at.uv[0] + B0 * at.uv[2] + B1 * at.uv[4];               // barycentric_interp, triangle.h:113-116
        h.v = B2 * at.uv[1] + B0 * at.uv[3] + B1 * at.uv[5];
        float3 n;
        if (!at.vertex_normals) n = cross3(e1, e0);                        // not normalised
        else n = f3(B2 * at.n[0] + B0 * at.n[3] + B1 * at.n[6], B2 * at.n[1] + B0 * at.n[4] + B1 * at.n[7],
                    B2 * at.n[2] + B0 * at.n[5] + B1 * at.n[8]);
        if (sv.semantics & RT_SEM_UNIT_TRIANGLE_NORMALS) n = unit3(n);
        set_face_normal(h, r.d, n);
    } else {
        float a, b;
        float3 outward_normal;
        if (type == RT_PRIM_XY_RECT) { a = r.o.x + t * r.d.x; b = r.o.y + t * r.d.y; outward_normal = f3(0, 0, 1); }
        else if (type == RT_PRIM_XZ_RECT) { a = r.o.x + t * r.d.x; b = r.o.z + t * r.d.z; outward_normal = f3(0, 1, 0); }
        else { a = r.o.y + t * r.d.y; b = r.o.z + t * r.d.z; outward_normal = f3(1, 0, 0); }
        h.u = (a - p.a.x) / (p.a.y - p.a.x);
        h.v = (b - p.a.z) / (p.a.w - p.a.z);
        set_face_normal(h, r.d, outward_normal);
    }
}

// Unwind a wrapper chain over an object-space hit record: translate::hit hittable.h:44-45, rotate_y::hit :128-140.
// dirs[k] = ray direction after applying ops[0..k] (the direction the wrapper at level k passed to its child).
__device__ __forceinline__ void unwind_chain(const SceneView& sv, uint32_t chain, const float3* dirs, HitAttr& h) {
    if (!chain) return;
    const rt_flat_chain c = sv.chains[chain];
    for (int k = (int)c.count - 1; k >= 0; k--) {
        const rt_flat_xop op = sv.xops[c.first + k];
        if (op.kind == RT_XOP_TRANSLATE) {
            h.p = h.p + f3(op.a, op.b, op.c);
            set_face_normal(h, dirs[k], h.normal);
        } else {
            const float s = op.a, cth = op.b;
            float3 p = h.p, n = h.normal;
            p.x = cth * h.p.x + s * h.p.z;
            p.z = -s * h.p.x + cth * h.p.z;
            n.x = cth * h.normal.x + s * h.normal.z;
            n.z = -s * h.normal.x + cth * h.normal.z;
            h.p = p;
            set_face_normal(h, dirs[k], n);     // against the ROTATED ray, as the reference does
        }
    }
}

// ray -> object space, remembering the per-level directions needed by unwind_chain
__device__ __forceinline__ Ray to_object_space_dirs(const SceneView& sv, uint32_t chain, Ray r, float3* dirs) {
    if (chain) {
        const rt_flat_chain c = sv.chains[chain];
        for (uint32_t k = 0; k < c.count; k++) { apply_xop(sv.xops[c.first + k], r); dirs[k] = r.d; }
    }
    return r;
}

// world-space attributes of a surface hit
__device__ __forceinline__ void prim_attr(const SceneView& sv, const DPrim& p, const Ray& world, float t, HitAttr& h) {
    float3 dirs[RT_MAX_CHAIN];
    const uint32_t chain = prim_chain(p);
    const Ray r = to_object_space_dirs(sv, chain, world, dirs);
    prim_attr_local(sv, p, r, t, h);
    unwind_chain(sv, chain, dirs, h);
}

// ---- constant_medium::hit — constant_medium.h:34-70 -------------------------------------------------
// The two boundary queries of constant_medium::hit (constant_medium.h:36-40): the closest hit of the boundary list in
// (-inf, inf), then the closest one from just behind it, (t1 + 0.0001, inf) — list semantics, hittable_list.h:23-39:
// a candidate is accepted unless t < t_min or t > closest-so-far, later objects win ties.
// Whether a primitive offers a candidate at all, and at which t, does not depend on the query range (a rect's bounds test,
// a triangle's barycentric tests, a sphere's near root), so each boundary primitive is intersected ONCE with the range
// wide open; both queries then replay the list's accept / shrink sequence over the stored candidates, comparison for
// comparison (NaN and infinite t included).  Half the primitive tests of a medium: Cornell smoke extend -13 %, step +9.3 % (the two
// box-bounded media there cost more than the BVH).  Falls back to intersecting per query where the candidate itself
// depends on the range (RT_SEM_SPHERE_FAR_ROOT) or the boundary has more than RT_BOUNDARY_CACHE primitives.
// The boundary primitives of one medium nearly always share ONE wrapper chain (a box under translate(rotate_y)), so the
// object-space ray is cached per chain across the primitives.
#define RT_BOUNDARY_CACHE 8
struct ChainCache { uint32_t chain; Ray ray; };

// Returns true and the hit distance when the medium scatters the ray before t_max.  `xi()` yields the uniform
// draw; it is only evaluated when the ray really crosses the boundary (constant_medium.h:57).
template <class XiFn>
__device__ __forceinline__ bool medium_t(const SceneView& sv, const DMedium& m, const Ray& world, float t_min, float t_max, XiFn&& xi, float& t) {
    const Ray r = to_object_space(sv, m.chain, world);
    ChainCache cc; cc.chain = 0; cc.ray = r;
    const bool per_query = (sv.semantics & RT_SEM_SPHERE_FAR_ROOT) != 0 || m.bprim_count > RT_BOUNDARY_CACHE;
    float cand[RT_BOUNDARY_CACHE];
    uint32_t offered = 0;
    // ONE instance of the boundary code, looped: k_extend's code size decides its instruction-cache hit rate on scenes
    // with media (profiles/r01_v6_*.md)
    float tq[2] = {0.f, 0.f};
    float lo = -INFINITY;
    #pragma unroll 1
    for (int q = 0; q < 2; q++) {
        bool any = false;
        float closest = INFINITY;
        for (uint32_t i = 0; i < m.bprim_count; i++) {
            float tt = 0.f;
            bool ok;
            if (q == 0 || per_query) {
                const DPrim p = sv.bprims[m.bprim_first + i];
                const uint32_t chain = prim_chain(p);
                if (chain != cc.chain) { cc.ray = to_object_space(sv, chain, r); cc.chain = chain; }
                ok = prim_t_local(p, cc.ray, per_query ? lo : -INFINITY, per_query ? closest : INFINITY, sv.semantics, tt);
                if (!per_query) { cand[i] = tt; offered |= (ok ? 1u : 0u) << i; }
            } else {
                tt = cand[i];
                ok = (offered >> i) & 1u;
            }
            if (!per_query) ok = ok && !(tt < lo) && !(tt > closest);       // the range test the primitive would have made
            if (ok) { any = true; closest = tt; }
        }
        if (!any) return false;
        tq[q] = closest;
        lo = closest + 0.0001f;
    }
    float t1 = tq[0], t2 = tq[1];
    if (t1 < t_min) t1 = t_min;
    if (t2 > t_max) t2 = t_max;
    if (t1 >= t2) return false;
    if (t1 < 0) t1 = 0;
    const float ray_length = len3(r.d);
    const float distance_inside_boundary = (t2 - t1) * ray_length;
    const float hit_distance = m.neg_inv_density * logf(xi());
    if (hit_distance > distance_inside_boundary) return false;
    t = t1 + hit_distance / ray_length;
    return true;
}

// attributes of a medium hit: p = r.at(t) in the medium's frame, normal (1,0,0), front_face true, then the
// wrappers above the medium unwind as for any other object
__device__ __forceinline__ void medium_attr(const SceneView& sv, const DMedium& m, const Ray& world, float t, HitAttr& h) {
    float3 dirs[RT_MAX_CHAIN];
    const Ray r = to_object_space_dirs(sv, m.chain, world, dirs);
    h.p = ray_at(r, t);
    h.normal = f3(1, 0, 0);
    h.front_face = true;
    h.u = 0.0f; h.v = 0.0f;
    unwind_chain(sv, m.chain, dirs, h);
}

} // namespace rtb
