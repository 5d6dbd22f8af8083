/* oracle/rt_oracle.c — TEST INFRASTRUCTURE: plain-C CPU restatement of the reference's path-tracing hot path.
 *
 * Consumes the flattened scene blob (include/rt_flat.h, produced by rt_scene_export) and restates, function by
 * function, the reference's intersection / shading / camera / texture arithmetic — every function cites the
 * reference file:line it follows (paths relative to the reference repo root).  Float expressions keep the
 * reference's operation order and are compiled with -ffp-contract=off, so results are IEEE-single exactly as
 * written.  The random stream is NOT the reference's per-pixel XORWOW (which is racy): it is the
 * product's Philox4x32-10 keyed on (pixel, sample, bounce), restated here independently from the published
 * Philox algorithm (Salmon et al., SC'11) and pinned by Random123's known-answer vectors in tests/.
 *
 * Parity status: PINNED — tests/test_oracle_vs_ref.py checks first-hit ids/t bit-exactly and converged images
 * statistically against oracle/_ref (the reference's own sources compiled for the host); the reference itself
 * ships no tests or golden vectors.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.  It is never
 * linked into or called from the product.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/rt_flat.h"

#define RTO_API __attribute__((visibility("default")))

typedef struct { float x, y, z; } v3;

/* ---- vec3.h ---------------------------------------------------------------------------------- */
static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }            /* vec3.h:97 */
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }            /* vec3.h:101 */
static inline v3 vmulv(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }           /* vec3.h:105 */
static inline v3 vmul(float t, v3 v) { return V(t * v.x, t * v.y, t * v.z); }               /* vec3.h:109 */
static inline v3 vdiv(v3 v, float t) { return vmul(1 / t, v); }                             /* vec3.h:117 */
static inline v3 vneg(v3 v) { return V(-v.x, -v.y, -v.z); }
static inline float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }          /* vec3.h:121 */
static inline v3 vcross(v3 u, v3 v) { return V(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); } /* vec3.h:125 */
static inline float vlen2(v3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }               /* vec3.h:58 */
static inline float vlen(v3 v) { return sqrtf(vlen2(v)); }                                  /* vec3.h:54 */
static inline v3 vunit(v3 v) { return vdiv(v, vlen(v)); }                                   /* vec3.h:129 */
static inline int vnear_zero(v3 v) { const float s = 1e-6f; return fabsf(v.x) < s && fabsf(v.y) < s && fabsf(v.z) < s; } /* vec3.h:72 */
static inline v3 vreflect(v3 v, v3 n) { return vsub(v, vmul(2 * vdot(v, n), n)); }          /* vec3.h:147 */
static inline v3 vrefract(v3 uv, v3 n, float etai_over_etat) {                               /* vec3.h:152 */
    const float cos_theta = fminf(vdot(vneg(uv), n), 1.0f);
    const v3 r_out_perp = vmul(etai_over_etat, vadd(uv, vmul(cos_theta, n)));
    const v3 r_out_parallel = vmul(-sqrtf(fabsf(1.0f - vlen2(r_out_perp))), n);
    return vadd(r_out_perp, r_out_parallel);
}

typedef struct { v3 o, d; float tm; } rayc;                                                  /* ray.h:5 */
static inline v3 ray_at(const rayc* r, float t) { return vadd(r->o, vmul(t, r->d)); }       /* ray.h:17 */

/* ---- Philox4x32-10 (counter-based RNG; replaces curand XORWOW, render.h:84-92) -------------------- */
#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u

RTO_API void rto_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        const uint64_t p0 = (uint64_t)PHILOX_M0 * c0, p1 = (uint64_t)PHILOX_M1 * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += PHILOX_W0; k1 += PHILOX_W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* uint32 -> (0,1], the mapping curand_uniform uses (curand_uniform.h:69-72), as one fused multiply-add */
static inline float u01(uint32_t x) { return fmaf((float)x, 2.3283064e-10f, 1.1641532e-10f); }
RTO_API float rto_u01(uint32_t x) { return u01(x); }

/* Stream ids (Philox key word 1) and counter layout: ctr = (pixel, sample, bounce, block). */
enum { RTO_STREAM_CAMERA = 0x43414d00u, RTO_STREAM_SCATTER = 0x53434100u, RTO_STREAM_MEDIUM = 0x4d454400u };

typedef struct { uint32_t pixel, sample, bounce, seed; const float* scripted_xi; } rng_ctx;

static void draw4(const rng_ctx* g, uint32_t stream, uint32_t block, float out[4]) {
    const uint32_t ctr[4] = {g->pixel, g->sample, g->bounce, block}, key[2] = {g->seed, stream};
    uint32_t r[4];
    rto_philox4x32_10(ctr, key, r);
    for (int i = 0; i < 4; i++) out[i] = u01(r[i]);
}

#define RTO_MAX_REJECT 64 /* rejection loops are bounded; P(overflow) < 1e-20 */

static inline float rf(float u, float mn, float mx) { return mn + (mx - mn) * u; }          /* vec3.h:11 */

/* random_in_unit_sphere (vec3.h:129-134): iteration k uses block k of the scatter stream */
static v3 random_in_unit_sphere(const rng_ctx* g) {
    v3 p = V(0, 0, 0);
    for (uint32_t k = 0; k < RTO_MAX_REJECT; k++) {
        float u[4];
        draw4(g, RTO_STREAM_SCATTER, k, u);
        p = V(rf(u[0], -1, 1), rf(u[1], -1, 1), rf(u[2], -1, 1));
        if (vlen2(p) < 1) return p;
    }
    return p;
}
static v3 random_unit_vector(const rng_ctx* g) { return vunit(random_in_unit_sphere(g)); }   /* vec3.h:143 */

/* ---- scene view over the blob ---------------------------------------------------------------- */
typedef struct rto_scene {
    const rt_flat_header* hdr;
    const rt_flat_prim *prims, *bprims;
    const rt_flat_tri_attr* tri_attrs;
    const rt_flat_xop* xops;
    const rt_flat_chain* chains;
    const rt_flat_medium* media;
    const rt_flat_material* materials;
    const rt_flat_texture* textures;
    const rt_flat_perlin* perlins;
    const rt_flat_image* images;
    const uint8_t* image_bytes;
    void* blob;
} rto_scene;

static size_t pad16(size_t n) { return (n + 15) & ~(size_t)15; }

RTO_API rto_scene* rto_scene_load(const void* blob, size_t size) {
    if (size < sizeof(rt_flat_header)) return NULL;
    rto_scene* s = (rto_scene*)calloc(1, sizeof(rto_scene));
    s->blob = malloc(size);
    memcpy(s->blob, blob, size);
    const uint8_t* p = (const uint8_t*)s->blob;
    s->hdr = (const rt_flat_header*)p;
    if (s->hdr->magic != RT_FLAT_MAGIC) { free(s->blob); free(s); return NULL; }
    p += sizeof(rt_flat_header);
    const rt_flat_header* h = s->hdr;
    s->prims = (const rt_flat_prim*)p; p += pad16(sizeof(rt_flat_prim) * h->n_prims);
    s->bprims = (const rt_flat_prim*)p; p += pad16(sizeof(rt_flat_prim) * h->n_bprims);
    s->tri_attrs = (const rt_flat_tri_attr*)p; p += pad16(sizeof(rt_flat_tri_attr) * h->n_tri_attrs);
    s->xops = (const rt_flat_xop*)p; p += pad16(sizeof(rt_flat_xop) * h->n_xops);
    s->chains = (const rt_flat_chain*)p; p += pad16(sizeof(rt_flat_chain) * h->n_chains);
    s->media = (const rt_flat_medium*)p; p += pad16(sizeof(rt_flat_medium) * h->n_media);
    s->materials = (const rt_flat_material*)p; p += pad16(sizeof(rt_flat_material) * h->n_materials);
    s->textures = (const rt_flat_texture*)p; p += pad16(sizeof(rt_flat_texture) * h->n_textures);
    s->perlins = (const rt_flat_perlin*)p; p += pad16(sizeof(rt_flat_perlin) * h->n_perlins);
    s->images = (const rt_flat_image*)p; p += pad16(sizeof(rt_flat_image) * h->n_images);
    s->image_bytes = p; p += pad16(h->image_bytes);
    if ((size_t)(p - (const uint8_t*)s->blob) > size) { free(s->blob); free(s); return NULL; }
    return s;
}

RTO_API void rto_scene_free(rto_scene* s) { if (s) { free(s->blob); free(s); } }

/* ---- hit_record (hittable.h:8-23) ------------------------------------------------------------- */
typedef struct {
    v3 p, normal;
    uint32_t material;
    float t;
    int front_face;
    float u, v;
    int32_t leaf_id;
} hitrec;

static inline void set_face_normal(hitrec* rec, const rayc* r, v3 outward_normal) {         /* hittable.h:17-22 */
    rec->front_face = vdot(r->d, outward_normal) < 0;
    rec->normal = rec->front_face ? outward_normal : vneg(outward_normal);
}

#define RTO_PI 3.1415926535897932385f                                                        /* common.h:43 */

/* sphere::hit — sphere.h:35-73 (the "second root" repeats the first: SURVEY Q1) */
static int hit_sphere(const rt_flat_prim* pr, const rayc* r, float t_min, float t_max, hitrec* rec) {
    const v3 center = V(pr->f[0], pr->f[1], pr->f[2]);
    const float radius = pr->f[3];
    const v3 oc = vsub(r->o, center);
    const float a = vlen2(r->d);
    const float half_b = vdot(oc, r->d);
    const float c = vlen2(oc) - radius * radius;
    const float discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return 0;
    const float sqrtd = sqrtf(discriminant);
    float root = (-half_b - sqrtd) / a;
    if (root < t_min || t_max < root) {
        root = (-half_b - sqrtd) / a;                                                       /* sphere.h:51 */
        if (root < t_min || t_max < root) return 0;
    }
    rec->t = root;
    rec->p = ray_at(r, rec->t);
    const v3 outward_normal = vdiv(vsub(rec->p, center), radius);
    set_face_normal(rec, r, outward_normal);
    /* get_sphere_uv — sphere.h:19-32 */
    const float theta = acosf(-outward_normal.y);
    const float phi = atan2f(-outward_normal.z, outward_normal.x) + RTO_PI;
    rec->u = phi / (2 * RTO_PI);
    rec->v = theta / RTO_PI;
    return 1;
}

/* moving_sphere::center / ::hit — moving_sphere.h:20-22, :26-59 (u, v not set: SURVEY Q15 -> 0) */
static inline v3 ms_center(const rt_flat_prim* pr, float time) {
    const v3 c0 = V(pr->f[0], pr->f[1], pr->f[2]), c1 = V(pr->f[3], pr->f[4], pr->f[5]);
    const float time0 = pr->f[6], time1 = pr->f[7];
    return vadd(c0, vmul((time - time0) / (time1 - time0), vsub(c1, c0)));
}
static int hit_moving_sphere(const rt_flat_prim* pr, const rayc* r, float t_min, float t_max, hitrec* rec) {
    const float radius = pr->f[8];
    const v3 oc = vsub(r->o, ms_center(pr, r->tm));
    const float a = vlen2(r->d);
    const float half_b = vdot(oc, r->d);
    const float c = vlen2(oc) - radius * radius;
    const float discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return 0;
    const float sqrtd = sqrtf(discriminant);
    float root = (-half_b - sqrtd) / a;
    if (root < t_min || t_max < root) {
        root = (-half_b - sqrtd) / a;
        if (root < t_min || t_max < root) return 0;
    }
    rec->t = root;
    rec->p = ray_at(r, rec->t);
    const v3 outward_normal = vdiv(vsub(rec->p, ms_center(pr, r->tm)), radius);
    set_face_normal(rec, r, outward_normal);
    rec->u = 0; rec->v = 0;
    return 1;
}

/* xy_rect / xz_rect / yz_rect ::hit — aarect.h:63-90, :94-120, :124-150 */
static int hit_rect(const rt_flat_prim* pr, const rayc* r, float t_min, float t_max, hitrec* rec) {
    const float a0 = pr->f[0], a1 = pr->f[1], b0 = pr->f[2], b1 = pr->f[3], k = pr->f[4];
    float t, a, b;
    v3 outward_normal;
    if (pr->type == RT_PRIM_XY_RECT) {
        t = (k - r->o.z) / r->d.z;
        if (t < t_min || t > t_max) return 0;
        a = r->o.x + t * r->d.x; b = r->o.y + t * r->d.y;
        if (a < a0 || a > a1 || b < b0 || b > b1) return 0;
        rec->u = (a - a0) / (a1 - a0); rec->v = (b - b0) / (b1 - b0);
        outward_normal = V(0, 0, 1);
    } else if (pr->type == RT_PRIM_XZ_RECT) {
        t = (k - r->o.y) / r->d.y;
        if (t < t_min || t > t_max) return 0;
        a = r->o.x + t * r->d.x; b = r->o.z + t * r->d.z;
        if (a < a0 || a > a1 || b < b0 || b > b1) return 0;
        rec->u = (a - a0) / (a1 - a0); rec->v = (b - b0) / (b1 - b0);
        outward_normal = V(0, 1, 0);
    } else { /* yz: a = y, b = z; u from y, v from z (aarect.h:139-140) */
        t = (k - r->o.x) / r->d.x;
        if (t < t_min || t > t_max) return 0;
        b = r->o.z + t * r->d.z; a = r->o.y + t * r->d.y;
        if (b < b0 || b > b1 || a < a0 || a > a1) return 0;
        rec->v = (b - b0) / (b1 - b0); rec->u = (a - a0) / (a1 - a0);
        outward_normal = V(1, 0, 0);
    }
    rec->t = t;
    set_face_normal(rec, r, outward_normal);
    rec->p = ray_at(r, t);
    return 1;
}

/* triangle::hit — triangle.h:120-178 (Möller–Trumbore, eps 1e-7, un-normalised normals: SURVEY Q13, Q14) */
static int hit_triangle(const rto_scene* s, const rt_flat_prim* pr, const rayc* r, float t_min, float t_max, hitrec* rec) {
    const float epsilon = 0.0000001f;
    const v3 vertex0 = V(pr->f[0], pr->f[1], pr->f[2]), e0 = V(pr->f[3], pr->f[4], pr->f[5]), e1 = V(pr->f[6], pr->f[7], pr->f[8]);
    const v3 h = vcross(r->d, e1);
    const float a = vdot(e0, h);
    if (a > -epsilon && a < epsilon) return 0;
    const float f = 1.0f / a;
    const v3 sv = vsub(r->o, vertex0);
    const float u = f * vdot(sv, h);
    if (u < 0.0f || u > 1.0f) return 0;
    const v3 q = vcross(sv, e0);
    const float v = f * vdot(r->d, q);
    if ((v < 0.0f) | (u + v > 1.0f)) return 0;
    const float t = f * vdot(e1, q);
    if (t < t_min || t > t_max || t < epsilon) return 0;
    rec->t = t;
    rec->p = ray_at(r, t);
    uint32_t ai; memcpy(&ai, &pr->f[9], 4);
    const rt_flat_tri_attr* at = &s->tri_attrs[ai];
    /* barycentric_coords — triangle.h:101-111 */
    const v3 v2 = vsub(rec->p, vertex0);
    const float d20 = vdot(v2, e0), d21 = vdot(v2, e1);
    const float B0 = (at->d11 * d20 - at->d01 * d21) * at->inv_denom;
    const float B1 = (at->d00 * d21 - at->d01 * d20) * at->inv_denom;
    const float B2 = 1.0f - B0 - B1;
    /* barycentric_interp — triangle.h:113-116: out = B2*i0 + B0*i1 + B1*i2 */
    rec->u = B2 * at->uv[0] + B0 * at->uv[2] + B1 * at->uv[4];
    rec->v = B2 * at->uv[1] + B0 * at->uv[3] + B1 * at->uv[5];
    if (!at->vertex_normals) {
        set_face_normal(rec, r, vcross(e1, e0));
    } else {
        v3 n;
        n.x = B2 * at->n[0] + B0 * at->n[3] + B1 * at->n[6];
        n.y = B2 * at->n[1] + B0 * at->n[4] + B1 * at->n[7];
        n.z = B2 * at->n[2] + B0 * at->n[5] + B1 * at->n[8];
        set_face_normal(rec, r, n);
    }
    return 1;
}

static int hit_leaf(const rto_scene* s, const rt_flat_prim* pr, const rayc* r, float t_min, float t_max, hitrec* rec) {
    int ok;
    switch (pr->type) {
    case RT_PRIM_SPHERE: ok = hit_sphere(pr, r, t_min, t_max, rec); break;
    case RT_PRIM_MOVING_SPHERE: ok = hit_moving_sphere(pr, r, t_min, t_max, rec); break;
    case RT_PRIM_TRIANGLE: ok = hit_triangle(s, pr, r, t_min, t_max, rec); break;
    default: ok = hit_rect(pr, r, t_min, t_max, rec); break;
    }
    if (ok) { rec->material = pr->material; rec->leaf_id = (int32_t)pr->leaf_id; }
    return ok;
}

/* A leaf-like object below a chain of translate / rotate_y wrappers.  `inner` hits the object in its own frame. */
typedef int (*inner_fn)(const rto_scene*, const void* obj, const rayc*, float, float, hitrec*, const rng_ctx*);

static int hit_chain(const rto_scene* s, const rt_flat_xop* ops, uint32_t n_ops, inner_fn inner, const void* obj,
                     const rayc* r, float t_min, float t_max, hitrec* rec, const rng_ctx* g) {
    if (n_ops == 0) return inner(s, obj, r, t_min, t_max, rec, g);
    const rt_flat_xop* op = ops;
    if (op->kind == RT_XOP_TRANSLATE) {                                                     /* translate::hit — hittable.h:37-49 */
        const v3 offset = V(op->a, op->b, op->c);
        rayc moved = *r;
        moved.o = vsub(r->o, offset);
        if (!hit_chain(s, ops + 1, n_ops - 1, inner, obj, &moved, t_min, t_max, rec, g)) return 0;
        rec->p = vadd(rec->p, offset);
        set_face_normal(rec, &moved, rec->normal);
        return 1;
    } else {                                                                                /* rotate_y::hit — hittable.h:112-143 */
        const float sin_theta = op->a, cos_theta = op->b;
        rayc rot = *r;
        rot.o.x = cos_theta * r->o.x - sin_theta * r->o.z;
        rot.o.z = sin_theta * r->o.x + cos_theta * r->o.z;
        rot.d.x = cos_theta * r->d.x - sin_theta * r->d.z;
        rot.d.z = sin_theta * r->d.x + cos_theta * r->d.z;
        if (!hit_chain(s, ops + 1, n_ops - 1, inner, obj, &rot, t_min, t_max, rec, g)) return 0;
        v3 p = rec->p, n = rec->normal;
        p.x = cos_theta * rec->p.x + sin_theta * rec->p.z;
        p.z = -sin_theta * rec->p.x + cos_theta * rec->p.z;
        n.x = cos_theta * rec->normal.x + sin_theta * rec->normal.z;
        n.z = -sin_theta * rec->normal.x + cos_theta * rec->normal.z;
        rec->p = p;
        set_face_normal(rec, &rot, n);          /* against the ROTATED ray, as the reference does */
        return 1;
    }
}

static int inner_leaf(const rto_scene* s, const void* obj, const rayc* r, float t_min, float t_max, hitrec* rec, const rng_ctx* g) {
    (void)g;
    return hit_leaf(s, (const rt_flat_prim*)obj, r, t_min, t_max, rec);
}

static int hit_prim(const rto_scene* s, const rt_flat_prim* pr, const rayc* r, float t_min, float t_max, hitrec* rec) {
    const rt_flat_chain* c = &s->chains[pr->chain];
    return hit_chain(s, s->xops + c->first, c->count, inner_leaf, pr, r, t_min, t_max, rec, NULL);
}

/* hittable_list::hit over a primitive range — hittable_list.h:23-39 (later object wins ties) */
static int hit_prim_list(const rto_scene* s, const rt_flat_prim* list, uint32_t n, const rayc* r, float t_min, float t_max, hitrec* rec) {
    hitrec temp;
    int hit_anything = 0;
    float closest_so_far = t_max;
    for (uint32_t i = 0; i < n; i++)
        if (hit_prim(s, &list[i], r, t_min, closest_so_far, &temp)) { hit_anything = 1; closest_so_far = temp.t; *rec = temp; }
    return hit_anything;
}

/* constant_medium::hit — constant_medium.h:34-70; the free-flight draw is stream MEDIUM, block = medium index */
static int inner_medium(const rto_scene* s, const void* obj, const rayc* r, float t_min, float t_max, hitrec* rec, const rng_ctx* g) {
    const rt_flat_medium* m = (const rt_flat_medium*)obj;
    const rt_flat_prim* b = s->bprims + m->bprim_first;
    hitrec rec1, rec2;
    if (!hit_prim_list(s, b, m->bprim_count, r, -INFINITY, INFINITY, &rec1)) return 0;
    if (!hit_prim_list(s, b, m->bprim_count, r, rec1.t + 0.0001f, INFINITY, &rec2)) return 0;
    if (rec1.t < t_min) rec1.t = t_min;
    if (rec2.t > t_max) rec2.t = t_max;
    if (rec1.t >= rec2.t) return 0;
    if (rec1.t < 0) rec1.t = 0;
    const float ray_length = vlen(r->d);
    const float distance_inside_boundary = (rec2.t - rec1.t) * ray_length;
    float xi;
    if (g->scripted_xi) xi = *g->scripted_xi;
    else { float u[4]; draw4(g, RTO_STREAM_MEDIUM, (uint32_t)(m - s->media), u); xi = u[0]; }
    const float hit_distance = m->neg_inv_density * logf(xi);
    if (hit_distance > distance_inside_boundary) return 0;
    rec->t = rec1.t + hit_distance / ray_length;
    rec->p = ray_at(r, rec->t);
    rec->normal = V(1, 0, 0);
    rec->front_face = 1;
    rec->material = m->material;
    rec->leaf_id = (int32_t)m->leaf_id;
    rec->u = 0; rec->v = 0;
    return 1;
}

/* The world query color_f makes (render.h:63): primitives and media merged in leaf-id order, i.e. in the
 * order a top-level hittable_list would visit them, with closest_so_far shrinking (hittable_list.h:29-35).
 * media_enabled = 0 skips media (first-hit queries without xi). */
static int world_hit(const rto_scene* s, const rayc* r, float t_min, float t_max, hitrec* rec, const rng_ctx* g, int media_enabled) {
    hitrec temp;
    int hit_anything = 0;
    float closest_so_far = t_max;
    uint32_t ip = 0, im = 0;
    const uint32_t np = s->hdr->n_prims, nm = media_enabled ? s->hdr->n_media : 0;
    while (ip < np || im < nm) {
        int ok;
        if (im >= nm || (ip < np && s->prims[ip].leaf_id < s->media[im].leaf_id)) {
            ok = hit_prim(s, &s->prims[ip], r, t_min, closest_so_far, &temp);
            ip++;
        } else {
            const rt_flat_medium* m = &s->media[im];
            const rt_flat_chain* c = &s->chains[m->chain];
            ok = hit_chain(s, s->xops + c->first, c->count, inner_medium, m, r, t_min, closest_so_far, &temp, g);
            im++;
        }
        if (ok) { hit_anything = 1; closest_so_far = temp.t; *rec = temp; }
    }
    return hit_anything;
}

/* rays: n x 7 floats (origin, direction, time); xi: NULL (media skipped) or n scripted free-flight draws */
RTO_API void rto_first_hit(const rto_scene* s, const float* rays, int n, float t_min, const float* xi, int32_t* out_id, float* out_t) {
    for (int i = 0; i < n; i++) {
        const float* q = rays + 7 * (size_t)i;
        rayc r; r.o = V(q[0], q[1], q[2]); r.d = V(q[3], q[4], q[5]); r.tm = q[6];
        rng_ctx g = {0, 0, 0, 0, xi ? xi + i : NULL};
        hitrec rec;
        if (world_hit(s, &r, t_min, INFINITY, &rec, &g, xi != NULL)) { out_id[i] = rec.leaf_id; out_t[i] = rec.t; }
        else { out_id[i] = -1; out_t[i] = INFINITY; }
    }
}

/* ---- perlin.h ---------------------------------------------------------------------------------- */
static float perlin_noise(const rt_flat_perlin* pn, v3 p) {                                   /* perlin.h:85-112 */
    const float u = p.x - floorf(p.x), v = p.y - floorf(p.y), w = p.z - floorf(p.z);
    const int i = (int)floorf(p.x), j = (int)floorf(p.y), k = (int)floorf(p.z);
    v3 c[2][2][2];
    for (int di = 0; di < 2; di++)
        for (int dj = 0; dj < 2; dj++)
            for (int dk = 0; dk < 2; dk++) {
                const int ix = pn->perm[0][(i + di) & 255] ^ pn->perm[1][(j + dj) & 255] ^ pn->perm[2][(k + dk) & 255];
                c[di][dj][dk] = V(pn->ranvec[3 * ix], pn->ranvec[3 * ix + 1], pn->ranvec[3 * ix + 2]);
            }
    /* perlin_interp — perlin.h:37-57 */
    const float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
    float accum = 0.0f;
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++)
            for (int cc = 0; cc < 2; cc++) {
                const v3 weight_v = V(u - a, v - b, w - cc);
                accum += (a * uu + (1 - a) * (1 - uu)) * (b * vv + (1 - b) * (1 - vv)) * (cc * ww + (1 - cc) * (1 - ww)) *
                         vdot(c[a][b][cc], weight_v);
            }
    return accum;
}

static float perlin_turb(const rt_flat_perlin* pn, v3 p, int depth) {                         /* perlin.h:114-126 (double accumulator) */
    double accum = 0.0, weight = 1.0;
    v3 temp_p = p;
    for (int i = 0; i < depth; i++) {
        accum += weight * perlin_noise(pn, temp_p);
        weight *= 0.5;
        temp_p = vmul(2, temp_p);
    }
    return (float)fabs(accum);
}

/* ---- texture.h --------------------------------------------------------------------------------- */
static inline float clamp_d(float x, float mn, float mx) { if (x < mn) return mn; if (x > mx) return mx; return x; } /* common.h:62 */

static v3 texture_value(const rto_scene* s, uint32_t ti, float u, float v, v3 p) {
    const rt_flat_texture* t = &s->textures[ti];
    switch (t->type) {
    case RT_TEX_SOLID: return V(t->rgb[0], t->rgb[1], t->rgb[2]);                             /* texture.h:17 */
    case RT_TEX_CHECKER: {                                                                    /* texture.h:37-45 */
        const float sines = sinf(10 * p.x) * sinf(10 * p.y) * sinf(10 * p.z);
        return texture_value(s, sines < 0 ? t->b : t->a, u, v, p);
    }
    case RT_TEX_NOISE: {                                                                      /* texture.h:57-59 */
        const float g = 0.5f * (1.0f + perlin_noise(&s->perlins[t->a], vmul(t->scale, p)));
        return V(g, g, g);
    }
    case RT_TEX_TURB: {                                                                       /* texture.h:74-76 */
        const float g = perlin_turb(&s->perlins[t->a], vmul(t->scale, p), t->depth);
        return V(g, g, g);
    }
    case RT_TEX_MARBLE: {                                                                     /* texture.h:88-90 */
        const float g = 0.5f * (1 + sinf(t->scale * p.z + 10 * perlin_turb(&s->perlins[t->a], vmul(t->scale, p), 7)));
        return V(g, g, g);
    }
    default: {                                                                                /* image_texture::value — texture.h:143-163 */
        const rt_flat_image* im = &s->images[t->a];
        const float uu = clamp_d(u, 0.0f, 1.0f);
        const double vv = 1.0 - clamp_d(v, 0.0f, 1.0f);
        int i = (int)(uu * im->width), j = (int)(vv * im->height);
        if (i >= (int)im->width) i = (int)im->width - 1;
        if (j >= (int)im->height) j = (int)im->height - 1;
        const float color_scale = 1.0f / 255.0f;
        const uint8_t* px = s->image_bytes + im->offset + (size_t)j * im->bytes_per_pixel * im->width + (size_t)i * im->bytes_per_pixel;
        return V(color_scale * px[0], color_scale * px[1], color_scale * px[2]);
    }
    }
}

/* ---- material.h -------------------------------------------------------------------------------- */
static v3 material_emitted(const rto_scene* s, const rt_flat_material* m, float u, float v, v3 p) {
    if (m->type == RT_MAT_DIFFUSE_LIGHT) return texture_value(s, m->texture, u, v, p);        /* material.h:119-121 */
    return V(0, 0, 0);                                                                        /* material.h:11-13 */
}

static int material_scatter(const rto_scene* s, const rt_flat_material* m, const rayc* ray_in, const hitrec* rec,
                            v3* attenuation, rayc* scattered, const rng_ctx* g) {
    switch (m->type) {
    case RT_MAT_LAMBERTIAN: {                                                                 /* material.h:25-35 */
        v3 scatter_direction = vadd(rec->normal, random_unit_vector(g));
        if (vnear_zero(scatter_direction)) scatter_direction = rec->normal;
        scattered->o = rec->p; scattered->d = scatter_direction; scattered->tm = ray_in->tm;
        *attenuation = texture_value(s, m->texture, rec->u, rec->v, rec->p);
        return 1;
    }
    case RT_MAT_METAL: {                                                                      /* material.h:50-55 */
        const v3 reflected = vreflect(vunit(ray_in->d), rec->normal);
        scattered->o = rec->p; scattered->d = vadd(reflected, vmul(m->param, random_in_unit_sphere(g))); scattered->tm = ray_in->tm;
        *attenuation = texture_value(s, m->texture, rec->u, rec->v, rec->p);
        return vdot(scattered->d, rec->normal) > 0;
    }
    case RT_MAT_DIELECTRIC: {                                                                 /* material.h:64-95 */
        const float ir = m->param;
        *attenuation = V(1.0f, 1.0f, 1.0f);
        const float refraction_ratio = rec->front_face ? (1.0f / ir) : ir;
        const v3 unit_direction = vunit(ray_in->d);
        const float cos_theta = fminf(vdot(vneg(unit_direction), rec->normal), 1.0f);
        const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
        const int cannot_refract = refraction_ratio * sin_theta > 1.0f;
        int reflect_it = cannot_refract;
        if (!reflect_it) {
            /* reflectance — material.h:98-103 (x^5 as four multiplies) */
            const float sqrt_r0 = (1.0f - refraction_ratio) / (1.0f + refraction_ratio);
            const float r0 = sqrt_r0 * sqrt_r0;
            const float x = 1.0f - cos_theta;
            const float refl = r0 + (1.0f - r0) * (x * x * x * x * x);
            float u[4];
            draw4(g, RTO_STREAM_SCATTER, 0, u);
            reflect_it = refl > u[0];
        }
        const v3 direction = reflect_it ? vreflect(unit_direction, rec->normal) : vrefract(unit_direction, rec->normal, refraction_ratio);
        scattered->o = rec->p; scattered->d = direction; scattered->tm = ray_in->tm;
        return 1;
    }
    case RT_MAT_DIFFUSE_LIGHT: return 0;                                                      /* material.h:115-117 */
    default: {                                                                                /* isotropic — material.h:133-137 */
        scattered->o = rec->p; scattered->d = random_in_unit_sphere(g); scattered->tm = ray_in->tm;
        *attenuation = texture_value(s, m->texture, rec->u, rec->v, rec->p);
        return 1;
    }
    }
}

/* ---- camera::get_ray — camera.h:49-58; draws: block 0 = (jitter u, jitter v, time, -), lens iteration k
 * uses words (2k mod 4, 2k+1 mod 4) of block 1 + k/2 ------------------------------------------------ */
static void camera_sample(const rto_scene* s, const rng_ctx* g, int i, int j, int max_x, int max_y, rayc* out) {
    const rt_flat_camera* c = &s->hdr->camera;
    float b0[4];
    draw4(g, RTO_STREAM_CAMERA, 0, b0);
    const float su = (float)(i + b0[0]) / max_x;                                              /* render.h:105 */
    const float sv = (float)(j + b0[1]) / max_y;                                              /* render.h:106 */
    v3 pd = V(0, 0, 0);
    for (uint32_t k = 0; k < RTO_MAX_REJECT; k++) {                                           /* random_in_unit_disk — vec3.h:136-141 */
        float u[4];
        draw4(g, RTO_STREAM_CAMERA, 1 + k / 2, u);
        pd = V(rf(u[2 * (k & 1)], -1, 1), rf(u[2 * (k & 1) + 1], -1, 1), 0);
        if (vlen2(pd) < 1) break;
    }
    const v3 rd = vmul(c->lens_radius, pd);
    const v3 cu = V(c->u[0], c->u[1], c->u[2]), cv = V(c->v[0], c->v[1], c->v[2]);
    const v3 offset = vadd(vmul(rd.x, cu), vmul(rd.y, cv));
    const v3 origin = V(c->origin[0], c->origin[1], c->origin[2]);
    const v3 llc = V(c->lower_left_corner[0], c->lower_left_corner[1], c->lower_left_corner[2]);
    const v3 hor = V(c->horizontal[0], c->horizontal[1], c->horizontal[2]), ver = V(c->vertical[0], c->vertical[1], c->vertical[2]);
    out->o = vadd(origin, offset);
    out->d = vsub(vsub(vadd(vadd(llc, vmul(su, hor)), vmul(sv, ver)), origin), offset);
    out->tm = rf(b0[2], c->time0, c->time1);
}

/* color_f — render.h:55-81.  Also counts path segments (world queries). */
static v3 color_f(const rto_scene* s, rayc r, rng_ctx* g, int depth, v3 background, uint64_t* rays) {
    rayc cur_ray = r;
    v3 cur_attenuation = V(1, 1, 1);
    for (int i = 0; i < depth; i++) {
        hitrec rec;
        g->bounce = (uint32_t)i;
        (*rays)++;
        if (!world_hit(s, &cur_ray, 0.001f, INFINITY, &rec, g, 1)) return vmulv(cur_attenuation, background);
        rayc scattered;
        v3 attenuation;
        const rt_flat_material* m = &s->materials[rec.material];
        const v3 emitted = material_emitted(s, m, rec.u, rec.v, rec.p);
        if (material_scatter(s, m, &cur_ray, &rec, &attenuation, &scattered, g)) {
            cur_attenuation = vmulv(cur_attenuation, attenuation);
            cur_ray = scattered;
        } else {
            return vmulv(cur_attenuation, emitted);
        }
    }
    return V(0, 0, 0);
}

/* Radiance of single samples (debug / per-sample parity): for each k, pixel index pix[k] and sample smp[k]. */
RTO_API void rto_sample_radiance(const rto_scene* s, int width, int height, int max_depth, uint32_t seed,
                                 const uint32_t* pix, const uint32_t* smp, int n, float* out_rgb, uint32_t* out_segments) {
    const v3 bg = V(s->hdr->background[0], s->hdr->background[1], s->hdr->background[2]);
    for (int k = 0; k < n; k++) {
        rng_ctx g = {pix[k], smp[k], 0, seed, NULL};
        rayc r;
        camera_sample(s, &g, (int)(pix[k] % (uint32_t)width), (int)(pix[k] / (uint32_t)width), width, height, &r);
        uint64_t segs = 0;
        const v3 c = color_f(s, r, &g, max_depth, bg, &segs);
        out_rgb[3 * k] = c.x; out_rgb[3 * k + 1] = c.y; out_rgb[3 * k + 2] = c.z;
        if (out_segments) out_segments[k] = (uint32_t)segs;
    }
}

/* render — render.h:94-113 restated with Philox keying.  Rows [row0, row1) of samples [s0, s0+ns);
 * out: width*height*3 SUMS over the samples (not means), row 0 = bottom.  Returns path segments traced.
 * Single-threaded; the caller (tests / bench) fans rows out over processes or threads. */
RTO_API uint64_t rto_render_rows(const rto_scene* s, int width, int height, int row0, int row1, uint32_t s0, uint32_t ns,
                                 int max_depth, uint32_t seed, float* out_sums) {
    const v3 bg = V(s->hdr->background[0], s->hdr->background[1], s->hdr->background[2]);
    uint64_t rays = 0;
    for (int j = row0; j < row1; j++)
        for (int i = 0; i < width; i++) {
            const uint32_t pixel = (uint32_t)j * (uint32_t)width + (uint32_t)i;
            v3 col = V(0, 0, 0);
            for (uint32_t sm = s0; sm < s0 + ns; sm++) {
                rng_ctx g = {pixel, sm, 0, seed, NULL};
                rayc r;
                camera_sample(s, &g, i, j, width, height, &r);
                col = vadd(col, color_f(s, r, &g, max_depth, bg, &rays));
            }
            out_sums[3 * (size_t)pixel] = col.x; out_sums[3 * (size_t)pixel + 1] = col.y; out_sums[3 * (size_t)pixel + 2] = col.z;
        }
    return rays;
}
