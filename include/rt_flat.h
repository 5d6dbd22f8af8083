/* rt_flat.h — the flattened scene: what rt_scene_flatten() produces from the recorded scene graph,
 * what rt_scene_export() serialises, what the CUDA side uploads to HBM and what the CPU oracle
 * (oracle/rt_oracle.c) consumes.  Plain C, fixed-size records, every record a multiple of 16 bytes
 * so the device can fetch them with 128-bit loads.
 *
 * Flattening resolves the reference's object graph (hittable_list / bvh_node / triangle_mesh / box /
 * translate / rotate_y / constant_medium) into
 *   - surface primitives  (rt_flat_prim, `prims`)      — members of the one LBVH,
 *   - instance chains     (rt_flat_chain + rt_flat_xop) — the translate / rotate_y wrappers above a leaf,
 *                                                         outermost first, applied to the RAY exactly as
 *                                                         hittable.h:37-49 and :112-143 do,
 *   - media               (rt_flat_medium)              — constant_medium objects, each with its boundary
 *                                                         primitives (`bprims`, chains relative to the medium),
 *   - materials, textures, Perlin tables, images, camera.
 * Leaf ids: depth-first pre-order over the world graph; every leaf primitive consumes one id, a medium
 * consumes one id for itself and then one per boundary leaf (boundary ids are never reported).
 */
#ifndef RT_FLAT_H
#define RT_FLAT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_FLAT_MAGIC 0x52544631u /* "RTF1" */

enum rt_prim_type {
    RT_PRIM_SPHERE = 0,        /* f[0..2] center, f[3] radius                                    (sphere.h) */
    RT_PRIM_MOVING_SPHERE = 1, /* f[0..2] center0, f[3..5] center1, f[6] time0, f[7] time1, f[8] radius (moving_sphere.h) */
    RT_PRIM_XY_RECT = 2,       /* f[0] x0, f[1] x1, f[2] y0, f[3] y1, f[4] k                     (aarect.h:8) */
    RT_PRIM_XZ_RECT = 3,       /* f[0] x0, f[1] x1, f[2] z0, f[3] z1, f[4] k                     (aarect.h:27) */
    RT_PRIM_YZ_RECT = 4,       /* f[0] y0, f[1] y1, f[2] z0, f[3] z1, f[4] k                     (aarect.h:45) */
    RT_PRIM_TRIANGLE = 5       /* f[0..2] vertex0, f[3..5] v0 = vertex1-vertex0, f[6..8] v1 = vertex2-vertex0,
                                  f[9] = tri_attr index (uint32 bits)                            (triangle.h) */
};

typedef struct rt_flat_prim { /* 64 B */
    uint32_t type;     /* rt_prim_type */
    uint32_t material; /* index into materials */
    uint32_t chain;    /* index into chains; 0 = identity */
    uint32_t leaf_id;  /* DFS id reported by first-hit queries */
    float f[12];
} rt_flat_prim;

typedef struct rt_flat_tri_attr { /* 96 B — only read when shading the closest hit (triangle.h:160-173) */
    float uv[6];        /* u_0 v_0 u_1 v_1 u_2 v_2 */
    float n[9];         /* normal0, normal1, normal2 (NOT normalised) */
    float d00, d01, d11, inv_denom;
    uint32_t vertex_normals;
    uint32_t pad[4];
} rt_flat_tri_attr;

enum rt_xop_kind { RT_XOP_TRANSLATE = 0 /* a,b,c = offset */, RT_XOP_ROTATE_Y = 1 /* a = sin_theta, b = cos_theta */ };
typedef struct rt_flat_xop { uint32_t kind; float a, b, c; } rt_flat_xop;     /* 16 B */
typedef struct rt_flat_chain { uint32_t first, count, pad0, pad1; } rt_flat_chain; /* 16 B; ops [first, first+count), outermost first */

typedef struct rt_flat_medium { /* 32 B — constant_medium.h:8-31 */
    uint32_t chain;        /* wrappers above the medium */
    uint32_t bprim_first;  /* boundary primitives: bprims[bprim_first .. +bprim_count) */
    uint32_t bprim_count;
    uint32_t material;     /* the isotropic phase function */
    uint32_t leaf_id;
    float neg_inv_density; /* -1/density, as stored by the reference */
    uint32_t pad[2];
} rt_flat_medium;

enum rt_material_type { RT_MAT_LAMBERTIAN = 0, RT_MAT_METAL = 1, RT_MAT_DIELECTRIC = 2, RT_MAT_DIFFUSE_LIGHT = 3, RT_MAT_ISOTROPIC = 4 };
typedef struct rt_flat_material { uint32_t type, texture; float param /* fuzz | ir */; uint32_t pad; } rt_flat_material; /* 16 B */

enum rt_texture_type { RT_TEX_SOLID = 0, RT_TEX_CHECKER = 1, RT_TEX_NOISE = 2, RT_TEX_TURB = 3, RT_TEX_MARBLE = 4, RT_TEX_IMAGE = 5 };
typedef struct rt_flat_texture { /* 32 B */
    uint32_t type;
    uint32_t a;      /* checker: even texture | noise/turb/marble: perlin table | image: image index */
    uint32_t b;      /* checker: odd texture */
    float scale;     /* noise/turb/marble */
    float rgb[3];    /* solid */
    int32_t depth;   /* turb */
} rt_flat_texture;

typedef struct rt_flat_perlin { float ranvec[256 * 3]; int32_t perm[3][256]; } rt_flat_perlin; /* 6144 B, perlin.h:9-15 */
typedef struct rt_flat_image { uint32_t width, height, bytes_per_pixel, offset; } rt_flat_image; /* offset into the byte pool */

typedef struct rt_flat_camera { /* camera.h:5-16 */
    float origin[3], lower_left_corner[3], horizontal[3], vertical[3], u[3], v[3], w[3];
    float lens_radius, time0, time1;
} rt_flat_camera; /* 96 B */

/* Export blob = this header, then the arrays in the order of the count fields, each array starting at a
 * 16-byte boundary: prims, bprims, tri_attrs, xops, chains, media, materials, textures, perlins, images,
 * image byte pool. */
typedef struct rt_flat_header { /* 176 B */
    uint32_t magic, num_leaf_ids;
    uint32_t n_prims, n_bprims, n_tri_attrs, n_xops, n_chains, n_media, n_materials, n_textures, n_perlins, n_images;
    uint32_t image_bytes, pad0, pad1, pad2;
    float background[3];
    float pad3;
    rt_flat_camera camera;
} rt_flat_header;

#ifdef __cplusplus
}
#endif
#endif /* RT_FLAT_H */
