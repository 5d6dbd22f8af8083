/* rt_program.h — "scene program": a flat uint32 word stream that serialises scene-description
 * calls (one record per constructor call of the reference's scene API).
 *
 * Used three ways:
 *   - rt_scene_replay() (include/rt_b200.h) replays a program into a scene through the same
 *     code path as the individual rt_* constructor calls;
 *   - oracle/_ref exports the object graph built by the reference's own constructors as a program,
 *     so the product and the reference see bit-identical scenes;
 *   - tests/golden/ stores small programs as fixtures.
 *
 * Layout: repeated records [opcode][payload words...]. Floats are stored as their IEEE-754 bit
 * pattern. Every record that creates a texture / material / hittable / perlin table / image gets
 * the next index of its kind (0-based, creation order); later records refer to those indices.
 * All payloads have a fixed length except RTP_PERLIN, RTP_IMAGE, RTP_LIST, RTP_BVH (length words
 * inside the payload, documented below).
 */
#ifndef RT_PROGRAM_H
#define RT_PROGRAM_H

#include <stdint.h>

#define RTP_MAGIC 0x52545031u /* "RTP1", first word of every program */

enum rtp_opcode {
    RTP_END = 0,
    /* --- resources --- */
    RTP_PERLIN = 1,        /* 768 floats ranvec[256][3], 768 ints perm_x|perm_y|perm_z (perlin.h:63-76) */
    RTP_IMAGE = 2,         /* w, h, bytes_per_pixel, nbytes, then ceil(nbytes/4) words of pixel bytes (texture.h:125) */
    /* --- textures (texture.h) --- */
    RTP_TEX_SOLID = 10,    /* r g b */
    RTP_TEX_CHECKER = 11,  /* even_tex odd_tex */
    RTP_TEX_NOISE = 12,    /* perlin scale */
    RTP_TEX_TURB = 13,     /* perlin scale depth */
    RTP_TEX_MARBLE = 14,   /* perlin scale */
    RTP_TEX_IMAGE = 15,    /* image */
    /* --- materials (material.h) --- */
    RTP_MAT_LAMBERTIAN = 20,    /* tex */
    RTP_MAT_METAL = 21,         /* tex fuzz */
    RTP_MAT_DIELECTRIC = 22,    /* ir */
    RTP_MAT_DIFFUSE_LIGHT = 23, /* tex */
    RTP_MAT_ISOTROPIC = 24,     /* tex */
    /* --- hittables --- */
    RTP_SPHERE = 30,        /* cx cy cz r mat                       (sphere.h:13) */
    RTP_MOVING_SPHERE = 31, /* c0xyz c1xyz t0 t1 r mat              (moving_sphere.h:13) */
    RTP_XY_RECT = 32,       /* x0 x1 y0 y1 k mat                    (aarect.h:13) */
    RTP_XZ_RECT = 33,       /* x0 x1 z0 z1 k mat                    (aarect.h:32) */
    RTP_YZ_RECT = 34,       /* y0 y1 z0 z1 k mat                    (aarect.h:50) */
    RTP_BOX = 35,           /* p0xyz p1xyz mat                      (box.h:14) */
    RTP_TRIANGLE = 36,      /* v0 v1 v2 (9), u0 v0 u1 v1 u2 v2, mat (triangle.h:19) */
    RTP_TRIANGLE_N = 37,    /* v0 v1 v2 (9), n0 n1 n2 (9), uv (6), mat (triangle.h:34) */
    RTP_TRANSLATE = 38,     /* child dx dy dz                       (hittable.h:35) */
    RTP_ROTATE_Y = 39,      /* child angle_degrees                  (hittable.h:78) */
    RTP_ROTATE_Y_SC = 40,   /* child sin_theta cos_theta            (exact form, used by exporters) */
    RTP_MEDIUM = 41,        /* boundary density tex                 (constant_medium.h:15) */
    RTP_MEDIUM_NID = 42,    /* boundary neg_inv_density tex         (exact form, used by exporters) */
    RTP_LIST = 43,          /* n, child[n]                          (hittable_list.h:12) */
    RTP_BVH = 44,           /* n, child[n]   — grouping hint only   (bvh.h:136; triangle_mesh.h:21) */
    /* --- scene --- */
    RTP_WORLD = 50,         /* hittable */
    RTP_BACKGROUND = 51,    /* r g b                                (scenes.h:69-77) */
    RTP_CAMERA = 52,        /* lookfrom(3) lookat(3) vup(3) vfov aspect aperture focus t0 t1 (camera.h:18) */
    RTP_CAMERA_RAW = 53     /* origin(3) lower_left(3) horizontal(3) vertical(3) u(3) v(3) w(3) lens_radius t0 t1 */
};

#endif /* RT_PROGRAM_H */
