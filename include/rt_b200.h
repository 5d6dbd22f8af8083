/* rt_b200.h — C-ABI of the B200-native path-tracing core (librt_b200.so).
 *
 * The reference (daRoyalCacti/Raytracing_GPU) has no library boundary: its de-facto operator interface is
 *     void draw(scene& curr_scene, render_settings settings)            render.h:118
 * fed by the device-side constructors that scenes.h calls.  Each entry point below names the
 * reference interface it replaces.  Plain pointers and sizes only; no torch / C++ types.
 *
 * Conventions: every function returns >= 0 on success (a resource index where one is created) and a
 * negative rt_status on failure; rt_last_error() returns the message of the calling thread's last failure.
 * The library never calls exit() (the reference's checkCudaErrors does: common.h:30-38).
 * A scene is not thread-safe; distinct scenes may be used from distinct threads.
 * Scene recording, flattening and export are host-only and work without a GPU; rt_scene_commit,
 * rt_first_hit and rt_render* need a CUDA device and fail with RT_ERR_CUDA when there is none —
 * there is no CPU fallback.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>
#include "rt_flat.h"
#include "rt_program.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_scene rt_scene;

enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,    /* bad argument / index / malformed program */
    RT_ERR_STATE = -2,      /* call order (e.g. render before commit) */
    RT_ERR_CUDA = -3,       /* CUDA runtime error or no device */
    RT_ERR_UNSUPPORTED = -4 /* graph shape outside what the flattener supports */
};

const char* rt_last_error(void);
const char* rt_version(void);

/* sizeof() of a public struct as this library was compiled (no counterpart in the reference): lets a binding that mirrors the
 * structs by hand (ctypes, cgo, JNI) check its layout at load time instead of corrupting memory.  Unknown id: 0. */
enum rt_struct_id { RT_STRUCT_RENDER_SETTINGS = 0, RT_STRUCT_RENDER_STATS = 1, RT_STRUCT_BUILD_SETTINGS = 2, RT_STRUCT_BUILD_STATS = 3,
                    RT_STRUCT_MULTI_SETTINGS = 4, RT_STRUCT_MULTI_STATS = 5, RT_STRUCT_DRAW_SETTINGS = 6, RT_STRUCT_CHECKPOINT_INFO = 7,
                    RT_STRUCT_CATALOG_PARAMS = 8 };
size_t rt_struct_size(int rt_struct_id);

/* ---- scene lifetime: replaces struct scene's ctor/dtor (scenes.h:36-67) ---- */
int rt_scene_create(rt_scene** out);
void rt_scene_destroy(rt_scene* s);

/* ---- textures (texture.h:12-164) ---- */
int rt_tex_solid(rt_scene* s, float r, float g, float b);                           /* solid_color            texture.h:14 */
int rt_tex_checker(rt_scene* s, int even_tex, int odd_tex);                         /* checker_texture        texture.h:31 */
int rt_perlin_tables(rt_scene* s, const float* ranvec768, const int32_t* perm768);  /* perlin(curandState*)   perlin.h:63 (tables made by the caller) */
int rt_tex_noise(rt_scene* s, int perlin, float scale);                             /* noise_texture          texture.h:53 */
int rt_tex_turbulent(rt_scene* s, int perlin, float scale, int depth);              /* turbulent_texture      texture.h:70 */
int rt_tex_marble(rt_scene* s, int perlin, float scale);                            /* marble_texture         texture.h:84 */
int rt_image(rt_scene* s, const uint8_t* pixels, int width, int height, int bytes_per_pixel); /* make_image texture.h:188 (decoded bytes; copied) */
int rt_tex_image(rt_scene* s, int image);                                           /* image_texture          texture.h:125 */

/* ---- materials (material.h) ---- */
int rt_mat_lambertian(rt_scene* s, int tex);            /* material.h:19-22 */
int rt_mat_metal(rt_scene* s, int tex, float fuzz);     /* material.h:45-48 */
int rt_mat_dielectric(rt_scene* s, float ir);           /* material.h:62 */
int rt_mat_diffuse_light(rt_scene* s, int tex);         /* material.h:110-113 */
int rt_mat_isotropic(rt_scene* s, int tex);             /* material.h:128-131 */

/* ---- hittables ---- */
int rt_sphere(rt_scene* s, float cx, float cy, float cz, float radius, int mat);                         /* sphere.h:13 */
int rt_moving_sphere(rt_scene* s, const float c0[3], const float c1[3], float t0, float t1, float radius, int mat); /* moving_sphere.h:13 */
int rt_xy_rect(rt_scene* s, float x0, float x1, float y0, float y1, float k, int mat);                   /* aarect.h:13 */
int rt_xz_rect(rt_scene* s, float x0, float x1, float z0, float z1, float k, int mat);                   /* aarect.h:32 */
int rt_yz_rect(rt_scene* s, float y0, float y1, float z0, float z1, float k, int mat);                   /* aarect.h:50 */
int rt_box(rt_scene* s, const float p0[3], const float p1[3], int mat);                                  /* box.h:14 */
int rt_triangle(rt_scene* s, const float v[9], const float uv[6], const float* normals9_or_null, int mat); /* triangle.h:19, :34 */
/* One mesh object as create_meshes_d builds it (triangle_mesh.h:147-204, single object): triangle k uses
 * vertices/normals/uvs of indices[3k..3k+2]; returns the triangle_mesh hittable (triangle_mesh.h:26). */
int rt_mesh(rt_scene* s, const float* vertices, int n_vertices, const uint32_t* indices, int n_triangles,
            const float* uvs, const float* normals, int mat);
int rt_translate(rt_scene* s, int child, float dx, float dy, float dz);       /* hittable.h:35 */
int rt_rotate_y(rt_scene* s, int child, float angle_degrees);                  /* hittable.h:78 */
int rt_rotate_y_sincos(rt_scene* s, int child, float sin_theta, float cos_theta);
int rt_constant_medium(rt_scene* s, int boundary, float density, int tex);     /* constant_medium.h:15 */
int rt_constant_medium_nid(rt_scene* s, int boundary, float neg_inv_density, int tex);
int rt_list(rt_scene* s, const int* children, int n);                          /* hittable_list.h:12 */
int rt_bvh(rt_scene* s, const int* children, int n);                           /* bvh.h:136 — grouping hint: everything lands in the one LBVH */

/* ---- world / camera / background ---- */
int rt_set_world(rt_scene* s, int hittable);                                   /* *d_world = ...            scenes.h:88 etc. */
int rt_set_background(rt_scene* s, float r, float g, float b);                 /* scene::set_background     scenes.h:69 */
int rt_set_camera(rt_scene* s, const float lookfrom[3], const float lookat[3], const float vup[3], float vfov_degrees,
                  float aspect_ratio, float aperture, float focus_dist, float time0, float time1); /* camera.h:18 */
int rt_set_camera_raw(rt_scene* s, const rt_flat_camera* cam);

/* Semantics switches.  The default, RT_SEM_REFERENCE, reproduces the reference including the quirks its
 * images depend on; the other bits repair one quirk each and are NOT covered by the parity contract (no reference output
 * exists for them).  Takes effect at the next rt_scene_commit.
 *   RT_SEM_SPHERE_FAR_ROOT       sphere / moving_sphere::hit try the far root (-half_b + sqrtd) / a when the near one is out
 *                                of range, as the book does — the reference repeats the near root (sphere.h:48-55,
 *                                moving_sphere.h:38-45).  Rays that start inside a sphere then hit its far side
 *                                (glass refracts on exit, hollow spheres work) and constant_medium with a sphere boundary
 *                                scatters (constant_medium.h:38-44; Q2) instead of being inert.
 *   RT_SEM_UNIT_TRIANGLE_NORMALS triangle face and interpolated vertex normals are normalised before set_face_normal
 *                                (triangle.h:166, :170-173; Q13).
 * Q3 (xz_rect bounding box, aarect.h:39) needs no switch: the LBVH computes its own boxes. */
enum rt_semantics {
    RT_SEM_REFERENCE = 0,
    RT_SEM_SPHERE_FAR_ROOT = 1,
    RT_SEM_UNIT_TRIANGLE_NORMALS = 2,
    RT_SEM_FIXED = 3
};
int rt_set_semantics(rt_scene* s, uint32_t semantics);

/* ---- bulk: replay a scene program (include/rt_program.h) through the calls above ---- */
int rt_scene_replay(rt_scene* s, const uint32_t* words, size_t n_words);

/* ---- flatten + export (host only) ---- */
int rt_scene_flatten(rt_scene* s);                          /* returns the number of leaf ids */
size_t rt_scene_export_size(rt_scene* s);                   /* 0 if not flattened */
int rt_scene_export(rt_scene* s, void* buf, size_t cap);    /* layout: rt_flat.h */
/* 64-bit FNV-1a of the exported blob and the semantics word: identifies the scene a checkpoint belongs to. */
int rt_scene_hash(rt_scene* s, uint64_t* out);

/* image_height = int(width / aspect) in double, as render_settings::calc_height (render.h:32) */
int rt_image_height(int width, float aspect);

/* ---- device side ---- */
enum rt_builder { RT_BUILDER_KARRAS = 0, RT_BUILDER_PLOC = 1, RT_BUILDER_HOST_SAH = 2, RT_BUILDER_GPU_SAH = 3 };

typedef struct rt_build_stats {
    uint32_t n_prims, n_nodes;
    float ms_upload, ms_morton, ms_sort, ms_hierarchy, ms_refit, ms_collapse, ms_total;
    uint32_t bvh_depth;      /* depth of the binary tree */
    uint32_t build_rounds;   /* PLOC builds: clustering rounds; SAH sweep on the GPU: levels */
    float sah_internal;      /* sum of internal-node surface areas / root area */
    uint32_t ploc_radius;    /* PLOC builds: neighbour search radius of the tree that was kept */
    uint32_t builder;        /* rt_builder: which hierarchy builder produced the tree */
    uint32_t n_wide_nodes;   /* > 0: the traversal walks the compressed 4-wide layout (this many 64-byte nodes) */
    float ms_host;           /* host wall clock of the whole rt_scene_commit call */
    float ms_wide;           /* collapse of the binary tree into wide nodes (CUDA events; not part of ms_total) */
    uint32_t wide_depth;     /* levels of the wide tree */
} rt_build_stats;

/* How rt_scene_commit_ex builds the acceleration structure.  Zero-initialise for the defaults.  (Round 1 read these
 * from environment variables; a library behind a C-ABI takes them as arguments.) */
enum rt_build_choice {
    RT_BUILD_AUTO = 0,       /* SAH sweep on the GPU up to sah_max_prims primitives, PLOC (GPU) with a SAH-swept top above */
    RT_BUILD_KARRAS = 1,     /* Morton-prefix hierarchy (Karras 2012) + bottom-up refit: A/B only */
    RT_BUILD_PLOC = 2,       /* parallel locally-ordered clustering on the GPU */
    RT_BUILD_HOST_SAH = 3,   /* round 1's host-thread sweep: kept as the checker of the GPU sweep (same tree, bit for bit) */
    RT_BUILD_GPU_SAH = 4     /* full-sweep SAH, top-down, on the GPU */
};
enum rt_wide_choice { RT_WIDE_AUTO = 0, RT_WIDE_OFF = 1, RT_WIDE_ON = 2 };
enum rt_build_flags { RT_BUILD_TRACE = 1 /* host wall clock of the commit phases on stderr */ };
typedef struct rt_build_settings {
    uint32_t builder;        /* rt_build_choice */
    uint32_t sah_max_prims;  /* RT_BUILD_AUTO: largest scene the SAH sweep takes; 0 = 65536 */
    uint32_t ploc_radius;    /* PLOC neighbour search radius; 0 = chosen per scene */
    uint32_t ploc_top;       /* PLOC hands the top of the tree to the SAH sweep once this many clusters are left; 0 = 32768 */
    uint32_t wide_nodes;     /* rt_wide_choice: collapse the binary tree into 4-wide nodes with 8-bit child boxes.
                                AUTO = scenes of 65 536 primitives and more (the binary tree no longer fits the L1) */
    uint32_t flags;          /* rt_build_flags */
    uint32_t reserved[10];
} rt_build_settings;

/* Flatten (if needed), upload to `device` and build the LBVH there.  Replaces the one-thread device
 * constructors create_*_world<<<1,1>>> + bvh_node::bvh_node (scenes.h:82..., bvh.h:270-346). */
int rt_scene_commit(rt_scene* s, int device, rt_build_stats* stats_or_null);
int rt_scene_commit_ex(rt_scene* s, int device, const rt_build_settings* settings_or_null, rt_build_stats* stats_or_null);

/* Test hook: the (nth + 1)-th device allocation the library makes from now on fails with an out-of-memory error (nth < 0:
 * off).  Error paths must then return RT_ERR_CUDA, leak nothing and leave the scene usable for a retry. */
void rt_debug_fail_allocation(int nth);

/* Copies the binary BVH of the committed scene to the host: rt_build_stats::n_nodes records of 64 bytes (16 words:
 * child 0 lo.x hi.x lo.y hi.y | child 1 lo.x hi.x lo.y hi.y | child 0 lo.z hi.z, child 1 lo.z hi.z | child codes, 2 pad words;
 * a child code >= 0 is a node index, < 0 the complement of a primitive's position in leaf order).  For tools and tests
 * (tree equality of two builders).  Returns the number of nodes. */
int rt_scene_bvh_nodes(rt_scene* s, void* buf, size_t cap_bytes);

/* Closest hit for a caller-supplied ray set — what color_f asks of the world at render.h:63.
 * rays: n x 7 floats (origin, direction, time), host memory.  xi: NULL, or n floats in (0,1]: the value used
 * for every constant_medium free-flight draw of that ray (constant_medium.h:57); with NULL media are skipped.
 * prim_id: DFS leaf id or -1; t: hit distance (+inf on miss). */
int rt_first_hit(rt_scene* s, const float* rays, int n, float t_min, const float* xi, int32_t* prim_id, float* t);

/* rt_first_hit without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and results as rt_first_hit, bit for bit; the device staging of the rays and of the results lives in a
 * file-scope __device__ array of RT_FIRST_HIT_STATIC_MAX rays (2.5 MiB, held by every GPU that loads the library until the
 * process exits) instead of a buffer the scene grows.  n > RT_FIRST_HIT_STATIC_MAX is rejected with RT_ERR_INVALID.  The
 * call makes no cudaMalloc / cudaFree (the scene must already be committed — rt_scene_commit allocates).  Every call, from
 * every scene on the same GPU, shares the one array, which keeps the last call's contents: calls that overlap on a GPU
 * race — serialise them (the call itself returns only after its results are in host memory). */
#define RT_FIRST_HIT_STATIC_MAX 65536
int rt_first_hit_static(rt_scene* s, const float* rays, int n, float t_min, const float* xi, int32_t* prim_id, float* t);

enum rt_render_flags {
    RT_RENDER_DEFAULT = 0,
    RT_RENDER_COUNTERS = 1, /* fill the traversal counters in rt_render_stats (slower) */
    RT_RENDER_TIMERS = 2,   /* CUDA-event time per kernel class (ms_extend / ms_shade / ms_generate / ms_finish) */
    RT_RENDER_ACCUMULATE = 4 /* rt_render only — progressive rendering: out_rgb holds SUMS; it is read, the samples of this call
                                are added, and it is written back un-normalised (divide by the samples rendered so far
                                to display).  Because every sample's random numbers are keyed by (pixel, sample index), a
                                frame rendered in slices [0,a) [a,b) ... equals one rendered in one call up to the order of
                                the float additions; together with rt_checkpoint_save/load (rt_output.h) a render resumes
                                after a restart */
};

#define RT_TILE_BLOCK_PIXELS 1024u

typedef struct rt_render_settings {  /* replaces render_settings (render.h:21-50) */
    uint32_t width, height;
    uint32_t samples_per_pixel;      /* total over all ranks: samples_per_pixel_per_fb * no_fb */
    uint32_t max_depth;              /* render.h:27, default 50 */
    uint32_t seed;                   /* Philox key; the reference hard-codes 1984 (render.h:91) */
    uint32_t sample_first;           /* this call renders sample indices [sample_first, sample_first+sample_count) */
    uint32_t sample_count;           /* 0 = all */
    uint32_t flags;
    uint32_t pool_paths;             /* wavefront pool size in paths; 0 = default: 32 Mi paths (4.5 GB; 16 Mi = 9.4 GB with the deferred
                                        extend), never more than the frame has samples, halved while the GPU lacks the memory */
    uint32_t tile_index, tile_count; /* tile-split partitioning: with tile_count > 1 this call renders only the film blocks
                                        b with b % tile_count == tile_index (blocks of RT_TILE_BLOCK_PIXELS consecutive
                                        pixels in film order = 32 tiles of 8x4); the other pixels of the buffer are left
                                        untouched, so per-rank buffers still sum with ONE reduce.  0 / 1 = whole frame */
    uint32_t extend_defer;           /* rt_defer_choice: deferred extend (stragglers of a warp are suspended and resumed 32 to
                                        a warp in the next iteration).  0 = automatic (on above 65 536 primitives), 1 = off,
                                        n >= 2 = on, a warp suspends once <= n - 1 of its rays are still walking */
    uint32_t traversal;              /* 0 = the default k_extend variant; RT_TRAVERSAL_FORCE | bits selects one for A/B:
                                        1 = stack entries carry their entry distance and are culled at pop, 2 = short stack in
                                        shared memory (trees up to 16 levels), 4 = min/max plane selection on the ALU pipe, 16 = two parked leaves per
                                        lane with the warp's leaf tests packed into its lowest lanes (alone); all but 0 only in the experiments build */
    uint32_t debug_flags;            /* rt_debug_flags */
    uint32_t reserved[2];
} rt_render_settings;
enum rt_defer_choice { RT_DEFER_AUTO = 0, RT_DEFER_OFF = 1 };
enum rt_traversal_bits { RT_TRAVERSAL_FORCE = 0x100, RT_TRAVERSAL_DEFAULT_BITS = 0 };
enum rt_debug_flags {
    RT_DEBUG_NO_GRAPH = 1,     /* direct kernel launches instead of the CUDA graph of 8 iterations */
    RT_DEBUG_RASTER_FILM = 2,  /* samples walk the film row by row instead of in 8x4 tiles */
    RT_DEBUG_WARP_SHADE = 4,   /* the barrier-free warp-tile shade kernel (measured slower than the block-tile default: A/B only) */
    RT_DEBUG_BLOCK_SHADE = 8   /* round 1's block-tile shade (a block sorts 256 paths by class, allocates with atomics, separate k_turnover) instead of
                                  the class-binned pipeline k_bin + k_shade_bins: an independent implementation of the same function, kept
                                  for A/B and as a cross-check in the tests (4-12 % slower on every BASELINE configuration) */
};

typedef struct rt_render_stats {
    uint64_t samples;       /* camera paths started */
    uint64_t rays;          /* path segments traced (closest-hit queries, render.h:63) */
    uint64_t iterations;    /* wavefront iterations */
    uint64_t kernel_launches;
    float ms_total;         /* CUDA events, first ray-gen launch -> accumulation complete */
    float ms_generate, ms_extend, ms_shade, ms_finish;
    uint64_t nodes_visited, prim_tests; /* only with RT_RENDER_COUNTERS */
    uint64_t node_slots;    /* with RT_RENDER_COUNTERS: sum over warps of 32 x the longest lane's node steps, i.e. the node
                               steps the SIMD machine paid for; nodes_visited / node_slots = lane utilisation lost to rays
                               of one warp finishing at different times */
    uint64_t pool_paths;    /* wavefront pool size this call used (the default halves when the GPU lacks the memory) */
    uint64_t reserved[2];
} rt_render_stats;

/* Render into host memory: out_rgb = width*height*3 floats, linear radiance, mean over the samples rendered
 * by this call, row 0 = bottom of the image (render.h:105-112, color.h:23).  Replaces draw() minus file output. */
int rt_render(rt_scene* s, const rt_render_settings* settings, float* out_rgb, rt_render_stats* stats_or_null);

/* Same, but accumulates SUMS (not means) into a caller-owned device buffer of width*height*3 floats on the
 * scene's device, on `cuda_stream` (a cudaStream_t cast to void*, NULL = default stream).  The buffer is added
 * to, not cleared.  Multi-GPU: each rank renders its sample range, then the caller sums the buffers (one NCCL
 * reduce) and divides by samples_per_pixel. */
int rt_render_device(rt_scene* s, const rt_render_settings* settings, float* d_accum_rgb, void* cuda_stream,
                     rt_render_stats* stats_or_null);

/* ---- multi-GPU behind the boundary: replaces draw()'s single `render<<<>>>` (render.h:118-162) ----
 * One host thread per device: the scene is committed on every device it is not current on (replicated: a deterministic
 * build is cheaper than a broadcast), every device renders its share into a zero-initialised W*H*3 float buffer — a
 * contiguous sample range (RT_SPLIT_SAMPLE) or the film blocks b with b % n == k (RT_SPLIT_TILE) — and the root device
 * (device_ids[0]) finishes with ONE reduce: a hand-written kernel that reads the other devices' buffers through NVLink
 * peer mappings and sums them in device order (bitwise reproducible for a given device count), scales by 1/samples and
 * writes the frame.  Without peer access the buffers are staged through cudaMemcpyPeer.  Philox keys make the sample set
 * independent of the split; float summation order differs with the device count (~1e-6 relative). */
enum rt_split_mode { RT_SPLIT_SAMPLE = 0, RT_SPLIT_TILE = 1 };
#define RT_MAX_DEVICES 16
typedef struct rt_multi_settings {
    uint32_t n_devices;                    /* 0 = every visible CUDA device, in order */
    int32_t device_ids[RT_MAX_DEVICES];    /* used when n_devices > 0; [0] is the root */
    uint32_t split_mode;                   /* rt_split_mode */
    uint32_t reserved[6];
} rt_multi_settings;
typedef struct rt_multi_stats {
    uint32_t n_devices;
    uint32_t peer_reduce;                  /* 1: the reduce kernel read peer memory directly; 0: staged copies */
    float ms_commit_max;                   /* slowest device's commit (0 where the device copy was current) */
    float ms_render[RT_MAX_DEVICES];       /* per device: first ray-gen launch -> its accumulation complete (CUDA events) */
    float ms_reduce;                       /* the reduce + scale kernel on the root (CUDA events) */
    float ms_wall;                         /* host wall clock of the call: commit + render + reduce + D2H */
    uint64_t samples, rays;                /* summed over the devices */
    uint64_t reserved[4];
} rt_multi_stats;
/* out_rgb: width*height*3 floats, host memory, linear means over the sample range [sample_first, sample_first + sample_count)
 * of `settings` (like rt_render: sample_count 0 = up to samples_per_pixel); that range is what is split over the devices.
 * settings->tile_index / tile_count must be 0.  multi_or_null == NULL: every visible device, sample-split. */
int rt_render_multi(rt_scene* s, const rt_render_settings* settings, const rt_multi_settings* multi_or_null, float* out_rgb,
                    rt_multi_stats* stats_or_null);

/* FP32 peak of `device` measured with an FMA microbenchmark (the denominator of bench.py's roofline_fp32). */
int rt_measure_fp32_peak(int device, double* tflops);

/* Reference-exact 8-bit output stage (color.h:19-49): gamma-2, clamp to 0.999, *256, row flip. Host only.
 * in: linear means, row 0 = bottom; out: width*height*3 bytes, row 0 = top. */
int rt_quantize_frame(const float* rgb_linear, int width, int height, uint8_t* out_rgb8);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
