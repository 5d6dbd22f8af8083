/* rt_catalog.h — library scenes: the reference's scene catalogue (scenes.h:82-621) plus the Next-Week
 * final scene it does not ship, recorded through the compat layer (rt_compat.h) into an rt_scene.
 * Part of librt_b200.so. */
#ifndef RT_CATALOG_H
#define RT_CATALOG_H
#include "rt_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_catalog_params {
    /* image for "earth" / "final" / "mesh" (decoded RGB8/RGBA8 bytes; the reference loads earthmap.jpg / Door_C.jpg) */
    const uint8_t* image; int image_width, image_height, image_bytes_per_pixel;
    /* "final": aspect ratio (Shirley: 1.0; BASELINE asks 16:9) */
    float aspect;
    /* "mesh": flat arrays as create_meshes_d takes them + camera (lookfrom, lookat, vup, vfov, aspect, aperture, focus, t0, t1) */
    const float* vertices; int n_vertices; const uint32_t* indices; int n_triangles; const float* uvs; const float* normals;
    const float* camera15;
    /* RNG-built scenes (big1; the Perlin tables of two_perlin / final): draws the reference passes as function ARGUMENTS
     * (scenes.h:152, vec3.h:63/:68) are taken in nvcc's left-to-right order by default, i.e. the scene the reference's
     * CUDA build constructs on the device; 1 = the host compiler's order (what a g++ build of the same sources gives,
     * used to pin the catalogue against oracle/_ref). */
    int host_argument_order;
} rt_catalog_params;

/* names: basic, first, big1, two_spheres, two_perlin, earth, cornell, cornell_smoke, triangle, triangles, mesh, final.
 * Fills `s` (world, camera, background); *aspect_out receives the scene's aspect ratio (scene::aspect). */
int rt_catalog_scene(rt_scene* s, const char* name, const rt_catalog_params* params_or_null, float* aspect_out);

#ifdef __cplusplus
}
#endif
#endif
