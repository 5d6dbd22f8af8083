/* rt_obj.h — Wavefront OBJ / MTL import (C-ABI, host only; part of librt_b200.so).
 *
 * Replaces the reference's assimp path: load_model / processNode / processMesh (triangle_mesh.h:51-143) and
 * the host half of create_meshes (triangle_mesh.h:208-352).  The reference asks assimp for
 * aiProcess_Triangulate | aiProcess_GenNormals (triangle_mesh.h:132); this loader does the same two things
 * itself:
 *   - faces with more than three corners are fan-triangulated (corner 0, i, i+1);
 *   - corners are unified: every distinct (v, vt, vn) index triple becomes one output vertex, so the four
 *     arrays (positions, uvs, normals; one entry per vertex) plus one index triple per triangle are exactly
 *     the flat arrays create_meshes_d consumes (triangle_mesh.h:147-204) and rt_mesh() takes;
 *   - when the file has no `vn` records, per-vertex normals are generated as the normalised sum of the
 *     (area-weighted) face normals of the faces that use the POSITION (smooth normals);
 *   - missing texture coordinates become (0, 0) (triangle_mesh.h:66-69);
 *   - the first `map_Kd` of the materials in the referenced .mtl files is reported as the diffuse map
 *     (the reference keeps one diffuse texture per object: triangle_mesh.h:275-281).
 * Deliberately NOT reproduced: the reference appends per-aiMesh indices without a base-vertex offset
 *  — a bug that depends on assimp's mesh splitting; here all objects / groups of the file land in
 * one consistent vertex array.
 * Supported records: v, vt, vn, f (v, v/vt, v//vn, v/vt/vn; negative = relative), mtllib, usemtl; o / g / s
 * and everything else are ignored.
 */
#ifndef RT_OBJ_H
#define RT_OBJ_H

#include <stdint.h>
#include "rt_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_obj rt_obj;

/* Parse `path`.  Returns RT_OK or a negative rt_status (rt_last_error() has the message). */
int rt_obj_load(const char* path, rt_obj** out);
/* Same, from memory (no .mtl lookup: mtl_text may be NULL). */
int rt_obj_parse(const char* obj_text, size_t obj_len, const char* mtl_text, size_t mtl_len, rt_obj** out);
void rt_obj_free(rt_obj* o);

uint32_t rt_obj_num_vertices(const rt_obj* o);
uint32_t rt_obj_num_triangles(const rt_obj* o);
const float* rt_obj_vertices(const rt_obj* o);     /* 3 per vertex */
const float* rt_obj_uvs(const rt_obj* o);          /* 2 per vertex */
const float* rt_obj_normals(const rt_obj* o);      /* 3 per vertex */
const uint32_t* rt_obj_indices(const rt_obj* o);   /* 3 per triangle */
int rt_obj_normals_generated(const rt_obj* o);     /* 1 when the file had no vn records */
const char* rt_obj_diffuse_map(const rt_obj* o);   /* "" when none; path as written in the .mtl */

/* The mesh as one triangle_mesh hittable of `scene` (= rt_mesh on the arrays above). */
int rt_obj_add_to_scene(rt_scene* scene, const rt_obj* o, int mat);

#ifdef __cplusplus
}
#endif
#endif /* RT_OBJ_H */
