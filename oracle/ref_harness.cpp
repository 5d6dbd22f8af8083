/* oracle/ref_harness.cpp — TEST INFRASTRUCTURE, not product code.
 *
 * Harness around the UNMODIFIED reference sources, which are compiled as host C++ straight from
 * /root/reference (see oracle/Makefile: -I/root/reference plus the shim in oracle/shim/; three
 * one-line build-time survivability patches are applied on the fly to throw-away copies, listed in
 * the Makefile).  The result, oracle/_ref/libref_oracle.so, is the authoritative CPU oracle:
 *   - scenes are built by the reference's own constructors (scenes.h bodies / this file's
 *     compositions for the scenes the reference does not ship),
 *   - first-hit queries go through the reference's own hittable::hit code (hittable_list.h:23,
 *     bvh.h:348, sphere.h:35, aarect.h:63..., triangle.h:120, constant_medium.h:34, hittable.h:37,112),
 *   - images go through the reference's own render()/color_f() (render.h:55-113), one emulated CUDA
 *     thread per pixel.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Everything in this file is harness code written for this repo; no reference code is copied.
 */
#include "scenes.h"
#include "render.h"
#include "../include/rt_program.h"

#include <map>
#include <thread>
#include <atomic>
#include <new>
#include <chrono>

#define REF_API extern "C" __attribute__((visibility("default")))

thread_local ref_idx3 ref_threadIdx, ref_blockIdx, ref_blockDim, ref_gridDim;
thread_local const float* ref_scripted_uniform = nullptr;

/* The reference frees Perlin tables and mesh textures that it keeps using and deletes
 * through non-virtual destructors.  Restating the intent ("tables stay alive"): every delete inside this
 * library is a no-op.  The operators are hidden + bound symbolically so the host process is unaffected. */
__attribute__((visibility("hidden"))) void operator delete(void*) noexcept {}
__attribute__((visibility("hidden"))) void operator delete[](void*) noexcept {}
__attribute__((visibility("hidden"))) void operator delete(void*, std::size_t) noexcept {}
__attribute__((visibility("hidden"))) void operator delete[](void*, std::size_t) noexcept {}

/* Wrapper material that carries the DFS leaf id of the primitive it was attached to; forwards
 * everything to the real material so rendering is unchanged. */
struct tag_material : public material {
    material* real;
    int leaf_id;
    tag_material(material* r, int id) : real(r), leaf_id(id) {}
    bool scatter(const ray& ray_in, const hit_record& rec, color& attenuation, ray& scattered, curandState* s) const override {
        return real->scatter(ray_in, rec, attenuation, scattered, s);
    }
    color emitted(const float u, const float v, const point3& p) const override { return real->emitted(u, v, p); }
};

struct ref_scene {
    scene* sc = nullptr;          /* reference scene object (never destroyed) */
    int num_leaves = 0;
    bool tagged = false;
    std::vector<uint32_t> program; /* exported before tagging */
};

static std::vector<ref_scene*> g_scenes;

static curandState* new_world_rng() {
    curandState* rs = new curandState;
    world_init(rs);               /* scenes.h:28 — curand_init(1984, 0, 0) */
    return rs;
}

/* A bare `scene` whose world/camera are filled by the caller (used for the compositions below). */
static scene* blank_scene(float aspect, background_color bc, int n_top) {
    scene* s = new scene(aspect, n_top, bc);
    return s;
}

/* ---- compositions the reference does not ship as a struct -------------------------------------- */

/* earth_scene (scenes.h:278-320) with the decoded image supplied by the caller instead of stb. */
static scene* build_earth(const unsigned char* img, int w, int h, int bpp) {
    scene* s = blank_scene(16.0f / 9.0f, background_color::black, 2);
    size_t nbytes = (size_t)w * h * bpp;
    unsigned char* data = (unsigned char*)malloc(nbytes);
    memcpy(data, img, nbytes);
    int* ws = new int[1]{w};
    int* hs = new int[1]{h};
    int* bs = new int[1]{bpp};
    create_earth_world(s->d_list, s->d_world, s->d_camera, data, ws, hs, bs);
    return s;
}

/* Next-Week final scene (absent from the reference; composed from reference constructors
 * following Shirley's published scene).  RNG order: box heights (i-major, then j), sphere positions
 * (x, y, z per sphere), then the Perlin tables — all from XORWOW(1984,0,0) — and only then the two
 * bvh_node constructors (whose axis picks, bvh.h:294, must not perturb the geometry draws). */
static scene* build_final(const unsigned char* img, int w, int h, int bpp, float aspect) {
    scene* s = blank_scene(aspect, background_color::black, 11);
    curandState* rs = new_world_rng();
    hittable** top = s->d_list;
    int k = 0;

    material* ground = new lambertian(color(0.48f, 0.83f, 0.53f));
    const int boxes_per_side = 20;
    hittable** boxes1 = new hittable*[boxes_per_side * boxes_per_side];
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
    hittable** boxes2 = new hittable*[ns];
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

    size_t nbytes = (size_t)w * h * bpp;
    unsigned char* data = (unsigned char*)malloc(nbytes);
    memcpy(data, img, nbytes);
    int* ws = new int[1]{w};
    int* hs = new int[1]{h};
    int* bs = new int[1]{bpp};
    top[k++] = new sphere(point3(400, 200, 400), 100, new lambertian(new image_texture(data, ws, hs, bs, 0)));

    top[k++] = new sphere(point3(220, 280, 300), 80, new lambertian(pertext));
    hittable* cluster = new bvh_node(boxes2, ns, 0, 1, rs);

    top[k++] = new translate(new rotate_y(cluster, 15), vec3(-100, 270, 395));

    *s->d_world = new hittable_list(top, k);
    *s->d_camera = new camera(vec3(478, 278, -600), vec3(278, 278, 0), vec3(0, 1, 0), 40, aspect, 0.0f, 10.0f, 0, 1);
    s->no_hittables = k;
    return s;
}

/* A composition for the media-ordering edge cases: two OVERLAPPING box-bounded media, a medium listed BEFORE a
 * surface that lies in front of its far side, and a medium listed before a NEARER medium — in a hittable_list a medium
 * only sees the objects listed before it and shrinks t_max for the ones after it (hittable_list.h:29-35).  Built from
 * reference constructors; not a scene the reference ships. */
static scene* build_media_mix() {
    scene* s = blank_scene(1.0f, background_color::sky, 8);
    hittable** top = s->d_list;
    int k = 0;
    material* white = new lambertian(color(0.73f, 0.73f, 0.73f));
    material* red = new lambertian(color(0.65f, 0.05f, 0.05f));
    /* far medium first, then the nearer one that overlaps it */
    hittable* far_box = new translate(new rotate_y(new box(point3(0, 0, 0), point3(200, 200, 200), white), 20), vec3(150, 50, 300));
    top[k++] = new constant_medium(far_box, 0.01f, color(0.1f, 0.1f, 0.1f));
    hittable* near_box = new translate(new box(point3(0, 0, 0), point3(200, 200, 200), white), vec3(200, 80, 150));
    top[k++] = new constant_medium(near_box, 0.006f, color(0.9f, 0.9f, 0.9f));
    /* a surface listed AFTER both media, standing inside the nearer one (in front of its far side) */
    top[k++] = new sphere(point3(300, 180, 250), 60, red);
    top[k++] = new xz_rect(-200, 800, -200, 900, 0, white);
    /* and a third medium listed last, in front of everything */
    hittable* front_box = new box(point3(180, 60, 20), point3(420, 300, 100), white);
    top[k++] = new constant_medium(front_box, 0.004f, color(0.2f, 0.4f, 0.9f));
    *s->d_world = new hittable_list(top, k);
    *s->d_camera = new camera(vec3(278, 278, -600), vec3(290, 160, 250), vec3(0, 1, 0), 40, 1.0f, 0.0f, 10.0f, 0, 1);
    s->no_hittables = k;
    return s;
}

/* door_scene / cup_scene shape (scenes.h:478-523, :576-621): one triangle_mesh + ground sphere, with the
 * flat arrays fed straight into the reference's create_meshes_d (triangle_mesh.h:147) instead of assimp. */
static scene* build_mesh(const float* verts, int nverts, const unsigned* idx, int ntris, const float* uvs,
                         const float* norms, const unsigned char* img, int w, int h, int bpp, const float* cam) {
    scene* s = blank_scene(cam[10], background_color::sky, 2);
    curandState* rs = new_world_rng();
    hittable** tris = new hittable*[ntris];
    unsigned* num = new unsigned[1]{(unsigned)ntris};
    size_t nbytes = (size_t)w * h * bpp;
    unsigned char* data = (unsigned char*)malloc(nbytes);
    memcpy(data, img, nbytes);
    int* ws = new int[1]{w};
    int* hs = new int[1]{h};
    int* bs = new int[1]{bpp};
    float* v = new float[3 * (size_t)nverts]; memcpy(v, verts, sizeof(float) * 3 * nverts);
    float* n = new float[3 * (size_t)nverts]; memcpy(n, norms, sizeof(float) * 3 * nverts);
    float* t = new float[2 * (size_t)nverts]; memcpy(t, uvs, sizeof(float) * 2 * nverts);
    unsigned* ix = new unsigned[3 * (size_t)ntris]; memcpy(ix, idx, sizeof(unsigned) * 3 * ntris);
    create_meshes_d(tris, num, 1, data, ws, hs, bs, v, ix, t, n);
    s->d_list[0] = new triangle_mesh(tris, num, 0, 1, rs, 0);
    s->d_list[1] = new sphere(vec3(0, -100, -1), 100, new lambertian(vec3(0, 1, 0)));
    *s->d_world = new hittable_list(s->d_list, 2);
    *s->d_camera = new camera(vec3(cam[0], cam[1], cam[2]), vec3(cam[3], cam[4], cam[5]), vec3(cam[6], cam[7], cam[8]),
                              cam[9], cam[10], cam[11], cam[12], cam[13], cam[14]);
    return s;
}

/* ---- object-graph walk: export as a scene program (include/rt_program.h) ---------------------- */

struct exporter {
    std::vector<uint32_t>& out;
    std::map<const void*, uint32_t> tex_ix, mat_ix, perlin_ix, image_ix;
    uint32_t n_tex = 0, n_mat = 0, n_hit = 0, n_perlin = 0, n_image = 0;
    explicit exporter(std::vector<uint32_t>& o) : out(o) { out.push_back(RTP_MAGIC); }

    void u(uint32_t v) { out.push_back(v); }
    void f(float v) { uint32_t b; memcpy(&b, &v, 4); out.push_back(b); }
    void v3(const vec3& v) { f(v.e[0]); f(v.e[1]); f(v.e[2]); }

    uint32_t perlin_tables(const perlin& p) {
        auto it = perlin_ix.find(p.ranvec);
        if (it != perlin_ix.end()) return it->second;
        u(RTP_PERLIN);
        for (int i = 0; i < 256; i++) v3(p.ranvec[i]);
        for (int i = 0; i < 256; i++) u((uint32_t)p.perm_x[i]);
        for (int i = 0; i < 256; i++) u((uint32_t)p.perm_y[i]);
        for (int i = 0; i < 256; i++) u((uint32_t)p.perm_z[i]);
        return perlin_ix[p.ranvec] = n_perlin++;
    }

    uint32_t image(const image_texture* t) {
        const unsigned char* base = t->data + t->index;
        auto it = image_ix.find(base);
        if (it != image_ix.end()) return it->second;
        const uint32_t nbytes = (uint32_t)t->bytes_per_scanline * (uint32_t)t->height;
        u(RTP_IMAGE); u(t->width); u(t->height); u(t->bytes_per_pixel); u(nbytes);
        const size_t at = out.size();
        out.resize(at + (nbytes + 3) / 4, 0u);
        memcpy(&out[at], base, nbytes);
        return image_ix[base] = n_image++;
    }

    uint32_t tex(const texturez* t) {
        auto it = tex_ix.find(t);
        if (it != tex_ix.end()) return it->second;
        if (auto p = dynamic_cast<const solid_color*>(t)) {
            const color c = p->value(0, 0, vec3(0, 0, 0));
            u(RTP_TEX_SOLID); v3(c);
        } else if (auto p = dynamic_cast<const checker_texture*>(t)) {
            const uint32_t e = tex(p->even), o = tex(p->odd);
            u(RTP_TEX_CHECKER); u(e); u(o);
        } else if (auto p = dynamic_cast<const noise_texture*>(t)) {
            const uint32_t k = perlin_tables(p->noise);
            u(RTP_TEX_NOISE); u(k); f(p->scale);
        } else if (auto p = dynamic_cast<const turbulent_texture*>(t)) {
            const uint32_t k = perlin_tables(p->noise);
            u(RTP_TEX_TURB); u(k); f(p->scale); u((uint32_t)p->depth);
        } else if (auto p = dynamic_cast<const marble_texture*>(t)) {
            const uint32_t k = perlin_tables(p->noise);
            u(RTP_TEX_MARBLE); u(k); f(p->scale);
        } else if (auto p = dynamic_cast<const image_texture*>(t)) {
            const uint32_t k = image(p);
            u(RTP_TEX_IMAGE); u(k);
        } else {
            fprintf(stderr, "ref_harness: unknown texture type\n"); abort();
        }
        return tex_ix[t] = n_tex++;
    }

    uint32_t mat(const material* m) {
        auto it = mat_ix.find(m);
        if (it != mat_ix.end()) return it->second;
        if (auto p = dynamic_cast<const lambertian*>(m)) { const uint32_t t = tex(p->albedo); u(RTP_MAT_LAMBERTIAN); u(t); }
        else if (auto p = dynamic_cast<const metal*>(m)) { const uint32_t t = tex(p->albedo); u(RTP_MAT_METAL); u(t); f(p->fuzz); }
        else if (auto p = dynamic_cast<const dielectric*>(m)) { u(RTP_MAT_DIELECTRIC); f(p->ir); }
        else if (auto p = dynamic_cast<const diffuse_light*>(m)) { const uint32_t t = tex(p->emit); u(RTP_MAT_DIFFUSE_LIGHT); u(t); }
        else if (auto p = dynamic_cast<const isotropic*>(m)) { const uint32_t t = tex(p->albedo); u(RTP_MAT_ISOTROPIC); u(t); }
        else { fprintf(stderr, "ref_harness: unknown material type\n"); abort(); }
        return mat_ix[m] = n_mat++;
    }

    uint32_t group(uint32_t op, hittable** objs, int n) {
        std::vector<uint32_t> kids(n);
        for (int i = 0; i < n; i++) kids[i] = hit(objs[i]);
        u(op); u((uint32_t)n);
        for (int i = 0; i < n; i++) u(kids[i]);
        return n_hit++;
    }

    /* no sharing map for hittables: every visit creates a fresh node, exactly as the flattener would */
    uint32_t hit(const hittable* h) {
        if (auto p = dynamic_cast<const sphere*>(h)) { const uint32_t m = mat(p->mat_ptr); u(RTP_SPHERE); v3(p->center); f(p->radius); u(m); return n_hit++; }
        if (auto p = dynamic_cast<const moving_sphere*>(h)) {
            const uint32_t m = mat(p->mat_ptr);
            u(RTP_MOVING_SPHERE); v3(p->center0); v3(p->center1); f(p->time0); f(p->time1); f(p->radius); u(m); return n_hit++;
        }
        if (auto p = dynamic_cast<const xy_rect*>(h)) { const uint32_t m = mat(p->mp); u(RTP_XY_RECT); f(p->x0); f(p->x1); f(p->y0); f(p->y1); f(p->k); u(m); return n_hit++; }
        if (auto p = dynamic_cast<const xz_rect*>(h)) { const uint32_t m = mat(p->mp); u(RTP_XZ_RECT); f(p->x0); f(p->x1); f(p->z0); f(p->z1); f(p->k); u(m); return n_hit++; }
        if (auto p = dynamic_cast<const yz_rect*>(h)) { const uint32_t m = mat(p->mp); u(RTP_YZ_RECT); f(p->y0); f(p->y1); f(p->z0); f(p->z1); f(p->k); u(m); return n_hit++; }
        if (auto p = dynamic_cast<const box*>(h)) {
            /* all six sides share one material (box.h:17-24) */
            const xy_rect* side = dynamic_cast<const xy_rect*>(p->temp_hittable[0]);
            const uint32_t m = mat(side->mp);
            u(RTP_BOX); v3(p->box_min); v3(p->box_max); u(m); return n_hit++;
        }
        if (auto p = dynamic_cast<const triangle*>(h)) {
            const uint32_t m = mat(p->mp);
            u(p->vertex_normals ? RTP_TRIANGLE_N : RTP_TRIANGLE);
            v3(p->vertex0); v3(p->vertex1); v3(p->vertex2);
            if (p->vertex_normals) { v3(p->normal0); v3(p->normal1); v3(p->normal2); }
            f(p->u_0); f(p->v_0); f(p->u_1); f(p->v_1); f(p->u_2); f(p->v_2); u(m); return n_hit++;
        }
        if (auto p = dynamic_cast<const translate*>(h)) { const uint32_t c = hit(p->ptr); u(RTP_TRANSLATE); u(c); v3(p->offset); return n_hit++; }
        if (auto p = dynamic_cast<const rotate_y*>(h)) { const uint32_t c = hit(p->ptr); u(RTP_ROTATE_Y_SC); u(c); f(p->sin_theta); f(p->cos_theta); return n_hit++; }
        if (auto p = dynamic_cast<const constant_medium*>(h)) {
            const uint32_t b = hit(p->boundary);
            const isotropic* iso = dynamic_cast<const isotropic*>(p->phase_function);
            const uint32_t t = tex(iso->albedo);
            u(RTP_MEDIUM_NID); u(b); f(p->neg_inv_density); u(t); return n_hit++;
        }
        if (auto p = dynamic_cast<const hittable_list*>(h)) return group(RTP_LIST, p->objects, p->list_size);
        if (auto p = dynamic_cast<const bvh_node*>(h)) return group(RTP_BVH, p->objs, p->n);
        if (auto p = dynamic_cast<const triangle_mesh*>(h)) return group(RTP_BVH, p->tris->objs, p->tris->n);
        fprintf(stderr, "ref_harness: unknown hittable type\n"); abort();
    }

    void finish(const scene* s) {
        const uint32_t w = hit(*s->d_world);
        u(RTP_WORLD); u(w);
        u(RTP_BACKGROUND); v3(s->background);
        const camera* c = *s->d_camera;
        u(RTP_CAMERA_RAW); v3(c->origin); v3(c->lower_left_corner); v3(c->horizontal); v3(c->vertical);
        v3(c->u); v3(c->v); v3(c->w); f(c->lens_radius); f(c->time0); f(c->time1);
        u(RTP_END);
    }
};

/* ---- object-graph walk: attach leaf ids (same DFS order as the product's flattener) ------------- */

struct tagger {
    int next = 0;
    material* wrap(material* m, bool visible) {
        const int id = next++;
        if (!visible || dynamic_cast<tag_material*>(m)) return m;   /* boundary leaves count but never report */
        return new tag_material(m, id);
    }
    void walk(hittable* h, bool visible) {
        if (auto p = dynamic_cast<sphere*>(h)) { p->mat_ptr = wrap(p->mat_ptr, visible); return; }
        if (auto p = dynamic_cast<moving_sphere*>(h)) { p->mat_ptr = wrap(p->mat_ptr, visible); return; }
        if (auto p = dynamic_cast<xy_rect*>(h)) { p->mp = wrap(p->mp, visible); return; }
        if (auto p = dynamic_cast<xz_rect*>(h)) { p->mp = wrap(p->mp, visible); return; }
        if (auto p = dynamic_cast<yz_rect*>(h)) { p->mp = wrap(p->mp, visible); return; }
        if (auto p = dynamic_cast<triangle*>(h)) { p->mp = wrap(p->mp, visible); return; }
        if (auto p = dynamic_cast<box*>(h)) { for (int i = 0; i < 6; i++) walk(p->temp_hittable[i], visible); return; }
        if (auto p = dynamic_cast<translate*>(h)) { walk(p->ptr, visible); return; }
        if (auto p = dynamic_cast<rotate_y*>(h)) { walk(p->ptr, visible); return; }
        if (auto p = dynamic_cast<constant_medium*>(h)) {
            p->phase_function = wrap(p->phase_function, visible);
            walk(p->boundary, false);
            return;
        }
        if (auto p = dynamic_cast<hittable_list*>(h)) { for (int i = 0; i < p->list_size; i++) walk(p->objects[i], visible); return; }
        if (auto p = dynamic_cast<bvh_node*>(h)) { for (int i = 0; i < p->n; i++) walk(p->objs[i], visible); return; }
        if (auto p = dynamic_cast<triangle_mesh*>(h)) { for (int i = 0; i < p->tris->n; i++) walk(p->tris->objs[i], visible); return; }
        fprintf(stderr, "ref_harness: unknown hittable type in tagger\n"); abort();
    }
};

static int register_scene(scene* s) {
    ref_scene* r = new ref_scene;
    r->sc = s;
    exporter ex(r->program);
    ex.finish(s);
    tagger tg;
    tg.walk(*s->d_world, true);
    r->num_leaves = tg.next;
    r->tagged = true;
    g_scenes.push_back(r);
    return (int)g_scenes.size() - 1;
}

/* ---- C entry points (loaded by tests / bench through ctypes) ------------------------------------- */

REF_API int ref_scene_named(const char* name) {
    const std::string n(name);
    scene* s = nullptr;
    if (n == "basic") s = new basic_scene;
    else if (n == "first") s = new first_scene;
    else if (n == "big1") s = new big_scene1;
    else if (n == "two_spheres") s = new two_spheres_scene;
    else if (n == "two_perlin") s = new two_perlin_spheres_scene;
    else if (n == "cornell") s = new cornell_box_scene;
    else if (n == "cornell_smoke") s = new cornell_smoke_box_scene;
    else if (n == "triangle") s = new triangle_scene;
    else if (n == "triangles") s = new triangles_scene;
    else if (n == "media_mix") s = build_media_mix();
    else return -1;
    return register_scene(s);
}

REF_API int ref_scene_earth(const unsigned char* img, int w, int h, int bpp) { return register_scene(build_earth(img, w, h, bpp)); }

REF_API int ref_scene_final(const unsigned char* img, int w, int h, int bpp, float aspect) {
    return register_scene(build_final(img, w, h, bpp, aspect));
}

/* cam: lookfrom(3) lookat(3) vup(3) vfov aspect aperture focus t0 t1 */
REF_API int ref_scene_mesh(const float* verts, int nverts, const unsigned* idx, int ntris, const float* uvs, const float* norms,
                           const unsigned char* img, int w, int h, int bpp, const float* cam) {
    return register_scene(build_mesh(verts, nverts, idx, ntris, uvs, norms, img, w, h, bpp, cam));
}

REF_API int ref_num_leaves(int h) { return g_scenes[h]->num_leaves; }
REF_API float ref_aspect(int h) { return g_scenes[h]->sc->aspect; }
REF_API void ref_background(int h, float* rgb) { const color b = g_scenes[h]->sc->background; rgb[0] = b.x(); rgb[1] = b.y(); rgb[2] = b.z(); }

/* image_height exactly as render_settings::calc_height (render.h:32-34) */
REF_API int ref_image_height(int h, int width) {
    render_settings st;
    st.image_width = (unsigned)width;
    st.aspect_ratio = g_scenes[h]->sc->aspect;
    st.calc_all();
    return (int)st.image_height;
}

REF_API size_t ref_program_words(int h) { return g_scenes[h]->program.size(); }
REF_API void ref_program_copy(int h, uint32_t* out) { memcpy(out, g_scenes[h]->program.data(), g_scenes[h]->program.size() * 4); }

/* First hit through the reference's own world->hit (what color_f calls at render.h:63).
 * rays: n x 7 floats (origin, direction, time).  xi: optional n floats, the value every
 * constant_medium free-flight draw returns for that ray (NULL = real XORWOW(1984, ray index, 0)).
 * out_id: DFS leaf id or -1; out_t: hit distance. */
REF_API void ref_first_hit(int h, const float* rays, int n, float t_min, const float* xi, int* out_id, float* out_t, int nthreads) {
    scene* s = g_scenes[h]->sc;
    hittable* world = *s->d_world;
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> pool;
    std::atomic<int> next(0);
    for (int w = 0; w < nthreads; w++)
        pool.emplace_back([&]() {
            for (;;) {
                const int base = next.fetch_add(256);
                if (base >= n) break;
                const int end = std::min(n, base + 256);
                for (int i = base; i < end; i++) {
                    const float* q = rays + 7 * (size_t)i;
                    const ray r(point3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), q[6]);
                    curandState st;
                    if (xi) ref_scripted_uniform = xi + i; else curand_init(1984, i, 0, &st);
                    hit_record rec;
                    if (world->hit(r, t_min, infinity, rec, &st)) {
                        const tag_material* tm = dynamic_cast<const tag_material*>(rec.mat_ptr);
                        out_id[i] = tm ? tm->leaf_id : -2;
                        out_t[i] = rec.t;
                    } else {
                        out_id[i] = -1;
                        out_t[i] = infinity;
                    }
                }
            }
            ref_scripted_uniform = nullptr;
        });
    for (auto& t : pool) t.join();
}

/* One frame-buffer batch through the reference's own render_init + render (render.h:84-113), one emulated
 * CUDA thread per pixel (blockDim = 1x1, blockIdx = pixel).  fb: width*height*3 floats, row 0 = bottom.
 * `batch_id` is render()'s `id` argument (render.h:101).  The camera's shared-state race is kept:
 * worker threads really do share rand_state[0]. Returns 0. */
static double g_last_init_s = 0, g_last_render_s = 0;
/* wall-clock seconds of the last ref_render call: [0] = render_init pass, [1] = render pass */
REF_API void ref_last_times(double* out) { out[0] = g_last_init_s; out[1] = g_last_render_s; }

REF_API int ref_render(int h, int width, int height, int ns, int max_depth, int batch_id, int nthreads, float* fb_out) {
    scene* s = g_scenes[h]->sc;
    const size_t npix = (size_t)width * height;
    curandState* states = (curandState*)malloc(npix * sizeof(curandState));
    vec3* fb = (vec3*)fb_out;
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next_row(0);
    auto run_rows = [&](bool init) {
        next_row = 0;
        std::vector<std::thread> pool;
        for (int w = 0; w < nthreads; w++)
            pool.emplace_back([&]() {
                ref_blockDim = {1, 1, 1};
                ref_threadIdx = {0, 0, 0};
                for (;;) {
                    const int j = next_row.fetch_add(1);
                    if (j >= height) break;
                    for (int i = 0; i < width; i++) {
                        ref_blockIdx = {(unsigned)i, (unsigned)j, 0};
                        if (init) render_init(width, height, states);
                        else render(fb, width, height, ns, s->d_camera, states, s->d_world, max_depth, batch_id, s->background);
                    }
                }
            });
        for (auto& t : pool) t.join();
    };
    const auto t0 = std::chrono::steady_clock::now();
    run_rows(true);
    const auto t1 = std::chrono::steady_clock::now();
    run_rows(false);
    const auto t2 = std::chrono::steady_clock::now();
    g_last_init_s = std::chrono::duration<double>(t1 - t0).count();
    g_last_render_s = std::chrono::duration<double>(t2 - t1).count();
    free(states);
    return 0;
}
