// catalog.cpp — library scenes, written against the compat layer (include/rt_compat.h) exactly as a user of the
// reference's API would write them.  Geometry, materials, cameras and the order in which the XORWOW(1984,0,0)
// stream is consumed follow scenes.h (line references per scene) so the recorded scenes equal the reference's.
#include "../../include/rt_catalog.h"
#include "../../include/rt_compat.h"

#include <string>

namespace rtb { void set_error(const std::string& e); }

namespace {

const color SKY(0.7f, 0.8f, 1.0f), BLACK(0.0f, 0.0f, 0.0f);      // scenes.h:69-77
const float WIDE = 16.0f / 9.0f;

struct Built { hittable* world; camera* cam; color background; float aspect; };

curandState* world_rng() { curandState* s = new curandState; curand_init(1984, 0, 0, s); return s; }   // scenes.h:28-32

hittable* list_of(std::vector<hittable*>& v) {
    hittable** arr = new hittable*[v.size() ? v.size() : 1];
    for (size_t i = 0; i < v.size(); i++) arr[i] = v[i];
    return new hittable_list(arr, (int)v.size());
}

camera* look(vec3 from, vec3 at, float vfov, float aspect, float aperture) {
    return new camera(from, at, vec3(0, 1, 0), vfov, aspect, aperture, 10.0f, 0, 1);
}

Built basic() {                                                   // scenes.h:82-91
    std::vector<hittable*> o{new sphere(vec3(0, 0, -1), 0.5, new lambertian(vec3(0, 1, 0))),
                             new sphere(vec3(0, -100.5, -1), 100, new lambertian(vec3(0, 0, 1)))};
    return {list_of(o), look(vec3(0, 0, -3), vec3(0, 0, 0), 40, WIDE, 0.0f), SKY, WIDE};
}

Built first() {                                                   // scenes.h:106-128
    material* ground = new lambertian(color(0.8, 0.8, 0));
    material* center = new lambertian(color(0.1, 0.2, 0.3));
    material* left = new dielectric(2.5);
    material* right = new metal(color(0.8, 0.6, 0.2), 0.2);
    material* front = new dielectric(2);
    std::vector<hittable*> o{new sphere(point3(0.0, -100.5, -1.0), 100.0, ground), new sphere(point3(0.0, 0.0, -1.0), 0.5, center),
                             new sphere(point3(-1.0, 0.0, -1.0), 0.5, left),        new sphere(point3(1.0, 0.0, -1.0), 0.5, right),
                             new sphere(point3(0.0, 1.0, -0.75), 0.25, front),      new sphere(point3(0.0, 1.0, -0.75), -0.25, front)};
    return {list_of(o), look(vec3(-2, 2, -3), vec3(0, 0, -1), 20, WIDE, 0.0f), SKY, WIDE};
}

Built big1() {                                                    // scenes.h:140-199 (RTIOW random spheres)
    curandState* rs = world_rng();
    hittable** all = new hittable*[488];
    int k = 0;
    all[k++] = new sphere(point3(0, -1000, 0), 1000, new lambertian(new checker_texture(color(0.2f, 0.3f, 0.1f), color(0.9f, 0.9f, 0.9f))));
    for (int a = -11; a < 11; a++)
        for (int b = -11; b < 11; b++) {
            const float choose_mat = random_float(rs);
            point3 center;
            if (rtc::device_argument_order()) {   // scenes.h:152 passes both draws as constructor arguments: nvcc takes x first
                const float cx = a + 0.9f * random_float(rs); const float cz = b + 0.9f * random_float(rs);
                center = point3(cx, 0.2f, cz);
            } else center = point3(a + 0.9f * random_float(rs), 0.2f, b + 0.9f * random_float(rs));   // the host compiler's order
            if ((center - point3(4, 0.2, 0)).length_squared() > 0.9f * 0.9f) {
                if (choose_mat < 0.8) {
                    const color albedo = color::random(rs) * color::random(rs);
                    material* m = new lambertian(albedo);
                    const point3 center2 = center + vec3(0, random_float(rs, 0, 0.5f), 0);
                    all[k++] = new moving_sphere(center, center2, 0.0f, 1.0f, 0.2f, m);
                } else if (choose_mat < 0.95) {
                    const color albedo = color::random(rs, 0.5f, 1);
                    const float fuzz = random_float(rs, 0, 0.5f);
                    all[k++] = new sphere(center, 0.2f, new metal(albedo, fuzz));
                } else {
                    all[k++] = new sphere(center, 0.2f, new dielectric(1.5f));
                }
            } else {   // cells under the big metal ball become a far-away speck (scenes.h:178)
                all[k++] = new sphere(vec3(10000, -10000, 10000), 0.00001f, new lambertian(vec3(0, 0, 0)));
            }
        }
    all[k++] = new sphere(point3(0, 1, 0), 1.0f, new dielectric(1.5f));
    all[k++] = new sphere(point3(-4, 1, 0), 1.0f, new lambertian(color(0.4f, 0.2f, 0.1f)));
    all[k++] = new sphere(point3(4, 1, 0), 1.0f, new metal(color(0.7, 0.6, 0.5), 0.0));
    std::vector<hittable*> top{new bvh_node(all, 488, 0, 1, rs)};
    return {list_of(top), look(vec3(13.0f, 2.0f, -3.0f), vec3(0, 0, 0), 20, WIDE, 0.1f), SKY, WIDE};
}

Built two_spheres() {                                             // scenes.h:225-234
    texturez* checker = new checker_texture(color(0.2, 0.3, 0.1), color(0.9, 0.9, 0.9));
    std::vector<hittable*> o{new sphere(point3(0, -10, 0), 10, new lambertian(checker)), new sphere(point3(0, 10, 0), 10, new lambertian(checker))};
    return {list_of(o), look(vec3(13.0f, 2.0f, 3.0f), vec3(0, 0, 0), 20, WIDE, 0.1f), SKY, WIDE};
}

Built two_perlin() {                                              // scenes.h:248-259
    curandState* rs = world_rng();
    texturez* marble = new marble_texture(rs, 4);
    texturez* turb = new turbulent_texture(rs, 5);
    std::vector<hittable*> o{new sphere(point3(0, -1000, 0), 1000, new lambertian(marble)), new sphere(point3(0, 2, 0), 2, new lambertian(turb))};
    return {list_of(o), look(vec3(13.0f, 2.0f, 3.0f), vec3(0, 0, 0), 20, WIDE, 0.1f), SKY, WIDE};
}

texturez* image_from(const rt_catalog_params* p) {
    if (!p || !p->image) throw std::runtime_error("this scene needs rt_catalog_params::image");
    int w = p->image_width, h = p->image_height, b = p->image_bytes_per_pixel;
    return new image_texture(const_cast<unsigned char*>(p->image), &w, &h, &b, 0);
}

Built earth(const rt_catalog_params* p) {                         // scenes.h:278-293
    std::vector<hittable*> o{new sphere(point3(0, 0, 0), 2, new lambertian(image_from(p))),
                             new xy_rect(-5, 5, -3, 3, 6, new diffuse_light(color(4.0, 4.0, 4.0)))};
    return {list_of(o), look(vec3(13.0f, 0.0f, 3.0f), vec3(0, 0, 0), 20, WIDE, 0.1f), BLACK, WIDE};
}

void cornell_walls(std::vector<hittable*>& o, material*& white, float lx0, float lx1, float lz0, float lz1) {   // scenes.h:323-338
    material* red = new lambertian(color(0.65, 0.05, 0.05));
    white = new lambertian(color(0.73, 0.73, 0.73));
    material* green = new lambertian(color(0.12, 0.45, 0.15));
    material* light = new diffuse_light(color(15, 15, 15));
    o.push_back(new yz_rect(0, 555, 0, 555, 555, green));
    o.push_back(new yz_rect(0, 555, 0, 555, 0, red));
    o.push_back(new xz_rect(lx0, lx1, lz0, lz1, 554, light));
    o.push_back(new xz_rect(0, 555, 0, 555, 0, white));
    o.push_back(new xz_rect(0, 555, 0, 555, 555, white));
    o.push_back(new xy_rect(0, 555, 0, 555, 555, white));
}

hittable* cornell_block(float height, float angle, vec3 at, material* white) {   // scenes.h:342-349
    return new translate(new rotate_y(new box(point3(0, 0, 0), point3(165, height, 165), white), angle), at);
}

Built cornell(bool smoke) {                                       // scenes.h:323-353, :365-396
    std::vector<hittable*> o;
    material* white = nullptr;
    if (smoke) cornell_walls(o, white, 113, 443, 127, 432); else cornell_walls(o, white, 213, 343, 227, 332);
    hittable* tall = cornell_block(330, 15, vec3(265, 0, 295), white);
    hittable* small = cornell_block(165, -18, vec3(130, 0, 65), white);
    if (smoke) {
        o.push_back(new constant_medium(tall, 0.01, color(0, 0, 0)));
        o.push_back(new constant_medium(small, 0.01, color(1, 1, 1)));
    } else {
        o.push_back(tall);
        o.push_back(small);
    }
    return {list_of(o), look(vec3(278, 278, -800), vec3(278, 278, 0), 40, 1.0f, 0.0f), BLACK, 1.0f};
}

Built one_triangle() {                                            // scenes.h:409-418
    std::vector<hittable*> o{new triangle(vec3(-0.5f, 0, 0), vec3(0, 1, 10), vec3(0.f, 0, 0), 0, 0, 0, 1, 1, 0, new lambertian(vec3(0, 1, 0))),
                             new sphere(vec3(0, -100.5, -1), 100, new lambertian(vec3(0, 0, 1)))};
    return {list_of(o), look(vec3(0, 0, -3), vec3(0, 0, 0), 40, WIDE, 0.0f), SKY, WIDE};
}

Built four_triangles() {                                          // scenes.h:432-451
    curandState* rs = world_rng();
    hittable** t = new hittable*[4];
    t[0] = new triangle(vec3(-0.5f, 0, 0), vec3(0, 1, 10), vec3(0.5f, 0, 0), 0, 0, 0, 1, 1, 0, new lambertian(vec3(0, 1, 0)));
    t[1] = new triangle(vec3(0.5f, 0, 0), vec3(0, 1, 10), vec3(0.5f, 1, 0), 0, 0, 0, 1, 1, 0, new lambertian(vec3(1, 1, 0)));
    t[2] = new triangle(vec3(1.5f, 0, 0), vec3(0, 2, 10), vec3(1.5f, 1, 0), 0, 0, 0, 1, 1, 0, new lambertian(vec3(1, 1, 1)));
    t[3] = new triangle(vec3(1.5f, 0, 0), vec3(1.5f, 1, 10), vec3(1.5f, 0, 2), 0, 0, 0, 1, 1, 0, new lambertian(vec3(1, 1, 1)));
    std::vector<hittable*> o{new triangle_mesh(t, 4, 0, 1, rs), new sphere(vec3(0, -100.5, -1), 100, new lambertian(vec3(0, 0, 1)))};
    return {list_of(o), look(vec3(0, 0, -3), vec3(0, 0, 0), 40, WIDE, 0.0f), SKY, WIDE};
}

// door_scene / cup_scene shape (scenes.h:478-487, :576-585): one textured triangle_mesh + the green ground sphere
Built mesh_scene(const rt_catalog_params* p) {
    if (!p || !p->vertices || !p->indices || !p->uvs || !p->normals || !p->camera15) throw std::runtime_error("mesh scene needs arrays + camera15");
    material* mat = new lambertian(image_from(p));
    const int mesh = rtc::ok(rt_mesh(rtc::sc(), p->vertices, p->n_vertices, p->indices, p->n_triangles, p->uvs, p->normals, mat->h));
    hittable* m = new hittable; m->h = mesh;
    std::vector<hittable*> o{m, new sphere(vec3(0, -100, -1), 100, new lambertian(vec3(0, 1, 0)))};
    const float* c = p->camera15;
    camera* cam = new camera(vec3(c[0], c[1], c[2]), vec3(c[3], c[4], c[5]), vec3(c[6], c[7], c[8]), c[9], c[10], c[11], c[12], c[13], c[14]);
    return {list_of(o), cam, SKY, c[10]};
}

// Next-Week final scene (absent from the reference; Shirley, "The Next Week", final scene).
// Draw order on XORWOW(1984,0,0): 400 box heights (i-major), 1000 sphere positions (x, y, z), Perlin tables.
Built final_scene(const rt_catalog_params* p) {
    const float aspect = (p && p->aspect > 0) ? p->aspect : 1.0f;
    curandState* rs = world_rng();
    material* ground = new lambertian(color(0.48f, 0.83f, 0.53f));
    hittable** floor_boxes = new hittable*[400];
    int nb = 0;
    for (int i = 0; i < 20; i++)
        for (int j = 0; j < 20; j++) {
            const float w = 100.0f, x0 = -1000.0f + i * w, z0 = -1000.0f + j * w;
            const float y1 = random_float(rs, 1, 101);
            floor_boxes[nb++] = new box(point3(x0, 0.0f, z0), point3(x0 + w, y1, z0 + w), ground);
        }
    material* white = new lambertian(color(0.73f, 0.73f, 0.73f));
    hittable** cluster = new hittable*[1000];
    for (int j = 0; j < 1000; j++) {
        const float x = random_float(rs, 0, 165);
        const float y = random_float(rs, 0, 165);
        const float z = random_float(rs, 0, 165);
        cluster[j] = new sphere(point3(x, y, z), 10, white);
    }
    texturez* pertext = new noise_texture(rs, 0.1f);

    std::vector<hittable*> o;
    o.push_back(new bvh_node(floor_boxes, nb, 0, 1, rs));
    o.push_back(new xz_rect(123, 423, 147, 412, 554, new diffuse_light(color(7, 7, 7))));
    const point3 c1(400, 400, 200);
    o.push_back(new moving_sphere(c1, c1 + vec3(30, 0, 0), 0, 1, 50, new lambertian(color(0.7f, 0.3f, 0.1f))));
    o.push_back(new sphere(point3(260, 150, 45), 50, new dielectric(1.5f)));
    o.push_back(new sphere(point3(0, 150, 145), 50, new metal(color(0.8f, 0.8f, 0.9f), 1.0f)));
    hittable* boundary = new sphere(point3(360, 150, 145), 70, new dielectric(1.5f));
    o.push_back(boundary);
    o.push_back(new constant_medium(boundary, 0.2f, color(0.2f, 0.4f, 0.9f)));       // inert under reference semantics
    o.push_back(new constant_medium(new sphere(point3(0, 0, 0), 5000, new dielectric(1.5f)), 0.0001f, color(1, 1, 1)));
    o.push_back(new sphere(point3(400, 200, 400), 100, new lambertian(image_from(p))));
    o.push_back(new sphere(point3(220, 280, 300), 80, new lambertian(pertext)));
    o.push_back(new translate(new rotate_y(new bvh_node(cluster, 1000, 0, 1, rs), 15), vec3(-100, 270, 395)));
    return {list_of(o), look(vec3(478, 278, -600), vec3(278, 278, 0), 40, aspect, 0.0f), BLACK, aspect};
}

} // namespace

extern "C" int rt_catalog_scene(rt_scene* s, const char* name, const rt_catalog_params* p, float* aspect_out) {
    if (!s || !name) { rtb::set_error("catalog: null argument"); return RT_ERR_INVALID; }
    try {
        rtc::scope bind(s);
        struct OrderGuard { bool prev; OrderGuard(bool v) : prev(rtc::device_argument_order()) { rtc::device_argument_order() = v; }
                            ~OrderGuard() { rtc::device_argument_order() = prev; } } order(!(p && p->host_argument_order));
        const std::string n(name);
        Built b;
        if (n == "basic") b = basic();
        else if (n == "first") b = first();
        else if (n == "big1") b = big1();
        else if (n == "two_spheres") b = two_spheres();
        else if (n == "two_perlin") b = two_perlin();
        else if (n == "earth") b = earth(p);
        else if (n == "cornell") b = cornell(false);
        else if (n == "cornell_smoke") b = cornell(true);
        else if (n == "triangle") b = one_triangle();
        else if (n == "triangles") b = four_triangles();
        else if (n == "mesh") b = mesh_scene(p);
        else if (n == "final") b = final_scene(p);
        else { rtb::set_error("catalog: unknown scene '" + n + "'"); return RT_ERR_INVALID; }
        rtc::finish(b.world, b.cam, b.background);
        if (aspect_out) *aspect_out = b.aspect;
        return RT_OK;
    } catch (const std::exception& e) {
        rtb::set_error(std::string("catalog: ") + e.what());
        return RT_ERR_INVALID;
    }
}
