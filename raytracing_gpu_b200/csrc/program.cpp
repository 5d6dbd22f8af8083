// program.cpp — rt_scene_replay: decode a scene program (include/rt_program.h) into rt_* calls.
#include "scene_host.h"

#include <cstring>
#include <string>
#include <vector>

namespace rtb { void set_error(const std::string& e); }

namespace {

struct Reader {
    const uint32_t* w;
    size_t n, pos = 0;
    bool ok = true;
    bool need(size_t k) { if (pos + k > n) { ok = false; return false; } return true; }
    uint32_t u() { if (!need(1)) return 0; return w[pos++]; }
    int i() { return (int)u(); }
    float f() { const uint32_t b = u(); float v; std::memcpy(&v, &b, 4); return v; }
    void fv(float* d, int k) { for (int j = 0; j < k; j++) d[j] = f(); }
};

} // namespace

extern "C" int rt_scene_replay(rt_scene* s, const uint32_t* words, size_t n_words) {
    if (!s || !words || n_words < 2) { rtb::set_error("replay: empty program"); return RT_ERR_INVALID; }
    Reader r{words, n_words};
    if (r.u() != RTP_MAGIC) { rtb::set_error("replay: bad magic"); return RT_ERR_INVALID; }
    // per-program index spaces -> scene indices (a scene may already hold resources)
    std::vector<int> tex, mat, hit, perlin, image;
    auto T = [&](int k) { return (k >= 0 && (size_t)k < tex.size()) ? tex[k] : -1; };
    auto M = [&](int k) { return (k >= 0 && (size_t)k < mat.size()) ? mat[k] : -1; };
    auto H = [&](int k) { return (k >= 0 && (size_t)k < hit.size()) ? hit[k] : -1; };
    auto P = [&](int k) { return (k >= 0 && (size_t)k < perlin.size()) ? perlin[k] : -1; };
    auto I = [&](int k) { return (k >= 0 && (size_t)k < image.size()) ? image[k] : -1; };
    int rc = RT_OK;
#define PUSH(vec, call) do { const int _r = (call); if (_r < 0) return _r; (vec).push_back(_r); } while (0)
    for (;;) {
        const uint32_t op = r.u();
        if (!r.ok) break;
        if (op == RTP_END) return RT_OK;
        switch (op) {
        case RTP_PERLIN: {
            if (!r.need(1536)) break;
            float rv[768]; int32_t pm[768];
            r.fv(rv, 768);
            for (int k = 0; k < 768; k++) pm[k] = r.i();
            PUSH(perlin, rt_perlin_tables(s, rv, pm));
            break;
        }
        case RTP_IMAGE: {
            const int w = r.i(), h = r.i(), bpp = r.i();
            const uint32_t nbytes = r.u();
            const size_t nw = ((size_t)nbytes + 3) / 4;
            if (!r.need(nw)) break;
            if ((size_t)w * h * bpp != nbytes) { rtb::set_error("replay: image size mismatch"); return RT_ERR_INVALID; }
            PUSH(image, rt_image(s, (const uint8_t*)(r.w + r.pos), w, h, bpp));
            r.pos += nw;
            break;
        }
        case RTP_TEX_SOLID: { float c[3]; r.fv(c, 3); PUSH(tex, rt_tex_solid(s, c[0], c[1], c[2])); break; }
        case RTP_TEX_CHECKER: { const int e = r.i(), o = r.i(); PUSH(tex, rt_tex_checker(s, T(e), T(o))); break; }
        case RTP_TEX_NOISE: { const int p = r.i(); const float sc = r.f(); PUSH(tex, rt_tex_noise(s, P(p), sc)); break; }
        case RTP_TEX_TURB: { const int p = r.i(); const float sc = r.f(); const int d = r.i(); PUSH(tex, rt_tex_turbulent(s, P(p), sc, d)); break; }
        case RTP_TEX_MARBLE: { const int p = r.i(); const float sc = r.f(); PUSH(tex, rt_tex_marble(s, P(p), sc)); break; }
        case RTP_TEX_IMAGE: { const int k = r.i(); PUSH(tex, rt_tex_image(s, I(k))); break; }
        case RTP_MAT_LAMBERTIAN: { const int t = r.i(); PUSH(mat, rt_mat_lambertian(s, T(t))); break; }
        case RTP_MAT_METAL: { const int t = r.i(); const float fz = r.f(); PUSH(mat, rt_mat_metal(s, T(t), fz)); break; }
        case RTP_MAT_DIELECTRIC: { const float ir = r.f(); PUSH(mat, rt_mat_dielectric(s, ir)); break; }
        case RTP_MAT_DIFFUSE_LIGHT: { const int t = r.i(); PUSH(mat, rt_mat_diffuse_light(s, T(t))); break; }
        case RTP_MAT_ISOTROPIC: { const int t = r.i(); PUSH(mat, rt_mat_isotropic(s, T(t))); break; }
        case RTP_SPHERE: { float a[4]; r.fv(a, 4); const int m = r.i(); PUSH(hit, rt_sphere(s, a[0], a[1], a[2], a[3], M(m))); break; }
        case RTP_MOVING_SPHERE: { float a[9]; r.fv(a, 9); const int m = r.i(); PUSH(hit, rt_moving_sphere(s, a, a + 3, a[6], a[7], a[8], M(m))); break; }
        case RTP_XY_RECT: { float a[5]; r.fv(a, 5); const int m = r.i(); PUSH(hit, rt_xy_rect(s, a[0], a[1], a[2], a[3], a[4], M(m))); break; }
        case RTP_XZ_RECT: { float a[5]; r.fv(a, 5); const int m = r.i(); PUSH(hit, rt_xz_rect(s, a[0], a[1], a[2], a[3], a[4], M(m))); break; }
        case RTP_YZ_RECT: { float a[5]; r.fv(a, 5); const int m = r.i(); PUSH(hit, rt_yz_rect(s, a[0], a[1], a[2], a[3], a[4], M(m))); break; }
        case RTP_BOX: { float a[6]; r.fv(a, 6); const int m = r.i(); PUSH(hit, rt_box(s, a, a + 3, M(m))); break; }
        case RTP_TRIANGLE: { float v[9], uv[6]; r.fv(v, 9); r.fv(uv, 6); const int m = r.i(); PUSH(hit, rt_triangle(s, v, uv, nullptr, M(m))); break; }
        case RTP_TRIANGLE_N: { float v[9], nn[9], uv[6]; r.fv(v, 9); r.fv(nn, 9); r.fv(uv, 6); const int m = r.i(); PUSH(hit, rt_triangle(s, v, uv, nn, M(m))); break; }
        case RTP_TRANSLATE: { const int c = r.i(); float d[3]; r.fv(d, 3); PUSH(hit, rt_translate(s, H(c), d[0], d[1], d[2])); break; }
        case RTP_ROTATE_Y: { const int c = r.i(); const float a = r.f(); PUSH(hit, rt_rotate_y(s, H(c), a)); break; }
        case RTP_ROTATE_Y_SC: { const int c = r.i(); const float sn = r.f(), cs = r.f(); PUSH(hit, rt_rotate_y_sincos(s, H(c), sn, cs)); break; }
        case RTP_MEDIUM: { const int b = r.i(); const float d = r.f(); const int t = r.i(); PUSH(hit, rt_constant_medium(s, H(b), d, T(t))); break; }
        case RTP_MEDIUM_NID: { const int b = r.i(); const float d = r.f(); const int t = r.i(); PUSH(hit, rt_constant_medium_nid(s, H(b), d, T(t))); break; }
        case RTP_LIST:
        case RTP_BVH: {
            const uint32_t cnt = r.u();
            if (!r.need(cnt)) break;
            std::vector<int> kids(cnt);
            for (uint32_t k = 0; k < cnt; k++) kids[k] = H(r.i());
            PUSH(hit, op == RTP_LIST ? rt_list(s, kids.data(), (int)cnt) : rt_bvh(s, kids.data(), (int)cnt));
            break;
        }
        case RTP_WORLD: { const int h = r.i(); rc = rt_set_world(s, H(h)); if (rc < 0) return rc; break; }
        case RTP_BACKGROUND: { float c[3]; r.fv(c, 3); rc = rt_set_background(s, c[0], c[1], c[2]); if (rc < 0) return rc; break; }
        case RTP_CAMERA: {
            float a[15]; r.fv(a, 15);
            rc = rt_set_camera(s, a, a + 3, a + 6, a[9], a[10], a[11], a[12], a[13], a[14]);
            if (rc < 0) return rc;
            break;
        }
        case RTP_CAMERA_RAW: {
            rt_flat_camera c;
            r.fv((float*)&c, 24);
            rc = rt_set_camera_raw(s, &c);
            if (rc < 0) return rc;
            break;
        }
        default:
            rtb::set_error("replay: unknown opcode " + std::to_string(op));
            return RT_ERR_INVALID;
        }
        if (!r.ok) break;
    }
#undef PUSH
    rtb::set_error("replay: truncated program");
    return RT_ERR_INVALID;
}
