// Forwarding header: lets sources written against the reference's "render.h" build against the B200 core.
// render_settings (render.h:21-50) keeps its fields and calc_* helpers; draw(scene&, render_settings) (render.h:118-174)
// hands the recorded world to the bound rt_scene and runs rt_draw (include/rt_output.h): no_fb batches, per-batch
// 8-bit quantisation, average_images' arithmetic, ./image.png — the file the reference writes, without the
// temporary PPM directory.  threads_x / threads_y are accepted and ignored (the wavefront kernels size themselves).
#pragma once
#include "rt_compat_cuda.h"
#include "../rt_output.h"

struct render_settings {
    double aspect_ratio = 0;
    unsigned image_width = 1200, image_height = 0;
    unsigned samples_per_pixel_per_fb = 100;
    unsigned no_fb = 10;
    unsigned threads_x = 8, threads_y = 8;
    unsigned max_depth = 50;

    unsigned rays_per_pixel = 0;
    unsigned num_pixels = 0;

    inline void calc_height() { image_height = static_cast<int>(image_width / aspect_ratio); }   // render.h:32-34
    inline void calc_rays_per_pixel() { rays_per_pixel = samples_per_pixel_per_fb * no_fb; }
    inline void calc_num_pixels() { num_pixels = image_width * image_height; }
    inline void calc_all() { calc_height(); calc_rays_per_pixel(); calc_num_pixels(); }
};

namespace rtc {
inline const char*& draw_output_path() { static const char* p = "image.png"; return p; }   // render.h:168 writes "image.png"
inline int& draw_device() { static int d = 0; return d; }
inline unsigned& draw_devices() { static unsigned n = RT_DRAW_ALL_DEVICES; return n; }   // draw() renders on every visible GPU by default
}

// Scene = the reference's `scene` (scenes.h:36-79) or anything with d_world / d_camera / background
template <class Scene>
void draw(Scene& curr_scene, render_settings settings) {
    rtc::finish(*curr_scene.d_world, *curr_scene.d_camera, curr_scene.background);
    rt_draw_settings st{};
    st.image_width = settings.image_width;
    st.image_height = settings.image_height;
    st.aspect_ratio = (float)settings.aspect_ratio;
    st.samples_per_pixel_per_fb = settings.samples_per_pixel_per_fb;
    st.no_fb = settings.no_fb;
    st.max_depth = settings.max_depth;
    st.seed = 1984;                                     // render.h:91
    st.device = (uint32_t)rtc::draw_device();
    st.n_devices = rtc::draw_devices();
    std::cerr << "\rRendering " << settings.no_fb << " frame buffers        " << std::flush;
    if (rt_draw(rtc::sc(), &st, rtc::draw_output_path(), nullptr, nullptr, nullptr) < 0) {
        std::cerr << "rt_b200 error: " << rt_last_error() << std::endl;        // the reference would exit(99) (common.h:30-38)
        throw std::runtime_error(rt_last_error());
    }
}
