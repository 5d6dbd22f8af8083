// Forwarding header: lets sources written against the reference's "color.h" build against the B200 core.
// write_frame_buffer (color.h:19-49) maps to rt_quantize_frame + rt_write_ppm; average_images is rt_average_frames
// inside rt_draw (include/rt_output.h).
#pragma once
#include "rt_compat_cuda.h"
#include "../rt_output.h"

inline void write_frame_buffer(std::string file_loc, const color* f, const int width, const int height, const int no_fb = 1) {
    std::vector<float> lin((size_t)width * height * 3);
    for (size_t k = 0; k < (size_t)width * height; k++) { lin[3 * k] = f[k].x() / no_fb; lin[3 * k + 1] = f[k].y() / no_fb; lin[3 * k + 2] = f[k].z() / no_fb; }
    std::vector<uint8_t> px(lin.size());
    rtc::ok(rt_quantize_frame(lin.data(), width, height, px.data()));
    rtc::ok(rt_write_ppm(file_loc.c_str(), px.data(), width, height));
}
