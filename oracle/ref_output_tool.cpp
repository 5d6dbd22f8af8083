/* oracle/ref_output_tool.cpp — TEST INFRASTRUCTURE, not product code.
 *
 * The reference's own output stage (render.h:152-168) run from its unmodified color.h: each frame buffer goes
 * through write_frame_buffer (color.h:19-49: text PPM, gamma 2, 8 bit) into <dir_in>, then average_images
 * (color.h:57-170) re-parses the PPMs, un-gammas, averages and re-quantises; the png++ stand-in of oracle/shim
 * dumps the final pixels to <file_out> ("RAW8 w h\n" + bytes, row 0 = top).
 * usage: ref_output_tool <n_frames> <width> <height> <frames.f32: n x h x w x 3 linear floats, row 0 = bottom>
 *                        <dir_in (empty, existing)> <file_out (outside dir_in)>
 * Only tests/ may execute it.  Harness code written for this repo; no reference code is copied. */
#include "color.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

thread_local ref_idx3 ref_threadIdx, ref_blockIdx, ref_blockDim, ref_gridDim;
thread_local const float* ref_scripted_uniform = nullptr;

int main(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s n w h frames.f32 dir_in file_out\n", argv[0]); return 2; }
    const int n = atoi(argv[1]), w = atoi(argv[2]), h = atoi(argv[3]);
    const size_t npix = (size_t)w * h;
    std::vector<float> frames((size_t)n * npix * 3);
    FILE* f = fopen(argv[4], "rb");
    if (!f || fread(frames.data(), sizeof(float), frames.size(), f) != frames.size()) { fprintf(stderr, "cannot read %s\n", argv[4]); return 2; }
    fclose(f);
    for (int i = 0; i < n; i++)
        write_frame_buffer(std::string(argv[5]) + "/" + std::to_string(i) + ".ppm", (const color*)(frames.data() + (size_t)i * npix * 3), w, h);
    average_images(argv[5], argv[6]);
    return 0;
}
