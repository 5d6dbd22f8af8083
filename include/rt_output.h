/* rt_output.h — the reference's image OUTPUT stage behind a C-ABI (host only; part of librt_b200.so).
 *
 * draw() (render.h:118-174) does not hand back radiance: per frame-buffer batch it writes an 8-bit text PPM
 * (write_frame_buffer, color.h:19-49: gamma 2, clamp to 0.999, *256, top row first), then average_images
 * (color.h:57-170) re-parses the PPMs, maps every byte back with (v*v)/(255*255), averages over the batches,
 * applies gamma 2 / clamp / *256 again and writes ./image.png.  A caller that switches to this library and wants
 * the file the reference would have written gets it from the functions below, byte for byte
 * (tests/test_output.py checks them against the reference's own color.h code through oracle/_ref).
 */
#ifndef RT_OUTPUT_H
#define RT_OUTPUT_H

#include <stdint.h>
#include "rt_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* average_images' arithmetic (color.h:125-165) on n_frames quantised frames (each width*height*3 bytes as produced
 * by rt_quantize_frame, row 0 = top), summed in frame order.  out: width*height*3 bytes. */
int rt_average_frames(const uint8_t* frames_rgb8, int n_frames, int width, int height, uint8_t* out_rgb8);

/* Text PPM exactly as write_frame_buffer emits it ("P3\n<w> <h>\n255\n" then "r g b\n" per pixel; color.h:22-45). */
int rt_write_ppm(const char* path, const uint8_t* rgb8, int width, int height);
/* 8-bit RGB PNG (what png::image<png::rgb_pixel>::write produces, color.h:141-168); self-contained encoder
 * (stored deflate blocks), no libpng / zlib dependency. */
int rt_write_png(const char* path, const uint8_t* rgb8, int width, int height);

/* Linear float frame as a NumPy .npy file (version 1.0, little-endian float32, shape (height, width, 3), C order; row 0 =
 * bottom as rt_render returns it): the lossless dump parity work compares, next to the 8-bit files above. */
int rt_write_npy(const char* path, const float* rgb_linear, int width, int height);

/* ---- progressive rendering checkpoints -------------------------------------------------------------
 * A checkpoint is the accumulation buffer of rt_render(..., RT_RENDER_ACCUMULATE) plus what is needed to continue it:
 * which scene (rt_scene_hash), frame size, seed, depth and how many samples per pixel the sums hold.  Resuming = load,
 * then keep rendering with sample_first = samples_done.  File: 64-byte little-endian header + width*height*3 floats
 * + FNV-1a of the floats. */
typedef struct rt_checkpoint_info {
    uint32_t width, height;
    uint32_t seed, max_depth;
    uint32_t samples_done;       /* the sums hold sample indices [0, samples_done) of every pixel */
    uint32_t reserved0;
    uint64_t scene_hash;         /* rt_scene_hash of the scene the samples belong to */
    uint64_t reserved[3];
} rt_checkpoint_info;
int rt_checkpoint_save(const char* path, const rt_checkpoint_info* info, const float* sums_rgb);
/* sums_rgb may be NULL to read the header only; capacity_floats must be >= width*height*3 otherwise. */
int rt_checkpoint_load(const char* path, rt_checkpoint_info* info, float* sums_rgb, size_t capacity_floats);

typedef struct rt_draw_settings {      /* render_settings, render.h:21-30 */
    uint32_t image_width;              /* default 1200 */
    uint32_t image_height;             /* 0 = int(image_width / aspect_ratio), render.h:32-34 (one row short at 16:9) */
    float aspect_ratio;
    uint32_t samples_per_pixel_per_fb; /* default 100 */
    uint32_t no_fb;                    /* default 10 */
    uint32_t max_depth;                /* default 50 */
    uint32_t seed;                     /* Philox key (the reference hard-codes 1984, render.h:91) */
    uint32_t device;                   /* CUDA device used when the scene is not committed yet (n_devices <= 1) */
    uint32_t n_devices;                /* 0 / 1: one GPU (`device`); RT_DRAW_ALL_DEVICES: every visible GPU; n: devices 0 .. n-1.
                                          Each batch is then sample-split over the GPUs and reduced on the first (rt_render_multi) */
    uint32_t reserved[7];
} rt_draw_settings;
#define RT_DRAW_ALL_DEVICES 0xffffffffu

/* draw(scene, settings) (render.h:118-174) minus the temp directory: renders no_fb batches of
 * samples_per_pixel_per_fb samples each (batch i = sample indices [i*spp, (i+1)*spp)), quantises every batch mean,
 * averages the quantised batches and writes the PNG to png_path (NULL = no file).  out_rgb8_or_null receives the
 * final width*height*3 bytes (row 0 = top); out_height_or_null the image height used.  Needs a CUDA device. */
int rt_draw(rt_scene* scene, const rt_draw_settings* settings, const char* png_path, uint8_t* out_rgb8_or_null,
            uint32_t* out_height_or_null, rt_render_stats* stats_or_null);

#ifdef __cplusplus
}
#endif
#endif /* RT_OUTPUT_H */
