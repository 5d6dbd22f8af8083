// output.cpp — the reference's image output stage (include/rt_output.h): color.h:19-49, :57-170, render.h:152-168.
#include "../../include/rt_output.h"
#include "device_scene.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

using namespace rtb;

namespace {

int fail(int code, const std::string& msg) { set_error(msg); return code; }

// CRC-32 table, built once during static initialisation (thread-safe: rt_write_png may run on several threads)
struct CrcTable {
    uint32_t t[256];
    CrcTable() { for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; t[i] = c; } }
};
const CrcTable crc_table;
uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n) {
    for (size_t i = 0; i < n; i++) crc = crc_table.t[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return crc;
}
void put_be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
void put_chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
    put_be32(out, (uint32_t)data.size());
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put_be32(out, crc32_update(0xffffffffu, out.data() + start, out.size() - start) ^ 0xffffffffu);
}

} // namespace

extern "C" {

int rt_average_frames(const uint8_t* frames, int n_frames, int width, int height, uint8_t* out) {
    if (!frames || !out || n_frames <= 0 || width <= 0 || height <= 0) return fail(RT_ERR_INVALID, "average_frames: bad argument");
    const size_t n = (size_t)width * height * 3;
    std::vector<float> fb(n, 0.0f);
    for (int f = 0; f < n_frames; f++) {
        const uint8_t* src = frames + (size_t)f * n;
        for (size_t k = 0; k < n; k++) {
            const int v = src[k];
            fb[k] += (v * v) / (255.0f * 255.0f);          // color.h:125 (int product, float quotient), :127
        }
    }
    for (size_t k = 0; k < n; k++) {
        float v = fb[k];
        v /= n_frames;                                      // color.h:150-152
        v = std::sqrt(v);                                   // color.h:156-158
        v = v < 0.0f ? 0.0f : (v > 0.999f ? 0.999f : v);    // color.h:161-163
        out[k] = (uint8_t)static_cast<int>(256 * v);
    }
    return RT_OK;
}

int rt_write_npy(const char* path, const float* rgb, int width, int height) {
    if (!path || !rgb || width <= 0 || height <= 0) return fail(RT_ERR_INVALID, "write_npy: bad argument");
    std::string dict = "{'descr': '<f4', 'fortran_order': False, 'shape': (" + std::to_string(height) + ", " + std::to_string(width) + ", 3), }";
    while ((10 + dict.size() + 1) % 64 != 0) dict += ' ';   // magic 6 + version 2 + length 2 + dict + newline: data starts 64-byte aligned
    dict += '\n';
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(RT_ERR_INVALID, std::string("write_npy: cannot open ") + path);
    const uint8_t head[10] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0, (uint8_t)(dict.size() & 0xff), (uint8_t)(dict.size() >> 8)};
    const size_t n = (size_t)width * height * 3;
    const bool ok = std::fwrite(head, 1, 10, f) == 10 && std::fwrite(dict.data(), 1, dict.size(), f) == dict.size() && std::fwrite(rgb, sizeof(float), n, f) == n;
    if (std::fclose(f) != 0 || !ok) return fail(RT_ERR_INVALID, std::string("write_npy: write failed: ") + path);
    return RT_OK;
}

// ---- checkpoints ---------------------------------------------------------------------------------------------------
namespace {
struct CheckpointHeader {            // 64 bytes, little-endian
    char magic[8];                   // "RTB2CKPT"
    uint32_t version, width, height, seed, max_depth, samples_done;
    uint64_t scene_hash;
    uint64_t payload_hash;           // FNV-1a of the float payload
    uint64_t reserved[2];
};
static_assert(sizeof(CheckpointHeader) == 64, "checkpoint header layout");
uint64_t fnv1a(const void* p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    const uint8_t* b = static_cast<const uint8_t*>(p);
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}
} // namespace

int rt_checkpoint_save(const char* path, const rt_checkpoint_info* info, const float* sums) {
    if (!path || !info || !sums || info->width == 0 || info->height == 0) return fail(RT_ERR_INVALID, "checkpoint_save: bad argument");
    const size_t n = (size_t)info->width * info->height * 3;
    CheckpointHeader h{};
    std::memcpy(h.magic, "RTB2CKPT", 8);
    h.version = 1; h.width = info->width; h.height = info->height; h.seed = info->seed; h.max_depth = info->max_depth;
    h.samples_done = info->samples_done; h.scene_hash = info->scene_hash;
    h.payload_hash = fnv1a(sums, n * sizeof(float));
    // written next to the target and renamed, so an interrupted save never leaves a truncated checkpoint behind
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return fail(RT_ERR_INVALID, "checkpoint_save: cannot open " + tmp);
    const bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && std::fwrite(sums, sizeof(float), n, f) == n;
    if (std::fclose(f) != 0 || !ok) { std::remove(tmp.c_str()); return fail(RT_ERR_INVALID, "checkpoint_save: write failed: " + tmp); }
    if (std::rename(tmp.c_str(), path) != 0) { std::remove(tmp.c_str()); return fail(RT_ERR_INVALID, std::string("checkpoint_save: cannot rename to ") + path); }
    return RT_OK;
}

int rt_checkpoint_load(const char* path, rt_checkpoint_info* info, float* sums, size_t capacity_floats) {
    if (!path || !info) return fail(RT_ERR_INVALID, "checkpoint_load: bad argument");
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(RT_ERR_INVALID, std::string("checkpoint_load: cannot open ") + path);
    CheckpointHeader h{};
    if (std::fread(&h, sizeof(h), 1, f) != 1 || std::memcmp(h.magic, "RTB2CKPT", 8) != 0 || h.version != 1 || h.width == 0 || h.height == 0) {
        std::fclose(f);
        return fail(RT_ERR_INVALID, std::string("checkpoint_load: not a checkpoint: ") + path);
    }
    std::memset(info, 0, sizeof(*info));
    info->width = h.width; info->height = h.height; info->seed = h.seed; info->max_depth = h.max_depth;
    info->samples_done = h.samples_done; info->scene_hash = h.scene_hash;
    if (sums) {
        const size_t n = (size_t)h.width * h.height * 3;
        if (capacity_floats < n) { std::fclose(f); return fail(RT_ERR_INVALID, "checkpoint_load: buffer too small"); }
        const bool ok = std::fread(sums, sizeof(float), n, f) == n;
        if (!ok || fnv1a(sums, n * sizeof(float)) != h.payload_hash) { std::fclose(f); return fail(RT_ERR_INVALID, std::string("checkpoint_load: truncated or corrupt: ") + path); }
    }
    std::fclose(f);
    return RT_OK;
}

int rt_write_ppm(const char* path, const uint8_t* rgb8, int width, int height) {
    if (!path || !rgb8 || width <= 0 || height <= 0) return fail(RT_ERR_INVALID, "write_ppm: bad argument");
    FILE* f = fopen(path, "w");
    if (!f) return fail(RT_ERR_INVALID, std::string("write_ppm: cannot open ") + path);
    fprintf(f, "P3\n%d %d\n255\n", width, height);
    const size_t npix = (size_t)width * height;
    for (size_t k = 0; k < npix; k++) fprintf(f, "%d %d %d\n", rgb8[3 * k], rgb8[3 * k + 1], rgb8[3 * k + 2]);
    fclose(f);
    return RT_OK;
}

int rt_write_png(const char* path, const uint8_t* rgb8, int width, int height) {
    if (!path || !rgb8 || width <= 0 || height <= 0) return fail(RT_ERR_INVALID, "write_png: bad argument");
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, (uint32_t)width); put_be32(ihdr, (uint32_t)height);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8-bit RGB, no interlace
    put_chunk(out, "IHDR", ihdr);
    // raw scanlines (filter byte 0 each) inside a zlib stream of stored deflate blocks
    const size_t stride = (size_t)width * 3 + 1;
    std::vector<uint8_t> raw(stride * height);
    for (int j = 0; j < height; j++) { raw[j * stride] = 0; memcpy(&raw[j * stride + 1], rgb8 + (size_t)j * width * 3, (size_t)width * 3); }
    std::vector<uint8_t> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (size_t off = 0; off < raw.size() || off == 0; ) {
        const size_t len = raw.size() - off < 65535 ? raw.size() - off : 65535;
        const bool last = off + len >= raw.size();
        z.push_back(last ? 1 : 0);
        z.push_back(len & 0xff); z.push_back(len >> 8); z.push_back(~len & 0xff); z.push_back((~len >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + len);
        for (size_t i = 0; i < len; i++) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
        off += len;
        if (last) break;
    }
    put_be32(z, (b << 16) | a);
    put_chunk(out, "IDAT", z);
    put_chunk(out, "IEND", {});
    FILE* f = fopen(path, "wb");
    if (!f) return fail(RT_ERR_INVALID, std::string("write_png: cannot open ") + path);
    const size_t w = fwrite(out.data(), 1, out.size(), f);
    fclose(f);
    return w == out.size() ? RT_OK : fail(RT_ERR_INVALID, "write_png: short write");
}

int rt_draw(rt_scene* scene, const rt_draw_settings* st, const char* png_path, uint8_t* out_rgb8, uint32_t* out_height, rt_render_stats* stats) {
    if (!scene || !st) return fail(RT_ERR_INVALID, "draw: bad argument");
    const uint32_t width = st->image_width ? st->image_width : 1200u;
    const uint32_t spp = st->samples_per_pixel_per_fb ? st->samples_per_pixel_per_fb : 100u;
    const uint32_t no_fb = st->no_fb ? st->no_fb : 10u;
    uint32_t height = st->image_height;
    if (!height) {
        if (!(st->aspect_ratio > 0.0f)) return fail(RT_ERR_INVALID, "draw: image_height or aspect_ratio must be given");
        const int h = rt_image_height((int)width, st->aspect_ratio);
        if (h <= 0) return fail(RT_ERR_INVALID, "draw: computed image height is not positive");
        height = (uint32_t)h;
    }
    const bool multi = st->n_devices > 1;
    rt_multi_settings mset{};
    if (multi && st->n_devices != RT_DRAW_ALL_DEVICES) {
        if (st->n_devices > RT_MAX_DEVICES) return fail(RT_ERR_INVALID, "draw: n_devices exceeds RT_MAX_DEVICES");
        mset.n_devices = st->n_devices;
        for (uint32_t k = 0; k < st->n_devices; k++) mset.device_ids[k] = (int32_t)k;
    }
    if (!multi && !device_scene_is_current(scene)) { const int rc = rt_scene_commit(scene, (int)st->device, nullptr); if (rc < 0) return rc; }   // draw() takes the scene as it is now
    const size_t n = (size_t)width * height * 3;
    std::vector<float> fb(n);
    std::vector<uint8_t> frames(n * no_fb), final_px(n);
    rt_render_stats total{}, one{};
    for (uint32_t i = 0; i < no_fb; i++) {                  // render.h:152-162
        rt_render_settings rs{};
        rs.width = width; rs.height = height; rs.samples_per_pixel = spp * no_fb; rs.max_depth = st->max_depth ? st->max_depth : 50u;
        rs.seed = st->seed; rs.sample_first = i * spp; rs.sample_count = spp;
        int rc;
        if (multi) {   // the batch's sample range is split over the GPUs and reduced on the first
            rt_multi_stats mst{};
            rc = rt_render_multi(scene, &rs, &mset, fb.data(), &mst);
            one = rt_render_stats{};
            one.samples = mst.samples; one.rays = mst.rays; one.ms_total = mst.ms_wall;
        } else rc = rt_render(scene, &rs, fb.data(), &one);                    // mean of this batch (render.h:112)
        if (rc < 0) return rc;
        total.samples += one.samples; total.rays += one.rays; total.iterations += one.iterations; total.kernel_launches += one.kernel_launches;
        total.ms_total += one.ms_total;
        const int qc = rt_quantize_frame(fb.data(), (int)width, (int)height, frames.data() + (size_t)i * n);   // color.h:19-49
        if (qc < 0) return qc;
    }
    int rc = rt_average_frames(frames.data(), (int)no_fb, (int)width, (int)height, final_px.data());           // color.h:57-170
    if (rc < 0) return rc;
    if (png_path && (rc = rt_write_png(png_path, final_px.data(), (int)width, (int)height)) < 0) return rc;
    if (out_rgb8) memcpy(out_rgb8, final_px.data(), n);
    if (out_height) *out_height = height;
    if (stats) *stats = total;
    return RT_OK;
}

} // extern "C"
