/* oracle/_ref shim: png++ is not installed.  color.h's PNG writer (color.h:141-168) compiles against this
 * stand-in, which keeps the pixels and writes them as a raw dump ("RAW8 w h\n" + w*h*3 bytes) so the harness can
 * read back exactly what the reference would have put into the PNG. */
#pragma once
#include <string>
#include <cstddef>
#include <cstdio>
#include <vector>
namespace png {
struct rgb_pixel { unsigned char red, green, blue; rgb_pixel(unsigned char r = 0, unsigned char g = 0, unsigned char b = 0) : red(r), green(g), blue(b) {} };
template <typename P> struct image {
    size_t w, h;
    std::vector<P> px;
    image(size_t w_, size_t h_) : w(w_), h(h_), px(w_ * h_) {}
    struct row { P* p; P& operator[](size_t i) { return p[i]; } };
    row operator[](size_t j) { return row{px.data() + j * w}; }
    size_t get_width() const { return w; } size_t get_height() const { return h; }
    void set_pixel(size_t x, size_t y, P v) { px[y * w + x] = v; }
    void write(const std::string& path) {
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) return;
        fprintf(f, "RAW8 %zu %zu\n", w, h);
        for (const P& v : px) { unsigned char b[3] = {v.red, v.green, v.blue}; fwrite(b, 1, 3, f); }
        fclose(f);
    }
};
}
