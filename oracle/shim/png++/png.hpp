/* oracle/_ref shim: png++ is not installed; color.h's PNG writer only needs to compile. */
#pragma once
#include <string>
#include <cstddef>
namespace png {
struct rgb_pixel { unsigned char red, green, blue; rgb_pixel(unsigned char r = 0, unsigned char g = 0, unsigned char b = 0) : red(r), green(g), blue(b) {} };
template <typename P> struct image {
    image(size_t, size_t) {}
    struct row { P dummy; P& operator[](size_t) { return dummy; } };
    row r; row& operator[](size_t) { return r; }
    size_t get_width() const { return 0; } size_t get_height() const { return 0; }
    void set_pixel(size_t, size_t, P) {}
    void write(const std::string&) {}
};
}
