/* oracle/_ref shim: minimal thrust::device_ptr on the host heap (test infrastructure). */
#pragma once
#include <cstdlib>
#include <cstring>
namespace thrust {
template <typename T> struct device_ptr {
    T* p = nullptr;
    device_ptr() {}
    explicit device_ptr(T* q) : p(q) {}
    T& operator[](size_t i) const { return p[i]; }
    T* get() const { return p; }
};
template <typename T> device_ptr<T> device_malloc(size_t n) { return device_ptr<T>((T*)calloc(n ? n : 1, sizeof(T))); }
template <typename T> void device_free(device_ptr<T>) { /* keep alive: the reference reads freed memory */ }
template <typename T> T* raw_pointer_cast(const device_ptr<T>& d) { return d.p; }
template <typename It, typename T> void copy(It b, It e, device_ptr<T> d) { size_t i = 0; for (; b != e; ++b) d.p[i++] = *b; }
}
