// rng.cuh — counter-based Philox4x32-10 keyed on (pixel, sample, bounce): replaces the reference's per-pixel
// curandState / curand_init(1984, pixel, 0) (render.h:84-92).  Because a draw is a pure function of
// (seed, stream, pixel, sample, bounce, block) the image does not depend on how samples are partitioned over
// wavefront slots, kernels or GPUs.
#pragma once
#include <cstdint>

namespace rtb {

enum : uint32_t { RT_STREAM_CAMERA = 0x43414d00u, RT_STREAM_SCATTER = 0x53434100u, RT_STREAM_MEDIUM = 0x4d454400u };

struct RngKey { uint32_t pixel, sample, bounce, seed; };

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    #pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}

// uint32 -> (0,1] exactly as curand_uniform maps it (curand_uniform.h:69-72), one fused multiply-add
__device__ __forceinline__ float u01(uint32_t x) { return fmaf((float)x, 2.3283064e-10f, 1.1641532e-10f); }

__device__ __forceinline__ float4 draw4(const RngKey& g, uint32_t stream, uint32_t block) {
    const uint4 r = philox4x32_10(make_uint4(g.pixel, g.sample, g.bounce, block), make_uint2(g.seed, stream));
    return make_float4(u01(r.x), u01(r.y), u01(r.z), u01(r.w));
}

// random_float(s, min, max) — vec3.h:11
__device__ __forceinline__ float rf(float u, float mn, float mx) { return mn + (mx - mn) * u; }

} // namespace rtb
