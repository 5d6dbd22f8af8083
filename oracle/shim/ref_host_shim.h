/* TEST INFRASTRUCTURE (oracle/_ref build only) — host shim that lets g++ compile the
 * reference's CUDA headers (/root/reference/*.h) as plain C++.  Force-included (-include)
 * before anything else.  cuRAND comes from the real toolkit header (host-compilable), so the
 * XORWOW stream is NVIDIA's own code, not a replica.
 * Nothing here is product code and nothing here is derived from the reference sources. */
#pragma once
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <string>
#include <vector>
#include <curand_kernel.h>   /* real header; defines __device__/__host__/__global__ empty under g++ */

#ifndef __device__
#define __device__
#endif
#ifndef __host__
#define __host__
#endif
#ifndef __global__
#define __global__
#endif

/* --- CUDA runtime subset used by scenes.h / render.h, mapped onto the host heap.
 * cuda_runtime.h (pulled in by curand_kernel.h) already declares the real entry points, so the
 * reference's calls are redirected by macro; libcudart is never linked. --- */
static inline cudaError_t ref_cudaMalloc(void** p, size_t n) { *p = calloc(1, n ? n : 1); return cudaSuccess; }
static inline cudaError_t ref_cudaFree(void* p) { (void)p; return cudaSuccess; }   /* leak: the reference frees dangling ptrs */
static inline cudaError_t ref_cudaOk() { return cudaSuccess; }
#define cudaMalloc(p, n) ref_cudaMalloc((void**)(p), (n))
#define cudaMallocManaged(p, n) ref_cudaMalloc((void**)(p), (n))
#define cudaFree(p) ref_cudaFree((void*)(p))
#define cudaGetLastError() ref_cudaOk()
#define cudaDeviceSynchronize() ref_cudaOk()
#define cudaDeviceReset() ref_cudaOk()

/* --- built-in launch variables; the harness sets them per emulated thread --- */
struct ref_idx3 { unsigned x, y, z; };
extern thread_local ref_idx3 ref_threadIdx, ref_blockIdx, ref_blockDim, ref_gridDim;
#define threadIdx ref_threadIdx
#define blockIdx ref_blockIdx
#define blockDim ref_blockDim
#define gridDim ref_gridDim

/* first-hit parity: constant_medium's free-flight draw (constant_medium.h:57) can be scripted per ray */
extern thread_local const float* ref_scripted_uniform;
static inline float ref_curand_uniform(curandState* s) { return ref_scripted_uniform ? *ref_scripted_uniform : curand_uniform(s); }
#define curand_uniform ref_curand_uniform

/* material.h:78 uses unqualified min() (CUDA device overload) */
using std::min;
using std::max;
