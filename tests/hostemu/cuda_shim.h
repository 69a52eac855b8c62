// TEST INFRASTRUCTURE ONLY: lets g++ compile the per-thread __device__ functions of parsip_b200/csrc
// (interp.cuh, cull_walk.cuh) for the host, so the interpreter and the pruning walk can be unit-tested and
// debugged without a GPU. Nothing in the product links this; kernels (shared memory, barriers, TMA) are not emulated.
#ifndef PSGPU_TEST_CUDA_SHIM_H
#define PSGPU_TEST_CUDA_SHIM_H
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __align__(n) __attribute__((aligned(n)))
#define __launch_bounds__(...)
#define __constant__ static const
struct float4 { float x, y, z, w; };
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline float __saturatef(float x) { return x != x ? 0.f : fminf(fmaxf(x, 0.f), 1.f); }
static inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
static inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
static inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
static inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline float __fdiv_rn(float a, float b) { volatile float r = a / b; return r; }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline float __fsqrt_rn(float x) { volatile float r = sqrtf(x); return r; }
using std::max;
using std::min;
typedef int cudaStream_t;
#define PSGPU_HOST_EMULATION 1
#endif
