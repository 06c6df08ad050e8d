// Shared helpers for the be_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "be_b200.h"

#define BE_SM_COUNT 148  // B200: 2 dies x 74 SMs; grids for streaming kernels are sized in multiples of this

void be_set_error(const char* fmt, ...);

#define BE_REQUIRE(cond, code, ...) \
    do {                            \
        if (!(cond)) {              \
            be_set_error(__VA_ARGS__); \
            return (code);          \
        }                           \
    } while (0)

#define BE_CUDA(call)                                                                        \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            be_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return BE_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

#define BE_LAUNCH_CHECK()                                                                     \
    do {                                                                                      \
        cudaError_t e_ = cudaGetLastError();                                                  \
        if (e_ != cudaSuccess) {                                                              \
            be_set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return BE_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)

static inline size_t be_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Streaming loads/stores: data touched once goes around L1 so it does not evict the small tables.
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_f4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w));
}

// numpy fp32 sigmoid 1/(1+exp(-x)): exp rounded to fp32 (computed in fp64 so the rounding is the correct
// one), then an fp32 add and an fp32 divide, none contracted.
__device__ __forceinline__ float be_sigmoid_f32(float x) {
    float e = (float)exp(-(double)x);
    return __fdiv_rn(1.0f, __fadd_rn(1.0f, e));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
