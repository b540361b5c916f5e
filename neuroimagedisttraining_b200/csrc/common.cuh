// Shared helpers for the sm_100a kernels of neuroimagedisttraining_b200.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

// Every extern "C" entry point returns 0 on success; on failure the message is kept
// per host thread and readable through nidt_last_error().
#define NIDT_API extern "C" __attribute__((visibility("default")))
void nidt_set_error(const char* fmt, ...);

#define NIDT_CUDA_CHECK(expr)                                                              \
    do {                                                                                   \
        cudaError_t e_ = (expr);                                                           \
        if (e_ != cudaSuccess) {                                                           \
            nidt_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); \
            return 1;                                                                      \
        }                                                                                  \
    } while (0)

#define NIDT_REQUIRE(cond, ...)                                                            \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            nidt_set_error(__VA_ARGS__);                                                   \
            return 2;                                                                      \
        }                                                                                  \
    } while (0)

#define NIDT_LAUNCH_CHECK() NIDT_CUDA_CHECK(cudaGetLastError())

namespace nidt {

constexpr int kNumSMs = 148;  // B200

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Deterministic block-wide sum (fixed tree); result valid in every thread.
template <int kThreads>
__device__ __forceinline__ float block_sum(float v, float* smem /* >= kThreads/32 floats */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kThreads / 32; ++i) t += smem[i];
    return t;
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace nidt
