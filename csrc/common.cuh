// Shared device/host helpers for the b200dp kernel library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/b200dp.h"

#define B200DP_SMS 148

namespace b200dp {

void set_error(const char* fmt, ...);

#define B200DP_CHECK_LAUNCH()                                                     \
    do {                                                                          \
        cudaError_t _e = cudaPeekAtLastError();                                   \
        if (_e != cudaSuccess) {                                                  \
            b200dp::set_error("%s:%d launch failed: %s", __FILE__, __LINE__,      \
                              cudaGetErrorString(_e));                            \
            return (int)_e;                                                       \
        }                                                                         \
    } while (0)

#define B200DP_CUDA(call)                                                         \
    do {                                                                          \
        cudaError_t _e = (call);                                                  \
        if (_e != cudaSuccess) {                                                  \
            b200dp::set_error("%s:%d %s: %s", __FILE__, __LINE__, #call,          \
                              cudaGetErrorString(_e));                            \
            return (int)_e;                                                       \
        }                                                                         \
    } while (0)

#define B200DP_REQUIRE(cond, ...)                                                 \
    do {                                                                          \
        if (!(cond)) {                                                            \
            b200dp::set_error(__VA_ARGS__);                                       \
            return B200DP_ERR_INVALID;                                            \
        }                                                                         \
    } while (0)

// ---- dtype dispatch -------------------------------------------------------
template <typename T> struct Cvt;
template <> struct Cvt<float> {
    __device__ __forceinline__ static float to_f(float v) { return v; }
    __device__ __forceinline__ static float from_f(float v) { return v; }
};
template <> struct Cvt<__half> {
    __device__ __forceinline__ static float to_f(__half v) { return __half2float(v); }
    __device__ __forceinline__ static __half from_f(float v) { return __float2half_rn(v); }
};
template <> struct Cvt<__nv_bfloat16> {
    __device__ __forceinline__ static float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
    __device__ __forceinline__ static __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
};

// 8 x 16-bit values in one 16-byte vector
template <typename T> struct alignas(16) Vec8 { T v[8]; };
template <typename T> struct alignas(16) Vec4f { T v[4]; };

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&out)[8]) {
    Vec8<T> r = *reinterpret_cast<const Vec8<T>*>(p);
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = Cvt<T>::to_f(r.v[i]);
}
template <typename T>
__device__ __forceinline__ void store8(T* p, const float (&in)[8]) {
    Vec8<T> r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[i] = Cvt<T>::from_f(in[i]);
    *reinterpret_cast<Vec8<T>*>(p) = r;
}
template <>
__device__ __forceinline__ void load8<float>(const float* p, float (&out)[8]) {
    float4 a = reinterpret_cast<const float4*>(p)[0];
    float4 b = reinterpret_cast<const float4*>(p)[1];
    out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w;
    out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
}
template <>
__device__ __forceinline__ void store8<float>(float* p, const float (&in)[8]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(in[0], in[1], in[2], in[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(in[4], in[5], in[6], in[7]);
}

// two packed 16-bit values -> float2 (the float instantiation exists only so templates compile)
template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t u);
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t u) {
    return __half22float2(*reinterpret_cast<const __half2*>(&u));
}
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t u) {
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
template <> __device__ __forceinline__ float2 unpack2<float>(uint32_t u) { return make_float2(__uint_as_float(u), 0.f); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32). `red` needs 32 floats of smem.
__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
    if (w == 0) v = warp_sum(v);
    if (threadIdx.x == 0) red[0] = v;
    __syncthreads();
    return red[0];
}
__device__ __forceinline__ float block_max(float v, float* red) {
    v = warp_max(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? red[threadIdx.x] : -INFINITY;
    if (w == 0) v = warp_max(v);
    if (threadIdx.x == 0) red[0] = v;
    __syncthreads();
    return red[0];
}

__device__ __forceinline__ bool is_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace b200dp

#define B200DP_DISPATCH_DTYPE(dt, T, ...)                                         \
    switch (dt) {                                                                 \
        case B200DP_F32: { using T = float; __VA_ARGS__; break; }                 \
        case B200DP_F16: { using T = __half; __VA_ARGS__; break; }                \
        case B200DP_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }        \
        default: b200dp::set_error("unsupported dtype %d", (int)(dt));            \
                 return B200DP_ERR_INVALID;                                       \
    }
#define B200DP_DISPATCH_HALF(dt, T, ...)                                          \
    switch (dt) {                                                                 \
        case B200DP_F16: { using T = __half; __VA_ARGS__; break; }                \
        case B200DP_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }        \
        default: b200dp::set_error("expected fp16/bf16, got dtype %d", (int)(dt));\
                 return B200DP_ERR_INVALID;                                       \
    }
