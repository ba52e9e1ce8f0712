// Shared device helpers for the sm_100a kernels (vector I/O, bf16 packing, warp reductions).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_imagenet.h"

#define B200_SM_COUNT 148  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

#define B200_CHECK_LAUNCH()                         \
    do {                                            \
        cudaError_t e__ = cudaPeekAtLastError();    \
        if (e__ != cudaSuccess) return (int)e__;    \
    } while (0)

static inline bool b200_misaligned(const void *p, size_t a = 16) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) != 0; }

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);  // .x = lo (low half), .y = hi
    return *reinterpret_cast<uint32_t *>(&v);
}

// Load/store VEC consecutive elements of T as floats.  Every access is one instruction:
// float x4 = 16 B, bf16 x8 = 16 B, bf16 x4 = 8 B.
template <typename T, int VEC>
struct VecIO;

template <>
struct VecIO<float, 4> {
    static __device__ __forceinline__ void load(const float *p, float (&v)[4]) {
        float4 t = *reinterpret_cast<const float4 *>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static __device__ __forceinline__ void load_stream(const float *p, float (&v)[4]) {
        float4 t = __ldcs(reinterpret_cast<const float4 *>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static __device__ __forceinline__ void store(float *p, const float (&v)[4]) {
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};

template <>
struct VecIO<__nv_bfloat16, 4> {
    static __device__ __forceinline__ void unpack(uint2 t, float (&v)[4]) {
        v[0] = bf16_lo(t.x); v[1] = bf16_hi(t.x); v[2] = bf16_lo(t.y); v[3] = bf16_hi(t.y);
    }
    static __device__ __forceinline__ void load(const __nv_bfloat16 *p, float (&v)[4]) {
        unpack(*reinterpret_cast<const uint2 *>(p), v);
    }
    static __device__ __forceinline__ void load_stream(const __nv_bfloat16 *p, float (&v)[4]) {
        unpack(__ldcs(reinterpret_cast<const uint2 *>(p)), v);
    }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, const float (&v)[4]) {
        *reinterpret_cast<uint2 *>(p) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
    }
};

template <>
struct VecIO<__nv_bfloat16, 8> {
    static __device__ __forceinline__ void unpack(uint4 t, float (&v)[8]) {
        v[0] = bf16_lo(t.x); v[1] = bf16_hi(t.x); v[2] = bf16_lo(t.y); v[3] = bf16_hi(t.y);
        v[4] = bf16_lo(t.z); v[5] = bf16_hi(t.z); v[6] = bf16_lo(t.w); v[7] = bf16_hi(t.w);
    }
    static __device__ __forceinline__ void load(const __nv_bfloat16 *p, float (&v)[8]) {
        unpack(*reinterpret_cast<const uint4 *>(p), v);
    }
    static __device__ __forceinline__ void load_stream(const __nv_bfloat16 *p, float (&v)[8]) {
        unpack(__ldcs(reinterpret_cast<const uint4 *>(p)), v);
    }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, const float (&v)[8]) {
        *reinterpret_cast<uint4 *>(p) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]),
                                                   pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    }
};

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
