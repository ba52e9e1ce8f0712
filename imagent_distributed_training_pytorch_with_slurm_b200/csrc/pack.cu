// K5: gradient packing - many weight-gradient tensors -> their slots in the flat float32 gradient arena,
// one launch.  Reference site: DDP's Reducer copies every parameter's gradient into its bucket view
// (reducer.hpp:347-406; `bucket_view.copy_(grad)` / `mul_out(grad, 1/W)`), one small kernel per parameter.
// Here cuDNN's wgrad outputs (bf16 or float32, K,R,S,C order = the arena's order) are cast-copied by
// ONE kernel per flush; the pointer table travels in the kernel's parameter space (<= 64 tensors per
// launch), so no device-side table has to be allocated or copied.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxTensors = 64;

struct PackTable {
    const void *src[kMaxTensors];
    long long off[kMaxTensors];
    long long cnt[kMaxTensors];
};

template <typename T>
__global__ void __launch_bounds__(kThreads) pack_grads_kernel(const __grid_constant__ PackTable tab,
                                                              float *__restrict__ dst, float scale) {
    const int t = blockIdx.y;
    const T *src = static_cast<const T *>(tab.src[t]);
    float *out = dst + tab.off[t];
    const long long n = tab.cnt[t];
    const long long n8 = n >> 3;
    for (long long v = (long long)blockIdx.x * kThreads + threadIdx.x; v < n8; v += (long long)gridDim.x * kThreads) {
        float e[8];
        if constexpr (sizeof(T) == 2) {
            VecIO<__nv_bfloat16, 8>::load(reinterpret_cast<const __nv_bfloat16 *>(src) + v * 8, e);
        } else {
            float a[4], b[4];
            VecIO<float, 4>::load(reinterpret_cast<const float *>(src) + v * 8, a);
            VecIO<float, 4>::load(reinterpret_cast<const float *>(src) + v * 8 + 4, b);
#pragma unroll
            for (int k = 0; k < 4; ++k) { e[k] = a[k]; e[4 + k] = b[k]; }
        }
        float a[4] = {e[0] * scale, e[1] * scale, e[2] * scale, e[3] * scale};
        float b[4] = {e[4] * scale, e[5] * scale, e[6] * scale, e[7] * scale};
        VecIO<float, 4>::store(out + v * 8, a);
        VecIO<float, 4>::store(out + v * 8 + 4, b);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
        const long long i = (n8 << 3) + threadIdx.x;
        float v;
        if constexpr (sizeof(T) == 2) v = __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(src)[i]);
        else v = reinterpret_cast<const float *>(src)[i];
        out[i] = v * scale;
    }
}

}  // namespace

extern "C" int b200_pack_grads(const void *const *srcs, const int64_t *offsets, const int64_t *counts, int n_tensors,
                               float *dst, float scale, int src_dtype, void *stream) {
    if (n_tensors < 0 || (n_tensors > 0 && (!srcs || !offsets || !counts || !dst))) return (int)cudaErrorInvalidValue;
    if (src_dtype != B200_F32 && src_dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int base = 0; base < n_tensors; base += kMaxTensors) {
        const int m = n_tensors - base < kMaxTensors ? n_tensors - base : kMaxTensors;
        PackTable tab;
        long long biggest = 0;
        for (int i = 0; i < m; ++i) {
            if (!srcs[base + i] || offsets[base + i] < 0 || counts[base + i] < 0) return (int)cudaErrorInvalidValue;
            if (b200_misaligned(srcs[base + i]) || (offsets[base + i] & 3)) return (int)cudaErrorMisalignedAddress;
            tab.src[i] = srcs[base + i];
            tab.off[i] = offsets[base + i];
            tab.cnt[i] = counts[base + i];
            if (counts[base + i] > biggest) biggest = counts[base + i];
        }
        for (int i = m; i < kMaxTensors; ++i) { tab.src[i] = nullptr; tab.off[i] = 0; tab.cnt[i] = 0; }
        if (b200_misaligned(dst)) return (int)cudaErrorMisalignedAddress;
        long long bx = (biggest / 8 + kThreads - 1) / kThreads;
        if (bx < 1) bx = 1;
        if (bx > 32) bx = 32;
        const dim3 grid((unsigned)bx, (unsigned)m);
        if (src_dtype == B200_BF16) pack_grads_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(tab, dst, scale);
        else pack_grads_kernel<float><<<grid, kThreads, 0, st>>>(tab, dst, scale);
        B200_CHECK_LAUNCH();
    }
    return 0;
}
