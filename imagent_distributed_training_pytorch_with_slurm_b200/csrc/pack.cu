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

// elements one CTA converts: STEPS vectors of 8 per thread (STEPS = 4: 16 KB in, 32 KB out)
constexpr int block_elems(int steps) { return kThreads * 8 * steps; }

struct PackTable {
    const void *src[kMaxTensors];
    long long off[kMaxTensors];
    long long cnt[kMaxTensors];
    int first_block[kMaxTensors + 1];   // prefix sum of ceil(cnt / kBlockElems): CTAs are dealt out in proportion to size
};

// The first version gave every tensor the same 32 CTAs: the 9.4 MB tensors of layer4 set the duration while the CTAs
// of the small ones idled (3.5 TB/s).  Now the grid is flat over all tensors' 8192-element blocks and a CTA finds its
// tensor by binary search in the prefix table (kernel parameter space = constant bank, uniform across the CTA).
template <typename T, int STEPS>
__global__ void __launch_bounds__(kThreads) pack_grads_kernel(const __grid_constant__ PackTable tab, int n_tensors,
                                                              float *__restrict__ dst, float scale) {
    constexpr int kBlockElems = kThreads * 8 * STEPS;
    int lo = 0, hi = n_tensors;   // largest t with first_block[t] <= blockIdx.x
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tab.first_block[mid] <= (int)blockIdx.x) lo = mid; else hi = mid;
    }
    const int t = lo;
    const T *src = static_cast<const T *>(tab.src[t]);
    float *out = dst + tab.off[t];
    const long long n = tab.cnt[t];
    const long long n8 = n >> 3;
    const long long v0 = (long long)((int)blockIdx.x - tab.first_block[t]) * (kBlockElems / 8);
    constexpr int kSteps = STEPS;
    // all of a thread's loads are issued before its first store (the sources are dead after this kernel: streaming
    // loads; the arena is read next by the all-reduce / the optimizer: ordinary stores, it stays in L2)
    float e[kSteps][8];
#pragma unroll
    for (int step = 0; step < kSteps; ++step) {
        const long long v = v0 + step * kThreads + threadIdx.x;
        if (v < n8) {
            if constexpr (sizeof(T) == 2) {
                VecIO<__nv_bfloat16, 8>::load_stream(reinterpret_cast<const __nv_bfloat16 *>(src) + v * 8, e[step]);
            } else {
                float a[4], b[4];
                VecIO<float, 4>::load_stream(reinterpret_cast<const float *>(src) + v * 8, a);
                VecIO<float, 4>::load_stream(reinterpret_cast<const float *>(src) + v * 8 + 4, b);
#pragma unroll
                for (int k = 0; k < 4; ++k) { e[step][k] = a[k]; e[step][4 + k] = b[k]; }
            }
        }
    }
#pragma unroll
    for (int step = 0; step < kSteps; ++step) {
        const long long v = v0 + step * kThreads + threadIdx.x;
        if (v < n8) {
            float a[4] = {e[step][0] * scale, e[step][1] * scale, e[step][2] * scale, e[step][3] * scale};
            float b[4] = {e[step][4] * scale, e[step][5] * scale, e[step][6] * scale, e[step][7] * scale};
            VecIO<float, 4>::store(out + v * 8, a);
            VecIO<float, 4>::store(out + v * 8 + 4, b);
        }
    }
    if ((int)blockIdx.x == tab.first_block[t] && threadIdx.x < (n & 7)) {   // ragged tail, by the tensor's first CTA
        const long long i = (n8 << 3) + threadIdx.x;
        float v;
        if constexpr (sizeof(T) == 2) v = __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(src)[i]);
        else v = reinterpret_cast<const float *>(src)[i];
        out[i] = v * scale;
    }
}

}  // namespace

// b200_tune_set("pack_steps", 1 | 2 | 4 | 8): 16-byte vectors per thread.  Measured on B200 (tools/probe_pack.py, ResNet-50's
// 53 weight gradients = 141 MB moved, L2 flushed): 28.7 / 30.7 / 32.8 / 36.9 us = 4.9 / 4.6 / 4.3 / 3.8 TB/s - with 53 ragged
// tensors the small blocks fill the tail of the launch better than deep per-thread batching helps.
int g_pack_steps = 1;

extern "C" int b200_pack_grads(const void *const *srcs, const int64_t *offsets, const int64_t *counts, int n_tensors,
                               float *dst, float scale, int src_dtype, void *stream) {
    if (n_tensors < 0 || (n_tensors > 0 && (!srcs || !offsets || !counts || !dst))) return (int)cudaErrorInvalidValue;
    if (src_dtype != B200_F32 && src_dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int steps = (g_pack_steps == 8 || g_pack_steps == 4 || g_pack_steps == 2) ? g_pack_steps : 1;
    const int kBlockElems = block_elems(steps);
    for (int base = 0; base < n_tensors; base += kMaxTensors) {
        const int m = n_tensors - base < kMaxTensors ? n_tensors - base : kMaxTensors;
        PackTable tab;
        long long blocks = 0;
        for (int i = 0; i < m; ++i) {
            if (!srcs[base + i] || offsets[base + i] < 0 || counts[base + i] < 0) return (int)cudaErrorInvalidValue;
            if (b200_misaligned(srcs[base + i]) || (offsets[base + i] & 3)) return (int)cudaErrorMisalignedAddress;
            tab.src[i] = srcs[base + i];
            tab.off[i] = offsets[base + i];
            tab.cnt[i] = counts[base + i];
            tab.first_block[i] = (int)blocks;
            const long long nb = (counts[base + i] + kBlockElems - 1) / kBlockElems;
            blocks += nb > 0 ? nb : 1;   // an empty tensor keeps one (idle) CTA so that the search stays well-defined
            if (blocks > 0x7fffffffLL) return (int)cudaErrorInvalidValue;
        }
        for (int i = m; i < kMaxTensors; ++i) { tab.src[i] = nullptr; tab.off[i] = 0; tab.cnt[i] = 0; }
        for (int i = m; i <= kMaxTensors; ++i) tab.first_block[i] = (int)blocks;
        if (b200_misaligned(dst)) return (int)cudaErrorMisalignedAddress;
#define B200_PACK(T, S) pack_grads_kernel<T, S><<<(unsigned)blocks, kThreads, 0, st>>>(tab, m, dst, scale)
#define B200_PACK_T(T)                                                                                             \
    do {                                                                                                           \
        if (steps == 8) B200_PACK(T, 8);                                                                           \
        else if (steps == 4) B200_PACK(T, 4);                                                                      \
        else if (steps == 2) B200_PACK(T, 2);                                                                      \
        else B200_PACK(T, 1);                                                                                      \
    } while (0)
        if (src_dtype == B200_BF16) B200_PACK_T(__nv_bfloat16);
        else B200_PACK_T(float);
#undef B200_PACK_T
#undef B200_PACK
        B200_CHECK_LAUNCH();
    }
    return 0;
}
