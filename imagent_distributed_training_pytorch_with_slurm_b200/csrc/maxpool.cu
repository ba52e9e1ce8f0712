// K6: 3x3 / stride 2 / pad 1 max-pooling on channels-last activations, forward and backward.
// Reference site: torchvision resnet.py:200,270 (`self.maxpool = nn.MaxPool2d(3, 2, 1)` after the
// stem) and its gradient inside loss.backward() (imagenet.py:128).  PyTorch's channels-last kernels
// keep an int64 argmax per output element (8 bytes per 2-byte activation: 411 MB written and read
// back at N=256) and scatter gradients through it; here the forward keeps ONE BYTE per element (the
// position 0..8 inside the window, first maximum in row-major window order wins, NaN counts as
// maximum - the same tie rule as ATen's max_pool_forward_nhwc) and the backward is a gather: every
// input pixel looks at the <= 4 windows that contain it and adds the gradients of those whose argmax
// it is.  No atomics, no zero-fill, float32 accumulation, one rounding.
//
// HBM traffic: forward reads x once (neighbouring windows hit L1/L2) and writes y + 1 B/element of
// argmax; backward reads dy and the argmax bytes and writes dx.  Vectors of 8 channels (16 bytes bf16,
// 2 x 16 bytes float32) per thread, channels fastest so every warp access is contiguous.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

template <typename T>
struct V8;  // 8 consecutive channels
template <>
struct V8<__nv_bfloat16> {
    static __device__ __forceinline__ void load(const __nv_bfloat16 *p, float (&v)[8]) { VecIO<__nv_bfloat16, 8>::load(p, v); }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, const float (&v)[8]) { VecIO<__nv_bfloat16, 8>::store(p, v); }
};
template <>
struct V8<float> {
    static __device__ __forceinline__ void load(const float *p, float (&v)[8]) {
        float a[4], b[4];
        VecIO<float, 4>::load(p, a);
        VecIO<float, 4>::load(p + 4, b);
#pragma unroll
        for (int k = 0; k < 4; ++k) { v[k] = a[k]; v[4 + k] = b[k]; }
    }
    static __device__ __forceinline__ void store(float *p, const float (&v)[8]) {
        float a[4] = {v[0], v[1], v[2], v[3]}, b[4] = {v[4], v[5], v[6], v[7]};
        VecIO<float, 4>::store(p, a);
        VecIO<float, 4>::store(p + 4, b);
    }
};

template <typename T>
__global__ void __launch_bounds__(kThreads)
maxpool_fwd_kernel(const T *__restrict__ x, T *__restrict__ y, uint8_t *__restrict__ arg, int64_t total, int c8, int h,
                   int w, int oh, int ow) {
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= total) return;
    const int cv = (int)(gid % c8);
    int64_t r = gid / c8;
    const int ox = (int)(r % ow); r /= ow;
    const int oy = (int)(r % oh);
    const int64_t n = r / oh;
    const int c = c8 * 8;
    float best[8];
    unsigned pos[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; pos[k] = 0; }
    bool first = true;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
        const int iy = oy * 2 - 1 + kh;
        if (iy < 0 || iy >= h) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
            const int ix = ox * 2 - 1 + kw;
            if (ix < 0 || ix >= w) continue;
            float v[8];
            V8<T>::load(x + ((n * h + iy) * w + ix) * c + cv * 8, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (first || v[k] > best[k] || v[k] != v[k]) { best[k] = v[k]; pos[k] = kh * 3 + kw; }
            }
            first = false;
        }
    }
    V8<T>::store(y + gid * 8, best);
    uint2 packed;
    packed.x = pos[0] | (pos[1] << 8) | (pos[2] << 16) | (pos[3] << 24);
    packed.y = pos[4] | (pos[5] << 8) | (pos[6] << 16) | (pos[7] << 24);
    *reinterpret_cast<uint2 *>(arg + gid * 8) = packed;
}

// Backward as a gather over 2x2 input patches: the patch {2p, 2p+1} x {2q, 2q+1} is touched by exactly
// the four windows (p..p+1) x (q..q+1), so a thread loads those four (gradient, argmax) vectors once and
// produces four output vectors.
template <typename T>
__global__ void __launch_bounds__(kThreads)
maxpool_bwd_kernel(const T *__restrict__ dy, const uint8_t *__restrict__ arg, T *__restrict__ dx, int64_t total, int c8,
                   int h, int w, int oh, int ow) {
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= total) return;
    const int ph = (h + 1) >> 1, pw = (w + 1) >> 1;
    const int cv = (int)(gid % c8);
    int64_t r = gid / c8;
    const int px = (int)(r % pw); r /= pw;
    const int py = (int)(r % ph);
    const int64_t n = r / ph;
    const int c = c8 * 8;
    float g[2][2][8];
    unsigned pos[2][2][8];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int oy = py + a, ox = px + b;
            if (oy < oh && ox < ow) {
                const int64_t o = (((n * oh + oy) * ow + ox) * c8 + cv) * 8;
                const uint2 p = *reinterpret_cast<const uint2 *>(arg + o);
                V8<T>::load(dy + o, g[a][b]);
#pragma unroll
                for (int k = 0; k < 8; ++k) pos[a][b][k] = ((k < 4 ? p.x : p.y) >> ((k & 3) * 8)) & 0xffu;
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) { g[a][b][k] = 0.f; pos[a][b][k] = 0xffu; }
            }
        }
    // pixel (2py+i, 2px+j): window (py+a, px+b) contains it at (kh, kw) = (i+1-2a, j+1-2b) when both are in 0..2
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int iy = 2 * py + i;
        if (iy >= h) continue;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int ix = 2 * px + j;
            if (ix >= w) continue;
            float acc[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const int kh = i + 1 - 2 * a;
                if (kh < 0 || kh > 2) continue;
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const int kw = j + 1 - 2 * b;
                    if (kw < 0 || kw > 2) continue;
                    const unsigned me = (unsigned)(kh * 3 + kw);
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        if (pos[a][b][k] == me) acc[k] += g[a][b][k];
                }
            }
            V8<T>::store(dx + ((n * h + iy) * w + ix) * c + cv * 8, acc);
        }
    }
}

}  // namespace

extern "C" int b200_maxpool3x3s2_fwd(const void *x, void *y, uint8_t *argmax, int64_t n, int c, int h, int w,
                                     int dtype, void *stream) {
    if (!x || !y || !argmax || n < 1 || c < 8 || (c % 8) || h < 1 || w < 1) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(x) || b200_misaligned(y) || b200_misaligned(argmax, 8)) return (int)cudaErrorMisalignedAddress;
    const int oh = (h + 2 - 3) / 2 + 1, ow = (w + 2 - 3) / 2 + 1;
    const int64_t total = n * oh * ow * (c / 8);
    const unsigned blocks = (unsigned)((total + kThreads - 1) / kThreads);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_BF16)
        maxpool_fwd_kernel<__nv_bfloat16><<<blocks, kThreads, 0, st>>>(static_cast<const __nv_bfloat16 *>(x),
                                                                      static_cast<__nv_bfloat16 *>(y), argmax, total,
                                                                      c / 8, h, w, oh, ow);
    else
        maxpool_fwd_kernel<float><<<blocks, kThreads, 0, st>>>(static_cast<const float *>(x), static_cast<float *>(y),
                                                              argmax, total, c / 8, h, w, oh, ow);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_maxpool3x3s2_bwd(const void *dy, const uint8_t *argmax, void *dx, int64_t n, int c, int h, int w,
                                     int dtype, void *stream) {
    if (!dy || !dx || !argmax || n < 1 || c < 8 || (c % 8) || h < 1 || w < 1) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(dy) || b200_misaligned(dx) || b200_misaligned(argmax, 8)) return (int)cudaErrorMisalignedAddress;
    const int oh = (h + 2 - 3) / 2 + 1, ow = (w + 2 - 3) / 2 + 1;
    const int64_t total = n * ((h + 1) / 2) * ((w + 1) / 2) * (c / 8);
    const unsigned blocks = (unsigned)((total + kThreads - 1) / kThreads);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_BF16)
        maxpool_bwd_kernel<__nv_bfloat16><<<blocks, kThreads, 0, st>>>(static_cast<const __nv_bfloat16 *>(dy), argmax,
                                                                      static_cast<__nv_bfloat16 *>(dx), total, c / 8, h,
                                                                      w, oh, ow);
    else
        maxpool_bwd_kernel<float><<<blocks, kThreads, 0, st>>>(static_cast<const float *>(dy), argmax,
                                                              static_cast<float *>(dx), total, c / 8, h, w, oh, ow);
    B200_CHECK_LAUNCH();
    return 0;
}
