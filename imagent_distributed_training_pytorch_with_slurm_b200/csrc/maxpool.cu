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

// Rounds to the activation type and back: the fused stem kernel must pool over exactly the values the separate
// BatchNorm+ReLU kernel would have STORED.
template <typename T>
__device__ __forceinline__ float round_to(float v);
template <>
__device__ __forceinline__ float round_to<float>(float v) { return v; }
template <>
__device__ __forceinline__ float round_to<__nv_bfloat16>(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// AFFINE: every tap is first mapped through relu(x * scale[ch] + shift[ch]) (the BatchNorm+ReLU in front of the pool,
// torchvision resnet.py:269-270), so the normalised 4x larger activation is never written or read.
template <typename T, bool AFFINE>
__global__ void __launch_bounds__(kThreads)
maxpool_fwd_kernel(const T *__restrict__ x, T *__restrict__ y, uint8_t *__restrict__ arg, int64_t total, int c8, int h,
                   int w, int oh, int ow, const float *__restrict__ scale_shift) {
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
    float sc[8], sf[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        best[k] = -INFINITY;
        pos[k] = 0;
        sc[k] = AFFINE ? scale_shift[cv * 8 + k] : 1.f;
        sf[k] = AFFINE ? scale_shift[c + cv * 8 + k] : 0.f;
    }
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
            if (AFFINE) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = round_to<T>(fmaxf(fmaf(v[k], sc[k], sf[k]), 0.f));   // bn_apply's form
            }
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
maxpool_bwd_kernel(const T *__restrict__ dy, const T *__restrict__ dy2, const uint8_t *__restrict__ arg,
                   T *__restrict__ dx, int64_t total, int c8, int h, int w, int oh, int ow) {
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
                if (dy2) {
                    float g2[8];
                    V8<T>::load(dy2 + o, g2);
#pragma unroll
                    for (int k = 0; k < 8; ++k) g[a][b][k] += g2[k];
                }
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

// ---- bf16 fast path ---------------------------------------------------------------------------
// The generic kernels above spend ~5 instructions per tap and channel (unpack, two compares, two selects) plus three
// 64-bit divisions per thread, and are issue-bound at 2.8 TB/s.  Here (a) the index is 32-bit and decoded with
// multiply-shift division by host-computed constants, (b) bf16 pairs stay packed: HSET2 (greater-than mask), HMNMX2 and
// one LOP3 per tap and PAIR of channels.  Taps outside the image read as -inf, which never wins a strict comparison,
// and the argmax starts at the first tap inside the image (ATen's initial index), so the tie rule is unchanged.  A
// window that contains a NaN is redone by the scalar rule ("NaN counts as maximum, the last one wins").
struct FastDiv {  // n / d for n < 2^31 (Granlund-Montgomery, round-up variant)
    unsigned d, mul, shr;
};
FastDiv make_fastdiv(unsigned d) {
    FastDiv f{d, 0u, 0u};
    while ((1ull << f.shr) < d) ++f.shr;
    f.mul = (unsigned)((((1ull << f.shr) - d) << 32) / d + 1ull);
    return f;
}
__device__ __forceinline__ unsigned fdiv(unsigned n, const FastDiv &f) { return (__umulhi(n, f.mul) + n) >> f.shr; }

__device__ __forceinline__ __nv_bfloat162 as_bf162(unsigned u) {
    __nv_bfloat162 r;
    *reinterpret_cast<unsigned *>(&r) = u;
    return r;
}
__device__ __forceinline__ unsigned as_u32(__nv_bfloat162 v) { return *reinterpret_cast<unsigned *>(&v); }

template <bool AFFINE>
__global__ void __launch_bounds__(kThreads)
maxpool_fwd_bf16_kernel(const __nv_bfloat16 *__restrict__ x, __nv_bfloat16 *__restrict__ y, uint8_t *__restrict__ arg,
                        unsigned total, FastDiv dc8, FastDiv dow, FastDiv doh, int h, int w,
                        const float *__restrict__ scale_shift) {
    const unsigned gid = blockIdx.x * kThreads + threadIdx.x;
    if (gid >= total) return;
    const int c8 = (int)dc8.d;
    unsigned r = fdiv(gid, dc8);
    const int cv = (int)(gid - r * dc8.d);
    unsigned q = fdiv(r, dow);
    const int ox = (int)(r - q * dow.d);
    const unsigned n = fdiv(q, doh);
    const int oy = (int)(q - n * doh.d);
    const int iy0 = oy * 2 - 1, ix0 = ox * 2 - 1;
    const uint4 *base = reinterpret_cast<const uint4 *>(x) + (((int64_t)n * h + iy0) * w + ix0) * c8 + cv;
    const int rs = w * c8;  // uint4 units; one image row always fits 32 bits here (total < 2^31)
    const unsigned ninf = 0xff80ff80u;
    uint4 v[9];
    const bool interior = iy0 >= 0 && ix0 >= 0 && iy0 + 2 < h && ix0 + 2 < w;
    if (interior) {
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) v[kh * 3 + kw] = __ldg(base + kh * rs + kw * c8);
    } else {
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const bool ok = (unsigned)(iy0 + kh) < (unsigned)h && (unsigned)(ix0 + kw) < (unsigned)w;
                v[kh * 3 + kw] = make_uint4(ninf, ninf, ninf, ninf);
                if (ok) v[kh * 3 + kw] = __ldg(base + kh * rs + kw * c8);
            }
    }
    if (AFFINE) {
        // relu(x * scale + shift) in float32 exactly as the BatchNorm apply kernel does it, rounded once to bf16 - the
        // values that kernel would have stored; taps outside the image stay -inf.  (fmaxf maps NaN to 0, as there.)
        float sc[8], sf[8];
        const int cch = c8 * 8;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            sc[k] = __ldg(scale_shift + cv * 8 + k);
            sf[k] = __ldg(scale_shift + cch + cv * 8 + k);
        }
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            const int kh = t / 3, kw = t - kh * 3;
            if (!interior && ((unsigned)(iy0 + kh) >= (unsigned)h || (unsigned)(ix0 + kw) >= (unsigned)w)) continue;
            unsigned u[4] = {v[t].x, v[t].y, v[t].z, v[t].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float lo = fmaxf(fmaf(bf16_lo(u[j]), sc[2 * j], sf[2 * j]), 0.f);
                const float hi = fmaxf(fmaf(bf16_hi(u[j]), sc[2 * j + 1], sf[2 * j + 1]), 0.f);
                u[j] = pack_bf16(lo, hi);
            }
            v[t] = make_uint4(u[0], u[1], u[2], u[3]);
        }
    }
    const unsigned first = (unsigned)((iy0 < 0 ? 3 : 0) + (ix0 < 0 ? 1 : 0)) * 0x00010001u;
    __nv_bfloat162 m[4] = {as_bf162(ninf), as_bf162(ninf), as_bf162(ninf), as_bf162(ninf)};
    unsigned pos[4] = {first, first, first, first};
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        const unsigned tt = (unsigned)t * 0x00010001u;
        const unsigned vv[4] = {v[t].x, v[t].y, v[t].z, v[t].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 e = as_bf162(vv[j]);
            const unsigned gt = __hgt2_mask(e, m[j]);
            m[j] = __hmax2_nan(m[j], e);
            pos[j] = (pos[j] & ~gt) | (tt & gt);
        }
    }
    const unsigned nan = __hneu2_mask(m[0], m[0]) | __hneu2_mask(m[1], m[1]) | __hneu2_mask(m[2], m[2]) |
                         __hneu2_mask(m[3], m[3]);
    if (AFFINE || nan == 0u) {   // (after the ReLU there is no NaN: fmaxf maps it to 0)
        *reinterpret_cast<uint4 *>(y + (size_t)gid * 8) = make_uint4(as_u32(m[0]), as_u32(m[1]), as_u32(m[2]), as_u32(m[3]));
        *reinterpret_cast<uint2 *>(arg + (size_t)gid * 8) =
            make_uint2(__byte_perm(pos[0], pos[1], 0x6420), __byte_perm(pos[2], pos[3], 0x6420));
        return;
    }
    // rare: redo this window channel by channel from memory with the scalar rule
    const unsigned short *xs = reinterpret_cast<const unsigned short *>(base);
    unsigned short *ys = reinterpret_cast<unsigned short *>(y) + (size_t)gid * 8;
#pragma unroll 1
    for (int k = 0; k < 8; ++k) {
        float b = -INFINITY;
        unsigned short bb = 0xff80;
        int p = (int)(first & 0xffu);
#pragma unroll 1
        for (int t = 0; t < 9; ++t) {
            const int kh = t / 3, kw = t - kh * 3;
            if ((unsigned)(iy0 + kh) >= (unsigned)h || (unsigned)(ix0 + kw) >= (unsigned)w) continue;
            const unsigned short raw = xs[(int64_t)(kh * rs + kw * c8) * 8 + k];
            const float f = __uint_as_float((unsigned)raw << 16);
            if (f > b || f != f) { b = f; bb = raw; p = t; }
        }
        ys[k] = bb;
        arg[(size_t)gid * 8 + k] = (uint8_t)p;
    }
}

// Backward, bf16: same patch decomposition as the generic kernel, 32-bit index, window by window: the eight argmax
// bytes and gradients of a window are unpacked once and compared against the position each pixel of the patch has in
// that window (one compare + one predicated add per contribution).
__global__ void __launch_bounds__(kThreads, 3)
maxpool_bwd_bf16_kernel(const __nv_bfloat16 *__restrict__ dy, const __nv_bfloat16 *__restrict__ dy2,
                        const uint8_t *__restrict__ arg, __nv_bfloat16 *__restrict__ dx, unsigned total, FastDiv dc8,
                        FastDiv dpw, FastDiv dph, int h, int w, int oh, int ow) {
    const unsigned gid = blockIdx.x * kThreads + threadIdx.x;
    if (gid >= total) return;
    const int c8 = (int)dc8.d;
    unsigned r = fdiv(gid, dc8);
    const int cv = (int)(gid - r * dc8.d);
    unsigned q = fdiv(r, dpw);
    const int px = (int)(r - q * dpw.d);
    const unsigned n = fdiv(q, dph);
    const int py = (int)(q - n * dph.d);
    uint4 g[2][2], g2[2][2];
    uint2 p[2][2];
    const unsigned o00 = ((n * (unsigned)oh + py) * ow + px) * c8 + cv;  // < total: fits 32 bits
    const uint2 *ap = reinterpret_cast<const uint2 *>(arg) + o00;
    const uint4 *gp = reinterpret_cast<const uint4 *>(dy) + o00;
    const uint4 *gp2 = dy2 ? reinterpret_cast<const uint4 *>(dy2) + o00 : nullptr;   // second gradient branch (fan-in)
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            p[a][b] = make_uint2(0xffffffffu, 0xffffffffu);
            g[a][b] = make_uint4(0u, 0u, 0u, 0u);
            g2[a][b] = make_uint4(0u, 0u, 0u, 0u);
            if (py + a < oh && px + b < ow) {
                const int off = (a * ow + b) * c8;
                p[a][b] = __ldg(ap + off);
                g[a][b] = __ldg(gp + off);
                if (gp2) g2[a][b] = __ldg(gp2 + off);
            }
        }
    float acc[2][2][8];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[i][j][k] = 0.f;
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const unsigned gw[4] = {g[a][b].x, g[a][b].y, g[a][b].z, g[a][b].w};
            const unsigned hw[4] = {g2[a][b].x, g2[a][b].y, g2[a][b].z, g2[a][b].w};
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const unsigned pk = __byte_perm(k < 4 ? p[a][b].x : p[a][b].y, 0u, 0x4440u | (k & 3));
                float f = __uint_as_float((k & 1) ? (gw[k >> 1] & 0xffff0000u) : (gw[k >> 1] << 16));
                if (gp2) f += __uint_as_float((k & 1) ? (hw[k >> 1] & 0xffff0000u) : (hw[k >> 1] << 16));
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int kh = i + 1 - 2 * a;
                    if (kh < 0) continue;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int kw = j + 1 - 2 * b;
                        if (kw < 0) continue;
                        asm("{\n\t.reg .pred q;\n\tsetp.eq.u32 q, %1, %2;\n\t@q add.f32 %0, %0, %3;\n\t}"
                            : "+f"(acc[i][j][k])
                            : "r"(pk), "r"((unsigned)(kh * 3 + kw)), "f"(f));  // compare + predicated add
                    }
                }
            }
        }
    uint4 *xp = reinterpret_cast<uint4 *>(dx) + (((size_t)n * h + 2 * py) * w + 2 * px) * c8 + cv;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        if (2 * py + i >= h) continue;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (2 * px + j >= w) continue;
            uint4 o;
            o.x = as_u32(__floats2bfloat162_rn(acc[i][j][0], acc[i][j][1]));
            o.y = as_u32(__floats2bfloat162_rn(acc[i][j][2], acc[i][j][3]));
            o.z = as_u32(__floats2bfloat162_rn(acc[i][j][4], acc[i][j][5]));
            o.w = as_u32(__floats2bfloat162_rn(acc[i][j][6], acc[i][j][7]));
            xp[(i * w + j) * c8] = o;
        }
    }
}

}  // namespace

static int maxpool_fwd_launch(const void *x, const float *scale_shift, void *y, uint8_t *argmax, int64_t n, int c,
                              int h, int w, int dtype, void *stream) {
    if (!x || !y || !argmax || n < 1 || c < 8 || (c % 8) || h < 1 || w < 1) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(x) || b200_misaligned(y) || b200_misaligned(argmax, 8)) return (int)cudaErrorMisalignedAddress;
    const int oh = (h + 2 - 3) / 2 + 1, ow = (w + 2 - 3) / 2 + 1;
    const int64_t total = n * oh * ow * (c / 8);
    const unsigned blocks = (unsigned)((total + kThreads - 1) / kThreads);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool fast = dtype == B200_BF16 && total < (1ll << 31);
#define B200_POOL_FWD(AFFINE)                                                                                       \
    do {                                                                                                            \
        if (fast)                                                                                                   \
            maxpool_fwd_bf16_kernel<AFFINE><<<blocks, kThreads, 0, st>>>(                                           \
                static_cast<const __nv_bfloat16 *>(x), static_cast<__nv_bfloat16 *>(y), argmax, (unsigned)total,     \
                make_fastdiv(c / 8), make_fastdiv(ow), make_fastdiv(oh), h, w, scale_shift);                        \
        else if (dtype == B200_BF16)                                                                                \
            maxpool_fwd_kernel<__nv_bfloat16, AFFINE><<<blocks, kThreads, 0, st>>>(                                 \
                static_cast<const __nv_bfloat16 *>(x), static_cast<__nv_bfloat16 *>(y), argmax, total, c / 8, h, w, \
                oh, ow, scale_shift);                                                                               \
        else                                                                                                        \
            maxpool_fwd_kernel<float, AFFINE><<<blocks, kThreads, 0, st>>>(                                         \
                static_cast<const float *>(x), static_cast<float *>(y), argmax, total, c / 8, h, w, oh, ow,         \
                scale_shift);                                                                                       \
    } while (0)
    if (scale_shift) B200_POOL_FWD(true);
    else B200_POOL_FWD(false);
#undef B200_POOL_FWD
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_maxpool3x3s2_fwd(const void *x, void *y, uint8_t *argmax, int64_t n, int c, int h, int w,
                                     int dtype, void *stream) {
    return maxpool_fwd_launch(x, nullptr, y, argmax, n, c, h, w, dtype, stream);
}

extern "C" int b200_bn_relu_maxpool3x3s2_fwd(const void *x, const float *scale_shift, void *y, uint8_t *argmax, int64_t n,
                                             int c, int h, int w, int dtype, void *stream) {
    if (!scale_shift) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(scale_shift, 4)) return (int)cudaErrorMisalignedAddress;
    return maxpool_fwd_launch(x, scale_shift, y, argmax, n, c, h, w, dtype, stream);
}

extern "C" int b200_maxpool3x3s2_bwd(const void *dy, const void *dy2, const uint8_t *argmax, void *dx, int64_t n, int c,
                                     int h, int w, int dtype, void *stream) {
    if (!dy || !dx || !argmax || n < 1 || c < 8 || (c % 8) || h < 1 || w < 1) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(dy) || b200_misaligned(dx) || b200_misaligned(argmax, 8) || (dy2 && b200_misaligned(dy2)))
        return (int)cudaErrorMisalignedAddress;
    const int oh = (h + 2 - 3) / 2 + 1, ow = (w + 2 - 3) / 2 + 1;
    const int64_t total = n * ((h + 1) / 2) * ((w + 1) / 2) * (c / 8);
    const unsigned blocks = (unsigned)((total + kThreads - 1) / kThreads);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_BF16 && total < (1ll << 31))
        maxpool_bwd_bf16_kernel<<<blocks, kThreads, 0, st>>>(static_cast<const __nv_bfloat16 *>(dy),
                                                            static_cast<const __nv_bfloat16 *>(dy2), argmax,
                                                            static_cast<__nv_bfloat16 *>(dx), (unsigned)total,
                                                            make_fastdiv(c / 8), make_fastdiv((w + 1) / 2),
                                                            make_fastdiv((h + 1) / 2), h, w, oh, ow);
    else if (dtype == B200_BF16)
        maxpool_bwd_kernel<__nv_bfloat16><<<blocks, kThreads, 0, st>>>(static_cast<const __nv_bfloat16 *>(dy),
                                                                      static_cast<const __nv_bfloat16 *>(dy2), argmax,
                                                                      static_cast<__nv_bfloat16 *>(dx), total, c / 8, h,
                                                                      w, oh, ow);
    else
        maxpool_bwd_kernel<float><<<blocks, kThreads, 0, st>>>(static_cast<const float *>(dy),
                                                              static_cast<const float *>(dy2), argmax,
                                                              static_cast<float *>(dx), total, c / 8, h, w, oh, ow);
    B200_CHECK_LAUNCH();
    return 0;
}
