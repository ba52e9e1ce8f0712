// K1: input stage.  uint8 NHWC batch -> per-sample crop box -> Pillow-exact antialiased bilinear
// resize -> horizontal flip -> ToTensor (/255) -> Normalize ((x-mean)/std) -> bf16/f32, one pass.
// Reference sites: transforms.Compose([Resize((448,448)), ToTensor(), Normalize(.5,.5)])
// (imagenet.py:280-283) executed per image on PIL images inside the DataLoader workers
// (imagenet.py:350-354); north-star mode adds torchvision RandomResizedCrop/RandomHorizontalFlip
// (transforms.py:928-981,712-722), whose PIL path is crop -> the same resize -> flip.
//
// The resize follows Pillow's ImagingResample for 8-bit images (Resample.c, BILINEAR):
//   precompute_coeffs   : double-precision triangle weights, support = max(scale,1), normalised
//   normalize_coeffs_8bpc: fixed point, 22 fractional bits, round half up
//   horizontal pass     : clip8((2^21 + sum pix*k) >> 22) into an 8-bit intermediate
//   vertical pass       : the same on the intermediate
// All double arithmetic uses the round-to-nearest intrinsics so that no multiply-add is contracted
// and the integer coefficients equal Pillow's bit for bit.
//
// HBM traffic: the crop box is read once (neighbouring tiles re-read a few rows out of L2), the
// output is written once; the normalisation is a 3x256 float look-up table in shared memory
// because the resized pixel is an 8-bit value.
#include "common.cuh"

int g_input_s2d_fast = 1;  // b200_tune_set("input_s2d_fast"): 0 sends the stem layout through the generic kernel

namespace {

constexpr int kThreads = 256;
constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kTileRows = 8;

__host__ __device__ inline int ksize_for(int src, int out) {
    // ceil(max(src/out, 1)) * 2 + 1, in integers
    int s = (src + out - 1) / out;
    if (s < 1) s = 1;
    return s * 2 + 1;
}

__device__ __forceinline__ double tri_weight(double x) {
    if (x < 0.0) x = -x;
    return x < 1.0 ? __dsub_rn(1.0, x) : 0.0;
}

// tables per sample: bounds_x[out_w][2], coef_x[out_w][KX], bounds_y[out_h][2], coef_y[out_h][KY]
struct Norm { float mean[3]; float std[3]; };

__global__ void __launch_bounds__(kThreads)
input_coeff_kernel(const int32_t *__restrict__ boxes, int64_t n, int out_h, int out_w, int KX, int KY,
                   int32_t *__restrict__ ws, float *__restrict__ lut, Norm nrm) {
    const int per_row = out_w + out_h;
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (lut && gid < 3 * 256) {
        // ToTensor: float32(v) / 255 ; Normalize: (x - mean) / std, all IEEE float32
        const int ch = (int)gid >> 8, v = (int)gid & 255;
        lut[gid] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), nrm.mean[ch]), nrm.std[ch]);
    }
    if (gid >= n * per_row) return;
    const int64_t s = gid / per_row;
    const int idx = (int)(gid - s * per_row);
    const bool vertical = idx >= out_w;
    const int xx = vertical ? idx - out_w : idx;
    const int in_size = vertical ? boxes[s * 4 + 2] : boxes[s * 4 + 3];
    const int out_size = vertical ? out_h : out_w;
    const int K = vertical ? KY : KX;
    const int64_t per_sample = (int64_t)out_w * (2 + KX) + (int64_t)out_h * (2 + KY);
    int32_t *base = ws + s * per_sample;
    int32_t *bounds = vertical ? base + (int64_t)out_w * (2 + KX) : base;
    int32_t *coef = bounds + 2 * out_size;

    const double scale = __ddiv_rn((double)in_size, (double)out_size);
    const double fs = scale < 1.0 ? 1.0 : scale;
    const double support = fs;  // bilinear support 1.0 * filterscale
    const double center = __dmul_rn(__dadd_rn((double)xx, 0.5), scale);
    int lo = __double2int_rz(__dadd_rn(__dsub_rn(center, support), 0.5));
    if (lo < 0) lo = 0;
    int hi = __double2int_rz(__dadd_rn(__dadd_rn(center, support), 0.5));
    if (hi > in_size) hi = in_size;
    int cnt = hi - lo;
    if (cnt > K) cnt = K;  // cannot happen for boxes inside the image; keeps the table in bounds
    if (cnt < 0) cnt = 0;
    const double ss = __ddiv_rn(1.0, fs);
    double ww = 0.0;
    for (int x = 0; x < cnt; ++x)
        ww = __dadd_rn(ww, tri_weight(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + lo), center), 0.5), ss)));
    for (int x = 0; x < K; ++x) {
        int kk = 0;
        if (x < cnt) {
            double w = tri_weight(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + lo), center), 0.5), ss));
            if (ww != 0.0) w = __ddiv_rn(w, ww);
            kk = __double2int_rz(__dadd_rn(0.5, __dmul_rn(w, (double)(1 << kPrecisionBits))));
        }
        coef[(int64_t)xx * K + x] = kk;
    }
    bounds[2 * xx] = lo;
    bounds[2 * xx + 1] = cnt;
}

__device__ __forceinline__ int clip8(int acc) {
    int v = acc >> kPrecisionBits;
    return v < 0 ? 0 : (v > 255 ? 255 : v);
}

template <typename OutT>
__device__ __forceinline__ OutT cast_out(float v);
template <>
__device__ __forceinline__ float cast_out<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 cast_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// One CTA = one tile of 8 output rows of one sample.  Output rows of a tile are assembled in shared
// memory in their final memory layout and leave with 16-byte vector stores (NHWC: 8 contiguous rows;
// space-to-depth: 4 complete rows of 16-channel cells, border and spare channels included - the tile
// origin is shifted by the 3-pixel padding so that cells are never split between CTAs).
template <typename OutT>
__global__ void __launch_bounds__(kThreads)
input_resample_kernel(const uint8_t *__restrict__ images, const int32_t *__restrict__ boxes,
                      const uint8_t *__restrict__ flips, OutT *__restrict__ out, const int32_t *__restrict__ ws,
                      int src_h, int src_w, int out_h, int out_w, int KX, int KY, int max_src_rows, Norm nrm,
                      int layout, int inter_bytes, const int64_t *__restrict__ tab) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *lut = reinterpret_cast<float *>(smem_raw);          // [3][256]
    uint8_t *inter = smem_raw + 3 * 256 * sizeof(float);       // [max_src_rows][out_w][3]
    OutT *otile = reinterpret_cast<OutT *>(inter + inter_bytes);  // staged output tile (NHWC / S2D only)

    const int64_t s = blockIdx.y;
    const bool s2d = layout == B200_LAYOUT_S2D_STEM;
    const int ya = (int)blockIdx.x * kTileRows - (s2d ? 3 : 0);   // first output row of the tile (may be < 0)
    const int y0 = ya < 0 ? 0 : ya;
    const int y1 = (ya + kTileRows < out_h) ? ya + kTileRows : out_h;
    const int sy = (out_h + 6) >> 1, sx = (out_w + 6) >> 1;
    for (int i = threadIdx.x; i < 3 * 256; i += kThreads) {
        const int ch = i >> 8, v = i & 255;
        // ToTensor: float32(v) / 255 ; Normalize: (x - mean) / std, all IEEE float32
        lut[i] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), nrm.mean[ch]), nrm.std[ch]);
    }
    // staged tile geometry
    const int cell_rows = s2d ? ((sy - (int)blockIdx.x * 4 < 4) ? sy - (int)blockIdx.x * 4 : 4) : 0;
    const int tile_elems = s2d ? cell_rows * sx * 16 : (layout == B200_LAYOUT_NHWC ? (y1 - y0) * out_w * 3 : 0);
    if (s2d) {  // border cells and the 4 spare channels read as zero
        uint32_t *z = reinterpret_cast<uint32_t *>(otile);
        const int words = tile_elems * (int)sizeof(OutT) / 4;
        for (int i = threadIdx.x; i < words; i += kThreads) z[i] = 0u;
    }
    const int64_t per_sample = (int64_t)out_w * (2 + KX) + (int64_t)out_h * (2 + KY);
    const int32_t *bx = ws + s * per_sample;
    const int32_t *cx = bx + 2 * out_w;
    const int32_t *by = bx + (int64_t)out_w * (2 + KX);
    const int32_t *cy = by + 2 * out_h;
    const int top = boxes[s * 4 + 0], left = boxes[s * 4 + 1];
    const bool flip = flips && flips[s] != 0;
    if (y1 > y0) {
        const int ylo = by[2 * y0];
        const int nrows = by[2 * (y1 - 1)] + by[2 * (y1 - 1) + 1] - ylo;
        if (nrows > max_src_rows) return;  // contract violation (box outside the stated limits)

        // horizontal pass: a thread owns an output column, keeps that column's taps in registers and
        // walks down the needed source rows -> 8-bit intermediate in shared memory
        // dense batch: sample s is image s of [n, src_h, src_w, 3]; ragged batch (tab != nullptr): sample s starts
        // tab[4s] bytes into the slab and its rows are tab[4s+3] bytes apart
        const int64_t pitch = tab ? tab[4 * s + 3] : (int64_t)src_w * 3;
        const int64_t base = tab ? tab[4 * s] : (int64_t)s * src_h * pitch;
        const uint8_t *img = images + base + (int64_t)(top + ylo) * pitch + (int64_t)left * 3;
        for (int xx = threadIdx.x; xx < out_w; xx += kThreads) {
            const int xmin = bx[2 * xx], cnt = bx[2 * xx + 1];
            const int32_t *k = cx + (int64_t)xx * KX;
            int kk[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) kk[t] = (t < cnt) ? k[t] : 0;
            const uint8_t *px = img + xmin * 3;
            if (cnt <= 8) {
                for (int rr = 0; rr < nrows; ++rr, px += pitch) {
                    int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        if (t < cnt) {
                            a0 += (int)px[3 * t + 0] * kk[t];
                            a1 += (int)px[3 * t + 1] * kk[t];
                            a2 += (int)px[3 * t + 2] * kk[t];
                        }
                    }
                    uint8_t *o = inter + (rr * out_w + xx) * 3;
                    o[0] = (uint8_t)clip8(a0);
                    o[1] = (uint8_t)clip8(a1);
                    o[2] = (uint8_t)clip8(a2);
                }
            } else {  // very strong down-scaling: taps stay in memory
                for (int rr = 0; rr < nrows; ++rr, px += pitch) {
                    int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
                    for (int t = 0; t < cnt; ++t) {
                        const int kt = k[t];
                        a0 += (int)px[3 * t + 0] * kt;
                        a1 += (int)px[3 * t + 1] * kt;
                        a2 += (int)px[3 * t + 2] * kt;
                    }
                    uint8_t *o = inter + (rr * out_w + xx) * 3;
                    o[0] = (uint8_t)clip8(a0);
                    o[1] = (uint8_t)clip8(a1);
                    o[2] = (uint8_t)clip8(a2);
                }
            }
        }
        __syncthreads();

        // vertical pass -> flip -> normalise -> staged tile (or straight to memory for NCHW)
        const int npix = (y1 - y0) * out_w;
        for (int idx = threadIdx.x; idx < npix; idx += kThreads) {
            const int ry = idx / out_w, xx = idx - ry * out_w;
            const int yy = y0 + ry;
            const int ymin = by[2 * yy] - ylo, cnt = by[2 * yy + 1];
            const int32_t *k = cy + (int64_t)yy * KY;
            int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
            for (int t = 0; t < cnt; ++t) {
                const int kt = k[t];
                const uint8_t *p = inter + ((ymin + t) * out_w + xx) * 3;
                a0 += (int)p[0] * kt;
                a1 += (int)p[1] * kt;
                a2 += (int)p[2] * kt;
            }
            const float f0 = lut[clip8(a0)], f1 = lut[256 + clip8(a1)], f2 = lut[512 + clip8(a2)];
            const int xd = flip ? (out_w - 1 - xx) : xx;
            if (layout == B200_LAYOUT_NHWC) {
                OutT *o = otile + (ry * out_w + xd) * 3;
                o[0] = cast_out<OutT>(f0);
                o[1] = cast_out<OutT>(f1);
                o[2] = cast_out<OutT>(f2);
            } else if (s2d) {
                // image zero-padded by 3, then 2x2 space-to-depth: pixel (py,px) -> cell (py/2,px/2),
                // channels (py%2*2 + px%2)*3 + ch of a 16-channel cell (12 used, 4 stay zero)
                const int py = yy + 3, px = xd + 3;
                OutT *o = otile + ((((py >> 1) - (int)blockIdx.x * 4) * sx + (px >> 1)) << 4) + ((py & 1) * 2 + (px & 1)) * 3;
                o[0] = cast_out<OutT>(f0);
                o[1] = cast_out<OutT>(f1);
                o[2] = cast_out<OutT>(f2);
            } else {
                const int64_t plane = (int64_t)out_h * out_w;
                OutT *o = out + (int64_t)s * 3 * plane + (int64_t)yy * out_w + xd;
                o[0] = cast_out<OutT>(f0);
                o[plane] = cast_out<OutT>(f1);
                o[2 * plane] = cast_out<OutT>(f2);
            }
        }
    }
    if (tile_elems == 0) return;
    __syncthreads();
    // copy the staged tile out: it is one contiguous span of the output tensor
    OutT *dst = s2d ? out + (((int64_t)s * sy + (int64_t)blockIdx.x * 4) * sx << 4)
                    : out + ((int64_t)s * out_h + y0) * out_w * 3;
    const int bytes = tile_elems * (int)sizeof(OutT);
    if (((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)bytes) & 15) == 0) {
        const uint4 *sv = reinterpret_cast<const uint4 *>(otile);
        uint4 *dv = reinterpret_cast<uint4 *>(dst);
        for (int i = threadIdx.x; i < bytes / 16; i += kThreads) dv[i] = sv[i];
    } else {
        for (int i = threadIdx.x; i < tile_elems; i += kThreads) dst[i] = otile[i];
    }
}

// ---- space-to-depth fast path -------------------------------------------------------------------
// The generic kernel above spends ~270 instructions per output pixel (byte loads from global memory in the horizontal
// pass, byte-wise shared-memory traffic and per-pixel index arithmetic in the vertical pass) and is issue-bound.  The
// stem layout - the one the model consumes - gets its own kernel built on 32-bit RGBx pixels in shared memory:
//   1. the needed rows of the crop box are staged as RGBx words (word loads + funnel shift, 4 pixels per thread step);
//   2. horizontal pass: one 32-bit shared load per tap, one 32-bit store per intermediate pixel;
//   3. vertical pass: a thread owns one 16-channel CELL (2x2 pixels): one 64-bit shared load per tap and row, 12
//      look-ups, and the finished cell leaves with 16-byte stores straight to global memory (a warp writes 1 KB
//      contiguous) - border cells and the 4 spare channels are written as zeros, so there is no staged tile.
// clip8 needs no clamp for the bilinear filter: weights are >= 0 and sum to 2^22 + (<= taps/2), so
// 2^21 + 255 * sum < 256 * 2^22 for up to 17 taps (the API's 8x down-scaling limit); the generic tap loops keep `min`.
constexpr int kS2dPadWords = 4;

template <typename OutT>
__device__ __forceinline__ void store_cell(OutT *dst, const float (&f)[12]);
template <>
__device__ __forceinline__ void store_cell<float>(float *dst, const float (&f)[12]) {
    float4 *d = reinterpret_cast<float4 *>(dst);
    d[0] = make_float4(f[0], f[1], f[2], f[3]);
    d[1] = make_float4(f[4], f[5], f[6], f[7]);
    d[2] = make_float4(f[8], f[9], f[10], f[11]);
    d[3] = make_float4(0.f, 0.f, 0.f, 0.f);
}
template <>
__device__ __forceinline__ void store_cell<__nv_bfloat16>(__nv_bfloat16 *dst, const float (&f)[12]) {
    unsigned w[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
        w[i] = *reinterpret_cast<const unsigned *>(&h);
    }
    uint4 *d = reinterpret_cast<uint4 *>(dst);
    d[0] = make_uint4(w[0], w[1], w[2], w[3]);
    d[1] = make_uint4(w[4], w[5], 0u, 0u);
}

template <typename OutT>
__global__ void __launch_bounds__(kThreads, 5)
input_s2d_kernel(const uint8_t *__restrict__ images, int64_t image_bytes, const int32_t *__restrict__ boxes,
                 const uint8_t *__restrict__ flips, OutT *__restrict__ out, const int32_t *__restrict__ ws,
                 const float *__restrict__ lut_g, int src_h, int src_w, int out_h, int out_w, int KX, int KY,
                 int max_src_rows, int src_stride, int inter_stride, const int64_t *__restrict__ tab) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *lut = reinterpret_cast<float *>(smem_raw);                        // [3][256]
    uint32_t *src = reinterpret_cast<uint32_t *>(smem_raw + 3 * 256 * sizeof(float));  // [max_src_rows][src_stride]
    uint32_t *inter = src + (size_t)max_src_rows * src_stride;              // [max_src_rows][inter_stride]

    const int64_t s = blockIdx.y;
    const int ya = (int)blockIdx.x * kTileRows - 3;   // first output row of the tile (may be < 0)
    const int y0 = ya < 0 ? 0 : ya;
    const int y1 = (ya + kTileRows < out_h) ? ya + kTileRows : out_h;
    const int sy = (out_h + 6) >> 1, sx = (out_w + 6) >> 1;
    for (int i = threadIdx.x; i < 3 * 256; i += kThreads) lut[i] = lut_g[i];   // made once by the coefficient kernel
    const int64_t per_sample = (int64_t)out_w * (2 + KX) + (int64_t)out_h * (2 + KY);
    const int32_t *bx = ws + s * per_sample;
    const int32_t *cx = bx + 2 * out_w;
    const int32_t *by = bx + (int64_t)out_w * (2 + KX);
    const int32_t *cy = by + 2 * out_h;
    const int top = boxes[s * 4 + 0], left = boxes[s * 4 + 1], bw = boxes[s * 4 + 3];
    const bool flip = flips && flips[s] != 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int ylo = 0;
    bool rows_ok = y1 > y0;
    if (rows_ok) {
        ylo = by[2 * y0];
        const int nrows = by[2 * (y1 - 1)] + by[2 * (y1 - 1) + 1] - ylo;
        rows_ok = nrows <= max_src_rows && bw <= src_stride - kS2dPadWords;   // else: contract violation
        if (rows_ok) {
            // 1. stage the rows: 4 pixels = 12 bytes = (up to) 4 aligned words per step
            const int ngroups = (bw + 3) >> 2;
            const int64_t pitch = tab ? tab[4 * s + 3] : (int64_t)src_w * 3;        // see input_resample_kernel
            const int64_t base = tab ? tab[4 * s] : (int64_t)s * src_h * pitch;
            for (int rr = warp; rr < nrows; rr += kThreads / 32) {
                const int64_t off = base + (int64_t)(top + ylo + rr) * pitch + (int64_t)left * 3;
                const uintptr_t addr = reinterpret_cast<uintptr_t>(images) + (uintptr_t)off;
                const unsigned sh = (unsigned)(addr & 3) * 8u;
                const uint8_t *a0 = reinterpret_cast<const uint8_t *>(addr & ~(uintptr_t)3);
                const int64_t o0 = off - (int64_t)(addr & 3);   // offset of a0 relative to `images` (may be -3..0)
                uint4 *drow = reinterpret_cast<uint4 *>(src + (size_t)rr * src_stride);
                const bool safe = o0 >= 0 && o0 + 12 * (int64_t)ngroups + 4 <= image_bytes;
                for (int j = lane; j < ngroups; j += 32) {
                    unsigned w[4];
                    if (safe) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) w[q] = __ldg(reinterpret_cast<const unsigned *>(a0 + 12 * j + 4 * q));
                    } else {  // first / last bytes of the whole batch: assemble from the bytes that exist
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int64_t o = o0 + 12 * (int64_t)j + 4 * q;
                            w[q] = 0u;
#pragma unroll
                            for (int b = 0; b < 4; ++b)
                                if (o + b >= 0 && o + b < image_bytes) w[q] |= (unsigned)images[o + b] << (8 * b);
                        }
                    }
                    const unsigned b0 = __funnelshift_r(w[0], w[1], sh), b1 = __funnelshift_r(w[1], w[2], sh),
                                   b2 = __funnelshift_r(w[2], w[3], sh);
                    drow[j] = make_uint4(b0, __byte_perm(b0, b1, 0x0543), __byte_perm(b1, b2, 0x0432), b2 >> 8);
                }
            }
            __syncthreads();
            // 2. horizontal pass: a thread owns an output column (taps in registers) and walks down the rows
            for (int xx = threadIdx.x; xx < out_w; xx += kThreads) {
                const int xmin = bx[2 * xx], cnt = bx[2 * xx + 1];
                const int32_t *k = cx + (int64_t)xx * KX;
                const uint32_t *px = src + xmin;
                uint32_t *o = inter + xx + 1;
                const int cmax = __reduce_max_sync(__activemask(), cnt);   // warp-uniform loop bound
                if (cmax <= 4) {
                    int kk[4];
#pragma unroll
                    for (int t = 0; t < 4; ++t) kk[t] = (t < cnt) ? k[t] : 0;   // zero taps read the row padding
                    for (int rr = 0; rr < nrows; ++rr, px += src_stride, o += inter_stride) {
                        int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            if (t < cmax) {
                                const unsigned w = px[t];
                                a0 += (int)__byte_perm(w, 0u, 0x4440) * kk[t];
                                a1 += (int)__byte_perm(w, 0u, 0x4441) * kk[t];
                                a2 += (int)__byte_perm(w, 0u, 0x4442) * kk[t];
                            }
                        }
                        *o = (unsigned)(a0 >> kPrecisionBits) | ((unsigned)(a1 >> kPrecisionBits) << 8) |
                             ((unsigned)(a2 >> kPrecisionBits) << 16);
                    }
                } else {  // stronger down-scaling: taps stay in memory
                    for (int rr = 0; rr < nrows; ++rr, px += src_stride, o += inter_stride) {
                        int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
                        for (int t = 0; t < cnt; ++t) {
                            const int kt = k[t];
                            const unsigned w = px[t];
                            a0 += (int)(w & 0xffu) * kt;
                            a1 += (int)((w >> 8) & 0xffu) * kt;
                            a2 += (int)((w >> 16) & 0xffu) * kt;
                        }
                        *o = (unsigned)min(a0 >> kPrecisionBits, 255) | ((unsigned)min(a1 >> kPrecisionBits, 255) << 8) |
                             ((unsigned)min(a2 >> kPrecisionBits, 255) << 16);
                    }
                }
            }
            __syncthreads();
        }
    } else {
        __syncthreads();   // the look-up table
    }
    // 3. vertical pass, one cell per thread step: 64 threads per cell row
    const int cyl = threadIdx.x >> 6;
    const int cyy = (int)blockIdx.x * 4 + cyl;
    if (cyy >= sy) return;
    const int yy0 = 2 * cyy - 3;
    OutT *orow = out + (((int64_t)s * sy + cyy) * sx << 4);
    int ymin[2], ycnt[2], ky[2][4];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int yy = yy0 + r;
        const bool ok = rows_ok && yy >= 0 && yy < out_h;
        ymin[r] = ok ? by[2 * yy] - ylo : 0;
        ycnt[r] = ok ? by[2 * yy + 1] : 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) ky[r][t] = (ok && t < ycnt[r]) ? cy[(int64_t)yy * KY + t] : 0;   // row-uniform taps
    }
    for (int cxx = threadIdx.x & 63; cxx < sx; cxx += 64) {
        const int xd0 = 2 * cxx - 3;
        float f[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) f[i] = 0.f;
        const bool v0 = xd0 >= 0 && xd0 < out_w, v1 = xd0 + 1 >= 0 && xd0 + 1 < out_w;
        if (v0 || v1) {
            // words x+1 of the two source columns are an aligned pair: (2cx-2, 2cx-1), mirrored when flipping
            const int pair = flip ? out_w + 2 - 2 * cxx : 2 * cxx - 2;
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                if (ycnt[r] == 0) continue;
                const uint32_t *p = inter + (size_t)ymin[r] * inter_stride + pair;
                int a[6];
#pragma unroll
                for (int i = 0; i < 6; ++i) a[i] = 1 << (kPrecisionBits - 1);
                if (ycnt[r] <= 4) {
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        if (t < ycnt[r]) {
                            const uint2 w = *reinterpret_cast<const uint2 *>(p + (size_t)t * inter_stride);
                            const int kt = ky[r][t];
                            a[0] += (int)__byte_perm(w.x, 0u, 0x4440) * kt;
                            a[1] += (int)__byte_perm(w.x, 0u, 0x4441) * kt;
                            a[2] += (int)__byte_perm(w.x, 0u, 0x4442) * kt;
                            a[3] += (int)__byte_perm(w.y, 0u, 0x4440) * kt;
                            a[4] += (int)__byte_perm(w.y, 0u, 0x4441) * kt;
                            a[5] += (int)__byte_perm(w.y, 0u, 0x4442) * kt;
                        }
                    }
                } else {
                    const int32_t *k = cy + (int64_t)(yy0 + r) * KY;
                    for (int t = 0; t < ycnt[r]; ++t, p += inter_stride) {
                        const int kt = k[t];
                        const uint2 w = *reinterpret_cast<const uint2 *>(p);
                        a[0] += (int)(w.x & 0xffu) * kt;
                        a[1] += (int)((w.x >> 8) & 0xffu) * kt;
                        a[2] += (int)((w.x >> 16) & 0xffu) * kt;
                        a[3] += (int)(w.y & 0xffu) * kt;
                        a[4] += (int)((w.y >> 8) & 0xffu) * kt;
                        a[5] += (int)((w.y >> 16) & 0xffu) * kt;
                    }
                }
                // a[0..2] = lower source column, a[3..5] = upper; destination order swaps them when flipping.
                // No clamp: see the note above the kernel (the sum is provably < 256 << 22 and >= 0).
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    const int e0 = flip ? a[3 + ch] : a[ch], e1 = flip ? a[ch] : a[3 + ch];
                    if (v0) f[r * 6 + ch] = lut[ch * 256 + (e0 >> kPrecisionBits)];
                    if (v1) f[r * 6 + 3 + ch] = lut[ch * 256 + (e1 >> kPrecisionBits)];
                }
            }
        }
        store_cell<OutT>(orow + ((int64_t)cxx << 4), f);
    }
}

// Ask for the largest shared-memory carve-out, so that shared memory does not cap the resident CTAs below what the
// registers allow (a per-device function attribute: set on every call, it is a cheap host-side operation).
template <typename K>
cudaError_t prefer_shared(K kern) {
    return cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

int max_src_rows_for(int src_h, int out_h) {
    const double s = (double)src_h / (double)out_h;
    const double sup = s < 1.0 ? 1.0 : s;
    return (int)((kTileRows - 1) * s + 2.0 * sup + 1.0) + 2;
}

}  // namespace

extern "C" size_t b200_input_workspace_bytes(int64_t n, int src_h, int src_w, int out_h, int out_w) {
    if (n < 1 || src_h < 1 || src_w < 1 || out_h < 1 || out_w < 1) return 0;
    const int KX = ksize_for(src_w, out_w), KY = ksize_for(src_h, out_h);
    // per-sample coefficient tables, then the 3x256 float normalisation table
    return sizeof(int32_t) * (size_t)n * ((size_t)out_w * (2 + KX) + (size_t)out_h * (2 + KY)) + 3 * 256 * sizeof(float) + 16;
}

// Shared host side of the dense and the ragged entry point.  Dense: tab == nullptr, images = [n, src_h, src_w, 3].
// Ragged: tab = int64 [n][4] on the device, src_h / src_w are upper bounds over the batch, image_bytes the slab size.
static int input_stage_launch(const uint8_t *images, int64_t image_bytes, const int64_t *tab, const int32_t *boxes,
                              const uint8_t *flips, void *out, void *workspace, size_t workspace_bytes, int64_t n,
                              int src_h, int src_w, int out_h, int out_w, const float *mean3, const float *std3,
                              int out_dtype, int out_layout, void *stream) {
    if (!images || !boxes || !out || !workspace || !mean3 || !std3) return (int)cudaErrorInvalidValue;
    if (n < 1 || src_h < 1 || src_w < 1 || out_h < 1 || out_w < 1) return (int)cudaErrorInvalidValue;
    if (out_dtype != B200_F32 && out_dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (out_layout != B200_LAYOUT_NHWC && out_layout != B200_LAYOUT_NCHW && out_layout != B200_LAYOUT_S2D_STEM)
        return (int)cudaErrorInvalidValue;
    if (out_layout == B200_LAYOUT_S2D_STEM && ((out_h | out_w) & 1)) return (int)cudaErrorInvalidValue;
    if ((int64_t)src_h > 8 * (int64_t)out_h || (int64_t)src_w > 8 * (int64_t)out_w) return (int)cudaErrorInvalidValue;
    if (n > 65535) return (int)cudaErrorInvalidValue;
    if (workspace_bytes < b200_input_workspace_bytes(n, src_h, src_w, out_h, out_w)) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(workspace) || b200_misaligned(out, 4)) return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int KX = ksize_for(src_w, out_w), KY = ksize_for(src_h, out_h);
    int32_t *ws = static_cast<int32_t *>(workspace);
    const int64_t items = n * (int64_t)(out_w + out_h);
    float *lut_g = reinterpret_cast<float *>(ws + n * ((int64_t)out_w * (2 + KX) + (int64_t)out_h * (2 + KY)));
    Norm nrm0;
    for (int i = 0; i < 3; ++i) { nrm0.mean[i] = mean3[i]; nrm0.std[i] = std3[i]; }
    const int64_t items_l = items > 3 * 256 ? items : 3 * 256;
    input_coeff_kernel<<<(unsigned)((items_l + kThreads - 1) / kThreads), kThreads, 0, st>>>(boxes, n, out_h, out_w, KX,
                                                                                             KY, ws, lut_g, nrm0);
    B200_CHECK_LAUNCH();
    const int max_rows = max_src_rows_for(src_h, out_h);
    const size_t esz = out_dtype == B200_F32 ? 4 : 2;
    if (out_layout == B200_LAYOUT_S2D_STEM && g_input_s2d_fast) {
        const int src_stride = (src_w + kS2dPadWords + 3) / 4 * 4, inter_stride = out_w + 2;
        const size_t smem2 = 3 * 256 * sizeof(float) + (size_t)max_rows * (src_stride + inter_stride) * sizeof(uint32_t);
        if (smem2 <= 96 * 1024) {   // larger boxes / stronger down-scaling use the generic kernel below
            const dim3 grid2((out_h + 6 + kTileRows - 1) / kTileRows, (unsigned)n);
            cudaError_t e = cudaSuccess;
            if (out_dtype == B200_F32) {
                auto kern = input_s2d_kernel<float>;
                if (smem2 > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
                if (e == cudaSuccess) e = prefer_shared(kern);
                if (e != cudaSuccess) return (int)e;
                kern<<<grid2, kThreads, smem2, st>>>(images, image_bytes, boxes, flips, static_cast<float *>(out), ws, lut_g,
                                                     src_h, src_w, out_h, out_w, KX, KY, max_rows, src_stride, inter_stride,
                                                     tab);
            } else {
                auto kern = input_s2d_kernel<__nv_bfloat16>;
                if (smem2 > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
                if (e == cudaSuccess) e = prefer_shared(kern);
                if (e != cudaSuccess) return (int)e;
                kern<<<grid2, kThreads, smem2, st>>>(images, image_bytes, boxes, flips, static_cast<__nv_bfloat16 *>(out), ws,
                                                     lut_g, src_h, src_w, out_h, out_w, KX, KY, max_rows, src_stride,
                                                     inter_stride, tab);
            }
            B200_CHECK_LAUNCH();
            return 0;
        }
    }
    const int inter_bytes = (int)(((size_t)max_rows * out_w * 3 + 15) / 16 * 16);
    size_t tile_bytes = 0;
    if (out_layout == B200_LAYOUT_NHWC) tile_bytes = (size_t)kTileRows * out_w * 3 * esz;
    if (out_layout == B200_LAYOUT_S2D_STEM) tile_bytes = (size_t)4 * ((out_w + 6) / 2) * 16 * esz;
    const size_t smem = 3 * 256 * sizeof(float) + (size_t)inter_bytes + tile_bytes;
    if (smem > 200 * 1024) return (int)cudaErrorInvalidValue;
    Norm nrm;
    for (int i = 0; i < 3; ++i) { nrm.mean[i] = mean3[i]; nrm.std[i] = std3[i]; }
    const int tiles = out_layout == B200_LAYOUT_S2D_STEM ? (out_h + 6 + kTileRows - 1) / kTileRows
                                                         : (out_h + kTileRows - 1) / kTileRows;
    const dim3 grid(tiles, (unsigned)n);
    if (out_dtype == B200_F32) {
        auto kern = input_resample_kernel<float>;
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return (int)e;
        }
        kern<<<grid, kThreads, smem, st>>>(images, boxes, flips, static_cast<float *>(out), ws, src_h, src_w, out_h,
                                           out_w, KX, KY, max_rows, nrm, out_layout, inter_bytes, tab);
    } else {
        auto kern = input_resample_kernel<__nv_bfloat16>;
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return (int)e;
        }
        kern<<<grid, kThreads, smem, st>>>(images, boxes, flips, static_cast<__nv_bfloat16 *>(out), ws, src_h, src_w,
                                           out_h, out_w, KX, KY, max_rows, nrm, out_layout, inter_bytes, tab);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_input_crop_resize_flip_norm(const uint8_t *images, const int32_t *boxes, const uint8_t *flips,
                                                void *out, void *workspace, size_t workspace_bytes, int64_t n,
                                                int src_h, int src_w, int out_h, int out_w, const float *mean3,
                                                const float *std3, int out_dtype, int out_layout, void *stream) {
    if (n < 1 || src_h < 1 || src_w < 1) return (int)cudaErrorInvalidValue;
    return input_stage_launch(images, n * (int64_t)src_h * src_w * 3, nullptr, boxes, flips, out, workspace,
                              workspace_bytes, n, src_h, src_w, out_h, out_w, mean3, std3, out_dtype, out_layout, stream);
}

extern "C" int b200_input_crop_resize_flip_norm_ragged(const uint8_t *slab, int64_t slab_bytes, const int64_t *src_table,
                                                       const int32_t *boxes, const uint8_t *flips, void *out,
                                                       void *workspace, size_t workspace_bytes, int64_t n,
                                                       int max_src_h, int max_src_w, int out_h, int out_w,
                                                       const float *mean3, const float *std3, int out_dtype,
                                                       int out_layout, void *stream) {
    if (!src_table || slab_bytes < 1) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(src_table, 8)) return (int)cudaErrorMisalignedAddress;
    return input_stage_launch(slab, slab_bytes, src_table, boxes, flips, out, workspace, workspace_bytes, n, max_src_h,
                              max_src_w, out_h, out_w, mean3, std3, out_dtype, out_layout, stream);
}

// ---- decode staging --------------------------------------------------------------------------------
// A JPEG decoder on the device (nvJPEG) delivers every image as three planes [3][h][w]; the input stage reads
// interleaved RGB rows.  One launch re-packs a whole batch of planar images of different sizes into the ragged slab:
// table row i = {address of image i's planes, h, w, byte offset of image i in the slab}; a thread moves 4 consecutive
// pixels (12 plane bytes in, three aligned 32-bit words out), image i's pixels are dealt to the CTAs of blockIdx.y = i.
namespace {
__global__ void __launch_bounds__(kThreads)
interleave_planar_kernel(const int64_t *__restrict__ table, uint8_t *__restrict__ slab) {
    const int64_t *row = table + 4 * (int64_t)blockIdx.y;
    const uint8_t *src = reinterpret_cast<const uint8_t *>(static_cast<uintptr_t>(row[0]));
    const int64_t npix = row[1] * row[2];
    uint8_t *dst = slab + row[3];
    const uint8_t *r = src, *g = src + npix, *b = src + 2 * npix;
    for (int64_t q = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4; q < npix; q += (int64_t)gridDim.x * kThreads * 4) {
        if (q + 4 <= npix) {
            unsigned v[12];
#pragma unroll
            for (int i = 0; i < 4; ++i) { v[3 * i] = r[q + i]; v[3 * i + 1] = g[q + i]; v[3 * i + 2] = b[q + i]; }
            unsigned *o = reinterpret_cast<unsigned *>(dst + 3 * q);   // slab offsets are multiples of 16, 3q of 12
#pragma unroll
            for (int w = 0; w < 3; ++w)
                o[w] = v[4 * w] | (v[4 * w + 1] << 8) | (v[4 * w + 2] << 16) | (v[4 * w + 3] << 24);
        } else {
            for (int64_t i = q; i < npix; ++i) { dst[3 * i] = r[i]; dst[3 * i + 1] = g[i]; dst[3 * i + 2] = b[i]; }
        }
    }
}
}  // namespace

extern "C" int b200_interleave_planar_u8(const int64_t *table, uint8_t *slab, int64_t n, int64_t max_pixels,
                                         void *stream) {
    if (!table || !slab || n < 1 || n > 65535 || max_pixels < 1) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(table, 8) || b200_misaligned(slab)) return (int)cudaErrorMisalignedAddress;
    int64_t per_image = (max_pixels + 4 * kThreads * 4 - 1) / (4 * kThreads * 4);   // >= 4 steps per thread
    if (per_image > 64) per_image = 64;
    if (per_image < 1) per_image = 1;
    interleave_planar_kernel<<<dim3((unsigned)per_image, (unsigned)n), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        table, slab);
    B200_CHECK_LAUNCH();
    return 0;
}
