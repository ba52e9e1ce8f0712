// K2: BatchNorm2d (+ residual add) + ReLU, forward (train / eval) and backward, on channels-last
// rows [M = N*H*W, C].
// Reference sites: torchvision resnet.py:93-103,147-161,240-243,269-270 (bn, relu, `out += identity`);
// arithmetic torch/nn/modules/batchnorm.py:160-206; gradients = batch_norm_backward +
// threshold_backward + the add's fan-out, executed by loss.backward() (imagenet.py:128).
//
// All kernels are HBM-bound streams over the activation:
//   fwd train : stats pass reads x (1 E) ; apply pass reads x (+z) and writes y (2-3 E)
//   bwd       : reduce pass reads dy,x (2 E) ; apply pass reads dy,x and writes dx (+dz) (3-4 E)
// plus O(C * chunks) floats of partial sums.  Layout of the work: a CTA of 256 threads is
// TX channel-vectors wide (a vector = VEC consecutive channels = one 8/16-byte access) and
// TY = 256/TX rows tall, so a warp always touches >= 256 contiguous bytes; each thread keeps its
// channels' constants in registers and walks down the rows of its chunk with U independent vector
// loads in flight per tensor.  Per-thread partial sums are combined with warp shuffles (lanes that
// share a channel), staged through shared memory across warps, and written as one partial per
// (channel, chunk); a one-warp-per-channel finalize kernel folds the chunks.  The grid is
// (channel tiles) x (row chunks) ~= 148 SMs x 8.
//
// Variance: each chunk accumulates sums shifted by the chunk's first row (sum(x-K), sum((x-K)^2)),
// converts to (mean_b, M2_b) and the finalize combines chunks with the pairwise-update formula
// (M2 = sum M2_b + n_b (mean_b - mean)^2), so there is no E[x^2]-E[x]^2 cancellation.
#include <cooperative_groups.h>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

// tuning knobs (b200_tune_set)
int g_bn_target_ctas = B200_SM_COUNT * 8;
int g_bn_vec_bf16 = 8;
int g_bn_stats_reverse = 1;  // stats/reduce passes walk chunks back to front (tail of the producer's
                             // output is the part most likely still in L2), apply passes front to back
int g_bn_apply_reverse = 0;
int g_bn_pipeline = 1;       // 1 = bulk-async (TMA) shared-memory ring kernels where eligible, 0 = direct loads
int g_bn_pipe_ctas_per_sm = 2;
int g_bn_pipe_stages = 0;    // 0 = automatic
int g_bn_pdl = 1;            // finalize / apply kernels of the bulk-async path use programmatic dependent launch
int g_bn_fused = 0;          // 1 = one cooperative kernel per site (bn_fused.cuh) where the shape allows, 0 = the chain.
                             // Off by default: measured SLOWER than the chain on every ResNet-50 shape (see bn_fused.cuh)

template <typename T, int VEC>
struct Unroll {
    static constexpr int U = (32 / (VEC * (int)sizeof(T))) < 1 ? 1 : (32 / (VEC * (int)sizeof(T)));
};

struct Geom {
    int cv;          // channel vectors per row
    int txl;         // log2(TX)
    int tiles;       // channel tiles = cv / TX
    int nchunk;      // row chunks
    int64_t rpc;     // rows per chunk
};

Geom make_geom(int64_t m, int c, int vec, int unroll) {
    Geom g;
    g.cv = c / vec;
    int tx = 1, l = 0;
    while (tx < 32 && (g.cv % (tx * 2)) == 0) { tx *= 2; ++l; }
    g.txl = l;
    g.tiles = g.cv / tx;
    const int ty = kThreads / tx;
    const int64_t step = (int64_t)ty * unroll;
    const int64_t max_chunks = (m + step - 1) / step;
    int64_t want = g_bn_target_ctas / g.tiles;
    if (want < 1) want = 1;
    if (want > max_chunks) want = max_chunks;
    int64_t rpc = (m + want - 1) / want;
    rpc = (rpc + step - 1) / step * step;
    g.rpc = rpc;
    g.nchunk = (int)((m + rpc - 1) / rpc);
    return g;
}

// Sum `val[0..NV)` over the threads of the CTA that share a channel vector.  On return threads with
// threadIdx.x < TX hold the CTA totals of their channel vector.
template <int NV>
__device__ __forceinline__ void cta_reduce_rows(float (&val)[NV], int tx, float *smem /*[kThreads][NV]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (tx <= 32) {
#pragma unroll
        for (int k = 0; k < NV; ++k)
            for (int o = 16; o >= tx; o >>= 1) val[k] += __shfl_xor_sync(0xffffffffu, val[k], o);
        // every warp covers all TX channel vectors: lanes < TX hold the warp's sums
        if (lane < tx) {
#pragma unroll
            for (int k = 0; k < NV; ++k) smem[(warp * 32 + lane) * NV + k] = val[k];
        }
        __syncthreads();
        if (threadIdx.x < tx) {
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                float s = 0.f;
#pragma unroll
                for (int w = 0; w < kWarps; ++w) s += smem[(w * 32 + threadIdx.x) * NV + k];
                val[k] = s;
            }
        }
    } else {
        // TX in {64, 128, 256}: threads t, t+TX, ... share a channel vector and sit in different warps
#pragma unroll
        for (int k = 0; k < NV; ++k) smem[threadIdx.x * NV + k] = val[k];
        __syncthreads();
        if (threadIdx.x < tx) {
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                float s = 0.f;
                for (int j = threadIdx.x; j < kThreads; j += tx) s += smem[j * NV + k];
                val[k] = s;
            }
        }
    }
}

#include "bn_pipe.cuh"
#include "bn_fused.cuh"

// Geometry of the bulk-async kernels: every CTA owns whole rows (TX = CV <= 256), tiles of kTileVecs
// vectors.  The number of chunks is one wave of resident CTAs (148 x g_bn_pipe_ctas_per_sm), capped
// so that the per-chunk partial sums stay under ~3% of the tensor (never below one CTA per SM); the
// stage count makes the resident CTAs of an SM keep ~200 KB in flight.
bool pipe_eligible(int c, int vec, int esize) {
    if (!g_bn_pipeline || vec * esize != 16) return false;
    const int cv = c / vec;
    return cv >= 2 && cv <= kThreads && (cv & (cv - 1)) == 0;
}

struct PipeGeom {
    Geom g;
    int stages;
    size_t smem;
};

PipeGeom make_geom_pipe(int64_t m, int c, int vec, int esize, int nt, bool mask, size_t static_smem) {
    PipeGeom pg;
    Geom &g = pg.g;
    g.cv = c / vec;
    int l = 0;
    while ((1 << l) < g.cv) ++l;
    g.txl = l;
    g.tiles = 1;
    const int64_t rpt = kTileVecs / g.cv;  // rows per tile
    const int64_t max_chunks = (m + rpt - 1) / rpt;
    int64_t want = (int64_t)B200_SM_COUNT * g_bn_pipe_ctas_per_sm;
    const int64_t by_partial = (int64_t)((double)m * c * esize * 0.03 / ((double)c * 8.0));
    if (want > by_partial) want = by_partial < B200_SM_COUNT ? B200_SM_COUNT : by_partial;
    if (want > max_chunks) want = max_chunks;
    if (want < 1) want = 1;
    int64_t rpc = (m + want - 1) / want;
    rpc = (rpc + rpt - 1) / rpt * rpt;
    g.rpc = rpc;
    g.nchunk = (int)((m + rpc - 1) / rpc);
    const int per_sm = (g.nchunk + B200_SM_COUNT - 1) / B200_SM_COUNT;
    const size_t budget = (size_t)200 * 1024 / per_sm;
    int st = (int)((budget > static_smem + kRingHeader ? budget - static_smem - kRingHeader : 0) / pipe_stage_bytes(nt, mask));
    if (st > kMaxStages) st = kMaxStages;
    if (st < 2) st = 2;
    if (g_bn_pipe_stages > 0) {   // forced depth (tuning knob), but never more than fits beside the other resident CTAs
        st = g_bn_pipe_stages;
        while (st > 2 && static_smem + pipe_smem_bytes(nt, mask, st) > (size_t)220 * 1024 / per_sm) --st;
    }
    pg.stages = st;
    pg.smem = pipe_smem_bytes(nt, mask, st);
    return pg;
}

template <typename K>
cudaError_t allow_smem(K kern, size_t bytes) {
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

// Launch `kern`; with `dependent` (and the knob on) it carries the programmatic-stream-serialization
// attribute, i.e. it may begin while the previous kernel of the stream is still draining (see pdl_wait).
template <typename... KArgs, typename... Args>
cudaError_t launch(void (*kern)(KArgs...), unsigned grid, size_t smem, cudaStream_t st, bool dependent, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (dependent && g_bn_pdl) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Cooperative launch (grid-wide barriers inside the kernel): the whole grid must be resident at once.
template <typename... KArgs, typename... Args>
cudaError_t launch_coop(void (*kern)(KArgs...), unsigned grid, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// How many CTAs of `kern` with `smem` dynamic bytes fit on the device at once (0 on error); cached per (kernel, smem).
// (The dynamic-shared-memory attribute it sets for the query is per kernel, not per size: callers set it again before
// every launch, as the chain's launches do.)
template <typename K>
int coop_capacity(K kern, size_t smem) {
    struct Entry { const void *k; size_t smem; int dev; int cap; };
    static Entry cache[64];
    static int used = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    const void *key = reinterpret_cast<const void *>(kern);
    for (int i = 0; i < used; ++i)
        if (cache[i].k == key && cache[i].smem == smem && cache[i].dev == dev) return cache[i].cap;
    if (allow_smem(kern, smem) != cudaSuccess) return 0;
    int per_sm = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    const int cap = per_sm * sms;
    if (used < 64) cache[used++] = Entry{key, smem, dev, cap};
    return cap;
}

// Geometry of a fused (two-pass, cooperative) site: one chunk per CTA for both passes, at most `cap` CTAs.
struct FusedGeom {
    Geom g;
    int stages1, stages2;
    size_t smem;
};

FusedGeom make_geom_fused(int64_t m, int c, int vec, int esize, int nt1, bool mask1, int nt2, bool mask2,
                          size_t static_smem) {
    FusedGeom fg;
    const bool first_larger = pipe_stage_bytes(nt1, mask1) >= pipe_stage_bytes(nt2, mask2);
    const PipeGeom big = make_geom_pipe(m, c, vec, esize, first_larger ? nt1 : nt2, first_larger ? mask1 : mask2, static_smem);
    fg.g = big.g;
    fg.smem = big.smem;
    const size_t data = big.smem - kRingHeader;
    int s1 = (int)(data / pipe_stage_bytes(nt1, mask1)), s2 = (int)(data / pipe_stage_bytes(nt2, mask2));
    fg.stages1 = s1 > kMaxStages ? kMaxStages : s1;
    fg.stages2 = s2 > kMaxStages ? kMaxStages : s2;
    return fg;
}

// ------------------------------------------------------------------------------------------------
// forward pass 1: per-(channel, chunk) mean and M2
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads)
bn_stats_kernel(const T *__restrict__ x, float2 *__restrict__ partial, int64_t m, int c, int txl, int nchunk,
                int64_t rpc, int reverse) {
    constexpr int U = Unroll<T, VEC>::U;
    __shared__ float smem[kWarps * 32 * 2 * VEC];
    const int tx = 1 << txl, ty = kThreads >> txl;
    const int txi = threadIdx.x & (tx - 1), tyi = threadIdx.x >> txl;
    const int cvi = blockIdx.x * tx + txi;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.y) : (int)blockIdx.y;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    const T *xp = x + (int64_t)cvi * VEC;
    float sh[VEC];
    VecIO<T, VEC>::load(xp + r0 * c, sh);
    float acc[2 * VEC];
#pragma unroll
    for (int k = 0; k < 2 * VEC; ++k) acc[k] = 0.f;
    for (int64_t r = r0 + tyi; r < r1; r += (int64_t)ty * U) {
        float v[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t rr = r + (int64_t)u * ty;
            if (rr < r1) {
                VecIO<T, VEC>::load(xp + rr * c, v[u]);
            } else {
#pragma unroll
                for (int k = 0; k < VEC; ++k) v[u][k] = sh[k];
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                const float d = v[u][k] - sh[k];
                acc[k] += d;
                acc[VEC + k] = fmaf(d, d, acc[VEC + k]);
            }
    }
    cta_reduce_rows<2 * VEC>(acc, tx, smem);
    if (threadIdx.x < tx) {
        const float nb = (float)(r1 - r0);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const float s1 = acc[k], s2 = acc[VEC + k];
            const float mean_b = sh[k] + s1 / nb;
            const float m2_b = fmaxf(s2 - s1 * s1 / nb, 0.f);
            partial[(int64_t)(cvi * VEC + k) * nchunk + chunk] = make_float2(mean_b, m2_b);
        }
    }
}

// forward pass 2: one warp per channel folds the chunks, updates the running statistics and
// writes saved = {mean, invstd, scale, shift}
__global__ void __launch_bounds__(kThreads)
bn_fwd_finalize_kernel(const float2 *__restrict__ partial, int nchunk, int64_t rpc, int64_t m, int c,
                       const float *__restrict__ gamma, const float *__restrict__ beta, float *__restrict__ rmean,
                       float *__restrict__ rvar, int64_t *__restrict__ nbt, float *__restrict__ saved, float eps,
                       float momentum) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ch = blockIdx.x * kWarps + warp;
    if (blockIdx.x == 0 && threadIdx.x == 0 && nbt) *nbt += 1;
    if (ch >= c) return;
    const float2 *p = partial + (int64_t)ch * nchunk;
    const float last_n = (float)(m - (int64_t)(nchunk - 1) * rpc);
    float s = 0.f;
    for (int i = lane; i < nchunk; i += 32) s = fmaf(i == nchunk - 1 ? last_n : (float)rpc, p[i].x, s);
    s = warp_sum(s);
    const float mean = s / (float)m;
    float q = 0.f;
    for (int i = lane; i < nchunk; i += 32) {
        const float2 v = p[i];
        const float d = v.x - mean;
        q += fmaf(i == nchunk - 1 ? last_n : (float)rpc, d * d, v.y);
    }
    q = warp_sum(q);
    if (lane == 0) {
        const float var = q / (float)m;
        const float invstd = 1.0f / sqrtf(var + eps);
        const float sc = gamma[ch] * invstd;
        saved[ch] = mean;
        saved[c + ch] = invstd;
        saved[2 * c + ch] = sc;
        saved[3 * c + ch] = fmaf(-mean, sc, beta[ch]);
        if (rmean) rmean[ch] = fmaf(momentum, mean, (1.f - momentum) * rmean[ch]);
        if (rvar) rvar[ch] = fmaf(momentum, q / (float)(m - 1), (1.f - momentum) * rvar[ch]);
    }
}

// forward pass 3 (and the whole of eval mode): y = act(x*scale + shift [+ z])
template <typename T, int VEC, int ACT, bool EVAL>
__global__ void __launch_bounds__(kThreads)
bn_apply_kernel(const T *__restrict__ x, const T *__restrict__ z, T *__restrict__ y, uint8_t *__restrict__ mask,
                const float *__restrict__ saved, const float *__restrict__ gamma, const float *__restrict__ beta,
                const float *__restrict__ rmean, const float *__restrict__ rvar, float eps, int64_t m, int c, int txl,
                int nchunk, int64_t rpc, int reverse) {
    constexpr int U = Unroll<T, VEC>::U;
    const int tx = 1 << txl, ty = kThreads >> txl;
    const int txi = threadIdx.x & (tx - 1), tyi = threadIdx.x >> txl;
    const int cvi = blockIdx.x * tx + txi;
    const int cv = c / VEC;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.y) : (int)blockIdx.y;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    const int64_t coff = (int64_t)cvi * VEC;
    float sc[VEC], sf[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int ch = cvi * VEC + k;
        if (EVAL) {
            const float invstd = 1.0f / sqrtf(rvar[ch] + eps);
            sc[k] = gamma[ch] * invstd;
            sf[k] = fmaf(-rmean[ch], sc[k], beta[ch]);
        } else {
            sc[k] = saved[2 * c + ch];
            sf[k] = saved[3 * c + ch];
        }
    }
    for (int64_t r = r0 + tyi; r < r1; r += (int64_t)ty * U) {
        float v[U][VEC], w[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t rr = r + (int64_t)u * ty;
            if (rr < r1) {
                VecIO<T, VEC>::load_stream(x + rr * c + coff, v[u]);
                if (ACT == B200_ACT_ADD_RELU) VecIO<T, VEC>::load_stream(z + rr * c + coff, w[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t rr = r + (int64_t)u * ty;
            if (rr < r1) {
                unsigned bits = 0;
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    float pre = fmaf(v[u][k], sc[k], sf[k]);
                    if (ACT == B200_ACT_ADD_RELU) pre += w[u][k];
                    if (ACT != B200_ACT_NONE) {
                        bits |= (pre > 0.f ? 1u : 0u) << k;
                        pre = fmaxf(pre, 0.f);
                    }
                    v[u][k] = pre;
                }
                VecIO<T, VEC>::store(y + rr * c + coff, v[u]);
                if (ACT == B200_ACT_ADD_RELU && mask) mask[rr * cv + cvi] = (uint8_t)bits;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// backward pass 1: per-(channel, chunk) sum(dyh) and sum(dyh * (x - mean))
template <typename T, int VEC, int ACT>
__global__ void __launch_bounds__(kThreads)
bn_bwd_reduce_kernel(const T *__restrict__ dy, const T *__restrict__ x, const uint8_t *__restrict__ mask,
                     const float *__restrict__ saved, float2 *__restrict__ partial, int64_t m, int c, int txl,
                     int nchunk, int64_t rpc, int reverse) {
    constexpr int U = Unroll<T, VEC>::U;
    __shared__ float smem[kWarps * 32 * 2 * VEC];
    const int tx = 1 << txl, ty = kThreads >> txl;
    const int txi = threadIdx.x & (tx - 1), tyi = threadIdx.x >> txl;
    const int cvi = blockIdx.x * tx + txi;
    const int cv = c / VEC;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.y) : (int)blockIdx.y;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    const int64_t coff = (int64_t)cvi * VEC;
    float mean[VEC], sc[VEC], sf[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int ch = cvi * VEC + k;
        mean[k] = saved[ch];
        if (ACT == B200_ACT_RELU) { sc[k] = saved[2 * c + ch]; sf[k] = saved[3 * c + ch]; }
    }
    float acc[2 * VEC];
#pragma unroll
    for (int k = 0; k < 2 * VEC; ++k) acc[k] = 0.f;
    for (int64_t r = r0 + tyi; r < r1; r += (int64_t)ty * U) {
        float g[U][VEC], v[U][VEC];
        unsigned bits[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t rr = r + (int64_t)u * ty;
            if (rr < r1) {
                VecIO<T, VEC>::load(dy + rr * c + coff, g[u]);
                VecIO<T, VEC>::load(x + rr * c + coff, v[u]);
                if (ACT == B200_ACT_ADD_RELU) bits[u] = mask[rr * cv + cvi];
            } else {
#pragma unroll
                for (int k = 0; k < VEC; ++k) { g[u][k] = 0.f; v[u][k] = mean[k]; }
                bits[u] = 0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                float gk = g[u][k];
                if (ACT == B200_ACT_RELU) gk = fmaf(v[u][k], sc[k], sf[k]) > 0.f ? gk : 0.f;
                if (ACT == B200_ACT_ADD_RELU) gk = ((bits[u] >> k) & 1u) ? gk : 0.f;
                acc[k] += gk;
                acc[VEC + k] = fmaf(gk, v[u][k] - mean[k], acc[VEC + k]);
            }
    }
    cta_reduce_rows<2 * VEC>(acc, tx, smem);
    if (threadIdx.x < tx) {
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            partial[(int64_t)(cvi * VEC + k) * nchunk + chunk] = make_float2(acc[k], acc[VEC + k]);
    }
}

// backward pass 2: fold chunks; dgamma, dbeta; coefficients of dx = c0*dyh + c1*(x-mean) + c2
__global__ void __launch_bounds__(kThreads)
bn_bwd_finalize_kernel(const float2 *__restrict__ partial, int nchunk, int64_t m, int c,
                       const float *__restrict__ gamma, const float *__restrict__ saved, float *__restrict__ dgamma,
                       float *__restrict__ dbeta, float *__restrict__ coef) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ch = blockIdx.x * kWarps + warp;
    if (ch >= c) return;
    const float2 *p = partial + (int64_t)ch * nchunk;
    float a = 0.f, b = 0.f;
    for (int i = lane; i < nchunk; i += 32) {
        const float2 v = p[i];
        a += v.x;
        b += v.y;
    }
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0) {
        const float invstd = saved[c + ch];
        const float gs = gamma[ch] * invstd;
        const float inv_m = 1.0f / (float)m;
        const float ps = b * invstd * invstd * inv_m;
        const float gm = a * inv_m;
        if (dgamma) dgamma[ch] = b * invstd;
        if (dbeta) dbeta[ch] = a;
        coef[ch] = gs;
        coef[c + ch] = -ps * gs;
        coef[2 * c + ch] = -gm * gs;
    }
}

// backward pass 3: dx (and dz) from dy, x
template <typename T, int VEC, int ACT>
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_kernel(const T *__restrict__ dy, const T *__restrict__ x, const uint8_t *__restrict__ mask,
                    const float *__restrict__ saved, const float *__restrict__ coef, T *__restrict__ dx,
                    T *__restrict__ dz, int64_t m, int c, int txl, int nchunk, int64_t rpc, int reverse) {
    constexpr int U = Unroll<T, VEC>::U;
    const int tx = 1 << txl, ty = kThreads >> txl;
    const int txi = threadIdx.x & (tx - 1), tyi = threadIdx.x >> txl;
    const int cvi = blockIdx.x * tx + txi;
    const int cv = c / VEC;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.y) : (int)blockIdx.y;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    const int64_t coff = (int64_t)cvi * VEC;
    float mean[VEC], c0[VEC], c1[VEC], c2[VEC], sc[VEC], sf[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int ch = cvi * VEC + k;
        mean[k] = saved[ch];
        c0[k] = coef[ch];
        c1[k] = coef[c + ch];
        c2[k] = coef[2 * c + ch];
        if (ACT == B200_ACT_RELU) { sc[k] = saved[2 * c + ch]; sf[k] = saved[3 * c + ch]; }
    }
    for (int64_t r = r0 + tyi; r < r1; r += (int64_t)ty * U) {
        float g[U][VEC], v[U][VEC];
        unsigned bits[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t rr = r + (int64_t)u * ty;
            if (rr < r1) {
                VecIO<T, VEC>::load_stream(dy + rr * c + coff, g[u]);
                VecIO<T, VEC>::load_stream(x + rr * c + coff, v[u]);
                if (ACT == B200_ACT_ADD_RELU) bits[u] = mask[rr * cv + cvi];
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t rr = r + (int64_t)u * ty;
            if (rr < r1) {
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    float gk = g[u][k];
                    if (ACT == B200_ACT_RELU) gk = fmaf(v[u][k], sc[k], sf[k]) > 0.f ? gk : 0.f;
                    if (ACT == B200_ACT_ADD_RELU) gk = ((bits[u] >> k) & 1u) ? gk : 0.f;
                    g[u][k] = gk;
                    v[u][k] = fmaf(gk, c0[k], fmaf(v[u][k] - mean[k], c1[k], c2[k]));
                }
                VecIO<T, VEC>::store(dx + rr * c + coff, v[u]);
                if (ACT == B200_ACT_ADD_RELU) VecIO<T, VEC>::store(dz + rr * c + coff, g[u]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
struct Ws {
    float2 *partial;
    float *coef;
};

// sized for the largest user: the dual backward keeps float4 partials and two coefficient sets
size_t ws_bytes_for(const Geom &g, int c) { return sizeof(float4) * (size_t)c * g.nchunk + sizeof(float) * 6 * (size_t)c + 64; }

Ws carve(void *ws, const Geom &g, int c) {
    Ws w;
    w.partial = static_cast<float2 *>(ws);
    size_t off = sizeof(float4) * (size_t)c * g.nchunk;
    off = (off + 15) / 16 * 16;
    w.coef = reinterpret_cast<float *>(static_cast<char *>(ws) + off);
    return w;
}

template <typename T, int VEC>
int fwd_train_impl(const void *x_, const void *z_, void *y_, uint8_t *mask, const float *gamma, const float *beta,
                   float *rm, float *rv, int64_t *nbt, float *saved, void *ws, size_t ws_bytes, int64_t m, int c,
                   float eps, float momentum, int act, cudaStream_t st, const float *saved2 = nullptr) {
    // y_ == nullptr: statistics + finalize only.  saved2 != nullptr: z is a raw tensor to be normalised
    // with saved2{scale, shift} before the add (bulk-async kernels only).
    const T *x = static_cast<const T *>(x_);
    const T *z = static_cast<const T *>(z_);
    T *y = static_cast<T *>(y_);
    if constexpr (VEC * sizeof(T) == 16) if (pipe_eligible(c, VEC, sizeof(T))) {
        const size_t red_bytes = sizeof(float) * kThreads * 2 * VEC;
        if (g_bn_fused) {
            // one cooperative kernel: statistics, fold and apply (bn_fused.cuh)
            const int nt2 = (y && act == B200_ACT_ADD_RELU) ? 2 : 1;
            const FusedGeom fg = make_geom_fused(m, c, VEC, sizeof(T), 1, false, nt2, false, red_bytes);
            const Geom &g = fg.g;
            if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
            const Ws w = carve(ws, g, c);
            cudaError_t e = cudaErrorUnknown;
#define B200_FWD_FUSED(ACT)                                                                                       \
    do {                                                                                                          \
        auto k = bn_fwd_fused_kernel<T, VEC, ACT>;                                                                \
        if (coop_capacity(k, fg.smem) >= g.nchunk && allow_smem(k, fg.smem) == cudaSuccess)                       \
            e = launch_coop(k, g.nchunk, fg.smem, st, x, z, y, mask, w.partial, gamma, beta, rm, rv, nbt, saved,  \
                            saved2, eps, momentum, m, c, g.txl, g.nchunk, g.rpc, fg.stages1, fg.stages2);         \
    } while (0)
            if (!y || act == B200_ACT_NONE) B200_FWD_FUSED(B200_ACT_NONE);
            else if (act == B200_ACT_RELU) B200_FWD_FUSED(B200_ACT_RELU);
            else B200_FWD_FUSED(B200_ACT_ADD_RELU);
#undef B200_FWD_FUSED
            if (e == cudaSuccess) return 0;
            if (e == cudaErrorCooperativeLaunchTooLarge) (void)cudaGetLastError();   // nothing was launched
            else if (e != cudaErrorUnknown) return (int)e;
            // grid does not fit at once on this device: the three-kernel chain below
        }
        const PipeGeom ps = make_geom_pipe(m, c, VEC, sizeof(T), 1, false, red_bytes);
        const Geom &g = ps.g;
        if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
        const Ws w = carve(ws, g, c);
        {
            auto k = bn_stats_pipe_kernel<T, VEC>;
            cudaError_t e = allow_smem(k, ps.smem);
            if (e != cudaSuccess) return (int)e;
            e = launch(k, g.nchunk, ps.smem, st, false, x, w.partial, m, c, g.txl, g.nchunk, g.rpc, g_bn_stats_reverse,
                       ps.stages);
            if (e != cudaSuccess) return (int)e;
        }
        {
            cudaError_t e = launch(bn_fwd_finalize_kernel, (c + kWarps - 1) / kWarps, 0, st, true, w.partial, g.nchunk,
                                   g.rpc, m, c, gamma, beta, rm, rv, nbt, saved, eps, momentum);
            if (e != cudaSuccess) return (int)e;
        }
#define B200_APPLY_PIPE(ACT, NT)                                                                                  \
    do {                                                                                                          \
        auto k = bn_apply_pipe_kernel<T, VEC, ACT, false>;                                                        \
        const PipeGeom pa = make_geom_pipe(m, c, VEC, sizeof(T), NT, false, 0);                                   \
        cudaError_t e = allow_smem(k, pa.smem);                                                                   \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(k, pa.g.nchunk, pa.smem, st, true, x, z, y, mask, (const float *)saved, (const float *)nullptr,  \
                   (const float *)nullptr, (const float *)nullptr, (const float *)nullptr, eps, m, c, pa.g.txl,   \
                   pa.g.nchunk, pa.g.rpc, g_bn_apply_reverse, pa.stages, saved2);                                 \
        if (e != cudaSuccess) return (int)e;                                                                      \
    } while (0)
        if (!y) return 0;
        if (act == B200_ACT_NONE) B200_APPLY_PIPE(B200_ACT_NONE, 1);
        else if (act == B200_ACT_RELU) B200_APPLY_PIPE(B200_ACT_RELU, 1);
        else B200_APPLY_PIPE(B200_ACT_ADD_RELU, 2);
#undef B200_APPLY_PIPE
        B200_CHECK_LAUNCH();
        return 0;
    }
    if (saved2) return (int)cudaErrorNotSupported;
    const Geom g = make_geom(m, c, VEC, Unroll<T, VEC>::U);
    if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
    const Ws w = carve(ws, g, c);
    const dim3 grid(g.tiles, g.nchunk);
    bn_stats_kernel<T, VEC><<<grid, kThreads, 0, st>>>(x, w.partial, m, c, g.txl, g.nchunk, g.rpc, g_bn_stats_reverse);
    B200_CHECK_LAUNCH();
    bn_fwd_finalize_kernel<<<(c + kWarps - 1) / kWarps, kThreads, 0, st>>>(w.partial, g.nchunk, g.rpc, m, c, gamma, beta,
                                                                          rm, rv, nbt, saved, eps, momentum);
    B200_CHECK_LAUNCH();
    if (!y) return 0;
    const int rev = g_bn_apply_reverse;
#define B200_APPLY(ACT)                                                                                         \
    bn_apply_kernel<T, VEC, ACT, false><<<grid, kThreads, 0, st>>>(x, z, y, mask, saved, nullptr, nullptr, nullptr, \
                                                                   nullptr, eps, m, c, g.txl, g.nchunk, g.rpc, rev)
    if (act == B200_ACT_NONE) B200_APPLY(B200_ACT_NONE);
    else if (act == B200_ACT_RELU) B200_APPLY(B200_ACT_RELU);
    else B200_APPLY(B200_ACT_ADD_RELU);
#undef B200_APPLY
    B200_CHECK_LAUNCH();
    return 0;
}

template <typename T, int VEC>
int fwd_eval_impl(const void *x_, const void *z_, void *y_, const float *gamma, const float *beta, const float *rm,
                  const float *rv, int64_t m, int c, float eps, int act, cudaStream_t st) {
    const T *x = static_cast<const T *>(x_);
    const T *z = static_cast<const T *>(z_);
    T *y = static_cast<T *>(y_);
    if constexpr (VEC * sizeof(T) == 16) if (pipe_eligible(c, VEC, sizeof(T))) {
#define B200_APPLY_PIPE(ACT, NT)                                                                                  \
    do {                                                                                                          \
        auto k = bn_apply_pipe_kernel<T, VEC, ACT, true>;                                                         \
        const PipeGeom pa = make_geom_pipe(m, c, VEC, sizeof(T), NT, false, 0);                                   \
        cudaError_t e = allow_smem(k, pa.smem);                                                                   \
        if (e != cudaSuccess) return (int)e;                                                                      \
        k<<<pa.g.nchunk, kThreads, pa.smem, st>>>(x, z, y, nullptr, nullptr, gamma, beta, rm, rv, eps, m, c,      \
                                                  pa.g.txl, pa.g.nchunk, pa.g.rpc, 0, pa.stages, nullptr);        \
    } while (0)
        if (act == B200_ACT_NONE) B200_APPLY_PIPE(B200_ACT_NONE, 1);
        else if (act == B200_ACT_RELU) B200_APPLY_PIPE(B200_ACT_RELU, 1);
        else B200_APPLY_PIPE(B200_ACT_ADD_RELU, 2);
#undef B200_APPLY_PIPE
        B200_CHECK_LAUNCH();
        return 0;
    }
    const Geom g = make_geom(m, c, VEC, Unroll<T, VEC>::U);
    const dim3 grid(g.tiles, g.nchunk);
#define B200_APPLY(ACT)                                                                                      \
    bn_apply_kernel<T, VEC, ACT, true><<<grid, kThreads, 0, st>>>(x, z, y, nullptr, nullptr, gamma, beta, rm, rv, \
                                                                  eps, m, c, g.txl, g.nchunk, g.rpc, 0)
    if (act == B200_ACT_NONE) B200_APPLY(B200_ACT_NONE);
    else if (act == B200_ACT_RELU) B200_APPLY(B200_ACT_RELU);
    else B200_APPLY(B200_ACT_ADD_RELU);
#undef B200_APPLY
    B200_CHECK_LAUNCH();
    return 0;
}

template <typename T, int VEC>
int bwd_impl(const void *dy_, const void *dy2_, const void *x_, const uint8_t *mask, const float *gamma, const float *saved, void *dx_,
             void *dz_, float *dgamma, float *dbeta, void *ws, size_t ws_bytes, int64_t m, int c, int act,
             cudaStream_t st) {
    const T *dy = static_cast<const T *>(dy_);
    const T *dy2 = static_cast<const T *>(dy2_);
    const T *x = static_cast<const T *>(x_);
    T *dx = static_cast<T *>(dx_);
    T *dz = static_cast<T *>(dz_);
    if constexpr (VEC * sizeof(T) == 16) if (pipe_eligible(c, VEC, sizeof(T))) {
        if (act == B200_ACT_ADD_RELU && b200_misaligned(mask)) return (int)cudaErrorMisalignedAddress;
        const size_t red_bytes = sizeof(float) * kThreads * 2 * VEC;
        const bool masked = act == B200_ACT_ADD_RELU;
        const int nt = dy2 ? 3 : 2;
        if (g_bn_fused) {
            const FusedGeom fg = make_geom_fused(m, c, VEC, sizeof(T), nt, masked, nt, masked, red_bytes);
            const Geom &g = fg.g;
            if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
            const Ws w = carve(ws, g, c);
            cudaError_t e = cudaErrorUnknown;
#define B200_BWD_FUSED2(ACT, DY2)                                                                                 \
    do {                                                                                                          \
        auto k = bn_bwd_fused_kernel<T, VEC, ACT, DY2>;                                                           \
        if (coop_capacity(k, fg.smem) >= g.nchunk && allow_smem(k, fg.smem) == cudaSuccess)                       \
            e = launch_coop(k, g.nchunk, fg.smem, st, dy, dy2, x, mask, gamma, saved, w.partial, w.coef, dgamma,  \
                            dbeta, dx, dz, m, c, g.txl, g.nchunk, g.rpc, fg.stages1);                             \
    } while (0)
#define B200_BWD_FUSED(ACT)                                                                                       \
    do {                                                                                                          \
        if (dy2) B200_BWD_FUSED2(ACT, true);                                                                      \
        else B200_BWD_FUSED2(ACT, false);                                                                         \
    } while (0)
            if (act == B200_ACT_NONE) B200_BWD_FUSED(B200_ACT_NONE);
            else if (act == B200_ACT_RELU) B200_BWD_FUSED(B200_ACT_RELU);
            else B200_BWD_FUSED(B200_ACT_ADD_RELU);
#undef B200_BWD_FUSED
#undef B200_BWD_FUSED2
            if (e == cudaSuccess) return 0;
            if (e == cudaErrorCooperativeLaunchTooLarge) (void)cudaGetLastError();   // nothing was launched: the chain
            else if (e != cudaErrorUnknown) return (int)e;
        }
        const PipeGeom pr = make_geom_pipe(m, c, VEC, sizeof(T), nt, masked, red_bytes);
        const PipeGeom pa = make_geom_pipe(m, c, VEC, sizeof(T), nt, masked, 0);
        const Geom &g = pr.g;
        if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
        const Ws w = carve(ws, g, c);
#define B200_BWD_PIPE(ACT)                                                                                        \
    do {                                                                                                          \
        if (dy2) B200_BWD_PIPE2(ACT, true);                                                                       \
        else B200_BWD_PIPE2(ACT, false);                                                                          \
    } while (0)
#define B200_BWD_PIPE2(ACT, DY2)                                                                                  \
    do {                                                                                                          \
        auto kr = bn_bwd_reduce_pipe_kernel<T, VEC, ACT, DY2>;                                                    \
        auto ka = bn_bwd_apply_pipe_kernel<T, VEC, ACT, DY2>;                                                     \
        cudaError_t e = allow_smem(kr, pr.smem);                                                                  \
        if (e == cudaSuccess) e = allow_smem(ka, pa.smem);                                                        \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(kr, g.nchunk, pr.smem, st, false, dy, dy2, x, mask, saved, w.partial, m, c, g.txl, g.nchunk,   \
                   g.rpc, g_bn_stats_reverse, pr.stages);                                                         \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(bn_bwd_finalize_kernel, (c + kWarps - 1) / kWarps, 0, st, true, w.partial, g.nchunk, m, c,     \
                   gamma, saved, dgamma, dbeta, w.coef);                                                          \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(ka, pa.g.nchunk, pa.smem, st, true, dy, dy2, x, mask, saved, (const float *)w.coef, dx, dz, m, \
                   c, pa.g.txl, pa.g.nchunk, pa.g.rpc, g_bn_apply_reverse, pa.stages);                            \
        if (e != cudaSuccess) return (int)e;                                                                      \
    } while (0)
        if (act == B200_ACT_NONE) B200_BWD_PIPE(B200_ACT_NONE);
        else if (act == B200_ACT_RELU) B200_BWD_PIPE(B200_ACT_RELU);
        else B200_BWD_PIPE(B200_ACT_ADD_RELU);
#undef B200_BWD_PIPE
#undef B200_BWD_PIPE2
        return 0;
    }
    if (dy2) return (int)cudaErrorNotSupported;  // two-branch gradients only in the bulk-async kernels
    const Geom g = make_geom(m, c, VEC, Unroll<T, VEC>::U);
    if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
    const Ws w = carve(ws, g, c);
    const dim3 grid(g.tiles, g.nchunk);
    const int rrev = g_bn_stats_reverse, arev = g_bn_apply_reverse;
#define B200_BWD(ACT)                                                                                            \
    do {                                                                                                         \
        bn_bwd_reduce_kernel<T, VEC, ACT><<<grid, kThreads, 0, st>>>(dy, x, mask, saved, w.partial, m, c, g.txl,     \
                                                                     g.nchunk, g.rpc, rrev);                     \
        B200_CHECK_LAUNCH();                                                                                     \
        bn_bwd_finalize_kernel<<<(c + kWarps - 1) / kWarps, kThreads, 0, st>>>(w.partial, g.nchunk, m, c, gamma,  \
                                                                              saved, dgamma, dbeta, w.coef);    \
        B200_CHECK_LAUNCH();                                                                                     \
        bn_bwd_apply_kernel<T, VEC, ACT><<<grid, kThreads, 0, st>>>(dy, x, mask, saved, w.coef, dx, dz, m, c, g.txl, \
                                                                    g.nchunk, g.rpc, arev);                      \
        B200_CHECK_LAUNCH();                                                                                     \
    } while (0)
    if (act == B200_ACT_NONE) B200_BWD(B200_ACT_NONE);
    else if (act == B200_ACT_RELU) B200_BWD(B200_ACT_RELU);
    else B200_BWD(B200_ACT_ADD_RELU);
#undef B200_BWD
    return 0;
}

template <typename T, int VEC>
int bwd_dual_impl(const void *dy_, const void *dy2_, const void *x_, const void *x2_, const uint8_t *mask,
                  const float *gamma, const float *gamma2, const float *saved, const float *saved2, void *dx_, void *dx2_,
                  float *dgamma, float *dbeta, float *dgamma2, float *dbeta2, void *ws, size_t ws_bytes, int64_t m, int c,
                  cudaStream_t st) {
    if constexpr (VEC * sizeof(T) == 16) if (pipe_eligible(c, VEC, sizeof(T))) {
        const T *dy = static_cast<const T *>(dy_), *dy2 = static_cast<const T *>(dy2_);
        const T *x = static_cast<const T *>(x_), *x2 = static_cast<const T *>(x2_);
        T *dx = static_cast<T *>(dx_), *dx2 = static_cast<T *>(dx2_);
        if (b200_misaligned(mask)) return (int)cudaErrorMisalignedAddress;
        const int nt = dy2 ? 4 : 3;
        const PipeGeom pr = make_geom_pipe(m, c, VEC, sizeof(T), nt, true, sizeof(float) * kThreads * 3 * VEC);
        const PipeGeom pa = make_geom_pipe(m, c, VEC, sizeof(T), nt, true, 0);
        const Geom &g = pr.g;
        if (ws_bytes < ws_bytes_for(g, c)) return (int)cudaErrorInvalidValue;
        const Ws w = carve(ws, g, c);
        float4 *partial = reinterpret_cast<float4 *>(w.partial);
        cudaError_t e;
#define B200_DUAL(DY2)                                                                                            \
    do {                                                                                                          \
        auto kr = bn_bwd_reduce_dual_pipe_kernel<T, VEC, DY2>;                                                    \
        auto ka = bn_bwd_apply_dual_pipe_kernel<T, VEC, DY2>;                                                     \
        e = allow_smem(kr, pr.smem);                                                                              \
        if (e == cudaSuccess) e = allow_smem(ka, pa.smem);                                                        \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(kr, g.nchunk, pr.smem, st, false, dy, dy2, x, x2, mask, saved, saved2, partial, m, c, g.txl,   \
                   g.nchunk, g.rpc, g_bn_stats_reverse, pr.stages);                                               \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(bn_bwd_finalize_dual_kernel, (c + kWarps - 1) / kWarps, 0, st, true, (const float4 *)partial,  \
                   g.nchunk, m, c, gamma, gamma2, saved, saved2, dgamma, dbeta, dgamma2, dbeta2, w.coef);         \
        if (e != cudaSuccess) return (int)e;                                                                      \
        e = launch(ka, pa.g.nchunk, pa.smem, st, true, dy, dy2, x, x2, mask, saved, saved2, (const float *)w.coef, \
                   dx, dx2, m, c, pa.g.txl, pa.g.nchunk, pa.g.rpc, g_bn_apply_reverse, pa.stages);                \
        if (e != cudaSuccess) return (int)e;                                                                      \
    } while (0)
        if (dy2) B200_DUAL(true);
        else B200_DUAL(false);
#undef B200_DUAL
        return 0;
    }
    return (int)cudaErrorNotSupported;
}

int check_common(int64_t m, int c, int act, int dtype, int vec) {
    if (m < 1 || c < 1) return (int)cudaErrorInvalidValue;
    if (act < B200_ACT_NONE || act > B200_ACT_ADD_RELU) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (c % vec != 0) return (int)cudaErrorInvalidValue;
    return 0;
}

}  // namespace

extern "C" int b200_bn_vec(int dtype) { return dtype == B200_BF16 ? g_bn_vec_bf16 : 4; }

extern int g_input_s2d_fast;  // input.cu
extern int g_pack_steps;      // pack.cu

extern "C" int b200_tune_set(const char *key, int value) {
    if (!key) return (int)cudaErrorInvalidValue;
    auto eq = [&](const char *s) { const char *a = key; while (*a && *a == *s) { ++a; ++s; } return *a == *s; };
    if (eq("bn_target_ctas")) { if (value < 1) return (int)cudaErrorInvalidValue; g_bn_target_ctas = value; return 0; }
    if (eq("bn_vec_bf16")) { if (value != 4 && value != 8) return (int)cudaErrorInvalidValue; g_bn_vec_bf16 = value; return 0; }
    if (eq("bn_stats_reverse")) { g_bn_stats_reverse = value != 0; return 0; }
    if (eq("bn_apply_reverse")) { g_bn_apply_reverse = value != 0; return 0; }
    if (eq("bn_pipeline")) { g_bn_pipeline = value != 0; return 0; }
    if (eq("bn_pdl")) { g_bn_pdl = value != 0; return 0; }
    if (eq("bn_fused")) { g_bn_fused = value != 0; return 0; }
    if (eq("input_s2d_fast")) { g_input_s2d_fast = value != 0; return 0; }
    if (eq("pack_steps")) { if (value != 1 && value != 2 && value != 4 && value != 8) return (int)cudaErrorInvalidValue; g_pack_steps = value; return 0; }
    if (eq("bn_pipe_stages")) { if (value < 0 || value > kMaxStages || value == 1) return (int)cudaErrorInvalidValue; g_bn_pipe_stages = value; return 0; }
    if (eq("bn_pipe_ctas_per_sm")) { if (value < 1 || value > 8) return (int)cudaErrorInvalidValue; g_bn_pipe_ctas_per_sm = value; return 0; }
    return (int)cudaErrorInvalidValue;
}

extern "C" size_t b200_bn_workspace_bytes(int64_t m, int c, int dtype) {
    if (m < 1 || c < 1) return 0;
    // upper bound over both vector widths and any tuning of the chunk count
    size_t best = 0;
    const int vecs[2] = {4, 8};
    for (int i = 0; i < 2; ++i) {
        const int vec = vecs[i];
        if (dtype != B200_BF16 && vec == 8) continue;
        if (c % vec) continue;
        const int unroll = dtype == B200_BF16 ? (vec == 8 ? 2 : 4) : 2;
        const Geom g = make_geom(m, c, vec, unroll);
        const size_t b = ws_bytes_for(g, c);
        if (b > best) best = b;
        const int cv = c / vec;
        if (cv >= 2 && cv <= kThreads && (cv & (cv - 1)) == 0) {
            const int saved_ctas = g_bn_pipe_ctas_per_sm;
            g_bn_pipe_ctas_per_sm = 8;  // bound over any tuning
            const PipeGeom gp = make_geom_pipe(m, c, vec, dtype == B200_BF16 ? 2 : 4, 1, false, 0);
            g_bn_pipe_ctas_per_sm = saved_ctas;
            const size_t bp = ws_bytes_for(gp.g, c);
            if (bp > best) best = bp;
        }
    }
    return best;
}

extern "C" int b200_bn_fwd_train(const void *x, const void *z, void *y, uint8_t *mask, const float *gamma,
                                 const float *beta, float *running_mean, float *running_var,
                                 int64_t *num_batches_tracked, float *saved, void *workspace, size_t workspace_bytes,
                                 int64_t m, int c, float eps, float momentum, int act, int dtype, void *stream) {
    const int vec = b200_bn_vec(dtype);
    int e = check_common(m, c, act, dtype, vec);
    if (e) return e;
    if (m < 2) return (int)cudaErrorInvalidValue;  // "Expected more than 1 value per channel when training"
    if (!x || !y || !gamma || !beta || !saved || !workspace) return (int)cudaErrorInvalidValue;
    if (act == B200_ACT_ADD_RELU && !z) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(x) || b200_misaligned(y) || b200_misaligned(z) || b200_misaligned(workspace))
        return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_F32)
        return fwd_train_impl<float, 4>(x, z, y, mask, gamma, beta, running_mean, running_var, num_batches_tracked, saved,
                                        workspace, workspace_bytes, m, c, eps, momentum, act, st);
    if (vec == 8)
        return fwd_train_impl<__nv_bfloat16, 8>(x, z, y, mask, gamma, beta, running_mean, running_var,
                                                num_batches_tracked, saved, workspace, workspace_bytes, m, c, eps,
                                                momentum, act, st);
    return fwd_train_impl<__nv_bfloat16, 4>(x, z, y, mask, gamma, beta, running_mean, running_var, num_batches_tracked,
                                            saved, workspace, workspace_bytes, m, c, eps, momentum, act, st);
}

extern "C" int b200_bn_fwd_eval(const void *x, const void *z, void *y, const float *gamma, const float *beta,
                                const float *running_mean, const float *running_var, int64_t m, int c, float eps,
                                int act, int dtype, void *stream) {
    const int vec = b200_bn_vec(dtype);
    int e = check_common(m, c, act, dtype, vec);
    if (e) return e;
    if (!x || !y || !gamma || !beta || !running_mean || !running_var) return (int)cudaErrorInvalidValue;
    if (act == B200_ACT_ADD_RELU && !z) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(x) || b200_misaligned(y) || b200_misaligned(z)) return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_F32) return fwd_eval_impl<float, 4>(x, z, y, gamma, beta, running_mean, running_var, m, c, eps, act, st);
    if (vec == 8)
        return fwd_eval_impl<__nv_bfloat16, 8>(x, z, y, gamma, beta, running_mean, running_var, m, c, eps, act, st);
    return fwd_eval_impl<__nv_bfloat16, 4>(x, z, y, gamma, beta, running_mean, running_var, m, c, eps, act, st);
}

extern "C" int b200_bn_bwd(const void *dy, const void *dy2, const void *x, const uint8_t *mask, const float *gamma, const float *saved,
                           void *dx, void *dz, float *dgamma, float *dbeta, void *workspace, size_t workspace_bytes,
                           int64_t m, int c, int act, int dtype, void *stream) {
    const int vec = b200_bn_vec(dtype);
    int e = check_common(m, c, act, dtype, vec);
    if (e) return e;
    if (!dy || !x || !gamma || !saved || !dx || !workspace) return (int)cudaErrorInvalidValue;
    if (act == B200_ACT_ADD_RELU && (!mask || !dz)) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(dy) || b200_misaligned(dy2) || b200_misaligned(x) || b200_misaligned(dx) || b200_misaligned(dz) ||
        b200_misaligned(workspace))
        return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_F32)
        return bwd_impl<float, 4>(dy, dy2, x, mask, gamma, saved, dx, dz, dgamma, dbeta, workspace, workspace_bytes, m, c, act, st);
    if (vec == 8)
        return bwd_impl<__nv_bfloat16, 8>(dy, dy2, x, mask, gamma, saved, dx, dz, dgamma, dbeta, workspace, workspace_bytes, m,
                                          c, act, st);
    return bwd_impl<__nv_bfloat16, 4>(dy, dy2, x, mask, gamma, saved, dx, dz, dgamma, dbeta, workspace, workspace_bytes, m, c,
                                      act, st);
}

extern "C" int b200_bn_dual_supported(int c, int dtype) {
    if (dtype != B200_F32 && dtype != B200_BF16) return 0;
    const int vec = b200_bn_vec(dtype);
    if (c < 1 || c % vec) return 0;
    return pipe_eligible(c, vec, dtype == B200_F32 ? 4 : 2) ? 1 : 0;
}

extern "C" int b200_bn_stats(const void *x, const float *gamma, const float *beta, float *running_mean,
                             float *running_var, int64_t *num_batches_tracked, float *saved, void *workspace,
                             size_t workspace_bytes, int64_t m, int c, float eps, float momentum, int dtype,
                             void *stream) {
    const int vec = b200_bn_vec(dtype);
    int e = check_common(m, c, B200_ACT_NONE, dtype, vec);
    if (e) return e;
    if (m < 2 || !x || !gamma || !beta || !saved || !workspace) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(x) || b200_misaligned(workspace)) return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_F32)
        return fwd_train_impl<float, 4>(x, nullptr, nullptr, nullptr, gamma, beta, running_mean, running_var,
                                        num_batches_tracked, saved, workspace, workspace_bytes, m, c, eps, momentum,
                                        B200_ACT_NONE, st);
    if (vec == 8)
        return fwd_train_impl<__nv_bfloat16, 8>(x, nullptr, nullptr, nullptr, gamma, beta, running_mean, running_var,
                                                num_batches_tracked, saved, workspace, workspace_bytes, m, c, eps,
                                                momentum, B200_ACT_NONE, st);
    return fwd_train_impl<__nv_bfloat16, 4>(x, nullptr, nullptr, nullptr, gamma, beta, running_mean, running_var,
                                            num_batches_tracked, saved, workspace, workspace_bytes, m, c, eps, momentum,
                                            B200_ACT_NONE, st);
}

extern "C" int b200_bn_fwd_train_dual(const void *x, const void *x2, void *y, uint8_t *mask, const float *gamma,
                                      const float *beta, float *running_mean, float *running_var,
                                      int64_t *num_batches_tracked, float *saved, const float *saved2, void *workspace,
                                      size_t workspace_bytes, int64_t m, int c, float eps, float momentum, int dtype,
                                      void *stream) {
    const int vec = b200_bn_vec(dtype);
    int e = check_common(m, c, B200_ACT_ADD_RELU, dtype, vec);
    if (e) return e;
    if (m < 2 || !x || !x2 || !y || !gamma || !beta || !saved || !saved2 || !workspace) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(x) || b200_misaligned(x2) || b200_misaligned(y) || b200_misaligned(workspace))
        return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_F32)
        return fwd_train_impl<float, 4>(x, x2, y, mask, gamma, beta, running_mean, running_var, num_batches_tracked, saved,
                                        workspace, workspace_bytes, m, c, eps, momentum, B200_ACT_ADD_RELU, st, saved2);
    if (vec == 8)
        return fwd_train_impl<__nv_bfloat16, 8>(x, x2, y, mask, gamma, beta, running_mean, running_var,
                                                num_batches_tracked, saved, workspace, workspace_bytes, m, c, eps,
                                                momentum, B200_ACT_ADD_RELU, st, saved2);
    return (int)cudaErrorNotSupported;
}

extern "C" int b200_bn_bwd_dual(const void *dy, const void *dy2, const void *x, const void *x2, const uint8_t *mask,
                                const float *gamma, const float *gamma2, const float *saved, const float *saved2,
                                void *dx, void *dx2, float *dgamma, float *dbeta, float *dgamma2, float *dbeta2,
                                void *workspace, size_t workspace_bytes, int64_t m, int c, int dtype, void *stream) {
    const int vec = b200_bn_vec(dtype);
    int e = check_common(m, c, B200_ACT_ADD_RELU, dtype, vec);
    if (e) return e;
    if (!dy || !x || !x2 || !mask || !gamma || !gamma2 || !saved || !saved2 || !dx || !dx2 || !workspace)
        return (int)cudaErrorInvalidValue;
    if (b200_misaligned(dy) || b200_misaligned(dy2) || b200_misaligned(x) || b200_misaligned(x2) || b200_misaligned(dx) ||
        b200_misaligned(dx2) || b200_misaligned(workspace))
        return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == B200_F32)
        return bwd_dual_impl<float, 4>(dy, dy2, x, x2, mask, gamma, gamma2, saved, saved2, dx, dx2, dgamma, dbeta, dgamma2,
                                       dbeta2, workspace, workspace_bytes, m, c, st);
    if (vec == 8)
        return bwd_dual_impl<__nv_bfloat16, 8>(dy, dy2, x, x2, mask, gamma, gamma2, saved, saved2, dx, dx2, dgamma, dbeta,
                                               dgamma2, dbeta2, workspace, workspace_bytes, m, c, st);
    return (int)cudaErrorNotSupported;
}
