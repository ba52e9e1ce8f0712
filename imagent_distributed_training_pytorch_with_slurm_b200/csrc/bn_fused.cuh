// One cooperative kernel per BatchNorm site instead of the statistics -> finalize -> apply chain of three.
// Included by batchnorm.cu after bn_pipe.cuh (same namespace, same tile ring, same per-tile arithmetic).
//
// Why: on the small late-stage activations (12-50 MB) a site costs 24-30 us as three launches although its bytes
// stream in ~10 us - ramp-up, drain and the dependent-launch hand-offs dominate, and the 98 finalize launches of a
// ResNet-50 step add ~0.4 ms of critical path.  Here the grid is one co-resident wave (cooperative launch), every CTA
// keeps its chunk of rows for both passes, and the two places where the passes need the whole tensor's sums are grid-wide
// barriers inside the kernel:
//     pass 1 over the CTA's rows -> per-(channel, chunk) partials -> [prefetch the first tiles of pass 2]
//     grid barrier -> channels dealt out over the CTAs, one warp folds one channel's partials, writes saved / coef
//     grid barrier -> pass 2 over the same rows (they are the most recently touched lines of L2)
// The arithmetic, the partial layout and the order of every floating-point sum are those of the three-kernel chain,
// so the results are bit-identical to it (tests compare the two).
//
// MEASURED RESULT (B200, tools/probe_bn_fused.py, profiles/r02_probe_bn_fused.jsonl): this form is SLOWER than the
// chain on every ResNet-50 shape - forward bn+relu, activation just written by a producer kernel, chain / fused:
// 12.8 MB 25.6 / 25.6 us, 25.7 MB 27.7 / 40.0 us, 51 MB 37.9 / 50.2 us, 103 MB 62.5 / 69.7 us, 411 MB 218 / 220 us;
// backward 25.7 MB 29.7 / 49.2 us, 103 MB 87 / 106 us.  A cooperative launch costs more than the two ordinary
// dependent launches it replaces, and each grid-wide barrier stalls every SM for several microseconds, whereas the
// chain's programmatic dependent launches let an SM that has finished its part of one pass start the next kernel's
// prologue and prefetch at once.  It therefore stays OFF by default (b200_tune_set("bn_fused", 1) selects it) and is
// kept as a cross-check of the chain and as the record of the experiment.
#pragma once
// (<cooperative_groups.h> is included by batchnorm.cu at file scope; this header sits inside its namespace)
namespace cg = cooperative_groups;

__device__ __forceinline__ float2 ldcg2(const float2 *p) { return __ldcg(p); }

// forward fold of one channel by one warp (body of bn_fwd_finalize_kernel)
__device__ __forceinline__ void fold_fwd_channel(const float2 *__restrict__ partial, int nchunk, int64_t rpc, int64_t m,
                                                 int c, int ch, const float *__restrict__ gamma,
                                                 const float *__restrict__ beta, float *__restrict__ rmean,
                                                 float *__restrict__ rvar, float *__restrict__ saved, float eps,
                                                 float momentum, int lane) {
    const float2 *p = partial + (int64_t)ch * nchunk;
    const float last_n = (float)(m - (int64_t)(nchunk - 1) * rpc);
    float s = 0.f;
    for (int i = lane; i < nchunk; i += 32) s = fmaf(i == nchunk - 1 ? last_n : (float)rpc, ldcg2(p + i).x, s);
    s = warp_sum(s);
    const float mean = s / (float)m;
    float q = 0.f;
    for (int i = lane; i < nchunk; i += 32) {
        const float2 v = ldcg2(p + i);
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

// ---- forward: statistics + fold + apply ------------------------------------------------------------------------
// y == nullptr: statistics and fold only (the downsample branch of the "dual" form).
// saved2 != nullptr (ADD_RELU): z is the raw downsample convolution output, normalised on the fly ("dual" form).
template <typename T, int VEC, int ACT>
__global__ void __launch_bounds__(kThreads, 2)
bn_fwd_fused_kernel(const T *__restrict__ x, const T *__restrict__ z, T *__restrict__ y, uint8_t *__restrict__ mask,
                    float2 *__restrict__ partial, const float *__restrict__ gamma, const float *__restrict__ beta,
                    float *__restrict__ rmean, float *__restrict__ rvar, int64_t *__restrict__ nbt,
                    float *__restrict__ saved, const float *__restrict__ saved2, float eps, float momentum, int64_t m,
                    int c, int txl, int nchunk, int64_t rpc, int nstages1, int nstages2) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ float red[kWarps * 32 * 2 * VEC];
    cg::grid_group grid = cg::this_grid();
    constexpr int NT = (ACT == B200_ACT_ADD_RELU) ? 2 : 1;
    const int tx = 1 << txl;
    const int cvi = threadIdx.x & (tx - 1);
    const int cv = c / VEC;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunk = (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    const int64_t coff = (int64_t)cvi * VEC;
    {   // ---- pass 1 (bn_stats_pipe_kernel)
        TileRing<1, false> ring;
        ring.src[0] = reinterpret_cast<const unsigned char *>(x);
        ring.msrc = nullptr;
        ring.setup(dyn_smem, txl, nstages1, (int64_t)c * sizeof(T), r0, r1);
        float sh[VEC];
        VecIO<T, VEC>::load(x + r0 * c + coff, sh);
        float acc[2 * VEC];
#pragma unroll
        for (int k = 0; k < 2 * VEC; ++k) acc[k] = 0.f;
        B200_RING_LOOP(ring, t, s, ph) {
            const int nrows = ring.wait(t, s, (uint32_t)ph);
            const uint4 *tile = reinterpret_cast<const uint4 *>(ring.stage(s, 0));
#pragma unroll
            for (int u = 0; u < kPipeU; ++u) {
                const int v = threadIdx.x + u * kThreads;
                if ((v >> txl) < nrows) {
                    float e[VEC];
                    unpack16<T, VEC>(tile[v], e);
#pragma unroll
                    for (int k = 0; k < VEC; ++k) {
                        const float d = e[k] - sh[k];
                        acc[k] += d;
                        acc[VEC + k] = fmaf(d, d, acc[VEC + k]);
                    }
                }
            }
            ring.release(t, s);
        }
        cta_reduce_rows<2 * VEC>(acc, tx, red);
        if (threadIdx.x < tx) {
            const float nb = (float)(r1 - r0);
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                const float s1 = acc[k], s2 = acc[VEC + k];
                partial[(int64_t)(cvi * VEC + k) * nchunk + chunk] =
                    make_float2(sh[k] + s1 / nb, fmaxf(s2 - s1 * s1 / nb, 0.f));
            }
        }
    }
    // pass 2's first tiles go in flight now (x and z do not depend on the statistics); second barrier set of the header
    TileRing<NT, false> ring;
    if (y) {
        ring.src[0] = reinterpret_cast<const unsigned char *>(x);
        if (NT == 2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(z);
        ring.msrc = nullptr;
        ring.setup(dyn_smem, txl, nstages2, (int64_t)c * sizeof(T), r0, r1, kRingHeader / 2);
    }
    __threadfence();
    grid.sync();
    {   // ---- fold: channels dealt out over the CTAs, one warp per channel
        const int per = (c + (int)gridDim.x - 1) / (int)gridDim.x;
        for (int i = warp; i < per; i += kWarps) {
            const int ch = (int)blockIdx.x * per + i;
            if (ch < c) fold_fwd_channel(partial, nchunk, rpc, m, c, ch, gamma, beta, rmean, rvar, saved, eps, momentum, lane);
        }
        if (blockIdx.x == 0 && threadIdx.x == 0 && nbt) *nbt += 1;
    }
    __threadfence();
    grid.sync();
    if (!y) return;
    // ---- pass 2 (bn_apply_pipe_kernel)
    float sc[VEC], sf[VEC], sc2[VEC], sf2[VEC];
    const bool dual = (ACT == B200_ACT_ADD_RELU) && saved2 != nullptr;
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int ch = cvi * VEC + k;
        sc[k] = __ldcg(saved + 2 * c + ch);
        sf[k] = __ldcg(saved + 3 * c + ch);
        sc2[k] = dual ? __ldcg(saved2 + 2 * c + ch) : 1.f;
        sf2[k] = dual ? __ldcg(saved2 + 3 * c + ch) : 0.f;
    }
    B200_RING_LOOP(ring, t, s, ph) {
        const int nrows = ring.wait(t, s, (uint32_t)ph);
        const uint4 *tx_ = reinterpret_cast<const uint4 *>(ring.stage(s, 0));
        const uint4 *tz_ = reinterpret_cast<const uint4 *>(ring.stage(s, NT - 1));
        const int64_t row0 = r0 + (int64_t)t * ring.rows_per_tile;
#pragma unroll
        for (int u = 0; u < kPipeU; ++u) {
            const int v = threadIdx.x + u * kThreads;
            const int r = v >> txl;
            if (r < nrows) {
                float e[VEC], w[VEC];
                unpack16<T, VEC>(tx_[v], e);
                if (ACT == B200_ACT_ADD_RELU) unpack16<T, VEC>(tz_[v], w);
                unsigned bits = 0;
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    float pre = fmaf(e[k], sc[k], sf[k]);
                    if (ACT == B200_ACT_ADD_RELU) pre += dual ? fmaf(w[k], sc2[k], sf2[k]) : w[k];
                    if (ACT != B200_ACT_NONE) {
                        bits |= (pre > 0.f ? 1u : 0u) << k;
                        pre = fmaxf(pre, 0.f);
                    }
                    e[k] = pre;
                }
                const int64_t rr = row0 + r;
                VecIO<T, VEC>::store(y + rr * c + coff, e);
                if (ACT == B200_ACT_ADD_RELU && mask) mask[rr * cv + cvi] = (uint8_t)bits;
            }
        }
        ring.release(t, s);
    }
}

// ---- backward: reduce + fold + apply ---------------------------------------------------------------------------
template <typename T, int VEC, int ACT, bool DY2>
__global__ void __launch_bounds__(kThreads, 2)
bn_bwd_fused_kernel(const T *__restrict__ dy, const T *__restrict__ dy2, const T *__restrict__ x,
                    const uint8_t *__restrict__ mask, const float *__restrict__ gamma, const float *__restrict__ saved,
                    float2 *__restrict__ partial, float *__restrict__ coef, float *__restrict__ dgamma,
                    float *__restrict__ dbeta, T *__restrict__ dx, T *__restrict__ dz, int64_t m, int c, int txl,
                    int nchunk, int64_t rpc, int nstages) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ float red[kWarps * 32 * 2 * VEC];
    cg::grid_group grid = cg::this_grid();
    constexpr bool MASK = (ACT == B200_ACT_ADD_RELU);
    constexpr int NT = DY2 ? 3 : 2;
    const int tx = 1 << txl;
    const int cvi = threadIdx.x & (tx - 1);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunk = (int)blockIdx.x;
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
    TileRing<NT, MASK> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(dy);
    ring.src[1] = reinterpret_cast<const unsigned char *>(x);
    if (DY2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(dy2);
    ring.msrc = mask;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);
    {   // ---- pass 1 (bn_bwd_reduce_pipe_kernel)
        float acc[2 * VEC];
#pragma unroll
        for (int k = 0; k < 2 * VEC; ++k) acc[k] = 0.f;
        B200_RING_LOOP(ring, t, s, ph) {
            const int nrows = ring.wait(t, s, (uint32_t)ph);
            const uint4 *tg = reinterpret_cast<const uint4 *>(ring.stage(s, 0));
            const uint4 *tv = reinterpret_cast<const uint4 *>(ring.stage(s, 1));
            const uint4 *tg2 = reinterpret_cast<const uint4 *>(ring.stage(s, NT - 1));
            const unsigned char *tm = ring.mstage(s);
#pragma unroll
            for (int u = 0; u < kPipeU; ++u) {
                const int v = threadIdx.x + u * kThreads;
                if ((v >> txl) < nrows) {
                    float g[VEC], e[VEC];
                    unpack16<T, VEC>(tg[v], g);
                    unpack16<T, VEC>(tv[v], e);
                    if (DY2) {
                        float g2[VEC];
                        unpack16<T, VEC>(tg2[v], g2);
#pragma unroll
                        for (int k = 0; k < VEC; ++k) g[k] += g2[k];
                    }
                    const unsigned bits = MASK ? tm[v] : 0u;
#pragma unroll
                    for (int k = 0; k < VEC; ++k) {
                        float gk = g[k];
                        if (ACT == B200_ACT_RELU) gk = fmaf(e[k], sc[k], sf[k]) > 0.f ? gk : 0.f;
                        if (ACT == B200_ACT_ADD_RELU) gk = ((bits >> k) & 1u) ? gk : 0.f;
                        acc[k] += gk;
                        acc[VEC + k] = fmaf(gk, e[k] - mean[k], acc[VEC + k]);
                    }
                }
            }
            ring.release(t, s);
        }
        cta_reduce_rows<2 * VEC>(acc, tx, red);
        if (threadIdx.x < tx) {
#pragma unroll
            for (int k = 0; k < VEC; ++k)
                partial[(int64_t)(cvi * VEC + k) * nchunk + chunk] = make_float2(acc[k], acc[VEC + k]);
        }
    }
    // pass 2 reads the same tensors again: restart the ring (second barrier set) so its first tiles are in flight
    // while the grid meets
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1, kRingHeader / 2);
    __threadfence();
    grid.sync();
    {   // ---- fold (bn_bwd_finalize_kernel), channels dealt out over the CTAs
        const int per = (c + (int)gridDim.x - 1) / (int)gridDim.x;
        for (int i = warp; i < per; i += kWarps) {
            const int ch = (int)blockIdx.x * per + i;
            if (ch >= c) continue;
            const float2 *p = partial + (int64_t)ch * nchunk;
            float a = 0.f, b = 0.f;
            for (int j = lane; j < nchunk; j += 32) {
                const float2 v = ldcg2(p + j);
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
    }
    __threadfence();
    grid.sync();
    // ---- pass 2 (bn_bwd_apply_pipe_kernel)
    float c0[VEC], c1[VEC], c2[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int ch = cvi * VEC + k;
        c0[k] = __ldcg(coef + ch);
        c1[k] = __ldcg(coef + c + ch);
        c2[k] = __ldcg(coef + 2 * c + ch);
    }
    B200_RING_LOOP(ring, t, s, ph) {
        const int nrows = ring.wait(t, s, (uint32_t)ph);
        const uint4 *tg = reinterpret_cast<const uint4 *>(ring.stage(s, 0));
        const uint4 *tv = reinterpret_cast<const uint4 *>(ring.stage(s, 1));
        const uint4 *tg2 = reinterpret_cast<const uint4 *>(ring.stage(s, NT - 1));
        const unsigned char *tm = ring.mstage(s);
        const int64_t row0 = r0 + (int64_t)t * ring.rows_per_tile;
#pragma unroll
        for (int u = 0; u < kPipeU; ++u) {
            const int v = threadIdx.x + u * kThreads;
            const int r = v >> txl;
            if (r < nrows) {
                float g[VEC], e[VEC];
                unpack16<T, VEC>(tg[v], g);
                unpack16<T, VEC>(tv[v], e);
                if (DY2) {
                    float g2[VEC];
                    unpack16<T, VEC>(tg2[v], g2);
#pragma unroll
                    for (int k = 0; k < VEC; ++k) g[k] += g2[k];
                }
                const unsigned bits = MASK ? tm[v] : 0u;
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    float gk = g[k];
                    if (ACT == B200_ACT_RELU) gk = fmaf(e[k], sc[k], sf[k]) > 0.f ? gk : 0.f;
                    if (ACT == B200_ACT_ADD_RELU) gk = ((bits >> k) & 1u) ? gk : 0.f;
                    g[k] = gk;
                    e[k] = fmaf(gk, c0[k], fmaf(e[k] - mean[k], c1[k], c2[k]));
                }
                const int64_t rr = row0 + r;
                VecIO<T, VEC>::store(dx + rr * c + coff, e);
                if (ACT == B200_ACT_ADD_RELU) VecIO<T, VEC>::store(dz + rr * c + coff, g);
            }
        }
        ring.release(t, s);
    }
}
