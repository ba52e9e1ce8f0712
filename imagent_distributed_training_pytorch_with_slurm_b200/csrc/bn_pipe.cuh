// Bulk-async (TMA, cp.async.bulk + mbarrier) streaming variants of the BatchNorm passes.
// Included by batchnorm.cu inside its anonymous namespace.
//
// Why: with plain per-thread loads a B200 SM cannot keep enough bytes in flight for a read-heavy
// pass (measured: ~2.8 TB/s of reads at ~40 KB/SM outstanding, i.e. ~2 us loaded latency).  Here one
// warp of each CTA issues cp.async.bulk copies of whole tiles into a 4-stage shared-memory ring
// (8 KB per tensor per stage); completion is signalled on an mbarrier per stage (complete_tx), the
// 256 threads consume a stage with conflict-free 16-byte shared loads, and the stage is refilled
// right after the CTA-wide barrier that ends its consumption.  Outstanding bytes are bounded by
// shared memory (3-4 CTAs x 32-68 KB per SM) instead of by registers.
//
// Tile geometry (same thread->channel mapping as the direct kernels): a CTA owns a contiguous range
// of rows with ALL channels (TX = C/VEC <= 256 vectors per row), so a tile of kTileVecs 16-byte
// vectors = R whole rows is one contiguous span of memory, fetched with ONE bulk copy per tensor.
// Vector v of a tile (v = r*TX + tx) lives at byte 16*v of the stage, so thread `tid` reads vectors
// tid and tid+256 and always sees the same channels.  The number of stages is chosen by the host so
// that the resident CTAs of an SM together keep ~200 KB in flight.
#pragma once

constexpr int kMaxStages = 16;    // 16 x 8 KB per tensor: what ONE resident CTA per SM needs to keep ~128 KB in flight
constexpr int kRingHeader = 256;  // two sets of kMaxStages mbarriers (8 bytes each)
constexpr int kTileVecs = 512;
constexpr int kPipeU = kTileVecs / kThreads;  // vectors per thread per tile

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "LAB_WAIT%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra LAB_DONE%=;\n"
        "bra LAB_WAIT%=;\n"
        "LAB_DONE%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization attribute
// may start (be scheduled, run its prologue, prefetch inputs that older kernels produced) while its
// predecessor drains; pdl_wait() blocks until the predecessor has completed and its writes are visible.
// Both are no-ops for ordinary launches.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

template <typename T, int VEC>
__device__ __forceinline__ void unpack16(const uint4 &t, float (&v)[VEC]);
template <>
__device__ __forceinline__ void unpack16<float, 4>(const uint4 &t, float (&v)[4]) {
    v[0] = __uint_as_float(t.x); v[1] = __uint_as_float(t.y); v[2] = __uint_as_float(t.z); v[3] = __uint_as_float(t.w);
}
template <>
__device__ __forceinline__ void unpack16<__nv_bfloat16, 8>(const uint4 &t, float (&v)[8]) {
    VecIO<__nv_bfloat16, 8>::unpack(t, v);
}

constexpr size_t pipe_stage_bytes(int nt, bool mask) { return (size_t)nt * kTileVecs * 16 + (mask ? kTileVecs : 0); }
constexpr size_t pipe_smem_bytes(int nt, bool mask, int stages) { return kRingHeader + (size_t)stages * pipe_stage_bytes(nt, mask); }

// The shared-memory ring.  All 256 threads call wait()/release(); thread 0 is the producer.
template <int NT, bool MASK>
struct TileRing {
    uint64_t *bars;
    unsigned char *data, *msk;
    const unsigned char *src[NT];
    const unsigned char *msrc;
    int64_t row_bytes, r0, r1;
    int cv, rows_per_tile, ntile, stages;

    __device__ __forceinline__ unsigned char *stage(int s, int tensor) const {
        return data + ((size_t)(s * NT + tensor) * kTileVecs) * 16;
    }
    __device__ __forceinline__ unsigned char *mstage(int s) const { return msk + (size_t)s * kTileVecs; }

    // bar_off: which of the two barrier sets of the header to use (0 or kRingHeader / 2) - a kernel that walks its rows
    // twice re-arms the ring on the second set instead of re-initialising barriers that may still be in use
    __device__ void setup(unsigned char *smem, int txl, int nstages, int64_t c_bytes, int64_t r0_, int64_t r1_,
                          int bar_off = 0) {
        bars = reinterpret_cast<uint64_t *>(smem + bar_off);
        data = smem + kRingHeader;
        stages = nstages;
        msk = data + (size_t)nstages * NT * kTileVecs * 16;
        cv = 1 << txl;
        row_bytes = c_bytes;
        r0 = r0_;
        r1 = r1_;
        rows_per_tile = kTileVecs >> txl;
        ntile = (int)((r1 - r0 + rows_per_tile - 1) / rows_per_tile);
        if (threadIdx.x == 0) {
            for (int s = 0; s < nstages; ++s) mbar_init(bars + s, 1);
            fence_mbar_init();
            fence_proxy_async();
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int t = 0; t < nstages && t < ntile; ++t) issue(t, t);
    }

    // producer thread only: fetch tile t into stage s
    __device__ __forceinline__ void issue(int t, int s) {
        const int64_t row0 = r0 + (int64_t)t * rows_per_tile;
        const int nrows = (int)((r1 - row0 < rows_per_tile) ? (r1 - row0) : rows_per_tile);
        const uint32_t bytes = (uint32_t)nrows * (uint32_t)row_bytes;
        const uint32_t mbytes = MASK ? (((uint32_t)nrows * (uint32_t)cv + 15u) & ~15u) : 0u;
        mbar_arrive_expect_tx(bars + s, NT * bytes + mbytes);
#pragma unroll
        for (int i = 0; i < NT; ++i) bulk_g2s(stage(s, i), src[i] + row0 * row_bytes, bytes, bars + s);
        if (MASK) bulk_g2s(mstage(s), msrc + row0 * cv, mbytes, bars + s);
    }

    // returns the number of valid rows of tile t (held by stage s, in its `phase`-th use) once landed
    __device__ __forceinline__ int wait(int t, int s, uint32_t phase) const {
        mbar_wait(bars + s, phase);
        const int64_t row0 = r0 + (int64_t)t * rows_per_tile;
        return (int)((r1 - row0 < rows_per_tile) ? (r1 - row0) : rows_per_tile);
    }

    // every thread is done reading tile t (stage s): refill the stage with tile t + stages
    __device__ __forceinline__ void release(int t, int s) {
        __syncthreads();
        if (threadIdx.x == 0 && t + stages < ntile) {
            fence_proxy_async();
            issue(t + stages, s);
        }
    }
};

// loop helper: stage index and phase bit advance without integer division
#define B200_RING_LOOP(ring, t, s, ph) \
    for (int t = 0, s = 0, ph = 0; t < (ring).ntile; ++t, (s + 1 == (ring).stages ? (s = 0, ph ^= 1) : ++s))

// ---- forward pass 1 ----------------------------------------------------------------------------
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads)
bn_stats_pipe_kernel(const T *__restrict__ x, float2 *__restrict__ partial, int64_t m, int c, int txl, int nchunk,
                     int64_t rpc, int reverse, int nstages) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ float red[kWarps * 32 * 2 * VEC];
    pdl_trigger();
    const int tx = 1 << txl;
    const int txi = threadIdx.x & (tx - 1);
    const int cv = c / VEC;
    const int cvi = txi;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.x) : (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    TileRing<1, false> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(x);
    ring.msrc = nullptr;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);
    float sh[VEC];
    VecIO<T, VEC>::load(x + r0 * c + (int64_t)cvi * VEC, sh);
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

// ---- forward pass 3 / eval ---------------------------------------------------------------------
template <typename T, int VEC, int ACT, bool EVAL>
__global__ void __launch_bounds__(kThreads)
bn_apply_pipe_kernel(const T *__restrict__ x, const T *__restrict__ z, T *__restrict__ y, uint8_t *__restrict__ mask,
                     const float *__restrict__ saved, const float *__restrict__ gamma, const float *__restrict__ beta,
                     const float *__restrict__ rmean, const float *__restrict__ rvar, float eps, int64_t m, int c,
                     int txl, int nchunk, int64_t rpc, int reverse, int nstages, const float *__restrict__ saved2) {
    // saved2 != nullptr (ADD_RELU only): z is the RAW output of the downsample convolution and is
    // normalised here with the downsample BatchNorm's {scale, shift} before the add ("dual" form)
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    constexpr int NT = (ACT == B200_ACT_ADD_RELU) ? 2 : 1;
    const int tx = 1 << txl;
    const int txi = threadIdx.x & (tx - 1);
    const int cv = c / VEC;
    const int cvi = txi;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.x) : (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    TileRing<NT, false> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(x);
    if (NT == 2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(z);
    ring.msrc = nullptr;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);  // x (and z) tiles are already in flight ...
    if (!EVAL) pdl_wait();                                               // ... before the statistics are final
    float sc[VEC], sf[VEC], sc2[VEC], sf2[VEC];
    const bool dual = (ACT == B200_ACT_ADD_RELU) && saved2 != nullptr;
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
        sc2[k] = dual ? saved2[2 * c + ch] : 1.f;
        sf2[k] = dual ? saved2[3 * c + ch] : 0.f;
    }
    const int64_t coff = (int64_t)cvi * VEC;
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

// ---- backward pass 1 ---------------------------------------------------------------------------
template <typename T, int VEC, int ACT, bool DY2>
__global__ void __launch_bounds__(kThreads)
bn_bwd_reduce_pipe_kernel(const T *__restrict__ dy, const T *__restrict__ dy2, const T *__restrict__ x,
                          const uint8_t *__restrict__ mask,
                          const float *__restrict__ saved, float2 *__restrict__ partial, int64_t m, int c, int txl,
                          int nchunk, int64_t rpc, int reverse, int nstages) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ float red[kWarps * 32 * 2 * VEC];
    pdl_trigger();
    constexpr bool MASK = (ACT == B200_ACT_ADD_RELU);
    const int tx = 1 << txl;
    const int txi = threadIdx.x & (tx - 1);
    const int cv = c / VEC;
    const int cvi = txi;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.x) : (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    constexpr int NT = DY2 ? 3 : 2;
    TileRing<NT, MASK> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(dy);
    ring.src[1] = reinterpret_cast<const unsigned char *>(x);
    if (DY2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(dy2);
    ring.msrc = mask;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);
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
                if (DY2) {  // the incoming gradient is the sum of two branches (residual fan-in)
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

// ---- backward pass 3 ---------------------------------------------------------------------------
template <typename T, int VEC, int ACT, bool DY2>
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_pipe_kernel(const T *__restrict__ dy, const T *__restrict__ dy2, const T *__restrict__ x,
                         const uint8_t *__restrict__ mask,
                         const float *__restrict__ saved, const float *__restrict__ coef, T *__restrict__ dx,
                         T *__restrict__ dz, int64_t m, int c, int txl, int nchunk, int64_t rpc, int reverse, int nstages) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    constexpr bool MASK = (ACT == B200_ACT_ADD_RELU);
    const int tx = 1 << txl;
    const int txi = threadIdx.x & (tx - 1);
    const int cv = c / VEC;
    const int cvi = txi;
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.x) : (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    constexpr int NT = DY2 ? 3 : 2;
    TileRing<NT, MASK> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(dy);
    ring.src[1] = reinterpret_cast<const unsigned char *>(x);
    if (DY2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(dy2);
    ring.msrc = mask;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);  // dy / x tiles in flight ...
    pdl_wait();                                                          // ... before the coefficients exist
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
    const int64_t coff = (int64_t)cvi * VEC;
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

// ---- "dual" backward: y = relu(bn(x) + bn2(x2)) (residual block with a downsample branch) ---------
// One reduce pass accumulates sum(g), sum(g*(x-mean)), sum(g*(x2-mean2)) with g = (dy [+dy2]) * mask;
// one apply pass writes both dx and dx2.  Versus running the two BatchNorm backwards separately this
// never writes nor re-reads dz (the masked gradient): 8 E instead of 11 E of traffic.
template <typename T, int VEC, bool DY2>
__global__ void __launch_bounds__(kThreads)
bn_bwd_reduce_dual_pipe_kernel(const T *__restrict__ dy, const T *__restrict__ dy2, const T *__restrict__ x,
                               const T *__restrict__ x2, const uint8_t *__restrict__ mask,
                               const float *__restrict__ saved, const float *__restrict__ saved2,
                               float4 *__restrict__ partial, int64_t m, int c, int txl, int nchunk, int64_t rpc,
                               int reverse, int nstages) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ float red[kThreads * 3 * VEC];
    pdl_trigger();
    constexpr int NT = DY2 ? 4 : 3;
    const int tx = 1 << txl;
    const int cvi = threadIdx.x & (tx - 1);
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.x) : (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    TileRing<NT, true> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(dy);
    ring.src[1] = reinterpret_cast<const unsigned char *>(x);
    ring.src[2] = reinterpret_cast<const unsigned char *>(x2);
    if (DY2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(dy2);
    ring.msrc = mask;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);
    float mean[VEC], mean2[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        mean[k] = saved[cvi * VEC + k];
        mean2[k] = saved2[cvi * VEC + k];
    }
    float acc[3 * VEC];
#pragma unroll
    for (int k = 0; k < 3 * VEC; ++k) acc[k] = 0.f;
    B200_RING_LOOP(ring, t, s, ph) {
        const int nrows = ring.wait(t, s, (uint32_t)ph);
        const uint4 *tg = reinterpret_cast<const uint4 *>(ring.stage(s, 0));
        const uint4 *tv = reinterpret_cast<const uint4 *>(ring.stage(s, 1));
        const uint4 *tv2 = reinterpret_cast<const uint4 *>(ring.stage(s, 2));
        const uint4 *tg2 = reinterpret_cast<const uint4 *>(ring.stage(s, NT - 1));
        const unsigned char *tm = ring.mstage(s);
#pragma unroll
        for (int u = 0; u < kPipeU; ++u) {
            const int v = threadIdx.x + u * kThreads;
            if ((v >> txl) < nrows) {
                float g[VEC], e[VEC], e2[VEC];
                unpack16<T, VEC>(tg[v], g);
                unpack16<T, VEC>(tv[v], e);
                unpack16<T, VEC>(tv2[v], e2);
                if (DY2) {
                    float h[VEC];
                    unpack16<T, VEC>(tg2[v], h);
#pragma unroll
                    for (int k = 0; k < VEC; ++k) g[k] += h[k];
                }
                const unsigned bits = tm[v];
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const float gk = ((bits >> k) & 1u) ? g[k] : 0.f;
                    acc[k] += gk;
                    acc[VEC + k] = fmaf(gk, e[k] - mean[k], acc[VEC + k]);
                    acc[2 * VEC + k] = fmaf(gk, e2[k] - mean2[k], acc[2 * VEC + k]);
                }
            }
        }
        ring.release(t, s);
    }
    cta_reduce_rows<3 * VEC>(acc, tx, red);
    if (threadIdx.x < tx) {
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            partial[(int64_t)(cvi * VEC + k) * nchunk + chunk] = make_float4(acc[k], acc[VEC + k], acc[2 * VEC + k], 0.f);
    }
}

__global__ void __launch_bounds__(kThreads)
bn_bwd_finalize_dual_kernel(const float4 *__restrict__ partial, int nchunk, int64_t m, int c,
                            const float *__restrict__ gamma, const float *__restrict__ gamma2,
                            const float *__restrict__ saved, const float *__restrict__ saved2,
                            float *__restrict__ dgamma, float *__restrict__ dbeta, float *__restrict__ dgamma2,
                            float *__restrict__ dbeta2, float *__restrict__ coef /*[6][C]*/) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ch = blockIdx.x * kWarps + warp;
    if (ch >= c) return;
    const float4 *p = partial + (int64_t)ch * nchunk;
    float a = 0.f, b = 0.f, b2 = 0.f;
    for (int i = lane; i < nchunk; i += 32) {
        const float4 v = p[i];
        a += v.x;
        b += v.y;
        b2 += v.z;
    }
    a = warp_sum(a);
    b = warp_sum(b);
    b2 = warp_sum(b2);
    if (lane == 0) {
        const float inv_m = 1.0f / (float)m;
        const float gm = a * inv_m;
        {
            const float invstd = saved[c + ch], gs = gamma[ch] * invstd;
            if (dgamma) dgamma[ch] = b * invstd;
            if (dbeta) dbeta[ch] = a;
            coef[ch] = gs;
            coef[c + ch] = -(b * invstd * invstd * inv_m) * gs;
            coef[2 * c + ch] = -gm * gs;
        }
        {
            const float invstd = saved2[c + ch], gs = gamma2[ch] * invstd;
            if (dgamma2) dgamma2[ch] = b2 * invstd;
            if (dbeta2) dbeta2[ch] = a;
            coef[3 * c + ch] = gs;
            coef[4 * c + ch] = -(b2 * invstd * invstd * inv_m) * gs;
            coef[5 * c + ch] = -gm * gs;
        }
    }
}

template <typename T, int VEC, bool DY2>
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_dual_pipe_kernel(const T *__restrict__ dy, const T *__restrict__ dy2, const T *__restrict__ x,
                              const T *__restrict__ x2, const uint8_t *__restrict__ mask,
                              const float *__restrict__ saved, const float *__restrict__ saved2,
                              const float *__restrict__ coef, T *__restrict__ dx, T *__restrict__ dx2, int64_t m, int c,
                              int txl, int nchunk, int64_t rpc, int reverse, int nstages) {
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    constexpr int NT = DY2 ? 4 : 3;
    const int tx = 1 << txl;
    const int cvi = threadIdx.x & (tx - 1);
    const int chunk = reverse ? (nchunk - 1 - (int)blockIdx.x) : (int)blockIdx.x;
    const int64_t r0 = (int64_t)chunk * rpc;
    const int64_t r1 = (r0 + rpc < m) ? (r0 + rpc) : m;
    TileRing<NT, true> ring;
    ring.src[0] = reinterpret_cast<const unsigned char *>(dy);
    ring.src[1] = reinterpret_cast<const unsigned char *>(x);
    ring.src[2] = reinterpret_cast<const unsigned char *>(x2);
    if (DY2) ring.src[NT - 1] = reinterpret_cast<const unsigned char *>(dy2);
    ring.msrc = mask;
    ring.setup(dyn_smem, txl, nstages, (int64_t)c * sizeof(T), r0, r1);
    pdl_wait();
    float mean[VEC], c0[VEC], c1[VEC], c2[VEC], mean2[VEC], d0[VEC], d1[VEC], d2[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int ch = cvi * VEC + k;
        mean[k] = saved[ch];
        c0[k] = coef[ch];
        c1[k] = coef[c + ch];
        c2[k] = coef[2 * c + ch];
        mean2[k] = saved2[ch];
        d0[k] = coef[3 * c + ch];
        d1[k] = coef[4 * c + ch];
        d2[k] = coef[5 * c + ch];
    }
    const int64_t coff = (int64_t)cvi * VEC;
    B200_RING_LOOP(ring, t, s, ph) {
        const int nrows = ring.wait(t, s, (uint32_t)ph);
        const uint4 *tg = reinterpret_cast<const uint4 *>(ring.stage(s, 0));
        const uint4 *tv = reinterpret_cast<const uint4 *>(ring.stage(s, 1));
        const uint4 *tv2 = reinterpret_cast<const uint4 *>(ring.stage(s, 2));
        const uint4 *tg2 = reinterpret_cast<const uint4 *>(ring.stage(s, NT - 1));
        const unsigned char *tm = ring.mstage(s);
        const int64_t row0 = r0 + (int64_t)t * ring.rows_per_tile;
#pragma unroll
        for (int u = 0; u < kPipeU; ++u) {
            const int v = threadIdx.x + u * kThreads;
            const int r = v >> txl;
            if (r < nrows) {
                float g[VEC], e[VEC], e2[VEC];
                unpack16<T, VEC>(tg[v], g);
                unpack16<T, VEC>(tv[v], e);
                unpack16<T, VEC>(tv2[v], e2);
                if (DY2) {
                    float h[VEC];
                    unpack16<T, VEC>(tg2[v], h);
#pragma unroll
                    for (int k = 0; k < VEC; ++k) g[k] += h[k];
                }
                const unsigned bits = tm[v];
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const float gk = ((bits >> k) & 1u) ? g[k] : 0.f;
                    e[k] = fmaf(gk, c0[k], fmaf(e[k] - mean[k], c1[k], c2[k]));
                    e2[k] = fmaf(gk, d0[k], fmaf(e2[k] - mean2[k], d1[k], d2[k]));
                }
                const int64_t rr = row0 + r;
                VecIO<T, VEC>::store(dx + rr * c + coff, e);
                VecIO<T, VEC>::store(dx2 + rr * c + coff, e2);
            }
        }
        ring.release(t, s);
    }
}
