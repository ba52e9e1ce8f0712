// K3: cross-entropy (mean) + top-1/top-5 precision in one kernel, and the loss backward.
// Reference sites: criterion(logits, target) imagenet.py:124 (nn.CrossEntropyLoss defaults,
// torch/nn/functional.py:3421-3504) and accuracy(logits, target, (1,5)) imagenet.py:63-79,134.
//
// One warp per row.  A row (1000 classes) is 2-4 KB, so the whole [N,1000] problem is ~1 MB:
// latency-bound, not bandwidth-bound.  The design goal is ONE launch instead of ~12 (log_softmax,
// nll_loss, topk, transpose, eq, 2x(reshape,float,sum,mul)) and a bit-reproducible sum.
#include <math.h>

#include <mutex>

#include "common.cuh"

namespace {

constexpr int kWarpsPerBlock = 8;

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
struct RowVec;  // 16-byte vector view of a row
template <>
struct RowVec<float> {
    static constexpr int N = 4;
    static __device__ __forceinline__ void load(const float *p, float (&v)[4]) { VecIO<float, 4>::load(p, v); }
    static __device__ __forceinline__ void store(float *p, const float (&v)[4]) { VecIO<float, 4>::store(p, v); }
};
template <>
struct RowVec<__nv_bfloat16> {
    static constexpr int N = 8;
    static __device__ __forceinline__ void load(const __nv_bfloat16 *p, float (&v)[8]) { VecIO<__nv_bfloat16, 8>::load(p, v); }
    static __device__ __forceinline__ void store(__nv_bfloat16 *p, const float (&v)[8]) { VecIO<__nv_bfloat16, 8>::store(p, v); }
};

struct OnlineLse {
    float m = -INFINITY, s = 0.f;
    __device__ __forceinline__ void add(float x) {
        if (x > m) {
            s = s * expf(m - x) + 1.f;  // m = -inf, s = 0 on the first element: 0*0 + 1
            m = x;
        } else {
            s += expf(x - m);
        }
    }
    __device__ __forceinline__ void merge(float om, float os) {
        float nm = fmaxf(m, om);
        if (nm == -INFINITY) return;
        s = s * expf(m - nm) + os * expf(om - nm);
        m = nm;
    }
};

// workspace layout: [0,16): ticket (uint32) ; then float rows[3*n] = {loss_n, hit1_n, hit5_n}
template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
ce_topk_fwd_kernel(const T *__restrict__ logits, const int64_t *__restrict__ target, float *__restrict__ out3,
                   float *__restrict__ row_lse, unsigned *__restrict__ ticket, float *__restrict__ rows, int64_t n,
                   int64_t c, float pct, bool vec_ok) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kWarpsPerBlock + warp;
    if (row < n) {
        const T *x = logits + row * c;
        const int64_t t = target[row];
        const bool t_ok = (t >= 0 && t < c);
        const float xt = t_ok ? to_f<T>(x[t]) : NAN;
        OnlineLse acc;
        int ahead = 0;
        constexpr int V = RowVec<T>::N;
        if (vec_ok) {
            const int64_t nv = c / V;
            for (int64_t v = lane; v < nv; v += 32) {
                float e[V];
                RowVec<T>::load(x + v * V, e);
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    const int64_t j = v * V + k;
                    acc.add(e[k]);
                    ahead += (e[k] > xt) || (e[k] == xt && j < t);
                }
            }
            for (int64_t j = nv * V + lane; j < c; j += 32) {
                float e = to_f<T>(x[j]);
                acc.add(e);
                ahead += (e > xt) || (e == xt && j < t);
            }
        } else {
            for (int64_t j = lane; j < c; j += 32) {
                float e = to_f<T>(x[j]);
                acc.add(e);
                ahead += (e > xt) || (e == xt && j < t);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float om = __shfl_xor_sync(0xffffffffu, acc.m, o);
            float os = __shfl_xor_sync(0xffffffffu, acc.s, o);
            acc.merge(om, os);
            ahead += __shfl_xor_sync(0xffffffffu, ahead, o);
        }
        if (lane == 0) {
            const float lse = acc.m + logf(acc.s);
            row_lse[row] = lse;
            rows[row] = lse - xt;
            rows[n + row] = (t_ok && ahead < 1) ? 1.f : 0.f;
            rows[2 * n + row] = (t_ok && ahead < 5) ? 1.f : 0.f;
            __threadfence();  // the row results must be visible device-wide before this block takes its ticket
        }
    }
    // ---- last block to arrive adds the rows up in a fixed order ---------------------------------
    __shared__ bool is_last;
    __shared__ float red[3][kWarpsPerBlock];
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned prev = atomicAdd(ticket, 1u);
        is_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (int64_t r = threadIdx.x; r < n; r += blockDim.x) {
        s0 += __ldcg(rows + r);
        s1 += __ldcg(rows + n + r);
        s2 += __ldcg(rows + 2 * n + r);
    }
    s0 = warp_sum(s0); s1 = warp_sum(s1); s2 = warp_sum(s2);
    if (lane == 0) { red[0][warp] = s0; red[1][warp] = s1; red[2][warp] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f, d = 0.f;
        for (int w = 0; w < kWarpsPerBlock; ++w) { a += red[0][w]; b += red[1][w]; d += red[2][w]; }
        out3[0] = a / (float)n;   // nll_loss 'mean': sum / N
        out3[1] = b * pct;        // count.float() * (100.0 / batch)   (imagenet.py:77-78)
        out3[2] = d * pct;
        *ticket = 0u;             // ready for the next call
    }
}

template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
ce_bwd_kernel(const T *__restrict__ logits, const int64_t *__restrict__ target, const float *__restrict__ row_lse,
              const float *__restrict__ grad_out, float scale, T *__restrict__ dlogits, int64_t n, int64_t c,
              bool vec_ok) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kWarpsPerBlock + warp;
    if (row >= n) return;
    const float g = (grad_out ? __ldg(grad_out) : 1.f) * scale / (float)n;
    const T *x = logits + row * c;
    T *d = dlogits + row * c;
    const int64_t t = target[row];
    const float lse = row_lse[row];
    constexpr int V = RowVec<T>::N;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t nv = c / V;
        for (int64_t v = lane; v < nv; v += 32) {
            float e[V];
            RowVec<T>::load(x + v * V, e);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const int64_t j = v * V + k;
                e[k] = (expf(e[k] - lse) - (j == t ? 1.f : 0.f)) * g;
            }
            RowVec<T>::store(d + v * V, e);
        }
        done = nv * V;
    }
    for (int64_t j = done + lane; j < c; j += 32) {
        float e = (expf(to_f<T>(x[j]) - lse) - (j == t ? 1.f : 0.f)) * g;
        if constexpr (sizeof(T) == 4) d[j] = e;
        else d[j] = __float2bfloat16_rn(e);
    }
}

// rank of the target class per row (the count behind accuracy() for ANY k): one warp per row, same tie rule as above
template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
target_rank_kernel(const T *__restrict__ logits, const int64_t *__restrict__ target, int32_t *__restrict__ rank,
                   int64_t n, int64_t c, bool vec_ok) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kWarpsPerBlock + warp;
    if (row >= n) return;
    const T *x = logits + row * c;
    const int64_t t = target[row];
    const bool t_ok = (t >= 0 && t < c);
    const float xt = t_ok ? to_f<T>(x[t]) : NAN;
    int ahead = 0;
    constexpr int V = RowVec<T>::N;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t nv = c / V;
        for (int64_t v = lane; v < nv; v += 32) {
            float e[V];
            RowVec<T>::load(x + v * V, e);
#pragma unroll
            for (int k = 0; k < V; ++k) ahead += (e[k] > xt) || (e[k] == xt && v * V + k < t);
        }
        done = nv * V;
    }
    for (int64_t j = done + lane; j < c; j += 32) {
        const float e = to_f<T>(x[j]);
        ahead += (e > xt) || (e == xt && j < t);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ahead += __shfl_xor_sync(0xffffffffu, ahead, o);
    if (lane == 0) rank[row] = t_ok ? ahead : (int32_t)c;   // an invalid target is never "in the top k"
}

}  // namespace

extern "C" int b200_target_rank(const void *logits, const int64_t *target, int32_t *rank, int64_t n, int64_t c,
                                int dtype, void *stream) {
    if (!logits || !target || !rank || n <= 0 || c <= 0) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(logits)) return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = (int)((n + kWarpsPerBlock - 1) / kWarpsPerBlock);
    const size_t esz = dtype == B200_F32 ? 4 : 2;
    const bool vec_ok = ((size_t)c * esz) % 16 == 0;
    if (dtype == B200_F32)
        target_rank_kernel<float><<<blocks, kWarpsPerBlock * 32, 0, st>>>(static_cast<const float *>(logits), target,
                                                                           rank, n, c, vec_ok);
    else
        target_rank_kernel<__nv_bfloat16><<<blocks, kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const __nv_bfloat16 *>(logits), target, rank, n, c, vec_ok);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" size_t b200_ce_topk_workspace_bytes(int64_t n) {
    if (n < 0) n = 0;
    return 16 + sizeof(float) * 3 * (size_t)n;
}

extern "C" int b200_ce_topk_fwd(const void *logits, const int64_t *target, float *out3, float *row_lse,
                                void *workspace, size_t workspace_bytes, int64_t n, int64_t c, int dtype,
                                void *stream) {
    if (!logits || !target || !out3 || !row_lse || !workspace || n <= 0 || c <= 0) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (workspace_bytes < b200_ce_topk_workspace_bytes(n)) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(logits) || b200_misaligned(workspace)) return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned *ticket = static_cast<unsigned *>(workspace);
    float *rows = reinterpret_cast<float *>(static_cast<char *>(workspace) + 16);
    const int blocks = (int)((n + kWarpsPerBlock - 1) / kWarpsPerBlock);
    const float pct = (float)(100.0 / (double)n);
    const size_t esz = dtype == B200_F32 ? 4 : 2;
    const bool vec_ok = ((size_t)c * esz) % 16 == 0;
    if (dtype == B200_F32)
        ce_topk_fwd_kernel<float><<<blocks, kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const float *>(logits), target, out3, row_lse, ticket, rows, n, c, pct, vec_ok);
    else
        ce_topk_fwd_kernel<__nv_bfloat16><<<blocks, kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const __nv_bfloat16 *>(logits), target, out3, row_lse, ticket, rows, n, c, pct, vec_ok);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_ce_bwd(const void *logits, const int64_t *target, const float *row_lse, const float *grad_out,
                           float scale, void *dlogits, int64_t n, int64_t c, int dtype, void *stream) {
    if (!logits || !target || !row_lse || !dlogits || n <= 0 || c <= 0) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (b200_misaligned(logits) || b200_misaligned(dlogits)) return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = (int)((n + kWarpsPerBlock - 1) / kWarpsPerBlock);
    const size_t esz = dtype == B200_F32 ? 4 : 2;
    const bool vec_ok = ((size_t)c * esz) % 16 == 0;
    if (dtype == B200_F32)
        ce_bwd_kernel<float><<<blocks, kWarpsPerBlock * 32, 0, st>>>(static_cast<const float *>(logits), target,
                                                                      row_lse, grad_out, scale,
                                                                      static_cast<float *>(dlogits), n, c, vec_ok);
    else
        ce_bwd_kernel<__nv_bfloat16><<<blocks, kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const __nv_bfloat16 *>(logits), target, row_lse, grad_out, scale,
            static_cast<__nv_bfloat16 *>(dlogits), n, c, vec_ok);
    B200_CHECK_LAUNCH();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Loss + top-1/top-5 for hosts that cannot hand in a workspace (see the header for who that is).
// The scratch the kernel needs (a 16-byte ticket and three floats per row) lives in this array, which
// every GPU that loads the library carries until the process exits: 16 + 12 * 16384 = 196,624 bytes.
static __device__ __align__(256) unsigned char g_ce_static_scratch[16 + sizeof(float) * 3 * B200_CE_STATIC_MAX_ROWS];

namespace {
constexpr int kMaxDevices = 64;
// In host code `g_ce_static_scratch` is only a handle for the runtime, not an address a kernel can
// use; the real device address differs per GPU and is looked up once per device, below.
void *g_ce_static_addr[kMaxDevices] = {};
std::mutex g_ce_static_mutex;
}  // namespace

extern "C" int b200_ce_topk_fwd_static(const void *logits, const int64_t *target, float *out3, float *row_lse,
                                       int64_t n, int64_t c, int dtype, void *stream) {
    if (!logits || !target || !out3 || !row_lse || n <= 0 || c <= 0) return (int)cudaErrorInvalidValue;
    if (dtype != B200_F32 && dtype != B200_BF16) return (int)cudaErrorInvalidValue;
    if (n > B200_CE_STATIC_MAX_ROWS) return (int)cudaErrorInvalidValue;  // more rows than the built-in scratch holds
    if (b200_misaligned(logits)) return (int)cudaErrorMisalignedAddress;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    void *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_ce_static_mutex);
        scratch = g_ce_static_addr[dev];
        if (!scratch) {  // first call on this GPU (it is the current device): resolve the array's address here
            e = cudaGetSymbolAddress(&scratch, g_ce_static_scratch);
            if (e != cudaSuccess) return (int)e;
            g_ce_static_addr[dev] = scratch;
        }
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // The array still holds whatever the previous call left.  Only the ticket has to start at zero (the
    // per-row slots of rows < n are all overwritten before they are read), so clear just that.
    e = cudaMemsetAsync(scratch, 0, 16, st);
    if (e != cudaSuccess) return (int)e;
    unsigned *ticket = static_cast<unsigned *>(scratch);
    float *rows = reinterpret_cast<float *>(static_cast<char *>(scratch) + 16);
    const int blocks = (int)((n + kWarpsPerBlock - 1) / kWarpsPerBlock);
    const float pct = (float)(100.0 / (double)n);
    const size_t esz = dtype == B200_F32 ? 4 : 2;
    const bool vec_ok = ((size_t)c * esz) % 16 == 0;
    if (dtype == B200_F32)
        ce_topk_fwd_kernel<float><<<blocks, kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const float *>(logits), target, out3, row_lse, ticket, rows, n, c, pct, vec_ok);
    else
        ce_topk_fwd_kernel<__nv_bfloat16><<<blocks, kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const __nv_bfloat16 *>(logits), target, out3, row_lse, ticket, rows, n, c, pct, vec_ok);
    B200_CHECK_LAUNCH();
    return 0;
}
