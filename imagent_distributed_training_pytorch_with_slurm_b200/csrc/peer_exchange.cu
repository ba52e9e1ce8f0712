// K4c: gradient exchange + optimizer step of one gradient bucket as ONE kernel over NVLink peer memory.
// Reference sites: DDP's bucket all-reduce (reducer.hpp:497-524, driven by loss.backward(), imagenet.py:128)
// followed by optimizer.step() (imagenet.py:131).  The default path of this package does those as an NCCL all-reduce
// on the bucket + the flat SGD kernel (K4b); this kernel is the peer-memory form:
//
//   every rank owns 1/W of the bucket.  For its shard it (1) reads the W gradient shards - its own and, through
//   peer-mapped pointers over NVLink, everyone else's - and adds them in rank order (= reduce-scatter), (2) applies
//   weight decay, momentum and the update exactly as K4b does (sgd.cu), (3) writes the new parameters and their bf16
//   shadow into EVERY rank's arena (= all-gather).  No intermediate buffer, one pass over the gradients instead of
//   NCCL's reduce-scatter + all-gather + a separate optimizer pass, and nothing but kernel launches - so a whole
//   multi-GPU iteration can sit in one CUDA graph.
//
// Hand-off between ranks uses 32-bit flags in a peer-mapped signal pad per rank, written with st.release.sys and read
// with ld.acquire.sys; the value is a step number that only grows, so nothing is ever reset:
//   A[bucket][src]  "src's gradients of this bucket are final"      written by src into every rank's pad at kernel start
//   B[bucket][src]  "src has finished its shard of this bucket"     written by src's last CTA into every rank's pad
// b200_peer_wait (a one-CTA kernel at the end of the backward) waits for all B flags: after it every parameter of the
// step has arrived and every peer has finished reading this rank's gradients, so the next iteration may start.
// A flag that does not arrive within ~20 s latches the error word (pad[0]) instead of hanging the GPU.  From then on
// this rank's kernels store NOTHING into any parameter, shadow or momentum arena and raise no completion flag, and
// b200_peer_wait leaves the iteration counter where it was - so the peers time out in turn, latch their own error
// word and stop updating too: replicas are left as they were before the failing iteration, never half-updated from
// gradients that were not final.  The host reads the word wherever it synchronises (FlatDDP.peer_error(), checked by
// train() at the end of an epoch).
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxRanks = 8;
constexpr int kMaxBuckets = 24;
constexpr int kSlotStep = 1;                                   // pad[1]: iteration counter (device-resident)
constexpr int kSlotCounter = 8;                                // pad[8 + bucket]: CTA arrival counter of a bucket
constexpr int kSlotFlags = 64;                                 // pad[64 + (bucket*2 + {A=0,B=1})*kMaxRanks + src]
constexpr long long kTimeoutCycles = 40000000000LL;            // ~20 s at 2 GHz: ranks may be seconds apart while cuDNN autotunes

struct PeerPtrs {
    const float *grad[kMaxRanks];
    float *param[kMaxRanks];
    __nv_bfloat16 *shadow[kMaxRanks];
    unsigned *pad[kMaxRanks];
};

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int flag_slot(int bucket, int which, int src) {
    return kSlotFlags + (bucket * 2 + which) * kMaxRanks + src;
}
// spin until *p has reached `want` (wrap-safe); false on timeout
__device__ __forceinline__ bool wait_flag(const unsigned *p, unsigned want) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(p) - want) < 0) {
        if (clock64() - t0 > kTimeoutCycles) return false;
        __nanosleep(100);
    }
    return true;
}

__global__ void __launch_bounds__(kThreads)
peer_exchange_sgd_kernel(const __grid_constant__ PeerPtrs P, float *__restrict__ momentum, int rank, int world,
                         int bucket, long long b4, long long e4, float lr, float weight_decay, float mu,
                         float grad_scale, int first_step) {
    unsigned *mypad = P.pad[rank];
    __shared__ int failed;
    if (threadIdx.x == 0) failed = (int)*reinterpret_cast<volatile unsigned *>(mypad);   // latched by an earlier kernel
    __syncthreads();
    if (failed) return;                           // no flags, no stores: the peers run into their own time-out
    const unsigned seq = mypad[kSlotStep] + 1u;   // only this rank's b200_peer_wait writes the counter, after this kernel
    if (blockIdx.x == 0 && threadIdx.x < world) st_release_sys(P.pad[threadIdx.x] + flag_slot(bucket, 0, rank), seq);
    if (threadIdx.x < world && !wait_flag(mypad + flag_slot(bucket, 0, threadIdx.x), seq)) {
        mypad[0] = 1u;
        failed = 1;
    }
    __syncthreads();
    if (failed) return;                           // some peer's gradients are not known to be final: touch nothing

    const float4 *myp = reinterpret_cast<const float4 *>(P.param[rank]);
    float4 *mom = reinterpret_cast<float4 *>(momentum);
    for (long long i = b4 + (long long)blockIdx.x * kThreads + threadIdx.x; i < e4; i += (long long)gridDim.x * kThreads) {
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p) {
            if (p < world) {   // rank order: the sum has the same bits whoever owns the shard
                const float4 t = __ldcg(reinterpret_cast<const float4 *>(P.grad[p]) + i);
                g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
            }
        }
        float4 w = myp[i];
        float4 m = mom[i];
        // torch/optim/sgd.py:343-380 (as sgd.cu): g' = g*scale + wd*p ; buf = first ? g' : mu*buf + g' ; p -= lr*buf
        const float gx = fmaf(weight_decay, w.x, g.x * grad_scale), gy = fmaf(weight_decay, w.y, g.y * grad_scale),
                    gz = fmaf(weight_decay, w.z, g.z * grad_scale), gw = fmaf(weight_decay, w.w, g.w * grad_scale);
        m.x = first_step ? gx : fmaf(mu, m.x, gx);
        m.y = first_step ? gy : fmaf(mu, m.y, gy);
        m.z = first_step ? gz : fmaf(mu, m.z, gz);
        m.w = first_step ? gw : fmaf(mu, m.w, gw);
        w.x = fmaf(-lr, m.x, w.x);
        w.y = fmaf(-lr, m.y, w.y);
        w.z = fmaf(-lr, m.z, w.z);
        w.w = fmaf(-lr, m.w, w.w);
        mom[i] = m;
        const __nv_bfloat162 lo = __floats2bfloat162_rn(w.x, w.y), hi = __floats2bfloat162_rn(w.z, w.w);
        const uint2 sh = make_uint2(*reinterpret_cast<const unsigned *>(&lo), *reinterpret_cast<const unsigned *>(&hi));
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p) {
            if (p < world) {
                __stcg(reinterpret_cast<float4 *>(P.param[p]) + i, w);
                if (P.shadow[p]) __stcg(reinterpret_cast<uint2 *>(P.shadow[p]) + i, sh);
            }
        }
    }
    // completion: the last CTA to arrive tells every rank that this shard is in place
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned *counter = mypad + kSlotCounter + bucket;
        if (atomicAdd(counter, 1u) == gridDim.x - 1) {
            *counter = 0u;
            __threadfence_system();
            for (int p = 0; p < world; ++p) st_release_sys(P.pad[p] + flag_slot(bucket, 1, rank), seq);
        }
    }
}

// Buffer broadcast (DDP's per-iteration _sync_buffers, distributed.py:1590-1591, 2214-2257): the root raises a flag,
// every other rank copies the root's buffer arena through its peer mapping and reports back; the root's
// b200_peer_wait waits for those reports before its next forward may touch the buffers.  Uses the flag slots of the
// last bucket index, which gradient buckets therefore may not use.
constexpr int kBcastSlot = kMaxBuckets - 1;

__global__ void __launch_bounds__(kThreads)
peer_bcast_kernel(const __grid_constant__ PeerPtrs P, const unsigned char *root_buf, unsigned char *my_buf,
                  long long nbytes, int rank, int world, int root) {
    unsigned *mypad = P.pad[rank];
    __shared__ int failed;
    if (threadIdx.x == 0) failed = (int)*reinterpret_cast<volatile unsigned *>(mypad);
    __syncthreads();
    if (failed) return;
    const unsigned seq = mypad[kSlotStep] + 1u;
    if (rank == root) {
        if (blockIdx.x == 0 && threadIdx.x < world) st_release_sys(P.pad[threadIdx.x] + flag_slot(kBcastSlot, 0, root), seq);
        return;
    }
    if (threadIdx.x == 0 && !wait_flag(mypad + flag_slot(kBcastSlot, 0, root), seq)) {
        mypad[0] = 1u;
        failed = 1;
    }
    __syncthreads();
    if (failed) return;                           // the root's statistics never arrived: keep this rank's own
    const long long n8 = nbytes >> 3;
    const uint2 *src = reinterpret_cast<const uint2 *>(root_buf);
    uint2 *dst = reinterpret_cast<uint2 *>(my_buf);
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n8; i += (long long)gridDim.x * kThreads)
        dst[i] = __ldcg(src + i);
    if (blockIdx.x == 0 && threadIdx.x < (nbytes & 7)) my_buf[(n8 << 3) + threadIdx.x] = root_buf[(n8 << 3) + threadIdx.x];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned *counter = mypad + kSlotCounter + kBcastSlot;
        if (atomicAdd(counter, 1u) == gridDim.x - 1) {
            *counter = 0u;
            st_release_sys(P.pad[root] + flag_slot(kBcastSlot, 1, rank), seq);   // "rank has read the root's buffers"
        }
    }
}

__global__ void __launch_bounds__(kThreads)
peer_wait_kernel(unsigned *mypad, int world, int nbuckets, int bcast_readers_of_me, int rank) {
    __shared__ int failed;
    if (threadIdx.x == 0) failed = (int)*reinterpret_cast<volatile unsigned *>(mypad);
    __syncthreads();
    if (failed) return;                              // an exchange kernel of this iteration gave up: do not count it
    const unsigned seq = mypad[kSlotStep] + 1u;
    for (int i = threadIdx.x; i < world * nbuckets; i += kThreads)
        if (!wait_flag(mypad + flag_slot(i / world, 1, i % world), seq)) {
            mypad[0] = 1u;
            failed = 1;
        }
    if (bcast_readers_of_me && threadIdx.x < world && threadIdx.x != rank &&
        !wait_flag(mypad + flag_slot(kBcastSlot, 1, threadIdx.x), seq)) {
        mypad[0] = 1u;
        failed = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0 && !failed) mypad[kSlotStep] = seq;   // the next iteration's flags use seq + 1
}

}  // namespace

extern "C" size_t b200_peer_pad_bytes(void) { return sizeof(unsigned) * (kSlotFlags + kMaxBuckets * 2 * kMaxRanks); }

extern "C" int b200_peer_exchange_sgd(const void *const *grad_ptrs, void *const *param_ptrs, void *const *shadow_ptrs,
                                      void *const *pad_ptrs, float *momentum, int rank, int world, int bucket,
                                      int64_t lo, int64_t hi, float lr, float weight_decay, float momentum_coef,
                                      float grad_scale, int first_step, int max_ctas, void *stream) {
    if (!grad_ptrs || !param_ptrs || !pad_ptrs || !momentum) return (int)cudaErrorInvalidValue;
    if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world || bucket < 0 || bucket >= kMaxBuckets - 1)
        return (int)cudaErrorInvalidValue;   // the last bucket index carries the buffer broadcast's flags
    if (lo < 0 || hi < lo || (lo & 3) || (hi & 3)) return (int)cudaErrorInvalidValue;
    PeerPtrs P;
    for (int p = 0; p < kMaxRanks; ++p) {
        const bool on = p < world;
        P.grad[p] = on ? static_cast<const float *>(grad_ptrs[p]) : nullptr;
        P.param[p] = on ? static_cast<float *>(param_ptrs[p]) : nullptr;
        P.shadow[p] = (on && shadow_ptrs) ? static_cast<__nv_bfloat16 *>(shadow_ptrs[p]) : nullptr;
        P.pad[p] = on ? static_cast<unsigned *>(pad_ptrs[p]) : nullptr;
        if (on && (!P.grad[p] || !P.param[p] || !P.pad[p])) return (int)cudaErrorInvalidValue;
        if (on && (b200_misaligned(P.grad[p]) || b200_misaligned(P.param[p]) || b200_misaligned(P.pad[p], 4) ||
                   (P.shadow[p] && b200_misaligned(P.shadow[p], 8))))
            return (int)cudaErrorMisalignedAddress;
    }
    if (b200_misaligned(momentum)) return (int)cudaErrorMisalignedAddress;
    // this rank's shard of the bucket, in float4 units (absolute positions in the arenas)
    const int64_t n4 = (hi - lo) / 4;
    const int64_t b4 = lo / 4 + n4 * rank / world, e4 = lo / 4 + n4 * (rank + 1) / world;
    int64_t ctas = (e4 - b4 + kThreads * 4 - 1) / (kThreads * 4);   // >= 4 float4 per thread
    if (max_ctas < 1) max_ctas = 32;
    if (ctas > max_ctas) ctas = max_ctas;
    if (ctas < 1) ctas = 1;   // an empty shard still takes part in the flag hand-off
    peer_exchange_sgd_kernel<<<(unsigned)ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        P, momentum, rank, world, bucket, b4, e4, lr, weight_decay, momentum_coef, grad_scale, first_step);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_peer_bcast(void *const *buf_ptrs, void *const *pad_ptrs, int64_t nbytes, int rank, int world, int root,
                               void *stream) {
    if (!buf_ptrs || !pad_ptrs || nbytes < 0 || world < 1 || world > kMaxRanks || rank < 0 || rank >= world || root < 0 ||
        root >= world)
        return (int)cudaErrorInvalidValue;
    PeerPtrs P;
    for (int p = 0; p < kMaxRanks; ++p) {
        P.grad[p] = nullptr; P.param[p] = nullptr; P.shadow[p] = nullptr;
        P.pad[p] = p < world ? static_cast<unsigned *>(pad_ptrs[p]) : nullptr;
        if (p < world && (!P.pad[p] || !buf_ptrs[p])) return (int)cudaErrorInvalidValue;
        if (p < world && (b200_misaligned(P.pad[p], 4) || b200_misaligned(buf_ptrs[p], 8))) return (int)cudaErrorMisalignedAddress;
    }
    int64_t ctas = (nbytes / 8 + kThreads - 1) / kThreads;
    if (ctas > 16) ctas = 16;
    if (ctas < 1) ctas = 1;
    peer_bcast_kernel<<<(unsigned)ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        P, static_cast<const unsigned char *>(buf_ptrs[root]), static_cast<unsigned char *>(buf_ptrs[rank]), nbytes, rank,
        world, root);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_peer_wait(void *my_pad, int world, int n_buckets, int rank, int bcast_readers_of_me, void *stream) {
    if (!my_pad || world < 1 || world > kMaxRanks || n_buckets < 1 || n_buckets > kMaxBuckets - 1 || rank < 0 ||
        rank >= world)
        return (int)cudaErrorInvalidValue;
    if (b200_misaligned(my_pad, 4)) return (int)cudaErrorMisalignedAddress;
    peer_wait_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<unsigned *>(my_pad), world,
                                                                            n_buckets, bcast_readers_of_me, rank);
    B200_CHECK_LAUNCH();
    return 0;
}
