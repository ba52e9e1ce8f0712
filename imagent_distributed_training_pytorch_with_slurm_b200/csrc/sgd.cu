// K4b: one-pass SGD with momentum and weight decay over a flat float32 buffer.
// Reference site: optimizer.step() (imagenet.py:131), SGD(lr, momentum=0.9, weight_decay=1e-4)
// (imagenet.py:325); per-element meaning torch/optim/sgd.py:343-380.
//
// HBM-bound streaming kernel: 12 B read + 8 B written per element (+2 B for the bf16 copy).
// 16-byte vector accesses, two independent vectors per thread per iteration, grid sized to a
// multiple of the SM count with a grid-stride loop.
#include "common.cuh"

namespace {

template <bool FIRST, bool SHADOW>
__device__ __forceinline__ void sgd_elem4(float4 &p, const float4 &g, float4 &b, float lr, float wd, float mom,
                                          float gs) {
    float pv[4] = {p.x, p.y, p.z, p.w}, gv[4] = {g.x, g.y, g.z, g.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float d = fmaf(wd, pv[k], gv[k] * gs);       // g' = g*scale + wd*p
        float m = FIRST ? d : fmaf(mom, bv[k], d);   // buf = mom*buf + g'
        bv[k] = m;
        pv[k] = fmaf(-lr, m, pv[k]);                 // p -= lr*buf
    }
    p = make_float4(pv[0], pv[1], pv[2], pv[3]);
    b = make_float4(bv[0], bv[1], bv[2], bv[3]);
}

template <bool FIRST, bool SHADOW>
__global__ void __launch_bounds__(256) sgd_flat_kernel(float *__restrict__ p, const float *__restrict__ g,
                                                       float *__restrict__ buf, __nv_bfloat16 *__restrict__ shadow,
                                                       int64_t n, float lr, float wd, float mom, float gs) {
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    float4 *p4 = reinterpret_cast<float4 *>(p);
    const float4 *g4 = reinterpret_cast<const float4 *>(g);
    float4 *b4 = reinterpret_cast<float4 *>(buf);
    uint2 *s4 = reinterpret_cast<uint2 *>(shadow);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < n4; i += 2 * stride) {
        const int64_t j = i + stride;
        float4 pa = p4[i], pb = p4[j];
        float4 ga = __ldcs(g4 + i), gb = __ldcs(g4 + j);
        float4 ba, bb;
        if (!FIRST) { ba = b4[i]; bb = b4[j]; }
        sgd_elem4<FIRST, SHADOW>(pa, ga, ba, lr, wd, mom, gs);
        sgd_elem4<FIRST, SHADOW>(pb, gb, bb, lr, wd, mom, gs);
        p4[i] = pa; p4[j] = pb;
        b4[i] = ba; b4[j] = bb;
        if (SHADOW) {
            s4[i] = make_uint2(pack_bf16(pa.x, pa.y), pack_bf16(pa.z, pa.w));
            s4[j] = make_uint2(pack_bf16(pb.x, pb.y), pack_bf16(pb.z, pb.w));
        }
    }
    if (i < n4) {
        float4 pa = p4[i];
        float4 ga = __ldcs(g4 + i);
        float4 ba;
        if (!FIRST) ba = b4[i];
        sgd_elem4<FIRST, SHADOW>(pa, ga, ba, lr, wd, mom, gs);
        p4[i] = pa;
        b4[i] = ba;
        if (SHADOW) s4[i] = make_uint2(pack_bf16(pa.x, pa.y), pack_bf16(pa.z, pa.w));
    }
    // scalar tail (n not a multiple of 4)
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const int64_t e = (n4 << 2) + threadIdx.x;
        float d = fmaf(wd, p[e], g[e] * gs);
        float m = FIRST ? d : fmaf(mom, buf[e], d);
        buf[e] = m;
        float pn = fmaf(-lr, m, p[e]);
        p[e] = pn;
        if (SHADOW) shadow[e] = __float2bfloat16_rn(pn);
    }
}

}  // namespace

extern "C" int b200_sgd_flat_f32(float *p, const float *g, float *buf, void *shadow_bf16, int64_t n, float lr,
                                 float weight_decay, float momentum, float grad_scale, int first_step,
                                 void *stream) {
    if (n < 0 || !p || !g || !buf) return (int)cudaErrorInvalidValue;
    if (n == 0) return 0;
    if (b200_misaligned(p) || b200_misaligned(g) || b200_misaligned(buf) || b200_misaligned(shadow_bf16, 8))
        return (int)cudaErrorMisalignedAddress;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int threads = 256;
    const int64_t n4 = n >> 2;
    // two vectors per thread per iteration; at most 8 resident CTAs per SM
    int64_t want = (n4 + 2 * threads - 1) / (2 * threads);
    int blocks = (int)(want < 1 ? 1 : (want > (int64_t)B200_SM_COUNT * 8 ? (int64_t)B200_SM_COUNT * 8 : want));
    __nv_bfloat16 *sh = static_cast<__nv_bfloat16 *>(shadow_bf16);
    if (first_step) {
        if (sh) sgd_flat_kernel<true, true><<<blocks, threads, 0, st>>>(p, g, buf, sh, n, lr, weight_decay, momentum, grad_scale);
        else sgd_flat_kernel<true, false><<<blocks, threads, 0, st>>>(p, g, buf, sh, n, lr, weight_decay, momentum, grad_scale);
    } else {
        if (sh) sgd_flat_kernel<false, true><<<blocks, threads, 0, st>>>(p, g, buf, sh, n, lr, weight_decay, momentum, grad_scale);
        else sgd_flat_kernel<false, false><<<blocks, threads, 0, st>>>(p, g, buf, sh, n, lr, weight_decay, momentum, grad_scale);
    }
    B200_CHECK_LAUNCH();
    return 0;
}
