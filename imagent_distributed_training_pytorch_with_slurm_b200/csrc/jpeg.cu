// Batched JPEG decode into the ragged input slab through nvJPEG (see include/b200_jpeg.h).
// Reference site: datasets.ImageNet -> pil_loader in the DataLoader workers (imagenet.py:293-296,350-359).
// nvJPEG does the decoding; this file partitions a batch between its hardware-engine backend (baseline single-scan files,
// no host work) and its GPU-assisted backend (everything else nvJPEG can read; Huffman stage on a pool of host threads),
// and points both at interleaved RGB rows inside the caller's slab.
#include <cuda_runtime.h>
#include <nvjpeg.h>
#include <stdint.h>

#include <vector>

#include "../../include/b200_jpeg.h"

struct b200_jpeg_decoder {
    nvjpegHandle_t hw = nullptr, gpu = nullptr;
    nvjpegJpegState_t hw_state = nullptr, gpu_state = nullptr;
    nvjpegJpegStream_t probe = nullptr;
    int cpu_threads = 1, hw_batch = 0, gpu_batch = 0, last_hw = 0;
    std::vector<const unsigned char *> ptr_hw, ptr_gpu;
    std::vector<size_t> len_hw, len_gpu;
    std::vector<nvjpegImage_t> out_hw, out_gpu;
};

namespace {
inline int code(nvjpegStatus_t s) { return s == NVJPEG_STATUS_SUCCESS ? 0 : 1000 + (int)s; }
}  // namespace

extern "C" int b200_jpeg_create(int cpu_threads, int prefer_hardware, b200_jpeg_decoder **out) {
    if (!out) return (int)cudaErrorInvalidValue;
    b200_jpeg_decoder *d = new b200_jpeg_decoder();
    d->cpu_threads = cpu_threads < 1 ? 1 : cpu_threads;
    nvjpegStatus_t s = nvjpegCreateEx(NVJPEG_BACKEND_GPU_HYBRID, nullptr, nullptr, 0, &d->gpu);
    if (s != NVJPEG_STATUS_SUCCESS) s = nvjpegCreateEx(NVJPEG_BACKEND_DEFAULT, nullptr, nullptr, 0, &d->gpu);
    if (s == NVJPEG_STATUS_SUCCESS) s = nvjpegJpegStateCreate(d->gpu, &d->gpu_state);
    if (s != NVJPEG_STATUS_SUCCESS) {
        b200_jpeg_destroy(d);
        return code(s);
    }
    if (prefer_hardware) {   // absent engine / driver support: carry on with the GPU-assisted backend alone
        if (nvjpegCreateEx(NVJPEG_BACKEND_HARDWARE, nullptr, nullptr, 0, &d->hw) == NVJPEG_STATUS_SUCCESS) {
            if (nvjpegJpegStateCreate(d->hw, &d->hw_state) != NVJPEG_STATUS_SUCCESS ||
                nvjpegJpegStreamCreate(d->hw, &d->probe) != NVJPEG_STATUS_SUCCESS) {
                if (d->probe) nvjpegJpegStreamDestroy(d->probe);
                if (d->hw_state) nvjpegJpegStateDestroy(d->hw_state);
                nvjpegDestroy(d->hw);
                d->hw = nullptr; d->hw_state = nullptr; d->probe = nullptr;
            }
        } else {
            d->hw = nullptr;
        }
    }
    *out = d;
    return 0;
}

extern "C" void b200_jpeg_destroy(b200_jpeg_decoder *d) {
    if (!d) return;
    if (d->probe) nvjpegJpegStreamDestroy(d->probe);
    if (d->hw_state) nvjpegJpegStateDestroy(d->hw_state);
    if (d->hw) nvjpegDestroy(d->hw);
    if (d->gpu_state) nvjpegJpegStateDestroy(d->gpu_state);
    if (d->gpu) nvjpegDestroy(d->gpu);
    delete d;
}

extern "C" int b200_jpeg_has_hardware(const b200_jpeg_decoder *d) { return d && d->hw ? 1 : 0; }
extern "C" int b200_jpeg_last_hardware_count(const b200_jpeg_decoder *d) { return d ? d->last_hw : 0; }

extern "C" int b200_jpeg_info(b200_jpeg_decoder *d, const uint8_t *data, size_t length, int *height, int *width,
                              int *components) {
    if (!d || !data || !height || !width || !components) return (int)cudaErrorInvalidValue;
    int nc = 0, w[NVJPEG_MAX_COMPONENT] = {0}, h[NVJPEG_MAX_COMPONENT] = {0};
    nvjpegChromaSubsampling_t ss;
    const nvjpegStatus_t s = nvjpegGetImageInfo(d->gpu, data, length, &nc, &ss, w, h);
    if (s != NVJPEG_STATUS_SUCCESS) return code(s);
    *height = h[0];
    *width = w[0];
    *components = nc;
    return 0;
}

extern "C" int b200_jpeg_decode_batch(b200_jpeg_decoder *d, const uint8_t *const *data, const size_t *lengths, int n,
                                      uint8_t *slab, const int64_t *table, void *stream) {
    if (!d || !data || !lengths || !slab || !table || n < 1) return (int)cudaErrorInvalidValue;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    d->ptr_hw.clear(); d->len_hw.clear(); d->out_hw.clear();
    d->ptr_gpu.clear(); d->len_gpu.clear(); d->out_gpu.clear();
    for (int i = 0; i < n; ++i) {
        if (!data[i] || table[4 * i] < 0 || table[4 * i + 3] < 3 * table[4 * i + 2]) return (int)cudaErrorInvalidValue;
        nvjpegImage_t img = {};
        img.channel[0] = slab + table[4 * i];
        img.pitch[0] = (size_t)table[4 * i + 3];
        bool on_engine = false;
        if (d->hw) {   // is_supported == 0 means the hardware backend takes this file
            int unsupported = 1;
            if (nvjpegJpegStreamParse(d->hw, data[i], lengths[i], 0, 0, d->probe) == NVJPEG_STATUS_SUCCESS &&
                nvjpegDecodeBatchedSupported(d->hw, d->probe, &unsupported) == NVJPEG_STATUS_SUCCESS)
                on_engine = unsupported == 0;
        }
        if (on_engine) {
            d->ptr_hw.push_back(data[i]); d->len_hw.push_back(lengths[i]); d->out_hw.push_back(img);
        } else {
            d->ptr_gpu.push_back(data[i]); d->len_gpu.push_back(lengths[i]); d->out_gpu.push_back(img);
        }
    }
    d->last_hw = (int)d->ptr_hw.size();
    nvjpegStatus_t s;
    if (!d->ptr_hw.empty()) {
        const int b = (int)d->ptr_hw.size();
        if (b != d->hw_batch) {   // the batched API is initialised for one batch size at a time
            s = nvjpegDecodeBatchedInitialize(d->hw, d->hw_state, b, 1, NVJPEG_OUTPUT_RGBI);
            if (s != NVJPEG_STATUS_SUCCESS) return code(s);
            d->hw_batch = b;
        }
        s = nvjpegDecodeBatched(d->hw, d->hw_state, d->ptr_hw.data(), d->len_hw.data(), d->out_hw.data(), st);
        if (s != NVJPEG_STATUS_SUCCESS) return code(s);
    }
    if (!d->ptr_gpu.empty()) {
        const int b = (int)d->ptr_gpu.size();
        if (b != d->gpu_batch) {
            s = nvjpegDecodeBatchedInitialize(d->gpu, d->gpu_state, b, d->cpu_threads, NVJPEG_OUTPUT_RGBI);
            if (s != NVJPEG_STATUS_SUCCESS) return code(s);
            d->gpu_batch = b;
        }
        s = nvjpegDecodeBatched(d->gpu, d->gpu_state, d->ptr_gpu.data(), d->len_gpu.data(), d->out_gpu.data(), st);
        if (s != NVJPEG_STATUS_SUCCESS) return code(s);
    }
    return 0;
}

extern "C" const char *b200_jpeg_error_string(int c) {
    if (c == 0) return "success";
    if (c < 1000) return cudaGetErrorString((cudaError_t)c);
    switch (c - 1000) {
        case NVJPEG_STATUS_NOT_INITIALIZED: return "nvJPEG: not initialized";
        case NVJPEG_STATUS_INVALID_PARAMETER: return "nvJPEG: invalid parameter";
        case NVJPEG_STATUS_BAD_JPEG: return "nvJPEG: bad JPEG";
        case NVJPEG_STATUS_JPEG_NOT_SUPPORTED: return "nvJPEG: JPEG not supported";
        case NVJPEG_STATUS_ALLOCATOR_FAILURE: return "nvJPEG: allocator failure";
        case NVJPEG_STATUS_EXECUTION_FAILED: return "nvJPEG: execution failed";
        case NVJPEG_STATUS_ARCH_MISMATCH: return "nvJPEG: architecture mismatch";
        case NVJPEG_STATUS_INTERNAL_ERROR: return "nvJPEG: internal error";
        case NVJPEG_STATUS_IMPLEMENTATION_NOT_SUPPORTED: return "nvJPEG: implementation not supported";
        case NVJPEG_STATUS_INCOMPLETE_BITSTREAM: return "nvJPEG: incomplete bitstream";
        default: return "nvJPEG: unknown status";
    }
}
