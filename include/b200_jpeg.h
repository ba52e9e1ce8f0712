/*
 * b200_jpeg.h - C ABI of the batched JPEG decode that feeds the ragged input stage (libb200_jpeg.so).
 *
 * Replaces the decode half of the reference's data path: datasets.ImageNet -> pil_loader (Image.open(...).convert("RGB"))
 * executed per image in the DataLoader worker processes (imagenet.py:293-296,350-359).  The arithmetic is nvJPEG's (a CUDA
 * toolkit LIBRARY, as cuDNN is for the convolutions): this file only drives its batched API so that a whole batch of JPEG
 * byte strings is decoded - on the GPU's JPEG engines where the files allow it, otherwise with nvJPEG's GPU-assisted
 * decoder and a pool of host threads for the Huffman stage - as INTERLEAVED RGB rows written straight into the byte
 * slab b200_input_crop_resize_flip_norm_ragged reads (include/b200_imagenet.h), with no planar intermediate.
 *
 * It lives in a library of its own so that libb200_imagenet.so keeps depending on libcudart only.  Unlike the kernels of
 * b200_imagenet.h, nvJPEG allocates device and pinned host memory internally (at creation and when a batch is larger than
 * any before).  Functions return 0 on success or 1000 + nvjpegStatus_t (1004 = a file nvJPEG does not support: decode that
 * one elsewhere), or a cudaError_t below 1000.  Not thread-safe per decoder object; one object per GPU and thread.
 */
#ifndef B200_JPEG_H_
#define B200_JPEG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_jpeg_decoder b200_jpeg_decoder;

/* Creates a decoder on the current device.  cpu_threads: host threads nvJPEG may use for the Huffman stage of files the
 * hardware engine does not take.  prefer_hardware: try the GPU's JPEG engines first (baseline, single-scan files). */
int b200_jpeg_create(int cpu_threads, int prefer_hardware, b200_jpeg_decoder **out);
void b200_jpeg_destroy(b200_jpeg_decoder *dec);
/* 1 if the hardware-engine backend could be created on this GPU, else 0. */
int b200_jpeg_has_hardware(const b200_jpeg_decoder *dec);
/* How many images of the last b200_jpeg_decode_batch call went through the hardware engine. */
int b200_jpeg_last_hardware_count(const b200_jpeg_decoder *dec);

/* Header only: height, width, number of components (1 = grayscale, 3 = YCbCr, 4 = CMYK-style: not decodable to RGB here)
 * of a JPEG byte string in HOST memory. */
int b200_jpeg_info(b200_jpeg_decoder *dec, const uint8_t *data, size_t length, int *height, int *width, int *components);

/* Decodes n JPEG byte strings (HOST pointers data[i], lengths[i]) into interleaved RGB: image i lands at
 * slab + table[4*i] (DEVICE memory) with rows table[4*i+3] bytes apart; table[4*i+1], table[4*i+2] = its height and
 * width as reported by b200_jpeg_info (table is a HOST array, the same numbers the ragged input stage takes on the
 * device).  GPU work is enqueued on `stream`; the byte strings must stay valid until the call returns. */
int b200_jpeg_decode_batch(b200_jpeg_decoder *dec, const uint8_t *const *data, const size_t *lengths, int n, uint8_t *slab,
                           const int64_t *table, void *stream);

const char *b200_jpeg_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif
