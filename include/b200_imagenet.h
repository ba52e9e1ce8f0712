/*
 * b200_imagenet.h - C ABI of the B200 (sm_100a) kernels behind the data-parallel training step of
 * the reference's imagenet.py (train(), imagenet.py:97-151).
 *
 * The reference is a pure PyTorch script with no FFI of its own; each entry point below names the
 * reference call site (and the PyTorch dispatch behind it) that it replaces.  Conventions, the same
 * for every function unless its comment says otherwise:
 *   - plain pointers and sizes only; all data pointers are DEVICE pointers on the current device;
 *   - returns 0 on success, otherwise a cudaError_t value (argument errors are reported as
 *     cudaErrorInvalidValue = 1, misaligned pointers as cudaErrorMisalignedAddress = 716);
 *     b200_error_string() turns the code into text;
 *   - launches on the stream handed in (a cudaStream_t passed as void*; NULL = legacy default
 *     stream), never synchronises the device, never allocates or frees device memory: scratch
 *     ("workspace") is passed in by the caller and sized with the matching *_workspace_bytes();
 *   - data pointers must be 16-byte aligned;
 *   - dtype: B200_F32 (float) or B200_BF16 (__nv_bfloat16) for activations; statistics, parameters,
 *     gradients of parameters and all accumulation are float32;
 *   - activations are "rows x channels": a channels-last [N,H,W,C] tensor seen as [M = N*H*W, C].
 */
#ifndef B200_IMAGENET_H_
#define B200_IMAGENET_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 2

enum { B200_F32 = 0, B200_BF16 = 1 };
enum { B200_ACT_NONE = 0, B200_ACT_RELU = 1, B200_ACT_ADD_RELU = 2 };
enum { B200_LAYOUT_NHWC = 0, B200_LAYOUT_NCHW = 1, B200_LAYOUT_S2D_STEM = 2 };

/* Library identity / diagnostics. */
int b200_abi_version(void);
const char *b200_error_string(int code);

/* Process-wide launch tuning (no reference counterpart; used by the benchmark sweeps).  Keys:
 *   "bn_target_ctas"   CTAs a BatchNorm pass aims for (default 148*8)
 *   "bn_vec_bf16"      channels per thread for bf16 activations, 4 or 8 (default 8); forward and
 *                      backward of one layer must run under the same value (it fixes the mask layout)
 *   "bn_stats_reverse" 1 = statistics/reduce passes walk the rows back to front (default 1)
 *   "bn_apply_reverse" 1 = apply passes walk the rows back to front (default 0)
 *   "bn_pipeline"      1 = bulk-async (cp.async.bulk + mbarrier) shared-memory ring kernels where the
 *                      shape allows (16-byte vectors, power-of-two C/vec), 0 = direct-load kernels
 *   "bn_pipe_ctas_per_sm" resident CTAs per SM the bulk-async grid is sized for (default 2)
 *   "bn_pdl"           1 = the finalize/apply kernels of the bulk-async path are launched with programmatic
 *                      dependent launch (they start while their predecessor drains), default 1
 *   "bn_fused"         1 = one cooperative kernel per BatchNorm site (statistics, fold and apply with grid-wide barriers
 *                      inside) instead of the three-kernel chain; same bits, measured slower on B200, default 0
 *   "bn_pipe_stages"   shared-memory ring depth of the bulk-async kernels, 2..16 (default 0 = automatic)
 *   "input_s2d_fast"   1 (default): B200_LAYOUT_S2D_STEM output is produced by the cell-per-thread kernel;
 *                      0: by the generic kernel (same bits, slower; kept as the in-library cross-check)
 *   "pack_steps"       16-byte vectors per thread of the gradient pack kernel: 1 (default), 2, 4 or 8
 * Returns cudaErrorInvalidValue for an unknown key or value. */
int b200_tune_set(const char *key, int value);

/* ------------------------------------------------------------------------------------------------
 * K4b  fused flat-buffer SGD.
 * Replaces optimizer.step() (imagenet.py:131; SGD built at imagenet.py:325 with momentum 0.9,
 * weight_decay 1e-4, dampening 0, nesterov False), i.e. the four _foreach passes of
 * torch/optim/sgd.py:425-473 whose per-element meaning is torch/optim/sgd.py:343-380:
 *     g' = g*grad_scale + weight_decay*p ;  buf = first_step ? g' : momentum*buf + g' ;
 *     p  = p - lr*buf
 * over n contiguous float32 elements in ONE pass (read p,g,buf; write p,buf = 20 B/element).
 * shadow_bf16, when not NULL, receives the updated p rounded to bfloat16 (the copy the bf16
 * convolutions read), +2 B/element.  grad_scale folds DDP's 1/world_size
 * (reducer.hpp:497-500) when the caller all-reduces un-divided sums; pass 1.0f otherwise.
 */
int b200_sgd_flat_f32(float *p, const float *g, float *buf, void *shadow_bf16, int64_t n, float lr,
                      float weight_decay, float momentum, float grad_scale, int first_step, void *stream);

/* ------------------------------------------------------------------------------------------------
 * K3  cross-entropy + top-1/top-5 in one kernel.
 * Replaces criterion(logits, target) (imagenet.py:124; nn.CrossEntropyLoss() defaults,
 * torch/nn/functional.py:3421-3504 = log_softmax + nll_loss mean) AND accuracy(logits, target,
 * (1,5)) (imagenet.py:63-79,134 = topk + eq + sum).  One warp per row:
 *     out3[0] = mean_n(logsumexp(x_n) - x_n[t_n])
 *     out3[1] = 100/N * #{n : rank_n < 1},  out3[2] = 100/N * #{n : rank_n < 5}
 * rank_n = #{j : x_nj > x_nt  or (x_nj == x_nt and j < t)}   (ties go to the lower class index).
 * row_lse[n] = logsumexp(x_n) is kept for the backward.  Rows are summed in a fixed order by the
 * last block to finish, so the result is bit-reproducible.  workspace must be zero-filled before
 * its FIRST use (the kernel leaves it ready for the next call); calls sharing a workspace must not
 * overlap on the GPU.  target values must lie in [0, c).
 */
size_t b200_ce_topk_workspace_bytes(int64_t n);
int b200_ce_topk_fwd(const void *logits, const int64_t *target, float *out3, float *row_lse, void *workspace,
                     size_t workspace_bytes, int64_t n, int64_t c, int dtype, void *stream);

/* The same loss + top-1/top-5 computation for callers that cannot supply a workspace.
 * (This entry point has no counterpart in the reference.)
 * Who needs it: host code that is not allowed to obtain device memory when it wants a loss value and has
 * no scratch buffer of its own to pass - for example a metrics or validation hook that runs after the
 * training framework has claimed all of HBM for its own pool, or a monitoring thread that must not call
 * cudaMalloc/cudaFree because those stall every stream of the device while they run.  The scratch comes
 * out of the library image instead: a fixed array of 16 + 12*B200_CE_STATIC_MAX_ROWS = 196,624 bytes that
 * every GPU loading the library holds until the process ends.  The call itself allocates and frees
 * nothing (no cudaMalloc/cudaFree, no allocating library routine); logits, target, out3 and row_lse are
 * still the caller's buffers, with the same meaning as in b200_ce_topk_fwd.
 *   - n must not exceed B200_CE_STATIC_MAX_ROWS; a larger batch is refused with cudaErrorInvalidValue
 *     before anything is launched (use b200_ce_topk_fwd with your own workspace for those).
 *   - The first call on each GPU looks up that GPU's address of the array (the GPU must be the current
 *     device, as for every call here); make that first call before starting any stream capture.
 *   - All calls on a GPU share the one array.  Each call re-zeroes the 16-byte ticket on `stream` and
 *     rewrites the per-row slots it uses, so back-to-back calls on one stream are fine; two calls that
 *     can overlap on the GPU (different streams, no dependency) race on the array and give garbage.
 */
#define B200_CE_STATIC_MAX_ROWS 16384
int b200_ce_topk_fwd_static(const void *logits, const int64_t *target, float *out3, float *row_lse, int64_t n,
                            int64_t c, int dtype, void *stream);

/* accuracy(output, target, topk) for ANY tuple of k (imagenet.py:63-79: output.topk(maxk, 1, True, True), eq, sum):
 * rank[n] = #{j : x_nj > x_nt or (x_nj == x_nt and j < t)} as int32, so precision@k = 100/N * #{n : rank[n] < k} for
 * every k from one pass (the reference's top-k select + transpose + eq per call).  Same tie rule as b200_ce_topk_fwd;
 * a target outside [0, c) gets rank c.  No workspace. */
int b200_target_rank(const void *logits, const int64_t *target, int32_t *rank, int64_t n, int64_t c, int dtype,
                     void *stream);

/* Backward of the loss above (loss.backward(), imagenet.py:128, first node):
 *     dlogits[n][j] = (exp(x_nj - row_lse[n]) - [j == t_n]) * scale * (grad_out ? *grad_out : 1) / N
 * grad_out is a DEVICE scalar (the upstream gradient, 1.0 for loss.backward()) or NULL; scale is a
 * host-side multiplier.  dlogits has the dtype of logits.
 */
int b200_ce_bwd(const void *logits, const int64_t *target, const float *row_lse, const float *grad_out,
                float scale, void *dlogits, int64_t n, int64_t c, int dtype, void *stream);

/* ------------------------------------------------------------------------------------------------
 * K2  BatchNorm2d (+ residual add) + ReLU on channels-last rows.
 * Replaces nn.BatchNorm2d in training mode followed by ReLU(inplace) and, in the residual blocks,
 * `out += identity; out = relu(out)`:
 *   B200_ACT_RELU     torchvision resnet.py:93-94,147-148,151-152,269-270
 *   B200_ACT_ADD_RELU torchvision resnet.py:97-103,155-161           (y = relu(bn(x) + z))
 *   B200_ACT_NONE     torchvision resnet.py:240-243 (downsample)
 * arithmetic of torch/nn/modules/batchnorm.py:160-206: biased batch variance for the
 * normalisation, running_var updated with the unbiased one, num_batches_tracked += 1.
 *
 * forward, training:  three launches (per-chunk shifted sums -> per-channel finalize -> apply).
 *   saved: float[4*C] out = {mean, invstd, scale = gamma*invstd, shift = beta - mean*scale}
 *   mask : B200_ACT_ADD_RELU only, out, M*C/vec bytes (one byte per `vec` channels, bit k set iff
 *          the pre-activation of channel k of the group is > 0); vec = b200_bn_vec(dtype).  The
 *          buffer must be 16-byte aligned and 16 bytes longer than that (the backward fetches it
 *          with 16-byte bulk copies)
 *   z    : B200_ACT_ADD_RELU only.  x,z,y: [M,C] of `dtype`.  Requires M >= 2, C % vec == 0.
 */
int b200_bn_vec(int dtype);
size_t b200_bn_workspace_bytes(int64_t m, int c, int dtype);
int b200_bn_fwd_train(const void *x, const void *z, void *y, uint8_t *mask, const float *gamma, const float *beta,
                      float *running_mean, float *running_var, int64_t *num_batches_tracked, float *saved,
                      void *workspace, size_t workspace_bytes, int64_t m, int c, float eps, float momentum, int act,
                      int dtype, void *stream);

/* forward, eval mode (validate(), imagenet.py:176,187): y = act(gamma*(x-rm)/sqrt(rv+eps)+beta [+z]),
 * one launch, no statistics. */
int b200_bn_fwd_eval(const void *x, const void *z, void *y, const float *gamma, const float *beta,
                     const float *running_mean, const float *running_var, int64_t m, int c, float eps, int act,
                     int dtype, void *stream);

/* backward (loss.backward(), imagenet.py:128; cudnn/native batch_norm_backward + threshold_backward
 * + the add's two-way gradient):  dyh = dy*[pre-activation > 0] (ACT_NONE: dyh = dy)
 *     dbeta = sum dyh ; dgamma = invstd * sum dyh*(x-mean)
 *     dx = gamma*invstd*(dyh - dbeta/M - (x-mean)*invstd^2*sum(dyh*(x-mean))/M) ;  dz = dyh
 * three launches (per-chunk sums -> finalize -> apply).  The ReLU mask is recomputed from x and
 * saved{scale,shift} for B200_ACT_RELU, read from `mask` for B200_ACT_ADD_RELU.  dgamma/dbeta are
 * written (not accumulated) as float32 - they may point straight into a flat gradient buffer.
 * dy2 (nullable): a second gradient tensor of the same shape; the incoming gradient is then dy + dy2
 * (summed in float32 on the fly).  This is the residual fan-in of `out += identity`
 * (resnet.py:102,160): the block output's gradient arrives from the next block's first convolution
 * AND from its identity branch, and autograd would add the two with a separate kernel.  Only the
 * bulk-async kernels take dy2; other shapes return cudaErrorNotSupported (add the tensors first).
 */
int b200_bn_bwd(const void *dy, const void *dy2, const void *x, const uint8_t *mask, const float *gamma, const float *saved,
                void *dx, void *dz, float *dgamma, float *dbeta, void *workspace, size_t workspace_bytes,
                int64_t m, int c, int act, int dtype, void *stream);

/* "Dual" form for residual blocks whose identity branch is itself conv + BatchNorm (the downsample
 * branch, torchvision resnet.py:97-103,155-161 with :240-243):  y = relu(bn(x) + bn2(x2)).
 * Running the two BatchNorms separately writes the normalised identity z = bn2(x2) and reads it back,
 * and in the backward writes the masked gradient dz and reads it back twice; the dual kernels never
 * materialise either (forward 3 E instead of 5 E for the pair of apply passes, backward 8 E instead of 11 E).
 *   b200_bn_stats          statistics + finalize of one BatchNorm, no output tensor (use it for bn2):
 *                          fills saved{mean,invstd,scale,shift}, updates the running statistics
 *   b200_bn_fwd_train_dual statistics of x, then y = relu(x*scale+shift + x2*saved2.scale+saved2.shift),
 *                          mask as in b200_bn_fwd_train(ADD_RELU)
 *   b200_bn_bwd_dual       g = (dy [+ dy2]) * mask;  dbeta = dbeta2 = sum g;  dgamma / dgamma2 from
 *                          sum g*(x-mean) / sum g*(x2-mean2);  dx and dx2 by the BatchNorm backward formula
 * Only the bulk-async kernels implement the dual forms (C/vec a power of two <= 256); other shapes return
 * cudaErrorNotSupported and the caller uses the separate entry points. */
int b200_bn_dual_supported(int c, int dtype); /* 1 when the dual entry points handle this channel count */
int b200_bn_stats(const void *x, const float *gamma, const float *beta, float *running_mean, float *running_var,
                  int64_t *num_batches_tracked, float *saved, void *workspace, size_t workspace_bytes, int64_t m, int c,
                  float eps, float momentum, int dtype, void *stream);
int b200_bn_fwd_train_dual(const void *x, const void *x2, void *y, uint8_t *mask, const float *gamma, const float *beta,
                           float *running_mean, float *running_var, int64_t *num_batches_tracked, float *saved,
                           const float *saved2, void *workspace, size_t workspace_bytes, int64_t m, int c, float eps,
                           float momentum, int dtype, void *stream);
int b200_bn_bwd_dual(const void *dy, const void *dy2, const void *x, const void *x2, const uint8_t *mask,
                     const float *gamma, const float *gamma2, const float *saved, const float *saved2, void *dx,
                     void *dx2, float *dgamma, float *dbeta, float *dgamma2, float *dbeta2, void *workspace,
                     size_t workspace_bytes, int64_t m, int c, int dtype, void *stream);

/* ------------------------------------------------------------------------------------------------
 * K1  input stage: crop box -> antialiased bilinear resize -> horizontal flip -> /255 ->
 * (x-mean)/std -> cast, from uint8 NHWC batches, one pass.
 * Replaces transforms.Compose([Resize((448,448)), ToTensor(), Normalize(.5,.5)]) run on PIL images
 * in the DataLoader workers (imagenet.py:280-283,350-354) - box = whole image, no flip - and, with
 * per-sample boxes/flips, torchvision RandomResizedCrop + RandomHorizontalFlip
 * (transforms.py:928-981,712-722).  The resize is Pillow's 8-bit BILINEAR resample (Resample.c:
 * double-precision triangle weights -> 22-bit fixed point, horizontal pass rounded to uint8, then
 * vertical pass rounded to uint8), reproduced bit for bit; the normalisation is evaluated in float32
 * with IEEE division exactly as ToTensor/Normalize do, then rounded once to the output dtype.
 *   images : uint8 [n, src_h, src_w, 3]
 *   boxes  : int32 [n,4] rows (top, left, height, width), inside the image, on the DEVICE
 *   flips  : uint8 [n] (non-zero = mirror), on the DEVICE
 *   out    : [n,out_h,out_w,3] (B200_LAYOUT_NHWC), [n,3,out_h,out_w] (B200_LAYOUT_NCHW), or
 *            [n,(out_h+6)/2,(out_w+6)/2,16] (B200_LAYOUT_S2D_STEM: the image zero-padded by 3 pixels and
 *            2x2 space-to-depth'ed, cell channel (py%2*2+px%2)*3+ch, 4 spare channels zero; out_h and
 *            out_w even).  With the 7x7 weights rearranged the same way, the stem convolution
 *            (7x7, stride 2, padding 3, 3->64; torchvision resnet.py:197) becomes a 4x4 stride-1
 *            convolution over 16 channels - the same sums, a shape the tensor cores like better
 *   mean3/std3 : HOST pointers to 3 floats each
 * two launches (coefficient tables into workspace, then the resample).  Down-scaling is limited to
 * src/out <= 8 per axis.
 */
size_t b200_input_workspace_bytes(int64_t n, int src_h, int src_w, int out_h, int out_w);
int b200_input_crop_resize_flip_norm(const uint8_t *images, const int32_t *boxes, const uint8_t *flips, void *out,
                                     void *workspace, size_t workspace_bytes, int64_t n, int src_h, int src_w,
                                     int out_h, int out_w, const float *mean3, const float *std3, int out_dtype,
                                     int out_layout, void *stream);

/* K1 on a RAGGED batch: the images of a real dataset have different sizes (the reference resizes each PIL image on
 * its own, imagenet.py:280-283,293-296), so the decoded batch is one byte slab with a table instead of a dense
 * [n,h,w,3] tensor:
 *   slab      : uint8, slab_bytes long; image i is interleaved RGB with rows src_table[4i+3] bytes apart
 *   src_table : int64 [n,4] on the DEVICE, rows (byte offset of the image in the slab, height, width, row pitch)
 *   boxes     : as above, each inside ITS image (whole image = (0, 0, height, width): the reference's Resize)
 *   max_src_h / max_src_w : upper bounds of the heights / widths in the batch (they size the tap tables and the
 *               shared-memory staging; workspace = b200_input_workspace_bytes(n, max_src_h, max_src_w, ...));
 *               the 8x down-scaling limit applies to them.
 * Same arithmetic, same kernels and the same output layouts as the dense entry point. */
int b200_input_crop_resize_flip_norm_ragged(const uint8_t *slab, int64_t slab_bytes, const int64_t *src_table,
                                            const int32_t *boxes, const uint8_t *flips, void *out, void *workspace,
                                            size_t workspace_bytes, int64_t n, int max_src_h, int max_src_w, int out_h,
                                            int out_w, const float *mean3, const float *std3, int out_dtype,
                                            int out_layout, void *stream);

/* Decode staging for the ragged input stage.  The reference decodes JPEGs with PIL inside the DataLoader workers
 * (datasets.ImageNet -> pil_loader, imagenet.py:293-296,350-359); here the bytes can be decoded on the device by
 * nvJPEG (a library), which returns every image as three planes [3,h,w].  This re-packs a batch of such planar images
 * into the interleaved slab b200_input_crop_resize_flip_norm_ragged reads, in one launch:
 *   table : int64 [n,4] on the DEVICE, rows (device address of the planar image, height, width, byte offset of the
 *           image in the slab - a multiple of 16; rows are packed, pitch = 3*width)
 *   max_pixels : largest height*width in the batch (sizes the grid). */
int b200_interleave_planar_u8(const int64_t *table, uint8_t *slab, int64_t n, int64_t max_pixels, void *stream);

/* ------------------------------------------------------------------------------------------------
 * K6  3x3 / stride 2 / pad 1 max-pooling, channels-last, forward and backward.
 * Replaces nn.MaxPool2d(kernel_size=3, stride=2, padding=1) after the stem (torchvision
 * resnet.py:200,270) and its gradient inside loss.backward() (imagenet.py:128), i.e. ATen's
 * max_pool2d_with_indices / max_pool2d_with_indices_backward (int64 argmax tensor).
 *   x      : [n,h,w,c] of dtype, c % 8 == 0        y : [n,oh,ow,c], oh = (h-1)/2+1, ow = (w-1)/2+1
 *   argmax : uint8 [n,oh,ow,c]: position 0..8 (row-major inside the window) of the maximum; the first
 *            maximum wins, NaN counts as maximum (ATen's rule)
 *   backward: dx[n,h,w,c] = sum of dy over the (at most 4) windows whose argmax is that pixel,
 *            accumulated in float32; every element of dx is written (no zero-fill needed).
 *            dy2 (may be NULL): a second gradient of the same shape that is added to dy on the fly - the
 *            fan-in sum autograd performs where the pooled tensor feeds both branches of the first
 *            residual block (torchvision resnet.py:143-161,240-243), without a separate add kernel.
 */
int b200_maxpool3x3s2_fwd(const void *x, void *y, uint8_t *argmax, int64_t n, int c, int h, int w, int dtype,
                          void *stream);
int b200_maxpool3x3s2_bwd(const void *dy, const void *dy2, const uint8_t *argmax, void *dx, int64_t n, int c, int h,
                          int w, int dtype, void *stream);
/* The stem's BatchNorm apply + ReLU + max-pool in one pass (torchvision resnet.py:269-270 then :200/:271:
 * `x = self.bn1(x); x = self.relu(x); x = self.maxpool(x)`): every tap is mapped through
 * relu(x*scale[ch] + shift[ch]) - float32 fma, rounded once to the activation dtype, i.e. exactly the values
 * b200_bn_fwd_train(act = RELU) would have stored - before the comparison, so y and argmax are bit-identical to running
 * the two kernels one after the other, while the normalised activation (4x the pooled one: 411 MB at N=256) is neither
 * written nor read.  Nothing downstream needs it: the BatchNorm backward recomputes the ReLU mask from x, the pool's
 * backward routes through argmax.  scale_shift: float32 [2,c] on the device = rows 2 and 3 of the `saved` tensor that
 * b200_bn_stats / b200_bn_fwd_train write (saved + 2*c).  x: the convolution output [n,h,w,c]; y, argmax as above. */
int b200_bn_relu_maxpool3x3s2_fwd(const void *x, const float *scale_shift, void *y, uint8_t *argmax, int64_t n, int c,
                                  int h, int w, int dtype, void *stream);

/* ------------------------------------------------------------------------------------------------
 * K4c  gradient exchange + optimizer step of one gradient bucket in ONE kernel over NVLink peer memory
 * (opt-in: FlatDDP(exchange="peer"); the default path is NCCL all-reduce + b200_sgd_flat_f32).
 * Replaces, per bucket, DDP's all-reduce (reducer.hpp:497-524, driven by loss.backward(), imagenet.py:128)
 * followed by optimizer.step() (imagenet.py:131): every rank owns 1/world of the bucket, sums the world gradient
 * shards in rank order through peer-mapped pointers (reduce-scatter), applies the update of b200_sgd_flat_f32, and
 * writes the new parameters (+ bf16 shadow) into every rank's arena (all-gather).
 *   grad_ptrs/param_ptrs/shadow_ptrs/pad_ptrs : HOST arrays of `world` DEVICE pointers, entry r = rank r's float32
 *       gradient arena / float32 parameter arena / bf16 shadow (shadow_ptrs may be NULL) / signal pad, all mapped into
 *       this process (cudaIpc / symmetric memory); entry `rank` is the local one.
 *   momentum : this rank's float32 momentum arena (only the owned shard is read and written).
 *   bucket   : 0..22, the same on every rank for the same [lo,hi) (element offsets, multiples of 4).
 *   max_ctas : CTAs the kernel may use (it runs beside the backward); <= 0 means 32.
 * Signal pads: b200_peer_pad_bytes() bytes per rank, zero before the first call, never touched by the host afterwards
 * except word 0, which a kernel sets to 1 if a peer's flag did not arrive within ~20 s.  The word latches: from then on
 * this rank's kernels store nothing into any parameter / shadow / momentum arena, raise no completion flag and do not
 * advance the iteration counter, so the peers time out in turn and stop updating too - replicas stay as they were
 * before the failing iteration instead of being half-updated.  The host reads the word after a synchronise.  Every rank must launch the same buckets in the same
 * order each iteration and finish the iteration with b200_peer_wait, which waits until every rank's shard of every
 * bucket has arrived (and every peer has finished reading this rank's gradients) and advances the iteration counter.
 */
size_t b200_peer_pad_bytes(void);
int b200_peer_exchange_sgd(const void *const *grad_ptrs, void *const *param_ptrs, void *const *shadow_ptrs,
                           void *const *pad_ptrs, float *momentum, int rank, int world, int bucket, int64_t lo,
                           int64_t hi, float lr, float weight_decay, float momentum_coef, float grad_scale,
                           int first_step, int max_ctas, void *stream);
/* b200_peer_bcast: DDP's per-iteration buffer broadcast (distributed.py:1590-1591) in the same style - every rank but
 * `root` copies the root's buffer arena (buf_ptrs: host array of `world` peer-mapped device pointers, nbytes each)
 * into its own after the root's flag, and reports back; the ROOT passes bcast_readers_of_me = 1 to b200_peer_wait
 * so that it does not touch its buffers again before everyone has read them.  Gradient buckets use indices 0..22. */
int b200_peer_bcast(void *const *buf_ptrs, void *const *pad_ptrs, int64_t nbytes, int rank, int world, int root,
                    void *stream);
int b200_peer_wait(void *my_pad, int world, int n_buckets, int rank, int bcast_readers_of_me, void *stream);

/* ------------------------------------------------------------------------------------------------
 * K5  gradient packing: many weight-gradient tensors -> their slots of one flat float32 buffer.
 * Replaces the per-parameter `bucket_view.copy_(grad)` / `mul_out(grad, 1/W)` of DDP's Reducer
 * (reducer.hpp:347-406,497-500; driven by loss.backward(), imagenet.py:128) with one launch per flush:
 * tensor i has counts[i] contiguous elements of `src_dtype` at srcs[i] (DEVICE pointer, 16-byte aligned)
 * and lands at dst + offsets[i] (offsets in elements, multiples of 4), multiplied by `scale`.
 * srcs / offsets / counts are HOST arrays of length n_tensors; they are copied into kernel parameters
 * (64 tensors per launch), so nothing is allocated and the arrays may be reused as soon as the call returns.
 */
int b200_pack_grads(const void *const *srcs, const int64_t *offsets, const int64_t *counts, int n_tensors, float *dst,
                    float scale, int src_dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_IMAGENET_H_ */
