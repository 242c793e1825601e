/* C ABI of the torchgpipe_b200 native library (libtgb200 / torchgpipe_b200/_C*.so).
 *
 * All tensors are raw device pointers; activations are NHWC ("channels last"), convolution
 * weights are KRSC. Every function enqueues work on `stream` and returns immediately:
 * 0 on success, non-zero on error (message via tgb_last_error()).
 */
#ifndef TGB200_H_
#define TGB200_H_

#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Text of the last error raised on the calling thread. */
const char* tgb_last_error(void);

/* Number of kernels this library has launched in this process (bench.py reports the delta). */
long long tgb_launch_count(void);
void tgb_reset_launch_count(void);

/* Stride-1 2-D convolution forward, TF32 tensor cores (tcgen05), FP32 accumulate.
 *   y[n,ho,wo,k] = sum_{r,s,c} x[n,ho+r-pad_h,wo+s-pad_w,c] * w[k,r,s,c]
 * x: [N,H,W,C]  w: [K,R,S,C]  y: [N,Ho,Wo,K], Ho = H+2*pad_h-R+1, Wo = W+2*pad_w-S+1.
 * Requires C % 4 == 0, K % 4 == 0, 16-byte aligned pointers.
 * stats (optional, may be NULL): FP32 accumulators that the epilogue ADDS per-channel sum and
 * sum of squares of y into; the caller zeroes them first.
 *   stats_mode 1: [N][K][2] per (image, channel)  - InstanceNorm statistics
 *   stats_mode 2: [K][2]    per channel           - BatchNorm statistics
 *   stats_mode -1 / -2: as 1 / 2, and the call zeroes the accumulators itself first (all
 *   forward entry points below accept the negative forms too)
 * Also computes the data gradient of such a convolution when called with dy as x, the
 * tap-flipped / channel-transposed weights and pad' = R-1-pad.
 * Reference call sites: nn.Conv2d forward/backward-data in benchmarks/models/{unet,resnet,
 * amoebanet} (the reference dispatches these to cuDNN). */
int tgb_conv2d_fprop_tf32(const float* x, const float* w, float* y, float* stats,
                          int N, int H, int W, int C, int K, int R, int S,
                          int pad_h, int pad_w, int stats_mode, cudaStream_t stream);

/* Weight gradient of the same convolution (TF32 tcgen05, both operands MN-major from TMA):
 *   dw[k,r,s,c] = sum_{n,ho,wo} dy[n,ho,wo,k] * x[n,ho+r-pad_h,wo+s-pad_w,c]
 * dy: [N,Ho,Wo,K]  x: [N,H,W,C]  dw: [K,R,S,C]. Requires C % 4 == 0, K % 4 == 0.
 * When the pixel range is split across CTAs (few channel tiles, many pixels) or accumulate != 0,
 * fp32 partials go to `scratch` and are summed in a fixed order by a second kernel;
 * tgb_conv2d_wgrad_scratch_bytes() returns how many bytes that needs (0 = none, <0 = error).
 * accumulate != 0 adds the result onto dw instead of overwriting it.
 * Reference call site: autograd of nn.Conv2d (cuDNN in the reference). */
long long tgb_conv2d_wgrad_scratch_bytes(int N, int H, int W, int C, int K, int R, int S,
                                         int pad_h, int pad_w);
int tgb_conv2d_wgrad_tf32(const float* dy, const float* x, float* dw, float* scratch,
                          long long scratch_bytes, int N, int H, int W, int C, int K,
                          int R, int S, int pad_h, int pad_w, int accumulate,
                          cudaStream_t stream);

/* Weight gradient without any allocation. (No counterpart in the reference: it has no native
 * code and lets PyTorch's allocator serve every temporary.)
 *
 * Same result as tgb_conv2d_wgrad_tf32, for callers that must not allocate device memory - a
 * serving or real-time loop running inside a fixed memory reservation, a C/C++ host without a
 * caching allocator (where cudaMalloc/cudaFree would stall the device), or code that records a
 * step for later replay. The split partials are kept in a buffer built into the library,
 * TGB_WGRAD_NOALLOC_SCRATCH_BYTES (32 MiB) per GPU, resident from library load until exit.
 *   - Problems whose partials do not fit are refused (non-zero return, tgb_last_error() names
 *     the need and the limit); use tgb_conv2d_wgrad_tf32 with your own scratch for those.
 *   - The call itself performs no cudaMalloc / cudaFree and uses no allocating library routine.
 *   - The first call on each device resolves the buffer's device address (and, like every entry
 *     point here, configures the kernel on first use). Make that first call during start-up,
 *     before any stream capture begins.
 *   - The one buffer is shared by all calls on a device and keeps whatever the previous call
 *     wrote; each call overwrites every partial before reading it, so nothing needs clearing.
 *     Two calls that can run on the GPU at the same time (different streams, no ordering
 *     between them) race on the buffer: serialise them.
 *   - Each GPU has its own copy of the buffer and the call uses the copy of the CURRENT device:
 *     the device that owns dy / x / dw and `stream` must be current (cudaSetDevice) when calling,
 *     as for every entry point here. A dw that lives on a different GPU than the current one is
 *     refused with an error. */
#define TGB_WGRAD_NOALLOC_SCRATCH_BYTES (32ll << 20)
int tgb_conv2d_wgrad_tf32_noalloc(const float* dy, const float* x, float* dw,
                                  int N, int H, int W, int C, int K, int R, int S,
                                  int pad_h, int pad_w, int accumulate, cudaStream_t stream);

/* Fused normalisation kernels. Tensors are [G][P][C] fp32 with C % 4 == 0: G = images and
 * P = H*W for InstanceNorm; G = 1 and P = N*H*W for BatchNorm-style per-channel statistics.
 *
 * tgb_moments: out[g][c][0..1] += (sum x, sum x^2). (Reference torchgpipe/batchnorm.py:51-57
 *   computes these with two separate full-tensor reductions.)
 * tgb_in_prepare: from the accumulated moments `stats` [G][C][2] and an optional per-(g,c)
 *   Dropout2d scale `drop_scale` (0 or 1/(1-p); NULL = no dropout) computes
 *   mean_coef[g][c] = (mean, ms / sqrt(ms^2 * var + eps)) - InstanceNorm applied to the
 *   channel-dropped tensor, folded into one multiply-add.
 * tgb_in_leaky_fwd: z = LeakyReLU(coef * (y - mean)).
 * tgb_in_leaky_bwd_sums: sums[g][c] += (sum dn, sum dn*u) with dn, u re-derived from (dz, z),
 *   so neither the conv output y nor the normalised tensor is kept for backward.
 * tgb_in_leaky_bwd: dy = coef * (dn - S1/P - u * S2/P).
 * Reference call sites: Dropout2d + InstanceNorm2d + LeakyReLU of
 * benchmarks/models/unet/__init__.py:42-48 (three ATen ops each way there). */
int tgb_moments(const float* x, float* out, int G, int P, int C, cudaStream_t stream);
int tgb_in_prepare(const float* stats, const float* drop_scale, float* mean_coef,
                   int G, int P, int C, float eps, cudaStream_t stream);
/* Same, with the Dropout2d mask made in the kernel: `uniform` [G][C] holds random numbers in
 * [0,1) (from the caller's generator, so checkpoint recomputation can reproduce them); channel
 * (g,c) is kept when uniform >= drop_p and scaled by 1/(1-drop_p). */
int tgb_in_prepare_dropout(const float* stats, const float* uniform, float drop_p,
                           float* mean_coef, int G, int P, int C, float eps,
                           cudaStream_t stream);
int tgb_in_leaky_fwd(const float* y, const float* mean_coef, float* z, int G, int P, int C,
                     float slope, cudaStream_t stream);
int tgb_in_leaky_bwd_sums(const float* dz, const float* z, float* out, int G, int P, int C,
                          float slope, cudaStream_t stream);
int tgb_in_leaky_bwd(const float* dz, const float* z, const float* mean_coef, const float* sums,
                     float* dy, int G, int P, int C, float slope, cudaStream_t stream);

/* Fused BatchNorm2d (training) + optional ReLU on [P = N*H*W][C] fp32, C % 4 == 0.
 * tgb_bn_prepare: from per-channel (sum, sum of squares) `stats` [C][2] - taken from the
 *   convolution epilogue (stats_mode 2) or tgb_moments - computes coef[c] = (a, b) with
 *   bn(y) = y*a + b and saved[c] = (mean, rstd); when running_mean/var are given it also folds
 *   the batch statistics into them exactly like nn.BatchNorm2d (unbiased running variance).
 *   gamma/beta may be NULL (no affine).
 * tgb_bn_act_fwd: z = act(y*a + b), act = ReLU if relu != 0.
 * tgb_bn_act_bwd: zeroed `sums` [C][2] receives (sum dn, sum dn*xhat) = (dbeta, dgamma), then
 *   dy = a * (dn - S1/P - xhat*S2/P); the ReLU mask and xhat are re-derived from y.
 * Reference call sites: nn.BatchNorm2d (+ nn.ReLU) in benchmarks/models/resnet and amoebanet,
 * F.batch_norm in torchgpipe/batchnorm.py:112-121 (cuDNN / ATen there). */
int tgb_bn_prepare(const float* stats, const float* gamma, const float* beta, float* coef,
                   float* saved, float* running_mean, float* running_var, int P, int C,
                   float eps, float momentum, cudaStream_t stream);
/* tgb_bn_prepare that also keeps nn.BatchNorm2d's num_batches_tracked (int64, on the device):
 * the counter is incremented by the kernel, and momentum < 0 selects the cumulative average
 * 1 / (count after the increment) - no separate counter kernel, no host read of the counter. */
int tgb_bn_prepare_tracked(const float* stats, const float* gamma, const float* beta,
                           float* coef, float* saved, float* running_mean, float* running_var,
                           long long* num_batches_tracked, int P, int C, float eps,
                           float momentum, cudaStream_t stream);
int tgb_bn_act_fwd(const float* y, const float* coef, float* z, int P, int C, int relu,
                   cudaStream_t stream);
int tgb_bn_act_bwd(const float* dz, const float* y, const float* coef, const float* saved,
                   float* sums, float* dy, int P, int C, int relu, cudaStream_t stream);

/* NHWC fp32 data movement of the U-Net encoder / decoder (C, C1, C2 multiples of 4).
 * tgb_maxpool2x2_fwd/bwd: MaxPool2d(2, stride 2); backward re-derives the arg-max from the saved
 *   input x (first maximum in row-major window order, like ATen), no index tensor.
 * tgb_upsample2x_cat_fwd: out[N,Hs,Ws,C1+C2] = concat(pad(nearest_x2(x[N,Hx,Wx,C1])), skip) in
 *   one pass (reference: nn.Upsample + F.pad + torch.cat, benchmarks/models/unet/__init__.py:
 *   27-39,131). _bwd writes dskip (contiguous) and dx (2x2 window sums) in one pass over dout.
 * SURVEY K7 / K9 / K10 (ATen kernels in the reference). */
int tgb_maxpool2x2_fwd(const float* x, float* y, int N, int H, int W, int C, cudaStream_t stream);
int tgb_maxpool2x2_bwd(const float* x, const float* dy, float* dx, int N, int H, int W, int C,
                       cudaStream_t stream);
int tgb_upsample2x_cat_fwd(const float* x, const float* skip, float* out, int N, int Hx, int Wx,
                           int C1, int Hs, int Ws, int C2, cudaStream_t stream);
int tgb_upsample2x_cat_bwd(const float* dout, float* dx, float* dskip, int N, int Hx, int Wx,
                           int C1, int Hs, int Ws, int C2, cudaStream_t stream);

/* What tgb_conv2d_fprop_tf32 would do for a problem, without launching anything (works without a
 * GPU: 148 SMs are assumed then). out[16] = BN, BW, BH, BNB (output tile: BN channels x a
 * BW x BH x BNB pixel box), tiles_w, tiles_h, tiles_n, tiles_k, splits of the reduction,
 * iterations per split, haloed-tile variant (0/1), filter bank resident in shared memory (0/1),
 * pipeline stages, reduction iterations per tile, Ho, Wo. stats_mode as for the forward call
 * (0 when no statistics are wanted). */
int tgb_conv2d_fprop_plan(int N, int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                          int stats_mode, int* out);

/* Weights of the data-gradient pass: wt[c][R-1-r][S-1-s][k] = w[k][r][s][c] (tap flip + channel
 * transpose, KRSC memory on both sides) in one coalesced pass; the dgrad of a stride-1
 * convolution is then tgb_conv2d_fprop_tf32(dy, wt). */
int tgb_dgrad_weight(const float* w, float* wt, int K, int C, int R, int S, cudaStream_t stream);

/* Strided convolutions (strides 1 or 2 per axis, any integer padding; negative = crop).
 *   y[n,ho,wo,k] = sum x[n, sh*ho+r-pad_h, sw*wo+s-pad_w, c] * w[k,r,s,c]
 * Ho, Wo = 0: output extent from the usual arithmetic; > 0: compute exactly that extent (padding
 * that differs between the near and the far edge; what lies outside x reads as zero).
 * y_pixel_stride (elements): K, or larger when y is a channel slice of a wider NHWC tensor - a
 * concatenation written in place. The epilogue statistics work as for tgb_conv2d_fprop_tf32.
 * Reference call sites: benchmarks/models/resnet/__init__.py:41 (1x1/2 projection),
 * resnet/bottleneck.py:22,28 (3x3/2), amoebanet/__init__.py:28 and operations.py:71-84 (1x7, 7x1,
 * 3x3, 1x1 with stride), operations.py:31-32 (FactorizedReduce: two offset 1x1/2, the second one
 * is pad = -1 with the extent of the first) - cuDNN in the reference. */
int tgb_conv2d_fprop_strided_tf32(const float* x, const float* w, float* y, float* stats,
                                  int N, int H, int W, int C, int K, int R, int S,
                                  int pad_h, int pad_w, int stride_h, int stride_w,
                                  int Ho, int Wo, long long y_pixel_stride, int stats_mode,
                                  cudaStream_t stream);

/* Data gradient of the convolution above: one stride-1 launch of the forward kernel per parity
 * class of input pixels (h mod sh, w mod sw), each reading only the filter taps that reach its
 * class and storing through a strided view of dx.
 * tgb_dgrad_weight_strided prepares the per-class filter banks (tap-flipped, channel-transposed,
 * class-major in one buffer of tgb_dgrad_weight_strided_elems floats); redo it when w changes.
 * dy_pixel_stride: K, or larger when dy is a channel slice of a wider NHWC gradient.
 * keep_untouched != 0: input pixels no tap reads keep their contents (default: zero-filled) -
 * lets two convolutions over disjoint pixel classes of one input share one dx.
 * Ho_given / Wo_given: as Ho / Wo above (0 = derive). */
long long tgb_dgrad_weight_strided_elems(int K, int C, int R, int S, int pad_h, int pad_w,
                                         int stride_h, int stride_w);
int tgb_dgrad_weight_strided(const float* w, float* out, int K, int C, int R, int S,
                             int pad_h, int pad_w, int stride_h, int stride_w,
                             cudaStream_t stream);
int tgb_conv2d_dgrad_tf32(const float* dy, const float* wt_classes, float* dx,
                          int N, int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                          int stride_h, int stride_w, int Ho_given, int Wo_given,
                          long long dy_pixel_stride, int keep_untouched, cudaStream_t stream);

/* Weight gradient of the strided convolution (per-tap kernel; arguments as
 * tgb_conv2d_wgrad_tf32 plus strides, optional fixed output extent and dy pixel stride). */
long long tgb_conv2d_wgrad_strided_scratch_bytes(int N, int H, int W, int C, int K, int R, int S,
                                                 int pad_h, int pad_w, int stride_h,
                                                 int stride_w, int Ho, int Wo,
                                                 long long dy_pixel_stride);
int tgb_conv2d_wgrad_strided_tf32(const float* dy, const float* x, float* dw, float* scratch,
                                  long long scratch_bytes, int N, int H, int W, int C, int K,
                                  int R, int S, int pad_h, int pad_w, int stride_h, int stride_w,
                                  int Ho, int Wo, long long dy_pixel_stride, int accumulate,
                                  cudaStream_t stream);

/* Input stem: convolution of an image with 1-4 channels and filter rows of at most 8 taps.
 * tgb_stem_pack copies src (any layout: element strides s_n, s_c, s_h, s_w) into
 * dst [N][Hp][Wp][4], Hp = H + 2*pad_h, Wp = W + 2*pad_w - S + 8, zero in the halo and in the
 * missing channels; relu != 0 stores max(x, 0) (AmoebaNet's stem). In that buffer every run of 8
 * pixels is one 128-byte K block, which an overlapping-window tensor map (pixel stride 16 bytes,
 * innermost extent 128 bytes) feeds to the ordinary tcgen05 kernels: an R x S filter becomes an
 * R x 1 filter over 32 "channels" 4*s + c, wv [K][R][32] (zero where s >= S or c >= C).
 * Forward: y [N][Ho][Wo][K] (+ statistics as above). Weight gradient: dwv [K][R][32], entries of
 * taps / channels that do not exist are meaningless.
 * Reference call sites: resnet/__init__.py:55 (7x7/2), amoebanet/__init__.py:54 (3x3/2),
 * unet/__init__.py:44 first cell (3x3) - cuDNN in the reference. */
int tgb_stem_pack(const float* src, float* dst, int N, int C, int H, int W,
                  long long s_n, long long s_c, long long s_h, long long s_w,
                  int pad_h, int pad_w, int Hp, int Wp, int relu, cudaStream_t stream);
int tgb_conv2d_stem_fprop_tf32(const float* xp, const float* wv, float* y, float* stats,
                               int N, int Hp, int Wp, int K, int R, int stride_h, int stride_w,
                               int stats_mode, cudaStream_t stream);
long long tgb_conv2d_stem_wgrad_scratch_bytes(int N, int Hp, int Wp, int K, int R,
                                              int stride_h, int stride_w);
int tgb_conv2d_stem_wgrad_tf32(const float* dy, const float* xp, float* dwv, float* scratch,
                               long long scratch_bytes, int N, int Hp, int Wp, int K, int R,
                               int stride_h, int stride_w, cudaStream_t stream);

/* 1x1 convolution to 1-4 output channels (C % 4 == 0), plain fp32: y[p][k] = x[p][:] . w[k][:]
 * + bias[k] over `pixels` = N*H*W NHWC pixels; backward writes dx (may be NULL), and overwrites
 * dw [K][C] and db [K] (either may be NULL; x may be NULL when dw is). A streaming kernel: one
 * pass over x forward, one pass over x and one write of dx backward.
 * Reference call site: unet/__init__.py:138 (64 -> 1 head; cuDNN in the reference). */
int tgb_head_fwd(const float* x, const float* w, const float* bias, float* y,
                 long long pixels, int C, int K, cudaStream_t stream);
int tgb_head_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db,
                 long long pixels, int C, int K, cudaStream_t stream);

/* NHWC fp32 pooling family and residual joins (C multiple of 4; n multiple of 4).
 * tgb_pool2d_fwd/bwd: k x k window, stride s, zero padding p (2p <= k), floor output size.
 *   mode 0 max, 1 average over the valid (non-padding) pixels (count_include_pad=False),
 *   2 average over k*k. Backward gathers over the covering windows (deterministic, no index
 *   tensor; max routes to the first maximum in row-major window order like ATen) and needs x only
 *   for mode 0. Reference call sites: nn.MaxPool2d(3,2,1) resnet/__init__.py:58, AvgPool2d /
 *   MaxPool2d of amoebanet/operations.py:30-62.
 * tgb_global_avgpool_fwd/bwd: y[N,C] = mean over HW of x[N,HW,C] (AdaptiveAvgPool2d(1)).
 * tgb_add_act_fwd: z = relu?(a + b); tgb_relu_mask: d = dz where z > 0 else 0 (the gradient of
 *   both summands; resnet/bottleneck.py:70-76 + the following nn.ReLU).
 * tgb_bn_add_act_fwd: z = relu?(y * a[c] + b[c] + res) with coef from tgb_bn_prepare;
 * tgb_bn_add_act_bwd: sums[C,2] (zero on entry) = (sum dn, sum dn*xhat), dn = dz masked by z > 0;
 *   dy = a * (dn - S1/P - xhat*S2/P); dres = dn. (bn3 -> residual -> relu3 as one kernel.) */
int tgb_pool2d_fwd(const float* x, float* y, int N, int H, int W, int C, int k, int s, int p,
                   int mode, cudaStream_t stream);
int tgb_pool2d_bwd(const float* x, const float* dy, float* dx, int N, int H, int W, int C, int k,
                   int s, int p, int mode, cudaStream_t stream);
int tgb_global_avgpool_fwd(const float* x, float* y, int N, int HW, int C, cudaStream_t stream);
int tgb_global_avgpool_bwd(const float* dy, float* dx, int N, int HW, int C, cudaStream_t stream);
int tgb_add_act_fwd(const float* a, const float* b, float* z, long long n, int relu,
                    cudaStream_t stream);
int tgb_relu_mask(const float* dz, const float* z, float* d, long long n, cudaStream_t stream);
int tgb_bn_add_act_fwd(const float* y, const float* coef, const float* res, float* z, int P,
                       int C, int relu, cudaStream_t stream);
int tgb_bn_add_act_bwd(const float* dz, const float* y, const float* z, const float* coef,
                       const float* saved, float* sums, float* dy, float* dres, int P, int C,
                       int relu, cudaStream_t stream);

/* Fused gradient all-reduce across data-parallel replicas: out = scale * sum over replicas, in
 * place in every replica's fp32 bucket, optionally also as a local bf16 copy. One kernel per
 * rank: barrier, reduce-scatter (P2P loads from the peers' buckets, or multimem.ld_reduce through
 * the NVSwitch when `multicast` is given), scale, all-gather (peer loads / multimem.st), barrier.
 *   bufs, pads : host arrays of `world` device pointers - every replica's bucket and signal pad
 *                as mapped on the calling GPU (symmetric memory); pads are zero-initialised
 *                uint32 arrays of tgb_allreduce_pad_bytes(max_blocks) bytes used only by this call
 *   n          : fp32 elements, multiple of 4*world
 *   launch_index : same on every rank, +1 per call on these buffers (derives barrier tickets)
 *   error_flag : device int set to 1 if a peer did not arrive within timeout_ms
 * All replicas must launch the call; it must not be wrapped with other kernels that wait on a
 * peer. Wire bytes per direction per GPU: P2P 2N(R-1)/R, multicast N(R+1)/R.
 * Reference: none - the reference has no data parallelism; NCCL all_reduce on the
 * same bucket is the baseline. */
int tgb_allreduce_scale_f32(void* const* bufs, void* const* pads, void* multicast,
                            void* cast_out_bf16, unsigned long long n, float scale,
                            int rank, int world, unsigned int launch_index,
                            int timeout_ms, int* error_flag, int max_blocks,
                            cudaStream_t stream);
unsigned long long tgb_allreduce_pad_bytes(int max_blocks);

#ifdef __cplusplus
}
#endif
#endif  /* TGB200_H_ */
