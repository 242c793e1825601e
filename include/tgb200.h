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
 * Requires C % 4 == 0, K % 32 == 0, 16-byte aligned pointers.
 * stats (optional, may be NULL): FP32 accumulators that the epilogue ADDS per-channel sum and
 * sum of squares of y into; the caller zeroes them first.
 *   stats_mode 1: [N][K][2] per (image, channel)  - InstanceNorm statistics
 *   stats_mode 2: [K][2]    per channel           - BatchNorm statistics
 * Also computes the data gradient of such a convolution when called with dy as x, the
 * tap-flipped / channel-transposed weights and pad' = R-1-pad.
 * Reference call sites: nn.Conv2d forward/backward-data in benchmarks/models/{unet,resnet,
 * amoebanet} (the reference dispatches these to cuDNN). */
int tgb_conv2d_fprop_tf32(const float* x, const float* w, float* y, float* stats,
                          int N, int H, int W, int C, int K, int R, int S,
                          int pad_h, int pad_w, int stats_mode, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif  /* TGB200_H_ */
