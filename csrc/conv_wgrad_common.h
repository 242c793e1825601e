// Shared between the per-tap (conv_wgrad.cu) and tap-fused (conv_wgrad_fused.cu) weight-gradient
// kernels: geometry of the fused kernel and the functions the C entry points dispatch to.
#pragma once
#include <cuda_runtime.h>

namespace tgb {

struct WgradFusedGeom {
  int N, H, W, C, K;
  int R, S, pad_h, pad_w;
  int Ho, Wo;
  int P, segs_per_line;          // pixels per block (multiple of 8, <= 32), blocks per line
  // A "slice" is one MMA per 8-pixel step. n_stack > 1: the slice covers filter rows
  // 0..n_stack-1 of one filter column s (their X segments sit at a regular stride in shared
  // memory, so one B descriptor spans them and N = n_stack * BN). n_stack == 1: slice = one tap.
  int n_stack, slices, slices_per_group, groups;
  int k_tiles, c_tiles;
  int max_splits;                // scratch holds max_splits partial copies of dW
  int group_splits[12];          // pixel-range splits of each group (heavier groups get more)
  int group_blocks[12];          // pixel blocks per split of each group
  int group_item_begin[13];      // prefix sums of k_tiles*c_tiles*group_splits
};

bool wgrad_fused_applicable(int C, int S, int Wo);
bool wgrad_fused_preferred(int N, int Ho, int Wo, int C, int S);
int plan_wgrad_fused(int N, int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                     WgradFusedGeom& g, int& BN);
void wgrad_fused_tap_splits(const WgradFusedGeom& g, int* tap_splits);
int run_wgrad_fused(const float* dy, const float* x, float* out, const WgradFusedGeom& g, int BN,
                    cudaStream_t stream);

}  // namespace tgb
