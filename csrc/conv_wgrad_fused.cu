// Tap-fused weight gradient for small-channel / large-image convolutions (sm_100a, tcgen05).
//
// The per-tap kernel in conv_wgrad.cu re-reads dY and X once per filter tap; with C <= 128 the
// tensor core then waits on L2 (measured 77-80 TFLOP/s on U-Net's 192x192 layers). This kernel
// loads each block ONCE per group of taps and keeps one TMEM accumulator per tap:
//
//   * a block is P (<= 32) consecutive pixels of one image line: dY[n, h, w0:w0+P, :]
//   * for every filter row r of the group it loads the X line segment
//     X[n, h+r-pad_h, w0-pad_w : w0-pad_w+P+S-1, :]  (P+S-1 pixels, zero-filled outside)
//   * the B operand of tap (r,s) is that segment starting s pixels in. Both operands are MN-major
//     (pixel = reduction index = 128-byte smem row, SWIZZLE_128B_BASE32B); tcgen05 applies the
//     swizzle to absolute shared-memory addresses, so a descriptor may start at any 128-byte row
//     of a TMA-written tile (verified on B200 by `selftest shift_probe`).
//
// Taps per group = floor(512 TMEM columns / BN): 9 for BN=32, 5 for BN=64, 3 (one filter row)
// for BN=128. Work item = (k-tile, c-tile, tap group, pixel split); partials of the splits go to
// scratch and are reduced in fixed order by wgrad_reduce_kernel (conv_wgrad.cu).
#include "sm100_ptx.cuh"
#include "tensormap.h"
#include "tgb_internal.h"
#include "conv_wgrad_common.h"

namespace tgb {

constexpr int kWfThreads = 256;
constexpr int kWfM = 128;
constexpr int kWfChunkA = 4096;        // 32 pixels x 128 B
constexpr int kWfChunkB = 5120;        // up to 40 pixels x 128 B (P + S - 1 <= 40), 1 KB aligned
constexpr int kWfMaxTaps = 16;

template <int BN>
struct WfCfg {
  static constexpr int kMaxRows = BN == 128 ? 1 : 3;      // filter rows resident per stage
  static constexpr int kABytes = 4 * kWfChunkA;
  static constexpr int kBBytes = kMaxRows * (BN / 32) * kWfChunkB;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (200 * 1024) / kStageBytes;
  static constexpr int kBarOffset = kStages * kStageBytes;
  static constexpr int kTotal = kBarOffset + (2 * kStages + 2) * 8 + 16;
};

template <int BN>
__global__ void __launch_bounds__(kWfThreads, 1)
conv_wgrad_fused_kernel(const __grid_constant__ CUtensorMap tmap_dy,
                        const __grid_constant__ CUtensorMap tmap_x, const WgradFusedGeom g,
                        float* __restrict__ out) {
  using L = WfCfg<BN>;
  constexpr int STAGES = L::kStages;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int taps = g.R * g.S;
  const int num_items = g.group_item_begin[g.groups];
  const int segs = g.segs_per_line;
  const int total_blocks = g.N * g.Ho * segs;
  const int halo = g.P + g.S - 1;
  const int n_mma = g.n_stack * BN;          // N of one MMA (columns of one slice in TMEM)
  const int per_group = g.k_tiles * g.c_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 4);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // item -> (group, kt, ct, split) and the pixel-block range of that split
  auto decode = [&](int item, int& grp, int& kt, int& ct, int& split, int& b_begin,
                    int& b_end) {
    grp = 0;
    while (item >= g.group_item_begin[grp + 1]) ++grp;
    int rest = item - g.group_item_begin[grp];
    kt = rest % g.k_tiles;
    rest /= g.k_tiles;
    ct = rest % g.c_tiles;
    split = rest / g.c_tiles;
    b_begin = split * g.group_blocks[grp];
    b_end = min(total_blocks, b_begin + g.group_blocks[grp]);
  };
  // slices [sl_begin, sl_end) of a group and the filter rows [r_begin, r_end] they read
  auto group_range = [&](int grp, int& sl_begin, int& sl_end, int& r_begin, int& r_end) {
    sl_begin = grp * g.slices_per_group;
    sl_end = min(g.slices, sl_begin + g.slices_per_group);
    if (g.n_stack > 1) {
      r_begin = 0;
      r_end = g.n_stack - 1;
    } else {
      r_begin = sl_begin / g.S;
      r_end = (sl_end - 1) / g.S;
    }
  };
  (void)per_group;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        int grp, kt, ct, split, b_begin, b_end, sl_begin, sl_end, r_begin, r_end;
        decode(item, grp, kt, ct, split, b_begin, b_end);
        group_range(grp, sl_begin, sl_end, r_begin, r_end);
        const int a_chunks = min(4, (g.K - kt * kWfM + 31) / 32);
        const int n_rows = r_end - r_begin + 1;
        const uint32_t tx = static_cast<uint32_t>(a_chunks * g.P * 128 +
                                                  n_rows * (BN / 32) * halo * 128);
        for (int blk = b_begin; blk < b_end; ++blk) {
          const int seg = blk % segs;
          const int rest = blk / segs;
          const int h = rest % g.Ho;
          const int n = rest / g.Ho;
          const int w0 = seg * g.P;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = smem + stage * L::kStageBytes;
          uint8_t* b_dst = a_dst + L::kABytes;
          mbar_expect_tx(&full_bar[stage], tx);
          for (int ch = 0; ch < a_chunks; ++ch)
            tma_load_4d(a_dst + ch * kWfChunkA, &tmap_dy, &full_bar[stage],
                        kt * kWfM + ch * 32, w0, h, n);
          for (int rl = 0; rl < n_rows; ++rl)
#pragma unroll
            for (int ch = 0; ch < BN / 32; ++ch)
              tma_load_4d(b_dst + (rl * (BN / 32) + ch) * kWfChunkB, &tmap_x, &full_bar[stage],
                          ct * BN + ch * 32, w0 - g.pad_w, h + r_begin + rl - g.pad_h, n);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc(/*tf32*/ 2, kWfM, n_mma, /*a MN-major*/ 1, /*b*/ 1);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_phase = 0;
      const int k_steps = g.P / 8;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        int grp, kt, ct, split, b_begin, b_end, sl_begin, sl_end, r_begin, r_end;
        decode(item, grp, kt, ct, split, b_begin, b_end);
        group_range(grp, sl_begin, sl_end, r_begin, r_end);
        mbar_wait(tmem_empty, acc_phase ^ 1);
        tc_fence_after();
        for (int blk = b_begin; blk < b_end; ++blk) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t b_addr = a_addr + L::kABytes;
          const uint64_t a_desc = make_smem_desc(a_addr, kWfChunkA, 512, kLayoutSw128Base32);
          const uint32_t accumulate = blk > b_begin ? 1u : 0u;
          for (int sl = sl_begin; sl < sl_end; ++sl) {
            // stacked: slice = filter column s, rows 0..n_stack-1 (regions at stride kWfChunkB
            // continue regularly from channel chunks into the next filter row);
            // flat: slice = tap (r,s)
            int r, sft;
            if (g.n_stack > 1) {
              r = 0;
              sft = sl;
            } else {
              r = sl / g.S;
              sft = sl - r * g.S;
            }
            const uint32_t region = b_addr + (r - r_begin) * (BN / 32) * kWfChunkB + sft * 128;
            const uint64_t b_desc = make_smem_desc(region, kWfChunkB, 512, kLayoutSw128Base32);
            const uint32_t d_tmem = tmem_base + static_cast<uint32_t>((sl - sl_begin) * n_mma);
            for (int k = 0; k < k_steps; ++k)
              mma_tf32_ss(d_tmem, a_desc + 64 * k, b_desc + 64 * k, idesc,
                          (accumulate | (k > 0)) ? 1u : 0u);
          }
          mma_commit(&empty_bar[stage]);
          if (blk == b_end - 1) mma_commit(tmem_full);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (b_begin >= b_end) mma_commit(tmem_full);
        acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int row = q * 32 + lane;
    uint32_t acc_phase = 0;
    const size_t k_stride = static_cast<size_t>(taps) * g.C;
    const size_t split_stride = static_cast<size_t>(g.K) * k_stride;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      int grp, kt, ct, split, b_begin, b_end, sl_begin, sl_end, r_begin, r_end;
      decode(item, grp, kt, ct, split, b_begin, b_end);
      group_range(grp, sl_begin, sl_end, r_begin, r_end);
      const bool empty = b_begin >= b_end;
      mbar_wait(tmem_full, acc_phase);
      tc_fence_after();
      const int kglob = kt * kWfM + row;
      float* item_base = out + split * split_stride + static_cast<size_t>(kglob) * k_stride +
                         static_cast<size_t>(ct) * BN;
      for (int sl = sl_begin; sl < sl_end; ++sl) {
#pragma unroll 1
        for (int j = 0; j < n_mma / 32; ++j) {
          // which tap do these 32 columns belong to?
          int tap, c_in;
          if (g.n_stack > 1) {
            const int r = (j * 32) / BN;
            c_in = (j * 32) % BN;
            tap = r * g.S + sl;
          } else {
            c_in = j * 32;
            tap = sl;
          }
          uint32_t v[32];
          const uint32_t taddr = tmem_base +
                                 static_cast<uint32_t>((sl - sl_begin) * n_mma + j * 32) +
                                 (static_cast<uint32_t>(q * 32) << 16);
          tmem_ld_32x32(taddr, v);
          tmem_ld_wait();
          if (kglob < g.K) {
            float* dst = item_base + static_cast<size_t>(tap) * g.C + c_in;
            const int c_base = ct * BN + c_in;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              if (c_base + c * 4 < g.C) {
                float4 o = empty ? make_float4(0, 0, 0, 0)
                                 : make_float4(__uint_as_float(v[4 * c]),
                                               __uint_as_float(v[4 * c + 1]),
                                               __uint_as_float(v[4 * c + 2]),
                                               __uint_as_float(v[4 * c + 3]));
                *reinterpret_cast<float4*>(dst + c * 4) = o;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty);
      acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int BN>
static int launch_fused(const CUtensorMap& mdy, const CUtensorMap& mx, const WgradFusedGeom& g,
                        float* out, int grid, cudaStream_t stream) {
  using L = WfCfg<BN>;
  constexpr int smem_bytes = L::kTotal + 1024;
  // the opt-in shared-memory size is a per-device attribute of the kernel: one process may
  // drive several GPUs (single-process GPipe), so remember it per device
  static bool configured_on[64] = {false};
  int dev_ = 0;
  cudaGetDevice(&dev_);
  if (dev_ < 0 || dev_ >= 64) dev_ = 0;
  bool& configured = configured_on[dev_];
  auto kern = conv_wgrad_fused_kernel<BN>;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         smem_bytes);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_wgrad_fused): %s",
                                           cudaGetErrorString(e));
    configured = true;
  }
  kern<<<grid, kWfThreads, smem_bytes, stream>>>(mdy, mx, g, out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("conv_wgrad_fused launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

bool wgrad_fused_applicable(int C, int S, int Wo) {
  const int c32 = (C + 31) / 32 * 32;
  return c32 <= 128 && S <= 9 && Wo >= 8;
}

// Measured on B200 (selftest unet): the fused kernel wins whenever C <= 64, and for C <= 128
// once the image is large enough that the per-tap kernel re-reads more than L2 can hide.
bool wgrad_fused_preferred(int N, int Ho, int Wo, int C, int S) {
  if (!wgrad_fused_applicable(C, S, Wo)) return false;
  if (C <= 64) return true;
  // 65..128 channels: large images (96x96 and up) at any micro-batch - 128->128 @96x96 at
  // micro-batch 16: 0.109 ms fused vs 0.157 ms per tap, at 32: 0.183 vs 0.300 - while 48x48
  // layers stay on the per-tap kernel (128->256 @48x48: 0.087 vs 0.126 ms) unless the batch
  // makes the per-tap kernel's re-reads outgrow the L2
  return static_cast<long long>(Ho) * Wo >= 6000 ||
         static_cast<long long>(N) * Ho * Wo >= 200000;
}

int plan_wgrad_fused(int N, int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                     WgradFusedGeom& g, int& BN) {
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K;
  g.R = R; g.S = S; g.pad_h = pad_h; g.pad_w = pad_w;
  g.Ho = H + 2 * pad_h - R + 1;
  g.Wo = W + 2 * pad_w - S + 1;
  const int c32 = (C + 31) / 32 * 32;
  BN = c32 > 64 ? 128 : c32 > 32 ? 64 : 32;
  // widest multiple-of-8 segment (<= 32) that wastes the fewest pixels of a line
  int best_p = 8;
  double best = -1;
  for (int p = 8; p <= 32; p += 8) {
    const int segs = (g.Wo + p - 1) / p;
    const double useful = static_cast<double>(g.Wo) / (segs * p);
    const double score = useful * (0.75 + 0.25 * p / 32.0);
    if (score > best) { best = score; best_p = p; }
  }
  g.P = best_p;
  g.segs_per_line = (g.Wo + g.P - 1) / g.P;
  const int taps = R * S;
  const int max_rows = BN == 128 ? 1 : 3;
  // Stack the filter rows of one filter column into one wide MMA when they fit: N = R*BN <= 256
  // keeps the tensor core ahead of the shared-memory operand reads (an N=64 MMA is smem-bound).
  const bool stack = R > 1 && R <= max_rows && R * BN <= 256 && getenv("TGB_WGRAD_NOSTACK") == nullptr;
  if (stack) {
    g.n_stack = R;
    g.slices = S;
    g.slices_per_group = 512 / (R * BN);
  } else {
    g.n_stack = 1;
    g.slices = taps;
    g.slices_per_group = 512 / BN > kWfMaxTaps ? kWfMaxTaps : 512 / BN;
    // a group may touch at most max_rows filter rows (that many X segments fit in a stage)
    while (g.slices_per_group > 1) {
      bool ok = true;
      for (int t0 = 0; t0 < taps; t0 += g.slices_per_group) {
        const int t1 = (t0 + g.slices_per_group < taps ? t0 + g.slices_per_group : taps) - 1;
        if (t1 / S - t0 / S + 1 > max_rows) ok = false;
      }
      if (ok) break;
      --g.slices_per_group;
    }
  }
  if (g.slices_per_group > g.slices) g.slices_per_group = g.slices;
  g.groups = (g.slices + g.slices_per_group - 1) / g.slices_per_group;
  if (g.groups > 12) return set_error("conv_wgrad_fused: too many tap groups (%d)", g.groups);
  g.k_tiles = (K + kWfM - 1) / kWfM;
  g.c_tiles = (C + BN - 1) / BN;
  const int total_blocks = N * g.Ho * g.segs_per_line;
  const int sms = sm_count();
  // Give every group a number of pixel splits proportional to its number of slices so that all
  // work items take about the same time, and the grid is about one item per SM.
  const int per_group = g.k_tiles * g.c_tiles;
  g.max_splits = 1;
  g.group_item_begin[0] = 0;
  for (int gi = 0; gi < g.groups; ++gi) {
    const int sl0 = gi * g.slices_per_group;
    const int sl1 = sl0 + g.slices_per_group < g.slices ? sl0 + g.slices_per_group : g.slices;
    const double share = static_cast<double>(sl1 - sl0) / g.slices;
    int splits = static_cast<int>(share * sms / per_group + 0.5);
    const int max_splits = total_blocks / 16 > 0 ? total_blocks / 16 : 1;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    g.group_blocks[gi] = (total_blocks + splits - 1) / splits;
    g.group_splits[gi] = (total_blocks + g.group_blocks[gi] - 1) / g.group_blocks[gi];
    if (g.group_splits[gi] > g.max_splits) g.max_splits = g.group_splits[gi];
    g.group_item_begin[gi + 1] = g.group_item_begin[gi] + per_group * g.group_splits[gi];
  }
  return 0;
}

// Number of valid partial copies for each tap (taps of a light group have fewer splits).
void wgrad_fused_tap_splits(const WgradFusedGeom& g, int* tap_splits) {
  const int taps = g.R * g.S;
  for (int t = 0; t < taps; ++t) {
    const int slice = g.n_stack > 1 ? t % g.S : t;
    tap_splits[t] = g.group_splits[slice / g.slices_per_group];
  }
}

int run_wgrad_fused(const float* dy, const float* x, float* out, const WgradFusedGeom& g, int BN,
                    cudaStream_t stream) {
  CUtensorMap mdy, mx;
  {
    uint64_t dims[4] = {(uint64_t)g.K, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)g.N};
    uint64_t strides[3] = {(uint64_t)g.K * 4, (uint64_t)g.Wo * g.K * 4,
                           (uint64_t)g.Ho * g.Wo * g.K * 4};
    uint32_t box[4] = {32, (uint32_t)g.P, 1, 1};
    if (make_tensor_map(&mdy, dy, 4, 4, dims, strides, box, 2))
      return set_error("conv_wgrad_fused: tensor map for dY failed");
  }
  {
    uint64_t dims[4] = {(uint64_t)g.C, (uint64_t)g.W, (uint64_t)g.H, (uint64_t)g.N};
    uint64_t strides[3] = {(uint64_t)g.C * 4, (uint64_t)g.W * g.C * 4,
                           (uint64_t)g.H * g.W * g.C * 4};
    uint32_t box[4] = {32, (uint32_t)(g.P + g.S - 1), 1, 1};
    if (make_tensor_map(&mx, x, 4, 4, dims, strides, box, 2))
      return set_error("conv_wgrad_fused: tensor map for X failed");
  }
  const int items = g.group_item_begin[g.groups];
  const int sms = sm_count();
  const int grid = items < sms ? items : sms;
  switch (BN) {
    case 128: return launch_fused<128>(mdy, mx, g, out, grid, stream);
    case 64:  return launch_fused<64>(mdy, mx, g, out, grid, stream);
    default:  return launch_fused<32>(mdy, mx, g, out, grid, stream);
  }
}

}  // namespace tgb
