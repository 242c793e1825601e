// Weight gradient of a 2-D convolution for sm_100a (tcgen05, TF32 -> FP32 in TMEM).
//
//   dW[k,r,s,c] = sum_{n,ho,wo} dY[n,ho,wo,k] * X[n, sh*ho+r-pad_h, sw*wo+s-pad_w, c]
//
// Strides work as in conv_fprop.cu: a tap's X box is a stride-1 box inside the view of X that
// holds the pixels of the tap's parity class (one tensor map per class).
//
// GEMM view per filter tap: M = K (output channels), N = C (input channels), and the reduction
// runs over all N*Ho*Wo output pixels. In NHWC memory the reduction index is the *outer*
// dimension of both operands, so both are fed to the tensor core as MN-major (transposed)
// operands straight from the TMA boxes. For 32-bit MN-major operands tcgen05 accepts exactly one
// shared-memory layout, SWIZZLE_128B_BASE32B: 128-byte rows (32 channels of one pixel) whose four
// 32-byte atoms are XOR-ed with (pixel % 4); TMA writes that layout with
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, so a box of P pixels x 32 channels lands ready to use and
// one MMA (K = 8 pixels) consumes two 4-pixel swizzle atoms of every channel chunk.
// The X box is shifted by the tap offset and zero-filled outside the image by TMA (= padding).
//
// Work item = (k-tile of 128, c-tile of BN, tap, split of the pixel range); persistent CTAs.
// Small-channel layers have millions of pixels and few (k,c,tap) tiles, so the pixel range is
// split across CTAs; each split writes an fp32 partial into scratch and a second kernel adds
// the partials in a fixed order (deterministic, no atomics). Large-channel layers use 1 split
// and store dW directly.
//
// Reference call site: autograd of every nn.Conv2d (cuDNN in the reference).
#include "sm100_ptx.cuh"
#include "tensormap.h"
#include "tgb_internal.h"
#include "conv_wgrad_common.h"
#include <cstdlib>

namespace tgb {

constexpr int kWgThreads = 256;
constexpr int kWgM = 128;
constexpr int kWgChunkBytes = 4096;      // 32 pixels x 128 B: one channel chunk of one stage
constexpr int kWgMaxPix = 32;

__host__ __device__ inline int wg_floor_div(int a, int b) {
  const int q = a / b;
  return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q;
}

struct WgradGeom {
  int N, H, W, C, K;
  int R, S, pad_h, pad_w;
  int stride_h, stride_w;
  int Ho, Wo;
  int BW, BH, BNB, P;                    // pixel box, P = BW*BH*BNB (multiple of 8, <= 32)
  int tiles_w, tiles_h, tiles_n;         // pixel blocks
  int k_tiles, c_tiles, splits;
  int blocks_per_split;
  int desc_lbo, desc_sbo, desc_layout;   // shared-memory descriptor fields (bytes, bytes, type)
  int dbg_shift, dbg_ksteps, dbg_baseoff; // bring-up probes (0 = off), see plan_wgrad
};

template <int BN, int STAGES>
struct WgSmem {
  static constexpr int kABytes = 4 * kWgChunkBytes;
  static constexpr int kBBytes = (BN / 32) * kWgChunkBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarOffset = STAGES * kStageBytes;
  static constexpr int kTotal = kBarOffset + (2 * STAGES + 4) * 8 + 16;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_tf32_kernel(const __grid_constant__ CUtensorMap tmap_dy,
                       const __grid_constant__ CUtensorMap tmap_x,
                       const __grid_constant__ CUtensorMap tmap_x1,
                       const __grid_constant__ CUtensorMap tmap_x2,
                       const __grid_constant__ CUtensorMap tmap_x3, const WgradGeom g,
                       float* __restrict__ out) {
  using L = WgSmem<BN, STAGES>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int taps = g.R * g.S;
  const int num_items = g.k_tiles * g.c_tiles * taps * g.splits;
  const int total_blocks = g.tiles_w * g.tiles_h * g.tiles_n;
  const uint32_t stage_tx = static_cast<uint32_t>((4 + BN / 32) * g.P * 128);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // item -> (kt, ct, tap, split); kt fastest so concurrent CTAs share the X boxes in L2
  auto decode = [&](int item, int& kt, int& ct, int& tap, int& split) {
    kt = item % g.k_tiles;
    int rest = item / g.k_tiles;
    ct = rest % g.c_tiles;
    rest /= g.c_tiles;
    tap = rest % taps;
    split = rest / taps;
  };

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        int kt, ct, tap, split;
        decode(item, kt, ct, tap, split);
        const int r = tap / g.S, s = tap - r * g.S;
        // tap offset = stride * q + p: a stride-1 box, q pixels in, of the parity-p view of X
        const int ar = r - g.pad_h, as = s - g.pad_w;
        const int qr = wg_floor_div(ar, g.stride_h), qs = wg_floor_div(as, g.stride_w);
        const int view = (ar - qr * g.stride_h) * g.stride_w + (as - qs * g.stride_w);
        const CUtensorMap* mx = view == 0 ? &tmap_x : view == 1 ? &tmap_x1
                              : view == 2 ? &tmap_x2 : &tmap_x3;
        const int b_begin = split * g.blocks_per_split;
        const int b_end = min(total_blocks, b_begin + g.blocks_per_split);
        for (int blk = b_begin; blk < b_end; ++blk) {
          const int tw = blk % g.tiles_w;
          const int rest = blk / g.tiles_w;
          const int th = rest % g.tiles_h;
          const int tn = rest / g.tiles_h;
          const int w0 = tw * g.BW, h0 = th * g.BH, n0 = tn * g.BNB;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = smem + stage * L::kStageBytes;
          uint8_t* b_dst = a_dst + L::kABytes;
          mbar_expect_tx(&full_bar[stage], stage_tx);
#pragma unroll
          for (int ch = 0; ch < 4; ++ch)
            tma_load_4d(a_dst + ch * kWgChunkBytes, &tmap_dy, &full_bar[stage],
                        kt * kWgM + ch * 32, w0, h0, n0);
#pragma unroll
          for (int ch = 0; ch < BN / 32; ++ch)
            tma_load_4d(b_dst + ch * kWgChunkBytes, mx, &full_bar[stage],
                        ct * BN + ch * 32, w0 + qs, h0 + qr, n0);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(/*tf32*/ 2, kWgM, BN, /*a MN-major*/ 1, /*b*/ 1);
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const int k_steps = g.dbg_ksteps > 0 ? g.dbg_ksteps : g.P / 8;
      const uint64_t dbg_add = static_cast<uint64_t>(g.dbg_shift * 8) |
                               (static_cast<uint64_t>(g.dbg_baseoff & 7) << 49);
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        int kt, ct, tap, split;
        decode(item, kt, ct, tap, split);
        const int b_begin = split * g.blocks_per_split;
        const int b_end = min(total_blocks, b_begin + g.blocks_per_split);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        bool first = true;
        for (int blk = b_begin; blk < b_end; ++blk) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t b_addr = a_addr + L::kABytes;
          // MN-major, 128B swizzle with 32-byte atoms: LBO = stride between 32-channel chunks,
          // SBO = stride between 4-pixel swizzle atoms (512 B)
          const uint64_t a_desc = make_smem_desc(a_addr, g.desc_lbo, g.desc_sbo, g.desc_layout);
          const uint64_t b_desc = make_smem_desc(b_addr, g.desc_lbo, g.desc_sbo, g.desc_layout);
          for (int k = 0; k < k_steps; ++k) {
            // next 8 pixels = next 1024-byte swizzle atom: +64 in the (addr >> 4) field
            mma_tf32_ss(d_tmem, a_desc + 64 * k + dbg_add, b_desc + 64 * k + dbg_add, idesc,
                        first ? 0u : 1u);
            first = false;
          }
          mma_commit(&empty_bar[stage]);
          if (blk == b_end - 1) mma_commit(&tmem_full[acc]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (b_begin >= b_end) {
          // empty split (cannot happen with the host's split choice, but never deadlock)
          mma_commit(&tmem_full[acc]);
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int row = q * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    const size_t tap_stride = static_cast<size_t>(g.C);
    const size_t k_stride = static_cast<size_t>(taps) * g.C;
    const size_t split_stride = static_cast<size_t>(g.K) * k_stride;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      int kt, ct, tap, split;
      decode(item, kt, ct, tap, split);
      const int b_begin = split * g.blocks_per_split;
      const bool empty = b_begin >= min(total_blocks, b_begin + g.blocks_per_split);
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int kglob = kt * kWgM + row;
      float* dst_row = out + split * split_stride + static_cast<size_t>(kglob) * k_stride +
                       tap * tap_stride + static_cast<size_t>(ct) * BN;
#pragma unroll 1
      for (int j = 0; j < BN / 32; ++j) {
        uint32_t v[32];
        const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * BN + j * 32) +
                               (static_cast<uint32_t>(q * 32) << 16);
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        if (kglob < g.K) {
          const int c_base = ct * BN + j * 32;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            if (c_base + c * 4 < g.C) {
              float4 o = empty ? make_float4(0, 0, 0, 0)
                               : make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]),
                                             __uint_as_float(v[4 * c + 2]),
                                             __uint_as_float(v[4 * c + 3]));
              *reinterpret_cast<float4*>(dst_row + j * 32 + c * 4) = o;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

struct TapSplits {
  int n[84];          // valid partial copies per filter tap (up to 9x9 filters)
};

// out[i] = sum_s partial[s][i] (fixed order), optionally accumulated onto out (beta = 1).
// Element i belongs to tap (i / c4) % taps; only the first ts.n[tap] partial copies are valid.
__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out,
                                    size_t n4, TapSplits ts, int taps, int c4,
                                    size_t split_stride4, int accumulate) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float4 acc = accumulate ? reinterpret_cast<const float4*>(out)[i] : make_float4(0, 0, 0, 0);
    const int splits = ts.n[(i / c4) % taps];
    for (int s = 0; s < splits; ++s) {
      const float4 v = reinterpret_cast<const float4*>(partial)[s * split_stride4 + i];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    reinterpret_cast<float4*>(out)[i] = acc;
  }
}

static void choose_wgrad_box(int N, int Ho, int Wo, int& BW, int& BH, int& BNB) {
  double best = -1.0;
  BW = 8; BH = 1; BNB = 1;
  for (int bw = 1; bw <= Wo && bw <= kWgMaxPix; ++bw)
    for (int bh = 1; bh <= Ho && bw * bh <= kWgMaxPix; ++bh) {
      const int nb_hi = (bw == Wo && bh == Ho) ? kWgMaxPix / (bw * bh) : 1;
      for (int nb = 1; nb <= nb_hi && nb <= N; ++nb) {
        const int p = bw * bh * nb;
        if (p % 8) continue;
        const double blocks = static_cast<double>((Wo + bw - 1) / bw) * ((Ho + bh - 1) / bh) *
                              ((N + nb - 1) / nb);
        const double useful = static_cast<double>(N) * Ho * Wo / (blocks * p);
        // fewer, fuller stages win: utilisation first, then box size, then width
        const double score = useful * (0.75 + 0.25 * p / kWgMaxPix) + 1e-6 * bw;
        if (score > best) {
          best = score;
          BW = bw; BH = bh; BNB = nb;
        }
      }
    }
}

static int wgrad_tma_swizzle() {
  if (const char* env = getenv("TGB_WGRAD_DESC")) {
    int a, b, c, d;
    if (sscanf(env, "%d,%d,%d,%d", &a, &b, &c, &d) == 4) return d;
  }
  return 2;
}

template <int BN, int STAGES>
static int launch_wgrad(const CUtensorMap& mdy, const CUtensorMap* mx, const WgradGeom& g,
                        float* out, int grid, cudaStream_t stream) {
  using L = WgSmem<BN, STAGES>;
  constexpr int smem_bytes = L::kTotal + 1024;
  // the opt-in shared-memory size is a per-device attribute of the kernel: one process may
  // drive several GPUs (single-process GPipe), so remember it per device
  static bool configured_on[64] = {false};
  int dev_ = 0;
  cudaGetDevice(&dev_);
  if (dev_ < 0 || dev_ >= 64) dev_ = 0;
  bool& configured = configured_on[dev_];
  auto kern = conv_wgrad_tf32_kernel<BN, STAGES>;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         smem_bytes);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_wgrad): %s",
                                           cudaGetErrorString(e));
    configured = true;
  }
  kern<<<grid, kWgThreads, smem_bytes, stream>>>(mdy, mx[0], mx[1], mx[2], mx[3], g, out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("conv_wgrad launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

// Fills the geometry and the number of splits for a problem; shared by the planning query and
// the launcher so both always agree.
static int plan_wgrad(int N, int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                      int stride_h, int stride_w, WgradGeom& g, int& BN, int Ho_given = 0,
                      int Wo_given = 0) {
  if (C % 4 != 0 || K % 4 != 0) return set_error("conv_wgrad: need C %% 4 == 0 and K %% 4 == 0");
  if (stride_h < 1 || stride_w < 1 || stride_h * stride_w > 4)
    return set_error("conv_wgrad: unsupported stride %dx%d", stride_h, stride_w);
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K;
  g.R = R; g.S = S; g.pad_h = pad_h; g.pad_w = pad_w;
  g.stride_h = stride_h; g.stride_w = stride_w;
  g.Ho = Ho_given > 0 ? Ho_given : (H + 2 * pad_h - R) / stride_h + 1;
  g.Wo = Wo_given > 0 ? Wo_given : (W + 2 * pad_w - S) / stride_w + 1;
  if (g.Ho <= 0 || g.Wo <= 0 || (Ho_given <= 0 && H + 2 * pad_h < R) ||
      (Wo_given <= 0 && W + 2 * pad_w < S))
    return set_error("conv_wgrad: empty output");
  choose_wgrad_box(N, g.Ho, g.Wo, g.BW, g.BH, g.BNB);
  g.P = g.BW * g.BH * g.BNB;
  g.tiles_w = (g.Wo + g.BW - 1) / g.BW;
  g.tiles_h = (g.Ho + g.BH - 1) / g.BH;
  g.tiles_n = (N + g.BNB - 1) / g.BNB;
  const int c32 = (C + 31) / 32 * 32;
  BN = c32 >= 256 ? 256 : c32 >= 128 ? 128 : c32 >= 64 ? 64 : 32;
  g.k_tiles = (K + kWgM - 1) / kWgM;
  g.c_tiles = (C + BN - 1) / BN;
  const int items = g.k_tiles * g.c_tiles * R * S;
  const int total_blocks = g.tiles_w * g.tiles_h * g.tiles_n;
  const int sms = sm_count();
  int splits = 1;
  if (items < sms) {
    splits = (sms + items - 1) / items;
    // keep at least 8 pixel blocks per split so the prologue/epilogue stay amortised
    const int max_splits = total_blocks / 8 > 0 ? total_blocks / 8 : 1;
    if (splits > max_splits) splits = max_splits;
  }
  g.blocks_per_split = (total_blocks + splits - 1) / splits;
  g.splits = (total_blocks + g.blocks_per_split - 1) / g.blocks_per_split;
  g.desc_lbo = kWgChunkBytes;
  g.desc_sbo = 512;
  g.desc_layout = kLayoutSw128Base32;
  g.dbg_shift = g.dbg_ksteps = g.dbg_baseoff = 0;
  // Bring-up probe: TGB_WGRAD_DBG="shift_rows,k_steps,base_offset" starts both operands
  // `shift_rows` pixels into the box (does tcgen05 swizzle by absolute address?).
  if (const char* env = getenv("TGB_WGRAD_DBG")) {
    int a, b, c;
    if (sscanf(env, "%d,%d,%d", &a, &b, &c) == 3) {
      g.dbg_shift = a; g.dbg_ksteps = b; g.dbg_baseoff = c;
    }
  }
  // Bring-up aid: TGB_WGRAD_DESC="lbo,sbo,layout,tma_swizzle" overrides the descriptor fields.
  if (const char* env = getenv("TGB_WGRAD_DESC")) {
    int a, b, c, d;
    if (sscanf(env, "%d,%d,%d,%d", &a, &b, &c, &d) == 4) {
      g.desc_lbo = a; g.desc_sbo = b; g.desc_layout = c;
    }
  }
  return 0;
}

}  // namespace tgb

using namespace tgb;

// Which kernel serves a problem: 0 = per-tap (this file), 1 = tap-fused (conv_wgrad_fused.cu).
// TGB_WGRAD_KERNEL=pertap|fused overrides the choice (bring-up / A-B timing).
static int choose_wgrad_kernel(int N, int Ho, int Wo, int C, int S) {
  if (const char* env = getenv("TGB_WGRAD_KERNEL")) {
    if (env[0] == 'p') return 0;
    if (env[0] == 'f' && wgrad_fused_applicable(C, S, Wo)) return 1;
  }
  return wgrad_fused_preferred(N, Ho, Wo, C, S) ? 1 : 0;
}

namespace tgb {

// Explicit addressing (strides in elements), as ConvProblem in conv_fprop.cu:
//   X[n][h][w][c]   at x  + n*x_sn + h*x_sh + w*x_sw + c
//   dY[n][ho][wo][k] at dy + (n*Ho*Wo + ho*Wo + wo) * dy_sw + k   (dy_sw > K: a channel slice)
struct WgradProblem {
  const float* dy; const float* x; float* dw; float* scratch; long long scratch_bytes;
  int N, H, W, C, K, R, S, pad_h, pad_w, stride_h, stride_w;
  long long x_sw, x_sh, x_sn, dy_sw;
  int accumulate;
  int Ho = 0, Wo = 0;               // > 0: output extent fixed by the caller (asymmetric padding)
};

static bool wgrad_is_plain(const WgradProblem& p) {
  return p.stride_h == 1 && p.stride_w == 1 && p.x_sw == p.C &&
         p.x_sh == static_cast<long long>(p.W) * p.C &&
         p.x_sn == static_cast<long long>(p.H) * p.W * p.C && p.dy_sw == p.K;
}

static long long wgrad_scratch_need(int N, int H, int W, int C, int K, int R, int S, int pad_h,
                                    int pad_w, int stride_h, int stride_w, bool plain,
                                    int Ho_given = 0, int Wo_given = 0) {
  if (C % 4 != 0 || K % 4 != 0) {
    set_error("conv_wgrad: need C %% 4 == 0 and K %% 4 == 0");
    return -1;
  }
  if (stride_h < 1 || stride_w < 1) {
    set_error("conv_wgrad: strides must be positive");
    return -1;
  }
  int splits;
  const int Wo = Wo_given > 0 ? Wo_given : (W + 2 * pad_w - S) / stride_w + 1;
  const int Ho = Ho_given > 0 ? Ho_given : (H + 2 * pad_h - R) / stride_h + 1;
  int BN;
  plain = plain && Ho_given <= 0 && Wo_given <= 0;
  if (plain && choose_wgrad_kernel(N, Ho, Wo, C, S)) {
    WgradFusedGeom g;
    if (plan_wgrad_fused(N, H, W, C, K, R, S, pad_h, pad_w, g, BN)) return -1;
    splits = g.max_splits;
  } else {
    WgradGeom g;
    if (plan_wgrad(N, H, W, C, K, R, S, pad_h, pad_w, stride_h, stride_w, g, BN, Ho_given,
                   Wo_given))
      return -1;
    splits = g.splits;
  }
  if (splits <= 1) return 0;
  return static_cast<long long>(splits) * K * R * S * C * 4;
}

static int run_wgrad(const WgradProblem& p, cudaStream_t stream) {
  const int N = p.N, H = p.H, W = p.W, C = p.C, K = p.K, R = p.R, S = p.S;
  if (C % 4 != 0 || K % 4 != 0)
    return set_error("conv_wgrad: need C %% 4 == 0 and K %% 4 == 0");
  if ((reinterpret_cast<uintptr_t>(p.x) | reinterpret_cast<uintptr_t>(p.dy) |
       reinterpret_cast<uintptr_t>(p.dw) | reinterpret_cast<uintptr_t>(p.scratch)) & 15)
    return set_error("conv_wgrad: tensors must be 16-byte aligned");
  if ((p.x_sw | p.x_sh | p.x_sn | p.dy_sw) & 3)
    return set_error("conv_wgrad: strides must be multiples of 4 elements (16 bytes)");
  if (p.stride_h < 1 || p.stride_w < 1) return set_error("conv_wgrad: strides must be positive");
  const int Ho = p.Ho > 0 ? p.Ho : (H + 2 * p.pad_h - R) / p.stride_h + 1;
  const int Wo = p.Wo > 0 ? p.Wo : (W + 2 * p.pad_w - S) / p.stride_w + 1;
  if (Ho <= 0 || Wo <= 0) return set_error("conv_wgrad: empty output");
  const bool fused = wgrad_is_plain(p) && p.Ho <= 0 && p.Wo <= 0 &&
                     choose_wgrad_kernel(N, Ho, Wo, C, S) != 0;
  WgradGeom g;
  WgradFusedGeom gf;
  int BN, splits;
  if (fused) {
    if (plan_wgrad_fused(N, H, W, C, K, R, S, p.pad_h, p.pad_w, gf, BN)) return 1;
    splits = gf.max_splits;
  } else {
    if (plan_wgrad(N, H, W, C, K, R, S, p.pad_h, p.pad_w, p.stride_h, p.stride_w, g, BN, p.Ho,
                   p.Wo))
      return 1;
    splits = g.splits;
  }
  const size_t dw_elems = static_cast<size_t>(K) * R * S * C;
  const bool direct = splits <= 1 && !p.accumulate;
  if (!direct) {
    const long long need = static_cast<long long>(splits) * dw_elems * 4;
    if (p.scratch == nullptr || p.scratch_bytes < need)
      return set_error("conv_wgrad: scratch too small (need %lld bytes, got %lld)", need,
                       p.scratch_bytes);
  }
  float* out = direct ? p.dw : p.scratch;
  const int sms = sm_count();
  int rc;
  if (fused) {
    rc = run_wgrad_fused(p.dy, p.x, out, gf, BN, stream);
  } else {
    CUtensorMap mdy, mx[4];
    {
      uint64_t dims[4] = {(uint64_t)K, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)N};
      uint64_t strides[3] = {(uint64_t)p.dy_sw * 4, (uint64_t)g.Wo * p.dy_sw * 4,
                             (uint64_t)g.Ho * g.Wo * p.dy_sw * 4};
      uint32_t box[4] = {32, (uint32_t)g.BW, (uint32_t)g.BH, (uint32_t)g.BNB};
      if (make_tensor_map(&mdy, p.dy, 4, 4, dims, strides, box, wgrad_tma_swizzle()))
        return set_error("conv_wgrad: tensor map for dY failed");
    }
    const int views = p.stride_h * p.stride_w;
    for (int v = 0; v < 4; ++v) {
      const int ph = v < views ? v / p.stride_w : 0, pw = v < views ? v % p.stride_w : 0;
      const long long Hv = (H - ph + p.stride_h - 1) / p.stride_h;
      const long long Wv = (W - pw + p.stride_w - 1) / p.stride_w;
      if (Hv <= 0 || Wv <= 0) {
        mx[v] = mx[0];
        continue;
      }
      uint64_t dims[4] = {(uint64_t)C, (uint64_t)Wv, (uint64_t)Hv, (uint64_t)N};
      uint64_t strides[3] = {(uint64_t)(p.x_sw * p.stride_w) * 4,
                             (uint64_t)(p.x_sh * p.stride_h) * 4, (uint64_t)p.x_sn * 4};
      uint32_t box[4] = {32, (uint32_t)g.BW, (uint32_t)g.BH, (uint32_t)g.BNB};
      if (make_tensor_map(&mx[v], p.x + ph * p.x_sh + pw * p.x_sw, 4, 4, dims, strides, box,
                          wgrad_tma_swizzle()))
        return set_error("conv_wgrad: tensor map for X (view %d) failed", v);
    }
    const int items = g.k_tiles * g.c_tiles * R * S * g.splits;
    const int grid = items < sms ? items : sms;
    switch (BN) {
      case 256: rc = launch_wgrad<256, 4>(mdy, mx, g, out, grid, stream); break;
      case 128: rc = launch_wgrad<128, 6>(mdy, mx, g, out, grid, stream); break;
      case 64:  rc = launch_wgrad<64, 8>(mdy, mx, g, out, grid, stream); break;
      default:  rc = launch_wgrad<32, 8>(mdy, mx, g, out, grid, stream); break;
    }
  }
  if (rc) return rc;
  if (!direct) {
    const size_t n4 = dw_elems / 4;
    size_t blocks = (n4 + 255) / 256;
    if (blocks > static_cast<size_t>(sms) * 8) blocks = static_cast<size_t>(sms) * 8;
    TapSplits ts;
    if (R * S > 84) return set_error("conv_wgrad: filters larger than 84 taps are unsupported");
    if (fused) {
      wgrad_fused_tap_splits(gf, ts.n);
    } else {
      for (int t = 0; t < R * S; ++t) ts.n[t] = splits;
    }
    wgrad_reduce_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(
        p.scratch, p.dw, n4, ts, R * S, C / 4, n4, p.accumulate);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("wgrad_reduce launch: %s", cudaGetErrorString(e));
    count_launch();
  }
  return 0;
}

}  // namespace tgb

extern "C" long long tgb_conv2d_wgrad_scratch_bytes(int N, int H, int W, int C, int K, int R,
                                                    int S, int pad_h, int pad_w) {
  return wgrad_scratch_need(N, H, W, C, K, R, S, pad_h, pad_w, 1, 1, true);
}

extern "C" long long tgb_conv2d_wgrad_strided_scratch_bytes(int N, int H, int W, int C, int K,
                                                            int R, int S, int pad_h, int pad_w,
                                                            int stride_h, int stride_w,
                                                            int Ho, int Wo,
                                                            long long dy_pixel_stride) {
  const bool plain = stride_h == 1 && stride_w == 1 &&
                     (dy_pixel_stride <= 0 || dy_pixel_stride == K);
  return wgrad_scratch_need(N, H, W, C, K, R, S, pad_h, pad_w, stride_h, stride_w, plain, Ho,
                            Wo);
}

extern "C" int tgb_conv2d_wgrad_tf32(const float* dy, const float* x, float* dw, float* scratch,
                                     long long scratch_bytes, int N, int H, int W, int C, int K,
                                     int R, int S, int pad_h, int pad_w, int accumulate,
                                     cudaStream_t stream) {
  WgradProblem p{dy, x, dw, scratch, scratch_bytes, N, H, W, C, K, R, S, pad_h, pad_w, 1, 1,
                 C, static_cast<long long>(W) * C, static_cast<long long>(H) * W * C, K,
                 accumulate};
  return run_wgrad(p, stream);
}

extern "C" int tgb_conv2d_wgrad_strided_tf32(const float* dy, const float* x, float* dw,
                                             float* scratch, long long scratch_bytes, int N,
                                             int H, int W, int C, int K, int R, int S, int pad_h,
                                             int pad_w, int stride_h, int stride_w, int Ho,
                                             int Wo, long long dy_pixel_stride, int accumulate,
                                             cudaStream_t stream) {
  WgradProblem p{dy, x, dw, scratch, scratch_bytes, N, H, W, C, K, R, S, pad_h, pad_w, stride_h,
                 stride_w, C, static_cast<long long>(W) * C, static_cast<long long>(H) * W * C,
                 dy_pixel_stride > 0 ? dy_pixel_stride : K, accumulate};
  p.Ho = Ho;
  p.Wo = Wo;
  return run_wgrad(p, stream);
}

// Weight gradient of the packed small-channel stem (conv_small.cu): dwv [K][R][32] with
// dwv[k][r][4*s+c] = dW[k][c][r][s]; columns of taps / channels that do not exist receive the
// correlation with whatever pixels the window covers there - the caller ignores them.
extern "C" long long tgb_conv2d_stem_wgrad_scratch_bytes(int N, int Hp, int Wp, int K, int R,
                                                         int stride_h, int stride_w) {
  return wgrad_scratch_need(N, Hp, Wp - 7, 32, K, R, 1, 0, 0, stride_h, stride_w, false);
}

extern "C" int tgb_conv2d_stem_wgrad_tf32(const float* dy, const float* xp, float* dwv,
                                          float* scratch, long long scratch_bytes, int N, int Hp,
                                          int Wp, int K, int R, int stride_h, int stride_w,
                                          cudaStream_t stream) {
  if (Wp < 8) return set_error("stem_wgrad: packed rows must hold at least one 8-pixel window");
  WgradProblem p{dy, xp, dwv, scratch, scratch_bytes, N, Hp, Wp - 7, 32, K, R, 1, 0, 0, stride_h,
                 stride_w, 4, static_cast<long long>(Wp) * 4,
                 static_cast<long long>(Hp) * Wp * 4, K, 0};
  return run_wgrad(p, stream);
}

// ---------------------------------------------------------------------------------------------
// Weight gradient for callers that are not allowed to allocate device memory.
//
// Some users of this library run where a device allocation is off the table: a serving or
// real-time loop whose memory was sized and reserved up front, a plain C/C++ host with no
// caching allocator behind it (so every cudaMalloc/cudaFree would synchronise the device), or a
// step that is being recorded for replay. tgb_conv2d_wgrad_tf32 makes such callers size, obtain
// and pass a scratch buffer themselves. This entry point needs none: the partial sums of the
// pixel splits go to a fixed buffer that is part of the library image.
//
// The buffer is the file-scope array below. In host code its name is only a handle; the device
// address is looked up with cudaGetSymbolAddress once per device (while that device is current)
// and that address is what the kernels get.
// ---------------------------------------------------------------------------------------------
namespace tgb {

constexpr long long kNoAllocScratchBytes = 32ll << 20;      // 32 MiB, see tgb200.h
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchBytes / 4];

// Device address of g_noalloc_scratch on the current device (resolved on first use per device).
static int noalloc_scratch_ptr(float** out) {
  static float* resolved[64] = {nullptr};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_error("wgrad_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return set_error("wgrad_noalloc: device ordinal %d out of range", dev);
  float* p = __atomic_load_n(&resolved[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* addr = nullptr;
    e = cudaGetSymbolAddress(&addr, g_noalloc_scratch);
    if (e != cudaSuccess || addr == nullptr)
      return set_error("wgrad_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                       cudaGetErrorString(e));
    p = static_cast<float*>(addr);
    __atomic_store_n(&resolved[dev], p, __ATOMIC_RELEASE);    // same value from any thread
  }
  *out = p;
  return 0;
}

}  // namespace tgb

extern "C" int tgb_conv2d_wgrad_tf32_noalloc(const float* dy, const float* x, float* dw, int N,
                                             int H, int W, int C, int K, int R, int S, int pad_h,
                                             int pad_w, int accumulate, cudaStream_t stream) {
  long long need = tgb_conv2d_wgrad_scratch_bytes(N, H, W, C, K, R, S, pad_h, pad_w);
  if (need < 0) return 1;                                     // message already set
  if (need == 0 && accumulate) need = static_cast<long long>(K) * R * S * C * 4;
  if (need > kNoAllocScratchBytes)
    return set_error("wgrad_noalloc: this problem needs %lld bytes of scratch but the built-in "
                     "buffer holds %lld; call tgb_conv2d_wgrad_tf32 with your own scratch instead",
                     need, kNoAllocScratchBytes);
  // The built-in buffer of the CURRENT device is used and the launch goes to the current device:
  // tensors that live on another GPU mean the caller forgot cudaSetDevice. Refuse that instead
  // of handing the kernels an address that belongs to a different GPU.
  {
    int dev = -1;
    cudaPointerAttributes attr;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaPointerGetAttributes(&attr, dw) == cudaSuccess &&
        attr.type == cudaMemoryTypeDevice && attr.device != dev)
      return set_error("wgrad_noalloc: dw lives on device %d but device %d is current; make the "
                       "device that owns the tensors (and the stream) current before the call",
                       attr.device, dev);
    cudaGetLastError();       // a failed attribute query must not stick
  }
  float* scratch = nullptr;
  if (noalloc_scratch_ptr(&scratch)) return 1;
  // Every partial the reduction reads is written by the kernel of this same call first, so
  // whatever an earlier call left in the buffer is never observed and nothing has to be cleared.
  return tgb_conv2d_wgrad_tf32(dy, x, dw, scratch, kNoAllocScratchBytes, N, H, W, C, K, R, S,
                               pad_h, pad_w, accumulate, stream);
}
