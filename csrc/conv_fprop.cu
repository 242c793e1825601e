// Implicit-GEMM 2-D convolution forward for sm_100a: tcgen05.mma (TF32 in, FP32 accumulate in
// TMEM), operands staged by TMA, NHWC activations, KRSC weights.
//
//   Y[n,ho,wo,k] = sum_{r,s,c} X[n, sh*ho+r-pad_h, sw*wo+s-pad_w, c] * Wt[k,r,s,c]
//
// Strides: a tap offset a = r - pad_h splits into a = sh*q + p (0 <= p < sh). All output rows of
// a tile then read input rows sh*(ho+q) + p, i.e. rows ho+q of the sub-image "rows congruent to p
// mod sh" - a plain strided view of X with its own tensor map (base moved by p rows, row stride
// times sh). There is one such view per (p_h, p_w) pair, at most four for stride 2, and a tap's
// TMA box in its view looks exactly like a stride-1 box: same shared-memory layout, same MMAs,
// zero fill outside the view = zero padding outside the image. Negative padding (a crop) falls
// out of the same arithmetic.
//
// GEMM view: M = N*Ho*Wo output pixels, N = K output channels, reduction = R*S*C.
// One CTA tile = 128 output pixels arranged as a BW x BH x BNB box of (w, h, n) x BN channels.
// For every filter tap (r,s) and every 32-channel block the A operand is ONE 4-D TMA box of the
// input shifted by the tap offset; pixels outside the image are zero-filled by TMA, which is
// exactly the convolution's zero padding - no im2col buffer, no index arithmetic on the SM.
// The same kernel computes the data gradient of a stride-1 convolution when it is given dY, the
// tap-flipped / channel-transposed weights and pad' = R-1-pad.
//
// Warp roles (256 threads, 1 CTA per SM, persistent over tiles):
//   warp 0   TMA producer            (one elected lane)
//   warp 1   tcgen05.mma issuer      (one elected lane)
//   warp 2   TMEM allocator
//   warps 4-7 epilogue: TMEM -> registers -> swizzled smem -> TMA store, plus optional fused
//            per-channel (BatchNorm) or per-(image, channel) (InstanceNorm) sum / sum-of-squares
//            so the normalisation that follows never has to re-read Y for its statistics.
// Pipelines: smem full/empty ring (TMA <-> MMA) and a 2-deep TMEM accumulator ring
// (MMA <-> epilogue) so the epilogue of tile t overlaps the main loop of tile t+1.
//
// Covers reference call sites: every nn.Conv2d of benchmarks/models/{unet,resnet,amoebanet}
// with stride 1 (forward: K1; backward-data: K2 in SURVEY 2.3b).
#include "sm100_ptx.cuh"
#include "tensormap.h"
#include "tgb_internal.h"
#include "../include/tgb200.h"
#include <cstdlib>

namespace tgb {

constexpr int kBlockM = 128;        // UMMA M
constexpr int kBlockKBytes = 128;   // one swizzle row: 32 fp32
constexpr int kBlockK = 32;
constexpr int kUmmaK = 8;           // tf32: 32 bytes per MMA
constexpr int kEpiChunk = 32;       // columns per epilogue chunk (= one 128B row of Y)
constexpr int kThreads = 256;

__host__ __device__ inline int floor_div(int a, int b) {
  const int q = a / b;
  return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q;
}

struct ConvGeom {
  int N, H, W, C, K;
  int R, S, pad_h, pad_w;
  int stride_h, stride_w;
  int Ho, Wo;
  int BW, BH, BNB;                  // pixel box; rows_valid = BW*BH*BNB <= 128
  int tiles_w, tiles_h, tiles_n, tiles_k;
  int c_blocks;
  int stats_mode;                   // 0 none, 1 per (n,k), 2 per k
  int splits, iters_per_split;      // split of the reduction (taps x channel blocks) over CTAs
};

// HALO variant (filters up to 3x3, BN <= 64): the tile is 8 pixels wide x 16 rows and a stage
// holds, for one 32-channel block, the input tile WITH its halo ((8+S-1) x (16+R-1) pixels, one
// TMA box) plus the weights of all R*S taps. The A operand of tap (r,s) is the same shared-memory
// tile read through a descriptor that starts (r*pitch + s) rows in and steps pitch rows per
// 8-row group (SBO = pitch * 128 B): 8 consecutive pixels of an image row are 8 consecutive smem
// rows, the next image row starts one pitch later. tcgen05 applies the 128B swizzle to absolute
// shared-memory addresses, so a descriptor may start at any 128-byte row of a TMA-written tile.
// The input is then read once per tile instead of once per tap (6.25x less A traffic for 3x3).
constexpr int kHaloW = 8, kHaloH = 16, kHaloMaxTaps = 9;
constexpr int kHaloABytes = ((kHaloW + 2) * (kHaloH + 2) * kBlockKBytes + 1023) / 1024 * 1024;

// RESW (with HALO, all output channels in one tile): the whole filter bank (<= 144 KB: 64->64,
// 128->32, 32->32 at 3x3) is loaded into shared memory once per CTA and stays there for all of
// its tiles; the stages then carry only the 23 KB haloed input tile.
constexpr int kResidentWBytes = 147456;

template <int BN, int STAGES, bool HALO = false, bool RESW = false>
struct SmemLayout {
  static constexpr int kABytes = HALO ? kHaloABytes : kBlockM * kBlockKBytes;   // 16 KB / 23 KB
  static constexpr int kBBytes = BN * kBlockKBytes;                             // one tap
  static constexpr int kStageBytes = kABytes + (RESW ? 0 : (HALO ? kHaloMaxTaps : 1) * kBBytes);
  static constexpr int kWOffset = STAGES * kStageBytes;        // resident filter bank (RESW)
  static constexpr int kEpiBytes = kBlockM * kBlockKBytes;     // 16 KB per staging buffer
  static constexpr int kEpiOffset = kWOffset + (RESW ? kResidentWBytes : 0);
  static constexpr int kStatOffset = kEpiOffset + 2 * kEpiBytes;   // 1 KB partials + 512 B table
  static constexpr int kBarOffset = kStatOffset + 1536;
  static constexpr int kTotal = kBarOffset + (2 * STAGES + 5) * 8 + 16;
};

template <int BN, int STAGES, bool HALO = false, bool RESW = false>
__global__ void __launch_bounds__(kThreads, 1)
conv_fprop_tf32_kernel(const __grid_constant__ CUtensorMap tmap_x,
                       const __grid_constant__ CUtensorMap tmap_x1,
                       const __grid_constant__ CUtensorMap tmap_x2,
                       const __grid_constant__ CUtensorMap tmap_x3,
                       const __grid_constant__ CUtensorMap tmap_w,
                       const __grid_constant__ CUtensorMap tmap_y,
                       const ConvGeom g, float* __restrict__ stats) {
  using L = SmemLayout<BN, STAGES, HALO, RESW>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem is only guaranteed 16B aligned: round up to the 1024B the swizzle needs
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));

  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* w_bar = tmem_empty + 2;                 // resident filter bank has landed (RESW)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rows_valid = g.BW * g.BH * g.BNB;
  const int num_tiles = g.tiles_w * g.tiles_h * g.tiles_n * g.tiles_k * g.splits;
  const int taps = g.R * g.S;
  // HALO: one iteration per channel block (all taps inside); otherwise per (tap, channel block)
  const int k_iters = HALO ? g.c_blocks : taps * g.c_blocks;
  const int halo_pitch = g.BW + g.S - 1;                        // pixels per haloed tile row
  const uint32_t stage_tx =
      HALO ? static_cast<uint32_t>(halo_pitch * (g.BH + g.R - 1) * kBlockKBytes +
                                   (RESW ? 0 : taps * L::kBBytes))
           : static_cast<uint32_t>(rows_valid * kBlockKBytes + L::kBBytes);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    if (g.stride_h * g.stride_w > 1) {
      tma_prefetch_desc(&tmap_x1);
      if (g.stride_h * g.stride_w > 2) {
        tma_prefetch_desc(&tmap_x2);
        tma_prefetch_desc(&tmap_x3);
      }
    }
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_y);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4);       // one arrival per epilogue warp
    }
    mbar_init(w_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 2 * BN);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // A tile id also carries the reduction split: [it_begin, it_end) of the (tap, channel block)
  // iterations. splits == 1 is the normal case.
  int it_begin = 0, it_end = k_iters;
  auto decode_tile = [&](int tile, int& kt, int& tw, int& th, int& tn) {
    const int ks = tile % g.splits;
    tile /= g.splits;
    it_begin = ks * g.iters_per_split;
    it_end = min(k_iters, it_begin + g.iters_per_split);
    kt = tile % g.tiles_k;
    int rest = tile / g.tiles_k;
    tw = rest % g.tiles_w;
    rest /= g.tiles_w;
    th = rest % g.tiles_h;
    tn = rest / g.tiles_h;
  };

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      if (RESW && static_cast<int>(blockIdx.x) < num_tiles) {
        // the filter bank, [channel block][tap][BN rows x 128 B], once per CTA
        mbar_expect_tx(w_bar, static_cast<uint32_t>(g.c_blocks * taps * L::kBBytes));
        for (int cb = 0; cb < g.c_blocks; ++cb)
          for (int t = 0; t < taps; ++t)
            tma_load_4d(smem + L::kWOffset + (cb * taps + t) * L::kBBytes, &tmap_w, w_bar,
                        cb * kBlockK, t, 0, 0);
      }
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int kt, tw, th, tn;
        decode_tile(tile, kt, tw, th, tn);
        const int w0 = tw * g.BW - g.pad_w;
        const int h0 = th * g.BH - g.pad_h;
        const int n0 = tn * g.BNB;
        const int k0 = kt * BN;
        if (HALO) {
          // one haloed input box and the weights of every tap per channel block
          for (int cb = it_begin; cb < it_end; ++cb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* a_dst = smem + stage * L::kStageBytes;
            uint8_t* b_dst = a_dst + L::kABytes;
            mbar_expect_tx(&full_bar[stage], stage_tx);
            tma_load_4d(a_dst, &tmap_x, &full_bar[stage], cb * kBlockK, w0, h0, n0);
            if (!RESW)
              for (int t = 0; t < taps; ++t)
                tma_load_4d(b_dst + t * L::kBBytes, &tmap_w, &full_bar[stage], cb * kBlockK, t,
                            k0, 0);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
          continue;
        }
        int tap = it_begin / g.c_blocks;
        int cb = it_begin - tap * g.c_blocks;
        // Per-tap quantities are recomputed only when the tap changes (the producer thread has
        // one MMA-iteration of time per stage; integer divisions here would bound narrow tiles).
        // tap offset = stride * q + p: box origin moves by q inside the parity-p view of X.
        const bool unit_stride = g.stride_h == 1 && g.stride_w == 1;
        int qr = 0, qs = 0;
        const CUtensorMap* mx = &tmap_x;
        auto enter_tap = [&](int t) {
          const int r = t / g.S;
          const int s = t - r * g.S;
          const int ar = r - g.pad_h, as = s - g.pad_w;
          if (unit_stride) {
            qr = ar;
            qs = as;
          } else {
            qr = floor_div(ar, g.stride_h);
            qs = floor_div(as, g.stride_w);
            const int view = (ar - qr * g.stride_h) * g.stride_w + (as - qs * g.stride_w);
            mx = view == 0 ? &tmap_x : view == 1 ? &tmap_x1 : view == 2 ? &tmap_x2 : &tmap_x3;
          }
        };
        enter_tap(tap);
        for (int it = it_begin; it < it_end; ++it) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = smem + stage * L::kStageBytes;
          uint8_t* b_dst = a_dst + L::kABytes;
          mbar_expect_tx(&full_bar[stage], stage_tx);
          tma_load_4d(a_dst, mx, &full_bar[stage], cb * kBlockK, tw * g.BW + qs,
                      th * g.BH + qr, n0);
          tma_load_4d(b_dst, &tmap_w, &full_bar[stage], cb * kBlockK, tap, k0, 0);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
          if (++cb == g.c_blocks) {
            cb = 0;
            enter_tap(++tap);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc(/*tf32*/ 2, kBlockM, BN, 0, 0);
    // One thread waits, issues and commits: tcgen05.commit tracks the MMAs of the issuing thread.
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      if (RESW && static_cast<int>(blockIdx.x) < num_tiles) {
        mbar_wait(w_bar, 0);
        tc_fence_after();
      }
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int kt_, tw_, th_, tn_;
        decode_tile(tile, kt_, tw_, th_, tn_);            // sets [it_begin, it_end)
        const int n_it = it_end - it_begin;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int it = 0; it < n_it; ++it) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t b_addr = a_addr + L::kABytes;
          if (HALO) {
            // tap (r,s) reads the haloed tile (r*pitch + s) rows in; 8-row groups (= 8 pixels of
            // one image row) are one haloed row apart
            const uint32_t sbo = static_cast<uint32_t>(halo_pitch * kBlockKBytes);
            for (int t = 0; t < taps; ++t) {
              const int r = t / g.S, sft = t - r * g.S;
              const uint64_t a_desc = make_smem_desc_sw128(
                  a_addr + static_cast<uint32_t>((r * halo_pitch + sft) * kBlockKBytes), 16, sbo);
              const uint32_t bt_addr =
                  RESW ? smem_u32(smem + L::kWOffset) + (it * taps + t) * L::kBBytes
                       : b_addr + t * L::kBBytes;
              const uint64_t b_desc = make_smem_desc_sw128(bt_addr, 16, 1024);
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k)
                mma_tf32_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc,
                            (it > 0 || t > 0 || k > 0) ? 1u : 0u);
            }
          } else {
            const uint64_t a_desc = make_smem_desc_sw128(a_addr, 16, 1024);
            const uint64_t b_desc = make_smem_desc_sw128(b_addr, 16, 1024);
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k) {
              // advance 32 bytes inside the 128B swizzle row: +2 in the (addr >> 4) field
              mma_tf32_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc,
                          (it > 0 || k > 0) ? 1u : 0u);
            }
          }
          mma_commit(&empty_bar[stage]);                  // smem slot free once MMAs retire
          if (it == n_it - 1) mma_commit(&tmem_full[acc]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue
    const int q = warp - 4;                       // TMEM lane quarter of this warp
    const int row = q * 32 + lane;                // tile row owned by this thread
    const int epi_tid = threadIdx.x - 128;
    uint8_t* epi = smem + L::kEpiOffset;
    float2* stat_part = reinterpret_cast<float2*>(smem + L::kStatOffset);     // [4 warps][32]
    int* row_img = reinterpret_cast<int*>(smem + L::kStatOffset + 1024);      // [128]
    int acc = 0;
    uint32_t acc_phase = 0;
    int chunk_counter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int kt, tw, th, tn;
      decode_tile(tile, kt, tw, th, tn);
      bool tile_uniform = false;
      if (g.stats_mode != 0) {
        tile_uniform = g.BNB == 1 && (tw + 1) * g.BW <= g.Wo && (th + 1) * g.BH <= g.Ho;
        if (!tile_uniform) {
          // image index of every tile row (or -1 when the row is outside the output); the
          // table is published by the first bar.sync of the chunk loop. The barrier keeps a
          // fast warp from overwriting the table a slower warp still reads for the last tile.
          asm volatile("bar.sync 2, 128;" ::: "memory");
          const int ww = row % g.BW;
          const int hh = (row / g.BW) % g.BH;
          const int nn = row / (g.BW * g.BH);
          const bool inside = row < rows_valid && tw * g.BW + ww < g.Wo &&
                              th * g.BH + hh < g.Ho && tn * g.BNB + nn < g.N;
          row_img[row] = inside ? (g.stats_mode == 1 ? tn * g.BNB + nn : 0) : -1;
        }
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int j = 0; j < BN / kEpiChunk; ++j, ++chunk_counter) {
        uint32_t v[32];
        const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * BN + j * kEpiChunk) +
                               (static_cast<uint32_t>(q * 32) << 16);
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        if (j == BN / kEpiChunk - 1) {
          // accumulator fully read: hand the TMEM stage back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        uint8_t* buf = epi + (chunk_counter & 1) * L::kEpiBytes;
        // the TMA store issued two chunks ago read from this buffer: wait for it
        if (epi_tid == 0) tma_store_wait_read<1>();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        {
          // row-major [128 rows][128 B] with the 128B swizzle TMA expects:
          // 16-byte chunk c of row r lives at chunk (c ^ (r & 7))
          uint8_t* rowp = buf + row * kBlockKBytes;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int pc = c ^ (row & 7);
            *reinterpret_cast<uint4*>(rowp + pc * 16) =
                make_uint4(v[4 * c + 0], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
          }
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (epi_tid == 0) {
          if (g.splits > 1)       // partial sum of this reduction split: add into the zeroed Y
            tma_reduce_add_4d(&tmap_y, buf, kt * BN + j * kEpiChunk, tw * g.BW, th * g.BH,
                              tn * g.BNB);
          else
            tma_store_4d(&tmap_y, buf, kt * BN + j * kEpiChunk, tw * g.BW, th * g.BH,
                         tn * g.BNB);
          tma_store_commit();
        }
        if (g.stats_mode != 0) {
          // Column sums over this warp's 32 rows, read back from the staging buffer:
          // lane = column, so each LDS hits 32 distinct words of one 128B row (no conflicts).
          const int col = lane;
          const int kglob = kt * BN + j * kEpiChunk + col;
          const int sw_hi = col >> 2, sw_lo = (col & 3) * 4;
          if (tile_uniform) {
            // whole tile inside one image and inside the output: no per-row bookkeeping
            float s1 = 0.f, s2 = 0.f;
            const int r_end = min(32, rows_valid - q * 32);
#pragma unroll 8
            for (int rr = 0; rr < r_end; ++rr) {
              const int r2 = q * 32 + rr;
              const float x = *reinterpret_cast<const float*>(
                  buf + r2 * kBlockKBytes + ((sw_hi ^ (r2 & 7)) << 4) + sw_lo);
              s1 += x;
              s2 = fmaf(x, x, s2);
            }
            stat_part[q * 32 + col] = make_float2(s1, s2);
            asm volatile("bar.sync 2, 128;" ::: "memory");
            if (q == 0 && kglob < g.K) {
              const float2 p0 = stat_part[col], p1 = stat_part[32 + col];
              const float2 p2 = stat_part[64 + col], p3 = stat_part[96 + col];
              const int img = (g.stats_mode == 1) ? tn : 0;
              float* dst = stats + (static_cast<size_t>(img) * g.K + kglob) * 2;
              atomicAdd(dst, (p0.x + p1.x) + (p2.x + p3.x));
              atomicAdd(dst + 1, (p0.y + p1.y) + (p2.y + p3.y));
            }
          } else {
            float s1 = 0.f, s2 = 0.f;
            int cur_img = -1;
            for (int rr = 0; rr < 32; ++rr) {
              const int r2 = q * 32 + rr;
              if (r2 >= rows_valid) break;
              const int img = row_img[r2];          // -1: row lies outside the output
              if (img < 0) continue;
              if (img != cur_img) {
                if (cur_img >= 0 && kglob < g.K) {
                  float* dst = stats + (static_cast<size_t>(cur_img) * g.K + kglob) * 2;
                  atomicAdd(dst, s1);
                  atomicAdd(dst + 1, s2);
                }
                cur_img = img;
                s1 = 0.f;
                s2 = 0.f;
              }
              const float x = *reinterpret_cast<const float*>(
                  buf + r2 * kBlockKBytes + ((sw_hi ^ (r2 & 7)) << 4) + sw_lo);
              s1 += x;
              s2 = fmaf(x, x, s2);
            }
            if (cur_img >= 0 && kglob < g.K) {
              float* dst = stats + (static_cast<size_t>(cur_img) * g.K + kglob) * 2;
              atomicAdd(dst, s1);
              atomicAdd(dst + 1, s2);
            }
          }
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (epi_tid == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ---------------------------------------------------------------- CTA-pair variant (BN = 256)
// Two CTAs of a cluster compute a 256-pixel x 256-channel tile with tcgen05.mma.cta_group::2
// (M = 256): each CTA loads its own 128-pixel A box and HALF of the B tile (128 of the 256
// output channels) - per k-block 32 KB of operands per SM instead of 48 KB, which is what the
// 1-CTA BN=256 kernel is bound by (shared-memory operand reads of the MMA and L2 -> SM fill).
// The leader (cluster rank 0) issues every MMA; its full barrier is credited by the TMA loads
// of both CTAs, its tcgen05.commit arrives on the empty / accumulator-ready barriers of both
// CTAs (multicast), and the epilogue warps of both CTAs release the accumulator stage on the
// leader's barrier. Every CTA runs its own epilogue on its own 128 TMEM lanes.
// All tensor maps cover this GPU's memory only; no load is multicast.
constexpr int kPairBN = 256;

template <int STAGES>
struct PairSmem {
  static constexpr int kABytes = kBlockM * kBlockKBytes;              // 16 KB
  static constexpr int kBBytes = (kPairBN / 2) * kBlockKBytes;        // 16 KB: this CTA's half
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = kBlockM * kBlockKBytes;
  static constexpr int kEpiOffset = STAGES * kStageBytes;
  static constexpr int kStatOffset = kEpiOffset + 2 * kEpiBytes;
  static constexpr int kBarOffset = kStatOffset + 1536;
  static constexpr int kTotal = kBarOffset + (2 * STAGES + 4) * 8 + 16;
};

template <int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_fprop_tf32_pair_kernel(const __grid_constant__ CUtensorMap tmap_x,
                            const __grid_constant__ CUtensorMap tmap_x1,
                            const __grid_constant__ CUtensorMap tmap_x2,
                            const __grid_constant__ CUtensorMap tmap_x3,
                            const __grid_constant__ CUtensorMap tmap_w,
                            const __grid_constant__ CUtensorMap tmap_y,
                            const ConvGeom g, float* __restrict__ stats) {
  using L = PairSmem<STAGES>;
  constexpr int BN = kPairBN;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // both CTAs of the pair must use identical offsets: the dynamic shared window starts at the
  // same address in every CTA of a kernel, so the same round-up applies to both
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));

  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);   // used in the leader
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;                                     // used in the leader
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int rows_valid = g.BW * g.BH * g.BNB;
  const int pix_tiles = g.tiles_w * g.tiles_h * g.tiles_n;
  const int pair_tiles = (pix_tiles + 1) / 2;
  const int num_items = pair_tiles * g.tiles_k * g.splits;
  const int k_iters = g.R * g.S * g.c_blocks;
  const uint32_t stage_tx = static_cast<uint32_t>(rows_valid * kBlockKBytes + L::kBBytes);
  const int first_item = static_cast<int>(blockIdx.x >> 1);
  const int item_stride = static_cast<int>(gridDim.x >> 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_y);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);          // the leader's own arrive.expect_tx (bytes of both)
      mbar_init(&empty_bar[s], 1);         // one multicast commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);         // one multicast commit
      mbar_init(&tmem_empty[s], 8);        // four epilogue warps in each CTA of the pair
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_pair(tmem_slot, 2 * BN);
  tc_fence_before();
  cluster_sync_all();                      // barriers of both CTAs exist before anyone signals
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  int it_begin = 0, it_end = k_iters;
  bool valid = true;
  auto decode = [&](int item, int& kt, int& tw, int& th, int& tn) {
    const int ks = item % g.splits;
    item /= g.splits;
    it_begin = ks * g.iters_per_split;
    it_end = min(k_iters, it_begin + g.iters_per_split);
    kt = item % g.tiles_k;
    const int p = 2 * (item / g.tiles_k) + static_cast<int>(rank);    // this CTA's pixel tile
    valid = p < pix_tiles;
    if (valid) {
      tw = p % g.tiles_w;
      const int rest = p / g.tiles_w;
      th = rest % g.tiles_h;
      tn = rest / g.tiles_h;
    } else {              // odd tile count: the partner of the last tile works on nothing -
      tw = 0;             // its boxes lie outside the tensor (zero-filled loads, clipped stores)
      th = 0;
      tn = g.tiles_n;
    }
  };

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = first_item; item < num_items; item += item_stride) {
        int kt, tw, th, tn;
        decode(item, kt, tw, th, tn);
        const int n0 = tn * g.BNB;
        const int k0 = kt * BN + static_cast<int>(rank) * (BN / 2);   // this CTA's half of B
        int tap = it_begin / g.c_blocks;
        int cb = it_begin - tap * g.c_blocks;
        const bool unit_stride = g.stride_h == 1 && g.stride_w == 1;
        int qr = 0, qs = 0;
        const CUtensorMap* mx = &tmap_x;
        auto enter_tap = [&](int t) {       // per-tap work only when the tap changes
          const int r = t / g.S;
          const int s = t - r * g.S;
          const int ar = r - g.pad_h, as = s - g.pad_w;
          if (unit_stride) {
            qr = ar;
            qs = as;
          } else {
            qr = floor_div(ar, g.stride_h);
            qs = floor_div(as, g.stride_w);
            const int view = (ar - qr * g.stride_h) * g.stride_w + (as - qs * g.stride_w);
            mx = view == 0 ? &tmap_x : view == 1 ? &tmap_x1 : view == 2 ? &tmap_x2 : &tmap_x3;
          }
        };
        enter_tap(tap);
        for (int it = it_begin; it < it_end; ++it) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = smem + stage * L::kStageBytes;
          uint8_t* b_dst = a_dst + L::kABytes;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * stage_tx);
          const uint32_t bar = mapa_u32(&full_bar[stage], 0);         // the leader's barrier
          tma_load_4d_pair(a_dst, mx, bar, cb * kBlockK, tw * g.BW + qs, th * g.BH + qr, n0);
          tma_load_4d_pair(b_dst, &tmap_w, bar, cb * kBlockK, tap, k0, 0);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
          if (++cb == g.c_blocks) {
            cb = 0;
            enter_tap(++tap);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader only)
    constexpr uint32_t idesc = make_idesc(/*tf32*/ 2, 2 * kBlockM, BN, 0, 0);
    if (leader && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int item = first_item; item < num_items; item += item_stride) {
        int kt_, tw_, th_, tn_;
        decode(item, kt_, tw_, th_, tn_);
        const int n_it = it_end - it_begin;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int it = 0; it < n_it; ++it) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * L::kStageBytes);
          const uint64_t a_desc = make_smem_desc_sw128(a_addr, 16, 1024);
          const uint64_t b_desc = make_smem_desc_sw128(a_addr + L::kABytes, 16, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            mma_tf32_ss_pair(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc,
                             (it > 0 || k > 0) ? 1u : 0u);
          mma_commit_pair(&empty_bar[stage], 0b11);
          if (it == n_it - 1) mma_commit_pair(&tmem_full[acc], 0b11);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (both CTAs)
    const int q = warp - 4;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 128;
    uint8_t* epi = smem + L::kEpiOffset;
    float2* stat_part = reinterpret_cast<float2*>(smem + L::kStatOffset);
    int* row_img = reinterpret_cast<int*>(smem + L::kStatOffset + 1024);
    int acc = 0;
    uint32_t acc_phase = 0;
    int chunk_counter = 0;
    for (int item = first_item; item < num_items; item += item_stride) {
      int kt, tw, th, tn;
      decode(item, kt, tw, th, tn);
      bool tile_uniform = false;
      if (g.stats_mode != 0) {
        tile_uniform = valid && g.BNB == 1 && (tw + 1) * g.BW <= g.Wo &&
                       (th + 1) * g.BH <= g.Ho;
        if (!tile_uniform) {
          asm volatile("bar.sync 2, 128;" ::: "memory");
          const int ww = row % g.BW;
          const int hh = (row / g.BW) % g.BH;
          const int nn = row / (g.BW * g.BH);
          const bool inside = valid && row < rows_valid && tw * g.BW + ww < g.Wo &&
                              th * g.BH + hh < g.Ho && tn * g.BNB + nn < g.N;
          row_img[row] = inside ? (g.stats_mode == 1 ? tn * g.BNB + nn : 0) : -1;
        }
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int j = 0; j < BN / kEpiChunk; ++j, ++chunk_counter) {
        uint32_t v[32];
        const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * BN + j * kEpiChunk) +
                               (static_cast<uint32_t>(q * 32) << 16);
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        if (j == BN / kEpiChunk - 1) {
          // accumulator fully read: release the TMEM stage on the LEADER's barrier
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_u32(&tmem_empty[acc], 0));
        }
        uint8_t* buf = epi + (chunk_counter & 1) * L::kEpiBytes;
        if (epi_tid == 0) tma_store_wait_read<1>();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        {
          uint8_t* rowp = buf + row * kBlockKBytes;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int pc = c ^ (row & 7);
            *reinterpret_cast<uint4*>(rowp + pc * 16) =
                make_uint4(v[4 * c + 0], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
          }
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (epi_tid == 0) {
          if (g.splits > 1)
            tma_reduce_add_4d(&tmap_y, buf, kt * BN + j * kEpiChunk, tw * g.BW, th * g.BH,
                              tn * g.BNB);
          else
            tma_store_4d(&tmap_y, buf, kt * BN + j * kEpiChunk, tw * g.BW, th * g.BH,
                         tn * g.BNB);
          tma_store_commit();
        }
        if (g.stats_mode != 0) {
          const int col = lane;
          const int kglob = kt * BN + j * kEpiChunk + col;
          const int sw_hi = col >> 2, sw_lo = (col & 3) * 4;
          if (tile_uniform) {
            float s1 = 0.f, s2 = 0.f;
            const int r_end = min(32, rows_valid - q * 32);
#pragma unroll 8
            for (int rr = 0; rr < r_end; ++rr) {
              const int r2 = q * 32 + rr;
              const float x = *reinterpret_cast<const float*>(
                  buf + r2 * kBlockKBytes + ((sw_hi ^ (r2 & 7)) << 4) + sw_lo);
              s1 += x;
              s2 = fmaf(x, x, s2);
            }
            stat_part[q * 32 + col] = make_float2(s1, s2);
            asm volatile("bar.sync 2, 128;" ::: "memory");
            if (q == 0 && kglob < g.K) {
              const float2 p0 = stat_part[col], p1 = stat_part[32 + col];
              const float2 p2 = stat_part[64 + col], p3 = stat_part[96 + col];
              const int img = (g.stats_mode == 1) ? tn : 0;
              float* dst = stats + (static_cast<size_t>(img) * g.K + kglob) * 2;
              atomicAdd(dst, (p0.x + p1.x) + (p2.x + p3.x));
              atomicAdd(dst + 1, (p0.y + p1.y) + (p2.y + p3.y));
            }
          } else {
            float s1 = 0.f, s2 = 0.f;
            int cur_img = -1;
            for (int rr = 0; rr < 32; ++rr) {
              const int r2 = q * 32 + rr;
              if (r2 >= rows_valid) break;
              const int img = row_img[r2];
              if (img < 0) continue;
              if (img != cur_img) {
                if (cur_img >= 0 && kglob < g.K) {
                  float* dst = stats + (static_cast<size_t>(cur_img) * g.K + kglob) * 2;
                  atomicAdd(dst, s1);
                  atomicAdd(dst + 1, s2);
                }
                cur_img = img;
                s1 = 0.f;
                s2 = 0.f;
              }
              const float x = *reinterpret_cast<const float*>(
                  buf + r2 * kBlockKBytes + ((sw_hi ^ (r2 & 7)) << 4) + sw_lo);
              s1 += x;
              s2 = fmaf(x, x, s2);
            }
            if (cur_img >= 0 && kglob < g.K) {
              float* dst = stats + (static_cast<size_t>(cur_img) * g.K + kglob) * 2;
              atomicAdd(dst, s1);
              atomicAdd(dst + 1, s2);
            }
          }
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (epi_tid == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  cluster_sync_all();                      // nobody leaves while the partner may still signal
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 2 * BN);
  }
}

// ---------------------------------------------------------------- host side
static void choose_pixel_box(int N, int Ho, int Wo, int stats_mode, int& BW, int& BH, int& BNB) {
  // Largest-utilisation box with BW*BH*BNB <= 128 that wastes the fewest rows on partial tiles.
  double best = -1.0;
  BW = 1; BH = 1; BNB = 1;
  for (int bw = 1; bw <= Wo && bw <= 128; ++bw) {
    for (int bh = 1; bh <= Ho && bw * bh <= 128; ++bh) {
      // Several images per tile: whole small images (6x6 -> 3 images), or the same row band of
      // several images when one image does not divide into 128-pixel boxes (12x12 = 144 pixels:
      // a 12x10 box leaves 44 % of the MMA rows empty, 12x2 x 5 images fills 120 of 128).
      // Bands keep full-width rows so every TMA row is one contiguous run.
      const int max_nb = 128 / (bw * bh);
      const int nb_hi = (bw == Wo) ? max_nb : 1;
      for (int nb = 1; nb <= nb_hi && nb <= N; ++nb) {
        const long tiles = static_cast<long>((Wo + bw - 1) / bw) * ((Ho + bh - 1) / bh) *
                           ((N + nb - 1) / nb);
        const double useful = static_cast<double>(N) * Ho * Wo / (static_cast<double>(tiles) * 128);
        // on ties prefer wide boxes (longer contiguous runs per TMA row), then few images per
        // tile (fewer statistics segments in the epilogue)
        // Per-image statistics (mode 1) cost one pair of global atomics per column and image
        // segment in the epilogue: there an extra image per tile must buy 2.5 % utilisation.
        const double score = useful + 1e-6 * bw - (stats_mode == 1 ? 0.025 : 1e-8) * (nb - 1);
        if (score > best) {
          best = score;
          BW = bw; BH = bh; BNB = nb;
        }
      }
    }
  }
}

struct XMaps {
  CUtensorMap m[4];
};

template <int BN, int STAGES, bool HALO = false, bool RESW = false>
static int launch_fprop(const XMaps& mx, const CUtensorMap& mw, const CUtensorMap& my,
                        const ConvGeom& g, float* stats, int sm_count, cudaStream_t stream) {
  using L = SmemLayout<BN, STAGES, HALO, RESW>;
  constexpr int smem_bytes = L::kTotal + 1024;      // + alignment slack
  // the opt-in shared-memory size is a per-device attribute of the kernel: one process may
  // drive several GPUs (single-process GPipe), so remember it per device
  static bool configured_on[64] = {false};
  int dev_ = 0;
  cudaGetDevice(&dev_);
  if (dev_ < 0 || dev_ >= 64) dev_ = 0;
  bool& configured = configured_on[dev_];
  auto kern = conv_fprop_tf32_kernel<BN, STAGES, HALO, RESW>;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         smem_bytes);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_fprop): %s",
                                           cudaGetErrorString(e));
    configured = true;
  }
  const int num_tiles = g.tiles_w * g.tiles_h * g.tiles_n * g.tiles_k * g.splits;
  const int grid = num_tiles < sm_count ? num_tiles : sm_count;
  kern<<<grid, kThreads, smem_bytes, stream>>>(mx.m[0], mx.m[1], mx.m[2], mx.m[3], mw, my, g,
                                               stats);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("conv_fprop launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

template <int STAGES>
static int launch_fprop_pair(const XMaps& mx, const CUtensorMap& mw, const CUtensorMap& my,
                             const ConvGeom& g, float* stats, int sm_count, cudaStream_t stream) {
  using L = PairSmem<STAGES>;
  constexpr int smem_bytes = L::kTotal + 1024;
  static bool configured_on[64] = {false};
  int dev_ = 0;
  cudaGetDevice(&dev_);
  if (dev_ < 0 || dev_ >= 64) dev_ = 0;
  bool& configured = configured_on[dev_];
  auto kern = conv_fprop_tf32_pair_kernel<STAGES>;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         smem_bytes);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_fprop_pair): %s",
                                           cudaGetErrorString(e));
    configured = true;
  }
  const int pix_tiles = g.tiles_w * g.tiles_h * g.tiles_n;
  const int items = (pix_tiles + 1) / 2 * g.tiles_k * g.splits;
  const int pairs = sm_count / 2;
  const int grid = 2 * (items < pairs ? items : pairs);
  kern<<<grid, kThreads, smem_bytes, stream>>>(mx.m[0], mx.m[1], mx.m[2], mx.m[3], mw, my, g,
                                               stats);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("conv_fprop_pair launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

// CTA pairs for BN = 256 tiles (default on; TGB_FPROP_2CTA=0 falls back to one CTA per tile).
// Measured on B200, U-Net layers at micro-batch 40 (build/selftest, TFLOP/s TF32, 1 CTA -> pair):
// 2048->2048 @6x6 714 -> 784, 1024->1024 @12x12 753 -> 770, 512->512 @24x24 711 -> 728,
// 256->256 @48x48 789 -> 792 (the last one already sits at the TF32 roof of about 800).
static bool use_cta_pairs(int BN, bool halo, const ConvGeom& g) {
  if (BN != kPairBN || halo) return false;
  if (g.tiles_w * g.tiles_h * g.tiles_n < 2) return false;
  const char* env = getenv("TGB_FPROP_2CTA");
  return env != nullptr ? atoi(env) != 0 : true;
}

// Everything that is decided on the host before a launch: tile shapes, kernel variant, split of
// the reduction. Shared by the launcher and by tgb_conv2d_fprop_plan (which lets callers and the
// CPU test-suite see the decisions without a GPU). `stats_mode` is the effective mode (0 when no
// statistics buffer is given). Ho / Wo <= 0: derive the output extent from the convolution
// arithmetic; otherwise compute exactly that extent (data-gradient launches ask for the extent
// of one output-parity class). `allow_split`: the split-K path zero-fills and reduce-adds into Y
// and so needs a dense Y.
static int plan_fprop(int N, int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                      int stride_h, int stride_w, int Ho, int Wo, bool allow_split,
                      int stats_mode, ConvGeom& g, int& BN_out, bool& halo_out,
                      bool& resident_out) {
  if (stride_h < 1 || stride_w < 1 || stride_h * stride_w > 4)
    return set_error("conv_fprop: strides up to 2x2 (or 1x4 / 4x1) are supported, got %dx%d",
                     stride_h, stride_w);
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K;
  g.R = R; g.S = S; g.pad_h = pad_h; g.pad_w = pad_w;
  g.stride_h = stride_h; g.stride_w = stride_w;
  g.Ho = Ho > 0 ? Ho : (H + 2 * pad_h - R) / stride_h + 1;
  g.Wo = Wo > 0 ? Wo : (W + 2 * pad_w - S) / stride_w + 1;
  if (g.Ho <= 0 || g.Wo <= 0 || (Ho <= 0 && H + 2 * pad_h < R) || (Wo <= 0 && W + 2 * pad_w < S))
    return set_error("conv_fprop: empty output");
  const bool unit_stride = stride_h == 1 && stride_w == 1;
  choose_pixel_box(N, g.Ho, g.Wo, stats_mode, g.BW, g.BH, g.BNB);
  // Output-channel tile: the widest of {256,128,64,32} that wastes at most 1/8 of its columns.
  // A partial last tile is fine: weight rows past K are zero-filled by TMA and the store of
  // columns past K is clipped.
  int BN = 32;
  for (int cand = 256; cand >= 32; cand /= 2) {
    const int tiles = (K + cand - 1) / cand;
    if ((tiles * cand - K) * 8 <= tiles * cand) { BN = cand; break; }
  }
  // Narrow output tiles (BN <= 64) are bound by L2 -> SM traffic when every tap re-reads the
  // input: use the haloed-tile variant (8 x 16 pixel tiles, input read once per tile) when the
  // filter is at most 3x3, the image tiles well into 8 x 16 boxes and there are plenty of tiles.
  bool halo = false;
  if (BN <= 64 && R <= 3 && S <= 3 && R * S > 1 && unit_stride && pad_h >= 0 && pad_w >= 0) {
    const long tw8 = (g.Wo + kHaloW - 1) / kHaloW, th16 = (g.Ho + kHaloH - 1) / kHaloH;
    const double useful = static_cast<double>(g.Ho) * g.Wo / (tw8 * kHaloW * th16 * kHaloH);
    halo = useful >= 0.85 && tw8 * th16 * N * ((K + BN - 1) / BN) >= 4L * sm_count();
    if (const char* env = getenv("TGB_FPROP_HALO")) {     // 0: never, 2: whenever it can run
      const int mode = atoi(env);
      halo = mode == 2 ? true : (mode != 0 && halo);
    }
    if (halo) {
      g.BW = kHaloW;
      g.BH = kHaloH;
      g.BNB = 1;
    }
  }
  g.tiles_w = (g.Wo + g.BW - 1) / g.BW;
  g.tiles_h = (g.Ho + g.BH - 1) / g.BH;
  g.tiles_n = (N + g.BNB - 1) / g.BNB;
  g.tiles_k = (K + BN - 1) / BN;
  g.c_blocks = (C + kBlockK - 1) / kBlockK;
  g.stats_mode = stats_mode;
  // Few output tiles and a long reduction (small images, many channels - e.g. 2048 ch at 6x6,
  // or any layer at a small micro-batch): split the (tap, channel-block) loop over CTAs so the
  // whole GPU works; partial sums are combined by TMA reduce-add into a zeroed Y.
  const int sms = sm_count();
  const int base_tiles = g.tiles_w * g.tiles_h * g.tiles_n * g.tiles_k;
  const int k_iters = halo ? g.c_blocks : R * S * g.c_blocks;
  g.splits = 1;
  if (!halo && allow_split && k_iters >= 32 && getenv("TGB_FPROP_NOSPLIT") == nullptr) {
    // Wave model: `items` equal work items run in ceil(items / sms) waves; an item costs its
    // reduction iterations plus an epilogue worth about kEpi iterations. Splitting the reduction
    // s ways shortens the items and fills partial waves (90 tiles on 148 SMs leave 39 % of the
    // machine idle) at the price of a zero-fill of Y, s reduce-add epilogues per tile and - when
    // statistics are wanted - one more pass over Y. Take the cheapest s, if it wins by > 5 %.
    constexpr int kEpi = 8;
    const double y_pass = static_cast<double>(N) * g.Ho * g.Wo * K * 4.0 / 6.0e12 / 0.4e-6;
    double best = static_cast<double>((base_tiles + sms - 1) / sms) * (k_iters + kEpi);
    for (int sp = 2; sp <= 8 && k_iters / sp >= 16; ++sp) {
      const int iters = (k_iters + sp - 1) / sp;
      const int eff = (k_iters + iters - 1) / iters;
      const long long items = static_cast<long long>(base_tiles) * eff;
      double cost = static_cast<double>((items + sms - 1) / sms) * (iters + kEpi);
      cost += 4.0 + y_pass;                                   // memset + launch
      if (g.stats_mode != 0) cost += 12.0 + y_pass;           // separate moments pass
      if (cost * 1.05 < best) {
        best = cost;
        g.splits = eff;
      }
    }
    if (const char* force = getenv("TGB_FPROP_SPLITS")) {
      const int f = atoi(force);
      if (f >= 1 && f <= 8) g.splits = f;
    }
  }
  g.iters_per_split = (k_iters + g.splits - 1) / g.splits;
  g.splits = (k_iters + g.iters_per_split - 1) / g.iters_per_split;
  bool resident = halo && g.tiles_k == 1 &&
                  static_cast<long>(R) * S * g.c_blocks * BN * kBlockKBytes <= kResidentWBytes;
  if (const char* env = getenv("TGB_FPROP_RESW")) resident = resident && atoi(env) != 0;
  BN_out = BN;
  halo_out = halo;
  resident_out = resident;
  return 0;
}

// A convolution with explicit addressing of X and Y (all strides in elements):
//   X[n][h][w][c] at x + n*x_sn + h*x_sh + w*x_sw + c,  0 <= h < H, 0 <= w < W, 0 <= c < C
//   Y[n][ho][wo][k] at y + n*y_sn + ho*y_sh + wo*y_sw + k
// Dense NHWC is (C, W*C, H*W*C). Other choices serve the data gradient of a strided convolution
// (Y = the pixels of one parity class of dX), channel-offset outputs (concatenation for free:
// y_sw larger than K) and the packed small-channel stem (x_sw smaller than C: every "pixel" is a
// window of several real pixels, see tgb_conv2d_stem_*).
struct ConvProblem {
  const float* x; const float* w; float* y; float* stats;
  int N, H, W, C, K, R, S, pad_h, pad_w, stride_h, stride_w;
  int Ho, Wo;                       // <= 0: from the convolution arithmetic
  long long x_sw, x_sh, x_sn;
  long long y_sw, y_sh, y_sn;
  int stats_mode;
};

static int run_fprop(const ConvProblem& p_in, cudaStream_t stream) {
  if (p_in.C % 4 != 0 || p_in.K % 4 != 0)
    return set_error("conv_fprop: need C %% 4 == 0 and K %% 4 == 0 (got C=%d K=%d)", p_in.C,
                     p_in.K);
  if ((reinterpret_cast<uintptr_t>(p_in.x) | reinterpret_cast<uintptr_t>(p_in.w) |
       reinterpret_cast<uintptr_t>(p_in.y)) & 15)
    return set_error("conv_fprop: tensors must be 16-byte aligned");
  if ((p_in.x_sw | p_in.x_sh | p_in.x_sn | p_in.y_sw | p_in.y_sh | p_in.y_sn) & 3)
    return set_error("conv_fprop: strides must be multiples of 4 elements (16 bytes)");
  // stats_mode < 0: same as -stats_mode, and the accumulators are zeroed here first (a memset
  // on the stream instead of a fill kernel issued by the caller)
  ConvProblem q = p_in;
  if (q.stats_mode < 0) {
    q.stats_mode = -q.stats_mode;
    if (q.stats != nullptr) {
      const size_t n = (q.stats_mode == 1 ? static_cast<size_t>(q.N) : 1) * q.K * 2 * 4;
      cudaError_t e = cudaMemsetAsync(q.stats, 0, n, stream);
      if (e != cudaSuccess) return set_error("conv_fprop: memset(stats): %s", cudaGetErrorString(e));
    }
  }
  const ConvProblem& p = q;
  ConvGeom g;
  int BN = 32;
  bool halo = false, resident = false;
  const int Ho_dense = (p.H + 2 * p.pad_h - p.R) / p.stride_h + 1;
  const int Wo_dense = (p.W + 2 * p.pad_w - p.S) / p.stride_w + 1;
  const int Ho = p.Ho > 0 ? p.Ho : Ho_dense, Wo = p.Wo > 0 ? p.Wo : Wo_dense;
  const bool y_dense = p.y_sw == p.K && p.y_sh == static_cast<long long>(Wo) * p.K &&
                       p.y_sn == static_cast<long long>(Ho) * Wo * p.K;
  if (int rc = plan_fprop(p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.pad_h, p.pad_w, p.stride_h,
                          p.stride_w, p.Ho, p.Wo, y_dense, p.stats ? p.stats_mode : 0, g, BN,
                          halo, resident))
    return rc;
  const int sms = sm_count();
  // Per-channel statistics (mode 2) from the epilogue are 2*K addresses - a handful of cache
  // lines - that EVERY pixel tile adds into; the L2 serialises atomics on one line (measured
  // about 0.7 us per pixel tile). With a long reduction per tile that stays hidden behind the
  // MMAs; with a short one it is exposed (ResNet 1x1/2 projection 512->1024 at micro-batch 110:
  // 0.048 ms without statistics, 0.169 ms with). Then one streaming pass over the freshly
  // written Y (mostly L2 hits) is cheaper. Per-image statistics (mode 1) spread over N times as
  // many lines and keep the fused path.
  const bool contended = g.stats_mode == 2 && !halo && y_dense &&
                         static_cast<long>(p.R) * p.S * g.c_blocks * g.tiles_k < 400 &&
                         g.tiles_w * g.tiles_h * g.tiles_n > 16 &&
                         getenv("TGB_FPROP_FUSED_STATS") == nullptr;
  const bool stats_after = (g.splits > 1 || contended) && g.stats_mode != 0;
  const int stats_mode_req = g.stats_mode;
  if (contended) g.stats_mode = 0;
  if (g.splits > 1) {
    g.stats_mode = 0;         // partial tiles cannot produce sum of squares: tgb_moments below
    cudaError_t e = cudaMemsetAsync(p.y, 0, static_cast<size_t>(p.N) * g.Ho * g.Wo * p.K * 4,
                                    stream);
    if (e != cudaSuccess) return set_error("conv_fprop: memset: %s", cudaGetErrorString(e));
  }

  XMaps mx;
  CUtensorMap mw, my;
  const bool pairs = use_cta_pairs(BN, halo, g);
  const int views = p.stride_h * p.stride_w;
  for (int v = 0; v < 4; ++v) {
    // view (ph, pw): the pixels with h = stride_h*hh + ph, w = stride_w*ww + pw. Views past
    // `views` are never touched by the kernel; give them the geometry of view 0.
    const int ph = v < views ? v / p.stride_w : 0, pw = v < views ? v % p.stride_w : 0;
    const long long Hv = (p.H - ph + p.stride_h - 1) / p.stride_h;
    const long long Wv = (p.W - pw + p.stride_w - 1) / p.stride_w;
    if (Hv <= 0 || Wv <= 0) {               // a parity class without pixels (1-pixel images)
      mx.m[v] = mx.m[0];
      continue;
    }
    uint64_t dims[4] = {(uint64_t)p.C, (uint64_t)Wv, (uint64_t)Hv, (uint64_t)p.N};
    uint64_t strides[3] = {(uint64_t)(p.x_sw * p.stride_w) * 4,
                           (uint64_t)(p.x_sh * p.stride_h) * 4, (uint64_t)p.x_sn * 4};
    uint32_t box[4] = {32, (uint32_t)g.BW, (uint32_t)g.BH, (uint32_t)g.BNB};
    if (halo) {               // the input tile with its halo in one box
      box[1] = (uint32_t)(g.BW + p.S - 1);
      box[2] = (uint32_t)(g.BH + p.R - 1);
    }
    if (make_tensor_map(&mx.m[v], p.x + ph * p.x_sh + pw * p.x_sw, 4, 4, dims, strides, box, 1))
      return set_error("conv_fprop: tensor map for X (view %d) failed", v);
  }
  {
    // weights [K][R*S][C] viewed as a 3-D (4-D with a unit dim) tensor so that a partial last
    // channel block is zero-filled instead of bleeding into the next tap
    uint64_t dims[4] = {(uint64_t)p.C, (uint64_t)(p.R * p.S), (uint64_t)p.K, 1};
    uint64_t strides[3] = {(uint64_t)p.C * 4, (uint64_t)p.R * p.S * p.C * 4,
                           (uint64_t)p.K * p.R * p.S * p.C * 4};
    // a CTA pair loads the B tile in two halves, one per CTA
    uint32_t box[4] = {32, 1, (uint32_t)(pairs ? BN / 2 : BN), 1};
    if (make_tensor_map(&mw, p.w, 4, 4, dims, strides, box, 1))
      return set_error("conv_fprop: tensor map for W failed");
  }
  {
    uint64_t dims[4] = {(uint64_t)p.K, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)p.N};
    uint64_t strides[3] = {(uint64_t)p.y_sw * 4, (uint64_t)p.y_sh * 4, (uint64_t)p.y_sn * 4};
    uint32_t box[4] = {32, (uint32_t)g.BW, (uint32_t)g.BH, (uint32_t)g.BNB};
    if (make_tensor_map(&my, p.y, 4, 4, dims, strides, box, 1))
      return set_error("conv_fprop: tensor map for Y failed");
  }
  int rc;
  if (pairs) {
    rc = launch_fprop_pair<5>(mx, mw, my, g, p.stats, sms, stream);
  } else if (halo && resident) {
    rc = BN == 64 ? launch_fprop<64, 2, true, true>(mx, mw, my, g, p.stats, sms, stream)
                  : launch_fprop<32, 2, true, true>(mx, mw, my, g, p.stats, sms, stream);
  } else if (halo) {
    rc = BN == 64 ? launch_fprop<64, 2, true>(mx, mw, my, g, p.stats, sms, stream)
                  : launch_fprop<32, 3, true>(mx, mw, my, g, p.stats, sms, stream);
  } else
  switch (BN) {
    case 256: rc = launch_fprop<256, 4>(mx, mw, my, g, p.stats, sms, stream); break;
    case 128: rc = launch_fprop<128, 6>(mx, mw, my, g, p.stats, sms, stream); break;
    case 64:  rc = launch_fprop<64, 8>(mx, mw, my, g, p.stats, sms, stream); break;
    default:  rc = launch_fprop<32, 8>(mx, mw, my, g, p.stats, sms, stream); break;
  }
  if (rc == 0 && stats_after) {
    // Y is [N][Ho*Wo][K]: per-image (mode 1) or per-channel (mode 2) moments in one more pass
    rc = stats_mode_req == 1 ? tgb_moments(p.y, p.stats, p.N, g.Ho * g.Wo, p.K, stream)
                             : tgb_moments(p.y, p.stats, 1, p.N * g.Ho * g.Wo, p.K, stream);
  }
  return rc;
}

static ConvProblem dense_problem(const float* x, const float* w, float* y, float* stats, int N,
                                 int H, int W, int C, int K, int R, int S, int pad_h, int pad_w,
                                 int stride_h, int stride_w, int stats_mode) {
  ConvProblem p;
  p.x = x; p.w = w; p.y = y; p.stats = stats;
  p.N = N; p.H = H; p.W = W; p.C = C; p.K = K; p.R = R; p.S = S;
  p.pad_h = pad_h; p.pad_w = pad_w; p.stride_h = stride_h; p.stride_w = stride_w;
  p.Ho = p.Wo = 0;
  p.x_sw = C; p.x_sh = static_cast<long long>(W) * C; p.x_sn = static_cast<long long>(H) * W * C;
  const long long Ho = (H + 2 * pad_h - R) / (stride_h > 0 ? stride_h : 1) + 1;
  const long long Wo = (W + 2 * pad_w - S) / (stride_w > 0 ? stride_w : 1) + 1;
  p.y_sw = K; p.y_sh = Wo * K; p.y_sn = Ho * Wo * K;
  p.stats_mode = stats_mode;
  return p;
}

}  // namespace tgb

using namespace tgb;

extern "C" int tgb_conv2d_fprop_tf32(const float* x, const float* w, float* y, float* stats,
                                     int N, int H, int W, int C, int K, int R, int S, int pad_h,
                                     int pad_w, int stats_mode, cudaStream_t stream) {
  return run_fprop(dense_problem(x, w, y, stats, N, H, W, C, K, R, S, pad_h, pad_w, 1, 1,
                                 stats_mode), stream);
}

extern "C" int tgb_conv2d_fprop_strided_tf32(const float* x, const float* w, float* y,
                                             float* stats, int N, int H, int W, int C, int K,
                                             int R, int S, int pad_h, int pad_w, int stride_h,
                                             int stride_w, int Ho, int Wo,
                                             long long y_pixel_stride, int stats_mode,
                                             cudaStream_t stream) {
  if (stride_h < 1 || stride_w < 1) return set_error("conv_fprop: strides must be positive");
  ConvProblem p = dense_problem(x, w, y, stats, N, H, W, C, K, R, S, pad_h, pad_w, stride_h,
                                stride_w, stats_mode);
  // Ho / Wo > 0: the caller fixes the output extent (asymmetric padding: with pad = -1 and the
  // extent of the unshifted convolution, the far edge is padded with zeros by TMA)
  long long ho = p.y_sn / p.y_sh, wo = p.y_sh / K;
  if (Ho > 0) { p.Ho = Ho; ho = Ho; }
  if (Wo > 0) { p.Wo = Wo; wo = Wo; }
  if (y_pixel_stride <= 0) y_pixel_stride = K;
  // y_pixel_stride > K: Y is a channel slice of a wider NHWC tensor (part of a concatenation)
  if (y_pixel_stride < K) return set_error("conv_fprop: y_pixel_stride smaller than K");
  p.y_sw = y_pixel_stride; p.y_sh = wo * y_pixel_stride; p.y_sn = ho * wo * y_pixel_stride;
  return run_fprop(p, stream);
}

// ---- small-channel stem on the packed buffer of tgb_stem_pack (see conv_small.cu) ------------
// xp: [N][Hp][Wp][4] zero-haloed image; wv: [K][R][32] with wv[k][r][4*s+c] = W[k][c][r][s]
// (zero where s >= S or c >= C). The convolution is R x 1 over the 32-"channel" window view.
extern "C" int tgb_conv2d_stem_fprop_tf32(const float* xp, const float* wv, float* y,
                                          float* stats, int N, int Hp, int Wp, int K, int R,
                                          int stride_h, int stride_w, int stats_mode,
                                          cudaStream_t stream) {
  if (Wp < 8) return set_error("stem_fprop: packed rows must hold at least one 8-pixel window");
  ConvProblem p = dense_problem(xp, wv, y, stats, N, Hp, Wp - 7, 32, K, R, 1, 0, 0, stride_h,
                                stride_w, stats_mode);
  p.x_sw = 4;                                   // next window starts one pixel (4 floats) later
  p.x_sh = static_cast<long long>(Wp) * 4;
  p.x_sn = static_cast<long long>(Hp) * Wp * 4;
  return run_fprop(p, stream);
}

// ---- data gradient of a (possibly strided) convolution -------------------------------------
//   dX[n,h,w,c] = sum over (r,s,k) with (h+pad_h-r) % sh == 0 and (w+pad_w-s) % sw == 0 of
//                 dY[n,(h+pad_h-r)/sh,(w+pad_w-s)/sw,k] * W[k,r,s,c]
// For the input pixels of one parity class (h = sh*hh + a, w = sw*ww + b) only the filter rows
// r = r0 + sh*j (r0 = (a+pad_h) mod sh) take part, and the sum becomes a stride-1 correlation of
// dY with that sub-filter: one launch of the forward kernel per class, its output tensor map
// being the strided view "pixels of class (a,b)" of dX. The sub-filters, tap-flipped and
// channel-transposed, are prepared by tgb_dgrad_weight_strided (class-major, [C][Rj*Si][K] each).
struct DgradClass {
  int r0, Rj, dh, s0, Si, dw;
  long long w_offset;         // elements from the start of the class-major weight buffer
};

static void dgrad_classes(int C, int K, int R, int S, int pad_h, int pad_w, int sh, int sw,
                          DgradClass* cls, long long* total) {
  long long off = 0;
  for (int a = 0; a < sh; ++a)
    for (int b = 0; b < sw; ++b) {
      DgradClass& c = cls[a * sw + b];
      c.r0 = ((a + pad_h) % sh + sh) % sh;
      c.s0 = ((b + pad_w) % sw + sw) % sw;
      c.Rj = c.r0 < R ? (R - c.r0 + sh - 1) / sh : 0;
      c.Si = c.s0 < S ? (S - c.s0 + sw - 1) / sw : 0;
      c.dh = (a + pad_h - c.r0) / sh;
      c.dw = (b + pad_w - c.s0) / sw;
      c.w_offset = off;
      off += static_cast<long long>(C) * c.Rj * c.Si * K;
    }
  *total = off;
}

extern "C" long long tgb_dgrad_weight_strided_elems(int K, int C, int R, int S, int pad_h,
                                                    int pad_w, int stride_h, int stride_w) {
  if (stride_h < 1 || stride_w < 1 || stride_h * stride_w > 4) return -1;
  DgradClass cls[4];
  long long total = 0;
  dgrad_classes(C, K, R, S, pad_h, pad_w, stride_h, stride_w, cls, &total);
  return total;
}

__global__ void dgrad_weight_class_kernel(const float* __restrict__ w, float* __restrict__ out,
                                          int K, int C, int R, int S, int r0, int Rj, int s0,
                                          int Si, int sh, int sw) {
  // out[c][j'][i'][k] = w[k][r0 + sh*(Rj-1-j')][s0 + sw*(Si-1-i')][c]; a 32x32 (k, c) tile goes
  // through shared memory so that both the read (c fastest) and the write (k fastest) coalesce
  __shared__ float tile[32][33];
  const int tap = blockIdx.z;
  const int jp = tap / Si, ip = tap - jp * Si;
  const int r = r0 + sh * (Rj - 1 - jp), s = s0 + sw * (Si - 1 - ip);
  const int k0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int k = k0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (k < K && c < C)
        ? w[((static_cast<size_t>(k) * R + r) * S + s) * C + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, k = k0 + threadIdx.x;
    if (c < C && k < K)
      out[(static_cast<size_t>(c) * Rj * Si + tap) * K + k] = tile[threadIdx.x][i];
  }
}

extern "C" int tgb_dgrad_weight_strided(const float* w, float* out, int K, int C, int R, int S,
                                        int pad_h, int pad_w, int stride_h, int stride_w,
                                        cudaStream_t stream) {
  if (stride_h < 1 || stride_w < 1 || stride_h * stride_w > 4)
    return set_error("dgrad_weight_strided: unsupported stride %dx%d", stride_h, stride_w);
  DgradClass cls[4];
  long long total = 0;
  dgrad_classes(C, K, R, S, pad_h, pad_w, stride_h, stride_w, cls, &total);
  for (int v = 0; v < stride_h * stride_w; ++v) {
    const DgradClass& c = cls[v];
    if (c.Rj * c.Si == 0) continue;
    dim3 grid((K + 31) / 32, (C + 31) / 32, c.Rj * c.Si), block(32, 8);
    dgrad_weight_class_kernel<<<grid, block, 0, stream>>>(w, out + c.w_offset, K, C, R, S, c.r0,
                                                          c.Rj, c.s0, c.Si, stride_h, stride_w);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("dgrad_weight_strided launch: %s",
                                           cudaGetErrorString(e));
    count_launch();
  }
  return 0;
}

// dy_pixel_stride: elements between consecutive pixels of dY (K for a dense tensor, more when dY
// is a channel slice of a wider NHWC gradient). keep_untouched != 0: input pixels that no filter
// tap reads are left as they are instead of being zero-filled (two convolutions that read
// disjoint pixel classes of one input - AmoebaNet's FactorizedReduce - write one dX in turn).
extern "C" int tgb_conv2d_dgrad_tf32(const float* dy, const float* wt_classes, float* dx, int N,
                                     int H, int W, int C, int K, int R, int S, int pad_h,
                                     int pad_w, int stride_h, int stride_w, int Ho_given,
                                     int Wo_given, long long dy_pixel_stride, int keep_untouched,
                                     cudaStream_t stream) {
  if (stride_h < 1 || stride_w < 1 || stride_h * stride_w > 4)
    return set_error("conv_dgrad: unsupported stride %dx%d", stride_h, stride_w);
  const int Ho = Ho_given > 0 ? Ho_given : (H + 2 * pad_h - R) / stride_h + 1;
  const int Wo = Wo_given > 0 ? Wo_given : (W + 2 * pad_w - S) / stride_w + 1;
  if (Ho <= 0 || Wo <= 0) return set_error("conv_dgrad: empty output");
  DgradClass cls[4];
  long long total = 0;
  dgrad_classes(C, K, R, S, pad_h, pad_w, stride_h, stride_w, cls, &total);
  const int views = stride_h * stride_w;
  bool any_empty = false;
  for (int v = 0; v < views; ++v) any_empty |= cls[v].Rj * cls[v].Si == 0;
  if (dy_pixel_stride <= 0) dy_pixel_stride = K;
  if (dy_pixel_stride < K || (dy_pixel_stride & 3))
    return set_error("conv_dgrad: bad dy_pixel_stride %lld for K=%d", dy_pixel_stride, K);
  if (any_empty && !keep_untouched) {   // pixels no filter tap ever reads (e.g. 1x1 stride 2)
    cudaError_t e = cudaMemsetAsync(dx, 0, static_cast<size_t>(N) * H * W * C * 4, stream);
    if (e != cudaSuccess) return set_error("conv_dgrad: memset: %s", cudaGetErrorString(e));
  }
  for (int v = 0; v < views; ++v) {
    const DgradClass& c = cls[v];
    const int a = v / stride_w, b = v % stride_w;
    const int Hc = (H - a + stride_h - 1) / stride_h, Wc = (W - b + stride_w - 1) / stride_w;
    if (c.Rj * c.Si == 0 || Hc <= 0 || Wc <= 0) continue;
    ConvProblem p;
    p.x = dy; p.w = wt_classes + c.w_offset; p.stats = nullptr; p.stats_mode = 0;
    p.y = dx + (static_cast<long long>(a) * W + b) * C;
    p.N = N; p.H = Ho; p.W = Wo; p.C = K; p.K = C; p.R = c.Rj; p.S = c.Si;
    p.pad_h = c.Rj - 1 - c.dh; p.pad_w = c.Si - 1 - c.dw;
    p.stride_h = p.stride_w = 1;
    p.Ho = Hc; p.Wo = Wc;
    p.x_sw = dy_pixel_stride; p.x_sh = static_cast<long long>(Wo) * dy_pixel_stride;
    p.x_sn = static_cast<long long>(Ho) * Wo * dy_pixel_stride;
    p.y_sw = static_cast<long long>(stride_w) * C;
    p.y_sh = static_cast<long long>(stride_h) * W * C;
    p.y_sn = static_cast<long long>(H) * W * C;
    if (int rc = run_fprop(p, stream)) return rc;
  }
  return 0;
}

// out[16]: BN, BW, BH, BNB, tiles_w, tiles_h, tiles_n, tiles_k, splits, iters_per_split, halo,
// resident filter bank, pipeline stages, reduction iterations per tile, Ho, Wo.
extern "C" int tgb_conv2d_fprop_plan(int N, int H, int W, int C, int K, int R, int S, int pad_h,
                                     int pad_w, int stats_mode, int* out) {
  if (C % 4 != 0 || K % 4 != 0)
    return set_error("conv_fprop: need C %% 4 == 0 and K %% 4 == 0 (got C=%d K=%d)", C, K);
  ConvGeom g;
  int BN = 32;
  bool halo = false, resident = false;
  if (int rc = plan_fprop(N, H, W, C, K, R, S, pad_h, pad_w, 1, 1, 0, 0, true, stats_mode, g, BN,
                          halo, resident))
    return rc;
  const int stages = halo ? (resident ? 2 : (BN == 64 ? 2 : 3))
                          : (BN == 256 ? 4 : BN == 128 ? 6 : 8);
  const int values[16] = {BN, g.BW, g.BH, g.BNB, g.tiles_w, g.tiles_h, g.tiles_n, g.tiles_k,
                          g.splits, g.iters_per_split, halo ? 1 : 0, resident ? 1 : 0, stages,
                          (halo ? 1 : R * S) * g.c_blocks, g.Ho, g.Wo};
  for (int i = 0; i < 16; ++i) out[i] = values[i];
  return 0;
}
