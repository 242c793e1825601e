// Convolutions whose channel counts are too small for a 32-channel tensor-core K block:
//
//  * the input stem (3 input channels: ResNet 7x7/2, AmoebaNet 3x3/2, U-Net 3x3/1 - reference
//    benchmarks/models/resnet/__init__.py:55, amoebanet/__init__.py:54, unet/__init__.py:44).
//    tgb_stem_pack copies the image into a zero-haloed NHWC buffer with 4 channels per pixel.
//    In that buffer 8 neighbouring pixels of a row are 32 consecutive floats, so "the 8 pixels
//    starting at column w" is a legal 128-byte K block for every w: a tensor map whose
//    pixel-to-pixel stride is 16 bytes while its innermost extent is 128 bytes (overlapping
//    windows) presents the image as a 32-channel tensor in which channel index = 4*s + c. One
//    filter ROW of up to 8 taps is then a single K block, a 7x7 stem is a 7x1 convolution over
//    32 "channels" (28 of them real) on the ordinary tcgen05 kernels, forward and weight
//    gradient alike; nothing is ever expanded in memory. Zero padding comes from the halo, not
//    from TMA bounds (the windows run across pixel boundaries), strides from the parity views.
//
//  * the 1x1 head with very few output channels (U-Net 64 -> 1, unet/__init__.py:138): a dot
//    product per pixel. There is no tile to fill; it is a memory-bound streaming kernel
//    (one pass over X forward, one pass over X plus one write of dX backward).
#include "tgb_internal.h"
#include "../include/tgb200.h"
#include <cstdint>

namespace tgb {

// dst[n][hp][wp][0..3] <- src[n][c][h][w] (any strides), zero in the halo and in channel >= C
__global__ void stem_pack_kernel(const float* __restrict__ src, float4* __restrict__ dst, int N,
                                 int C, int H, int W, long long s_n, long long s_c,
                                 long long s_h, long long s_w, int pad_h, int pad_w, int Hp,
                                 int Wp, int relu) {
  const long long total = static_cast<long long>(N) * Hp * Wp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int wp = static_cast<int>(i % Wp);
    const long long t = i / Wp;
    const int hp = static_cast<int>(t % Hp);
    const int n = static_cast<int>(t / Hp);
    const int h = hp - pad_h, w = wp - pad_w;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (h >= 0 && h < H && w >= 0 && w < W) {
      const float* p = src + n * s_n + h * s_h + w * s_w;
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (c < C) {
          const float t = __ldg(p + c * s_c);
          v[c] = relu ? fmaxf(t, 0.f) : t;          // AmoebaNet's stem applies ReLU to the image
        }
    }
    dst[i] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// ---- 1x1 convolution with K <= 4 output channels ------------------------------------------
// G lanes share a pixel (G = power of two >= C/4, at most 32); lane g of a group owns the
// channel quads g, g+G, ...
template <int KMAX>
__global__ void head_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                const float* __restrict__ bias, float* __restrict__ y,
                                long long pixels, int C, int K, int G) {
  const int lane = threadIdx.x & 31;
  const int per_warp = 32 / G;
  const int sub = lane / G, g = lane % G;
  const long long warp_id = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long warps = (gridDim.x * static_cast<long long>(blockDim.x)) >> 5;
  const int c4n = C >> 2;
  for (long long p0 = warp_id * per_warp; p0 < pixels; p0 += warps * per_warp) {
    const long long p = p0 + sub;
    float acc[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) acc[k] = 0.f;
    if (p < pixels) {
      const float4* xp = reinterpret_cast<const float4*>(x + p * C);
      for (int c4 = g; c4 < c4n; c4 += G) {
        const float4 v = xp[c4];
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < K) {
            const float4 ww = __ldg(reinterpret_cast<const float4*>(w + k * C) + c4);
            acc[k] += v.x * ww.x + v.y * ww.y + v.z * ww.z + v.w * ww.w;
          }
      }
    }
#pragma unroll
    for (int k = 0; k < KMAX; ++k)
      for (int off = G >> 1; off > 0; off >>= 1)
        acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], off);
    if (p < pixels && g == 0) {
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (k < K) y[p * K + k] = acc[k] + (bias ? bias[k] : 0.f);
    }
  }
}

// dX[p][c] = sum_k dY[p][k] W[k][c];  dW[k][c] += sum_p dY[p][k] X[p][c];  db[k] += sum_p dY[p][k]
template <int KMAX>
__global__ void head_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                const float* __restrict__ w, float* __restrict__ dx,
                                float* __restrict__ dw, float* __restrict__ db, long long pixels,
                                int C, int K, int G) {
  extern __shared__ float red[];            // [KMAX][C] partial dW of this block, then [KMAX] db
  const int lane = threadIdx.x & 31;
  const int per_warp = 32 / G;
  const int sub = lane / G, g = lane % G;
  const long long warp_id = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long warps = (gridDim.x * static_cast<long long>(blockDim.x)) >> 5;
  const int c4n = C >> 2;
  for (int i = threadIdx.x; i < KMAX * C + KMAX; i += blockDim.x) red[i] = 0.f;
  __syncthreads();
  // each lane keeps partial dW for the first two channel quads it owns in registers (C <= 8*G
  // covers every model here); further quads go straight to shared atomics
  float4 part[KMAX][2];
  float dbp[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    part[k][0] = part[k][1] = make_float4(0.f, 0.f, 0.f, 0.f);
    dbp[k] = 0.f;
  }
  for (long long p0 = warp_id * per_warp; p0 < pixels; p0 += warps * per_warp) {
    const long long p = p0 + sub;
    if (p >= pixels) continue;
    float d[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) d[k] = k < K ? dy[p * K + k] : 0.f;
    if (g == 0) {
#pragma unroll
      for (int k = 0; k < KMAX; ++k) dbp[k] += d[k];
    }
    const float4* xp = x ? reinterpret_cast<const float4*>(x + p * C) : nullptr;
    float4* dxp = dx ? reinterpret_cast<float4*>(dx + p * C) : nullptr;
    int slot = 0;
    for (int c4 = g; c4 < c4n; c4 += G, ++slot) {
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 v = xp ? xp[c4] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (k < K) {
          if (dxp) {
            const float4 ww = __ldg(reinterpret_cast<const float4*>(w + k * C) + c4);
            o.x += d[k] * ww.x; o.y += d[k] * ww.y; o.z += d[k] * ww.z; o.w += d[k] * ww.w;
          }
          if (xp) {
            if (slot < 2) {
              float4& a = part[k][slot];
              a.x += d[k] * v.x; a.y += d[k] * v.y; a.z += d[k] * v.z; a.w += d[k] * v.w;
            } else {
              float* r = red + k * C + c4 * 4;
              atomicAdd(r, d[k] * v.x); atomicAdd(r + 1, d[k] * v.y);
              atomicAdd(r + 2, d[k] * v.z); atomicAdd(r + 3, d[k] * v.w);
            }
          }
        }
      if (dxp) dxp[c4] = o;
    }
  }
#pragma unroll
  for (int k = 0; k < KMAX; ++k)
    if (k < K) {
#pragma unroll
      for (int slot = 0; slot < 2; ++slot) {
        const int c4 = g + slot * G;
        if (c4 < c4n && x) {
          float* r = red + k * C + c4 * 4;
          atomicAdd(r, part[k][slot].x); atomicAdd(r + 1, part[k][slot].y);
          atomicAdd(r + 2, part[k][slot].z); atomicAdd(r + 3, part[k][slot].w);
        }
      }
      if (g == 0) atomicAdd(red + KMAX * C + k, dbp[k]);
    }
  __syncthreads();
  if (dw)
    for (int i = threadIdx.x; i < K * C; i += blockDim.x) atomicAdd(dw + i, red[i]);
  if (db)
    for (int k = threadIdx.x; k < K; k += blockDim.x) atomicAdd(db + k, red[KMAX * C + k]);
}

static int group_lanes(int C) {
  int g = 1;
  while (g < 32 && g * 4 < C) g <<= 1;
  return g;
}

}  // namespace tgb

using namespace tgb;

extern "C" int tgb_stem_pack(const float* src, float* dst, int N, int C, int H, int W,
                             long long s_n, long long s_c, long long s_h, long long s_w,
                             int pad_h, int pad_w, int Hp, int Wp, int relu,
                             cudaStream_t stream) {
  if (C < 1 || C > 4) return set_error("stem_pack: 1..4 input channels, got %d", C);
  if (Hp < H + pad_h || Wp < W + pad_w || pad_h < 0 || pad_w < 0)
    return set_error("stem_pack: padded extent %dx%d too small for %dx%d + pad", Hp, Wp, H, W);
  if (reinterpret_cast<uintptr_t>(dst) & 15) return set_error("stem_pack: dst must be 16-byte aligned");
  const long long total = static_cast<long long>(N) * Hp * Wp;
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  stem_pack_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(
      src, reinterpret_cast<float4*>(dst), N, C, H, W, s_n, s_c, s_h, s_w, pad_h, pad_w, Hp, Wp,
      relu);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("stem_pack launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

extern "C" int tgb_head_fwd(const float* x, const float* w, const float* bias, float* y,
                            long long pixels, int C, int K, cudaStream_t stream) {
  if (C % 4 != 0 || K < 1 || K > 4)
    return set_error("head_fwd: need C %% 4 == 0 and 1 <= K <= 4 (got C=%d K=%d)", C, K);
  const int G = group_lanes(C);
  const long long warps_needed = (pixels + (32 / G) - 1) / (32 / G);
  long long blocks = (warps_needed + 7) / 8;
  const long long cap = static_cast<long long>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  head_fwd_kernel<4><<<static_cast<int>(blocks), 256, 0, stream>>>(x, w, bias, y, pixels, C, K, G);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("head_fwd launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

extern "C" int tgb_head_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw,
                            float* db, long long pixels, int C, int K, cudaStream_t stream) {
  if (C % 4 != 0 || K < 1 || K > 4)
    return set_error("head_bwd: need C %% 4 == 0 and 1 <= K <= 4 (got C=%d K=%d)", C, K);
  if (C > 2048) return set_error("head_bwd: at most 2048 input channels");
  const int G = group_lanes(C);
  const long long warps_needed = (pixels + (32 / G) - 1) / (32 / G);
  long long blocks = (warps_needed + 7) / 8;
  const long long cap = static_cast<long long>(sm_count()) * 4;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaError_t e;
  if (dw) {
    e = cudaMemsetAsync(dw, 0, static_cast<size_t>(K) * C * 4, stream);
    if (e != cudaSuccess) return set_error("head_bwd memset: %s", cudaGetErrorString(e));
  }
  if (db) {
    e = cudaMemsetAsync(db, 0, static_cast<size_t>(K) * 4, stream);
    if (e != cudaSuccess) return set_error("head_bwd memset: %s", cudaGetErrorString(e));
  }
  const size_t smem = static_cast<size_t>(4 * C + 4) * 4;
  head_bwd_kernel<4><<<static_cast<int>(blocks), 256, smem, stream>>>(dy, dw ? x : nullptr, w, dx,
                                                                     dw, db, pixels, C, K, G);
  e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("head_bwd launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
