// NHWC fp32 data-movement kernels of the U-Net encoder / decoder (memory-bound, 16-byte accesses).
//
//  * MaxPool2d(2, stride 2) forward, and backward without an index tensor (the arg-max is
//    re-derived from the saved input).
//  * Nearest x2 upsample + pad-to-match + channel concat in ONE pass ("PopCat" of the reference
//    U-Net, benchmarks/models/unet/__init__.py:27-39,131): out[..., :C1] = up(x), out[..., C1:] =
//    skip. The reference runs Upsample, F.pad and torch.cat as three kernels with two
//    intermediate tensors. Backward splits the gradient into d(skip) (contiguous copy) and d(x)
//    (2x2 window sum) in one pass over d(out).
// SURVEY K7 / K9 / K10.
#include <cuda_runtime.h>
#include <cstdint>
#include "tgb_internal.h"

namespace tgb {

constexpr int kPrThreads = 256;

__device__ __forceinline__ float4 pr_ld(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
__device__ __forceinline__ void pr_st(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 f4max(float4 a, float4 b) {
  return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}

// x [N,H,W,C] -> y [N,H/2,W/2,C]
__global__ void __launch_bounds__(kPrThreads)
maxpool2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int N, int H, int W,
                    int C) {
  const int Ho = H >> 1, Wo = W >> 1, Q = C >> 2;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * Q;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    size_t r = i / Q;
    const int wo = static_cast<int>(r % Wo);
    r /= Wo;
    const int ho = static_cast<int>(r % Ho);
    const int n = static_cast<int>(r / Ho);
    const float* p = x + ((static_cast<size_t>(n) * H + 2 * ho) * W + 2 * wo) * C + q * 4;
    const float4 a = pr_ld(p), b = pr_ld(p + C);
    const float4 c = pr_ld(p + static_cast<size_t>(W) * C), d = pr_ld(p + static_cast<size_t>(W) * C + C);
    pr_st(y + i * 4, f4max(f4max(a, b), f4max(c, d)));
  }
}

// dx [N,H,W,C] from dy [N,H/2,W/2,C]: the gradient goes to the first maximum of each window
// (row-major window order, like ATen). Rows / columns an odd H / W leaves uncovered get zero.
__global__ void __launch_bounds__(kPrThreads)
maxpool2_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                    float* __restrict__ dx, int N, int H, int W, int C) {
  const int Ho = H >> 1, Wo = W >> 1, Q = C >> 2;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * Q;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    size_t r = i / Q;
    const int wo = static_cast<int>(r % Wo);
    r /= Wo;
    const int ho = static_cast<int>(r % Ho);
    const int n = static_cast<int>(r / Ho);
    const size_t base = ((static_cast<size_t>(n) * H + 2 * ho) * W + 2 * wo) * C + q * 4;
    const size_t row = static_cast<size_t>(W) * C;
    const float4 a = pr_ld(x + base), b = pr_ld(x + base + C);
    const float4 c = pr_ld(x + base + row), d = pr_ld(x + base + row + C);
    const float4 g = pr_ld(dy + i * 4);
    float4 ga, gb, gc, gd;
#define TGB_POOL_ROUTE(f)                                                     \
  {                                                                           \
    const float m = fmaxf(fmaxf(a.f, b.f), fmaxf(c.f, d.f));                  \
    const bool ta = a.f == m, tb = !ta && b.f == m, tc = !ta && !tb && c.f == m; \
    const bool td = !ta && !tb && !tc;                                        \
    ga.f = ta ? g.f : 0.f;                                                    \
    gb.f = tb ? g.f : 0.f;                                                    \
    gc.f = tc ? g.f : 0.f;                                                    \
    gd.f = td ? g.f : 0.f;                                                    \
  }
    TGB_POOL_ROUTE(x) TGB_POOL_ROUTE(y) TGB_POOL_ROUTE(z) TGB_POOL_ROUTE(w)
#undef TGB_POOL_ROUTE
    pr_st(dx + base, ga);
    pr_st(dx + base + C, gb);
    pr_st(dx + base + row, gc);
    pr_st(dx + base + row + C, gd);
  }
}

// out [N,Hs,Ws,C1+C2]: channels [0,C1) = x[n, h/2, w/2] (zero where 2*Hx <= h or 2*Wx <= w:
// the reference pads the upsampled map at the bottom / right to the skip's size), [C1,C1+C2) = skip
__global__ void __launch_bounds__(kPrThreads)
upcat_fwd_kernel(const float* __restrict__ x, const float* __restrict__ skip,
                 float* __restrict__ out, int N, int Hx, int Wx, int C1, int Hs, int Ws, int C2) {
  const int Ct = C1 + C2, Q = Ct >> 2, Q1 = C1 >> 2;
  const size_t total = static_cast<size_t>(N) * Hs * Ws * Q;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    size_t r = i / Q;
    const int w = static_cast<int>(r % Ws);
    r /= Ws;
    const int h = static_cast<int>(r % Hs);
    const int n = static_cast<int>(r / Hs);
    float4 v;
    if (q < Q1) {
      const int hx = h >> 1, wx = w >> 1;
      v = (hx < Hx && wx < Wx)
              ? pr_ld(x + ((static_cast<size_t>(n) * Hx + hx) * Wx + wx) * C1 + q * 4)
              : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      v = pr_ld(skip + ((static_cast<size_t>(n) * Hs + h) * Ws + w) * C2 + (q - Q1) * 4);
    }
    pr_st(out + i * 4, v);
  }
}

// dskip = dout[..., C1:]  (one thread per skip element quad)
// dx[n,hx,wx,c] = sum of dout[n, 2hx+{0,1}, 2wx+{0,1}, c] inside [Hs,Ws]
__global__ void __launch_bounds__(kPrThreads)
upcat_bwd_kernel(const float* __restrict__ dout, float* __restrict__ dx,
                 float* __restrict__ dskip, int N, int Hx, int Wx, int C1, int Hs, int Ws, int C2) {
  const int Ct = C1 + C2, Q1 = C1 >> 2, Q2 = C2 >> 2;
  const size_t n_skip = static_cast<size_t>(N) * Hs * Ws * Q2;
  const size_t n_x = static_cast<size_t>(N) * Hx * Wx * Q1;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_skip + n_x;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    if (i < n_skip) {
      const int q = static_cast<int>(i % Q2);
      const size_t pix = i / Q2;
      pr_st(dskip + i * 4, pr_ld(dout + pix * Ct + C1 + q * 4));
    } else {
      const size_t j = i - n_skip;
      const int q = static_cast<int>(j % Q1);
      size_t r = j / Q1;
      const int wx = static_cast<int>(r % Wx);
      r /= Wx;
      const int hx = static_cast<int>(r % Hx);
      const int n = static_cast<int>(r / Hx);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int dh = 0; dh < 2; ++dh)
#pragma unroll
        for (int dw = 0; dw < 2; ++dw) {
          const int h = 2 * hx + dh, w = 2 * wx + dw;
          if (h < Hs && w < Ws) {
            const float4 v = pr_ld(dout + ((static_cast<size_t>(n) * Hs + h) * Ws + w) * Ct + q * 4);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
          }
        }
      pr_st(dx + j * 4, acc);
    }
  }
}

// Weights for the data-gradient pass: wt[c][R-1-r][S-1-s][k] = w[k][r][s][c] (both KRSC-style
// channels-last memory). One 32x32 (k x c) tile of one filter tap per block, transposed through
// shared memory so that reads (along c) and writes (along k) are both coalesced. Replaces
// flip + permute + contiguous + copy (three passes, one of them a strided transpose) by one pass.
__global__ void __launch_bounds__(256)
dgrad_weight_kernel(const float* __restrict__ w, float* __restrict__ wt, int K, int C, int R,
                    int S) {
  __shared__ float tile[32][33];
  const int tap = blockIdx.z, r = tap / S, s = tap % S;
  const int tap_out = (R - 1 - r) * S + (S - 1 - s);
  const int c0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
  const int RS = R * S;
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int k = k0 + ty + j, c = c0 + tx;
    tile[ty + j][tx] = (k < K && c < C)
                           ? w[(static_cast<size_t>(k) * RS + tap) * C + c] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j, k = k0 + tx;
    if (c < C && k < K) wt[(static_cast<size_t>(c) * RS + tap_out) * K + k] = tile[tx][ty + j];
  }
}

static int pr_grid(size_t n) {
  size_t blocks = (n + kPrThreads - 1) / kPrThreads;
  const size_t cap = static_cast<size_t>(sm_count()) * 16;
  return static_cast<int>(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace tgb

using namespace tgb;

#define TGB_PR_CHECK(name)                                                           \
  do {                                                                               \
    cudaError_t e_ = cudaGetLastError();                                             \
    if (e_ != cudaSuccess) return set_error(name ": %s", cudaGetErrorString(e_));    \
    count_launch();                                                                  \
  } while (0)

extern "C" int tgb_maxpool2x2_fwd(const float* x, float* y, int N, int H, int W, int C,
                                  cudaStream_t stream) {
  if (C % 4) return set_error("maxpool2x2: C %% 4 != 0");
  const size_t n = static_cast<size_t>(N) * (H / 2) * (W / 2) * (C / 4);
  maxpool2_fwd_kernel<<<pr_grid(n), kPrThreads, 0, stream>>>(x, y, N, H, W, C);
  TGB_PR_CHECK("maxpool2x2_fwd");
  return 0;
}

extern "C" int tgb_maxpool2x2_bwd(const float* x, const float* dy, float* dx, int N, int H,
                                  int W, int C, cudaStream_t stream) {
  if (C % 4) return set_error("maxpool2x2: C %% 4 != 0");
  if ((H & 1) || (W & 1)) {
    cudaError_t e = cudaMemsetAsync(dx, 0, static_cast<size_t>(N) * H * W * C * 4, stream);
    if (e != cudaSuccess) return set_error("maxpool2x2_bwd memset: %s", cudaGetErrorString(e));
  }
  const size_t n = static_cast<size_t>(N) * (H / 2) * (W / 2) * (C / 4);
  maxpool2_bwd_kernel<<<pr_grid(n), kPrThreads, 0, stream>>>(x, dy, dx, N, H, W, C);
  TGB_PR_CHECK("maxpool2x2_bwd");
  return 0;
}

extern "C" int tgb_upsample2x_cat_fwd(const float* x, const float* skip, float* out, int N,
                                      int Hx, int Wx, int C1, int Hs, int Ws, int C2,
                                      cudaStream_t stream) {
  if ((C1 % 4) || (C2 % 4)) return set_error("upsample2x_cat: channels must be multiples of 4");
  if (Hs < 2 * Hx || Ws < 2 * Wx)
    return set_error("upsample2x_cat: skip (%dx%d) smaller than the upsampled map (%dx%d)", Hs, Ws,
                     2 * Hx, 2 * Wx);
  const size_t n = static_cast<size_t>(N) * Hs * Ws * ((C1 + C2) / 4);
  upcat_fwd_kernel<<<pr_grid(n), kPrThreads, 0, stream>>>(x, skip, out, N, Hx, Wx, C1, Hs, Ws, C2);
  TGB_PR_CHECK("upsample2x_cat_fwd");
  return 0;
}

extern "C" int tgb_upsample2x_cat_bwd(const float* dout, float* dx, float* dskip, int N, int Hx,
                                      int Wx, int C1, int Hs, int Ws, int C2,
                                      cudaStream_t stream) {
  if ((C1 % 4) || (C2 % 4)) return set_error("upsample2x_cat: channels must be multiples of 4");
  const size_t n = static_cast<size_t>(N) * Hs * Ws * (C2 / 4) +
                   static_cast<size_t>(N) * Hx * Wx * (C1 / 4);
  upcat_bwd_kernel<<<pr_grid(n), kPrThreads, 0, stream>>>(dout, dx, dskip, N, Hx, Wx, C1, Hs, Ws,
                                                          C2);
  TGB_PR_CHECK("upsample2x_cat_bwd");
  return 0;
}

extern "C" int tgb_dgrad_weight(const float* w, float* wt, int K, int C, int R, int S,
                                cudaStream_t stream) {
  if (K < 1 || C < 1 || R < 1 || S < 1 || R * S > 65535)
    return set_error("dgrad_weight: bad filter shape");
  const dim3 grid((C + 31) / 32, (K + 31) / 32, R * S);
  if (grid.y > 65535) return set_error("dgrad_weight: K too large");
  dgrad_weight_kernel<<<grid, 256, 0, stream>>>(w, wt, K, C, R, S);
  TGB_PR_CHECK("dgrad_weight");
  return 0;
}
