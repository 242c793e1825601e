// NHWC fp32 pooling family and residual-join kernels (ledger W14 / W15).
//
//  * tgb_pool2d_fwd / _bwd: square-window max / average pooling with stride and zero padding
//    (MaxPool2d(3,2,1) of the ResNet stem, AvgPool2d(3, s, 1, count_include_pad=False) and
//    MaxPool2d(2 / 3, s) of the AmoebaNet-D operation set, reference
//    benchmarks/models/amoebanet/operations.py:30-62, resnet/__init__.py:58). Backward is a gather
//    over the output windows that cover an input pixel: no index tensor, no atomics,
//    deterministic; the max route goes to the first maximum in row-major window order (ATen's
//    choice).
//  * tgb_global_avgpool_fwd / _bwd: AdaptiveAvgPool2d(1) in front of the classifiers.
//  * tgb_add_act_fwd / _bwd: z = relu?(a + b) - the residual join of a bottleneck
//    (resnet/bottleneck.py:70-76 adds, then a separate ReLU layer runs); backward is one masked
//    copy shared by both branches.
//  * tgb_bn_add_act_fwd / _bwd: z = relu?(BatchNorm(y) + r) in one pass from the coefficient table
//    tgb_bn_prepare made, and its backward (per-channel sums, dy, and the masked gradient for the
//    shortcut branch) - bn3 -> residual -> relu3 of every bottleneck as one kernel each way.
// All kernels are memory-bound: 16-byte accesses, grid-stride loops, <= 16 blocks per SM.
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdint>
#include "bn_reduce.cuh"
#include "tgb_internal.h"

namespace tgb {

constexpr int kPeThreads = 256;

__device__ __forceinline__ float4 pe_ld(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
__device__ __forceinline__ void pe_st(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

struct PoolGeom {
  int N, H, W, C, Ho, Wo, k, s, p, mode;   // mode 0 max, 1 avg over valid pixels, 2 avg over k*k
};

// One thread per (n, ho, wo, channel quad).
__global__ void __launch_bounds__(kPeThreads)
pool2d_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, PoolGeom g) {
  const int Q = g.C >> 2;
  const size_t total = static_cast<size_t>(g.N) * g.Ho * g.Wo * Q;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    size_t r = i / Q;
    const int wo = static_cast<int>(r % g.Wo);
    r /= g.Wo;
    const int ho = static_cast<int>(r % g.Ho);
    const int n = static_cast<int>(r / g.Ho);
    const int h0 = max(ho * g.s - g.p, 0), h1 = min(ho * g.s - g.p + g.k, g.H);
    const int w0 = max(wo * g.s - g.p, 0), w1 = min(wo * g.s - g.p + g.k, g.W);
    float4 acc = g.mode == 0 ? make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX)
                             : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int h = h0; h < h1; ++h)
      for (int w = w0; w < w1; ++w) {
        const float4 v = pe_ld(x + ((static_cast<size_t>(n) * g.H + h) * g.W + w) * g.C + q * 4);
        if (g.mode == 0) {
          acc.x = fmaxf(acc.x, v.x); acc.y = fmaxf(acc.y, v.y);
          acc.z = fmaxf(acc.z, v.z); acc.w = fmaxf(acc.w, v.w);
        } else {
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
    if (g.mode != 0) {
      const int cnt = g.mode == 1 ? (h1 - h0) * (w1 - w0) : g.k * g.k;
      const float inv = 1.f / static_cast<float>(max(cnt, 1));
      acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
    }
    pe_st(y + i * 4, acc);
  }
}

// One thread per (n, h, w, channel quad) of the INPUT: sums the contributions of every output
// window that covers the pixel.
__global__ void __launch_bounds__(kPeThreads)
pool2d_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                  float* __restrict__ dx, PoolGeom g) {
  const int Q = g.C >> 2;
  const size_t total = static_cast<size_t>(g.N) * g.H * g.W * Q;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    size_t r = i / Q;
    const int w = static_cast<int>(r % g.W);
    r /= g.W;
    const int h = static_cast<int>(r % g.H);
    const int n = static_cast<int>(r / g.H);
    // windows [o*s - p, o*s - p + k) containing h:  (h + p - k + 1) / s <= o <= (h + p) / s
    const int ho_lo = max((h + g.p - g.k + g.s) / g.s, 0);      // ceil((h+p-k+1)/s), numerator >= 0 ok
    const int ho_hi = min((h + g.p) / g.s, g.Ho - 1);
    const int wo_lo = max((w + g.p - g.k + g.s) / g.s, 0);
    const int wo_hi = min((w + g.p) / g.s, g.Wo - 1);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* xn = x + static_cast<size_t>(n) * g.H * g.W * g.C + q * 4;
    for (int ho = ho_lo; ho <= ho_hi; ++ho)
      for (int wo = wo_lo; wo <= wo_hi; ++wo) {
        const float4 gy =
            pe_ld(dy + ((static_cast<size_t>(n) * g.Ho + ho) * g.Wo + wo) * g.C + q * 4);
        const int h0 = max(ho * g.s - g.p, 0), h1 = min(ho * g.s - g.p + g.k, g.H);
        const int w0 = max(wo * g.s - g.p, 0), w1 = min(wo * g.s - g.p + g.k, g.W);
        if (g.mode == 0) {
          float4 best = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX);
          bool mx = false, my = false, mz = false, mw = false;    // "first maximum is this pixel"
          for (int hh = h0; hh < h1; ++hh)
            for (int ww = w0; ww < w1; ++ww) {
              const float4 v = pe_ld(xn + (static_cast<size_t>(hh) * g.W + ww) * g.C);
              const bool me = hh == h && ww == w;
              if (v.x > best.x) { best.x = v.x; mx = me; }
              if (v.y > best.y) { best.y = v.y; my = me; }
              if (v.z > best.z) { best.z = v.z; mz = me; }
              if (v.w > best.w) { best.w = v.w; mw = me; }
            }
          if (mx) acc.x += gy.x;
          if (my) acc.y += gy.y;
          if (mz) acc.z += gy.z;
          if (mw) acc.w += gy.w;
        } else {
          const int cnt = g.mode == 1 ? (h1 - h0) * (w1 - w0) : g.k * g.k;
          const float inv = 1.f / static_cast<float>(max(cnt, 1));
          acc.x += gy.x * inv; acc.y += gy.y * inv; acc.z += gy.z * inv; acc.w += gy.w * inv;
        }
      }
    pe_st(dx + i * 4, acc);
  }
}

// y[n, c] = mean over H*W of x[n, :, :, c]. One block per (n, 64-channel slab): threads are laid
// out 16 quads x 16 pixel lanes, partial sums meet in shared memory.
__global__ void __launch_bounds__(kPeThreads)
global_avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int HW, int C) {
  const int slabs = (C + 63) / 64;
  const int n = blockIdx.x / slabs, slab = blockIdx.x % slabs;
  const int ql = threadIdx.x & 15, lane = threadIdx.x >> 4;
  const int c = slab * 64 + ql * 4;
  __shared__ float4 part[16][16];
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c < C) {
    const float* base = x + static_cast<size_t>(n) * HW * C + c;
    for (int p = lane; p < HW; p += 16) {
      const float4 v = pe_ld(base + static_cast<size_t>(p) * C);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  part[lane][ql] = acc;
  __syncthreads();
  if (lane == 0 && c < C) {
    for (int l = 1; l < 16; ++l) {
      const float4 v = part[l][ql];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    const float inv = 1.f / static_cast<float>(HW);
    pe_st(y + static_cast<size_t>(n) * C + c, make_float4(acc.x * inv, acc.y * inv, acc.z * inv,
                                                          acc.w * inv));
  }
}

__global__ void __launch_bounds__(kPeThreads)
global_avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int N, int HW,
                          int C) {
  const int Q = C >> 2;
  const size_t total = static_cast<size_t>(N) * HW * Q;
  const float inv = 1.f / static_cast<float>(HW);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const size_t n = i / Q / HW;
    const float4 g = pe_ld(dy + n * C + q * 4);
    pe_st(dx + i * 4, make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv));
  }
}

// z = relu?(a + b)
__global__ void __launch_bounds__(kPeThreads)
add_act_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b,
                   float* __restrict__ z, size_t n4, int relu) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float4 u = pe_ld(a + i * 4), v = pe_ld(b + i * 4);
    float4 o = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
    if (relu) {
      o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f);
    }
    pe_st(z + i * 4, o);
  }
}

// d = dz where z > 0 else 0 (the gradient of both summands)
__global__ void __launch_bounds__(kPeThreads)
relu_mask_kernel(const float* __restrict__ dz, const float* __restrict__ z,
                 float* __restrict__ d, size_t n4) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float4 g = pe_ld(dz + i * 4), v = pe_ld(z + i * 4);
    pe_st(d + i * 4, make_float4(v.x > 0.f ? g.x : 0.f, v.y > 0.f ? g.y : 0.f,
                                 v.z > 0.f ? g.z : 0.f, v.w > 0.f ? g.w : 0.f));
  }
}

// z = relu?(y * a[c] + b[c] + r)      coef[c] = (a, b) from tgb_bn_prepare
__global__ void __launch_bounds__(kPeThreads)
bn_add_act_fwd_kernel(const float* __restrict__ y, const float* __restrict__ coef,
                      const float* __restrict__ res, float* __restrict__ z, size_t n4, int C,
                      int relu) {
  const int Q = C >> 2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const float4 c01 = pe_ld(coef + q * 8), c23 = pe_ld(coef + q * 8 + 4);
    const float4 v = pe_ld(y + i * 4), r = pe_ld(res + i * 4);
    float4 o;
    o.x = fmaf(v.x, c01.x, c01.y) + r.x; o.y = fmaf(v.y, c01.z, c01.w) + r.y;
    o.z = fmaf(v.z, c23.x, c23.y) + r.z; o.w = fmaf(v.w, c23.z, c23.w) + r.w;
    if (relu) {
      o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f);
    }
    pe_st(z + i * 4, o);
  }
}

// dy = a * (dn - S1/P - xhat * S2/P),  dres = dn
__global__ void __launch_bounds__(kPeThreads)
bn_add_act_bwd_apply_kernel(const float* __restrict__ dz, const float* __restrict__ y,
                            const float* __restrict__ z, const float* __restrict__ coef,
                            const float* __restrict__ saved, const float* __restrict__ sums,
                            float* __restrict__ dy, float* __restrict__ dres, size_t n4, int C,
                            float inv_p, int relu) {
  const int Q = C >> 2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const float4 g4 = pe_ld(dz + i * 4), v4 = pe_ld(y + i * 4), z4 = pe_ld(z + i * 4);
    const float4 c01 = pe_ld(coef + q * 8), c23 = pe_ld(coef + q * 8 + 4);
    const float4 m01 = pe_ld(saved + q * 8), m23 = pe_ld(saved + q * 8 + 4);
    const float4 t01 = pe_ld(sums + q * 8), t23 = pe_ld(sums + q * 8 + 4);
    const float g[4] = {g4.x, g4.y, g4.z, g4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
    const float zz[4] = {z4.x, z4.y, z4.z, z4.w};
    const float a[4] = {c01.x, c01.z, c23.x, c23.z};
    const float mu[4] = {m01.x, m01.z, m23.x, m23.z}, rs[4] = {m01.y, m01.w, m23.y, m23.w};
    const float S1[4] = {t01.x, t01.z, t23.x, t23.z}, S2[4] = {t01.y, t01.w, t23.y, t23.w};
    float o[4], d[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float dn = (!relu || zz[k] > 0.f) ? g[k] : 0.f;
      const float xh = (v[k] - mu[k]) * rs[k];
      d[k] = dn;
      o[k] = a[k] * (dn - S1[k] * inv_p - xh * S2[k] * inv_p);
    }
    pe_st(dy + i * 4, make_float4(o[0], o[1], o[2], o[3]));
    pe_st(dres + i * 4, make_float4(d[0], d[1], d[2], d[3]));
  }
}

static int pe_grid(size_t n) {
  size_t blocks = (n + kPeThreads - 1) / kPeThreads;
  const size_t cap = static_cast<size_t>(sm_count()) * 16;
  return static_cast<int>(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace tgb

using namespace tgb;

#define TGB_PE_CHECK(name)                                                           \
  do {                                                                               \
    cudaError_t e_ = cudaGetLastError();                                             \
    if (e_ != cudaSuccess) return set_error(name ": %s", cudaGetErrorString(e_));    \
    count_launch();                                                                  \
  } while (0)

static int pool_geom(PoolGeom* g, int N, int H, int W, int C, int k, int s, int p, int mode) {
  if (C % 4) return set_error("pool2d: C %% 4 != 0");
  if (k < 1 || k > 7 || s < 1 || p < 0 || 2 * p > k || mode < 0 || mode > 2)
    return set_error("pool2d: unsupported window k=%d s=%d p=%d mode=%d", k, s, p, mode);
  if (H + 2 * p < k || W + 2 * p < k) return set_error("pool2d: window larger than padded input");
  *g = PoolGeom{N, H, W, C, (H + 2 * p - k) / s + 1, (W + 2 * p - k) / s + 1, k, s, p, mode};
  return 0;
}

extern "C" int tgb_pool2d_fwd(const float* x, float* y, int N, int H, int W, int C, int k, int s,
                              int p, int mode, cudaStream_t stream) {
  PoolGeom g;
  if (int rc = pool_geom(&g, N, H, W, C, k, s, p, mode)) return rc;
  const size_t n = static_cast<size_t>(N) * g.Ho * g.Wo * (C / 4);
  pool2d_fwd_kernel<<<pe_grid(n), kPeThreads, 0, stream>>>(x, y, g);
  TGB_PE_CHECK("pool2d_fwd");
  return 0;
}

extern "C" int tgb_pool2d_bwd(const float* x, const float* dy, float* dx, int N, int H, int W,
                              int C, int k, int s, int p, int mode, cudaStream_t stream) {
  PoolGeom g;
  if (int rc = pool_geom(&g, N, H, W, C, k, s, p, mode)) return rc;
  const size_t n = static_cast<size_t>(N) * H * W * (C / 4);
  pool2d_bwd_kernel<<<pe_grid(n), kPeThreads, 0, stream>>>(x, dy, dx, g);
  TGB_PE_CHECK("pool2d_bwd");
  return 0;
}

extern "C" int tgb_global_avgpool_fwd(const float* x, float* y, int N, int HW, int C,
                                      cudaStream_t stream) {
  if (C % 4) return set_error("global_avgpool: C %% 4 != 0");
  const int slabs = (C + 63) / 64;
  global_avgpool_fwd_kernel<<<N * slabs, kPeThreads, 0, stream>>>(x, y, HW, C);
  TGB_PE_CHECK("global_avgpool_fwd");
  return 0;
}

extern "C" int tgb_global_avgpool_bwd(const float* dy, float* dx, int N, int HW, int C,
                                      cudaStream_t stream) {
  if (C % 4) return set_error("global_avgpool: C %% 4 != 0");
  const size_t n = static_cast<size_t>(N) * HW * (C / 4);
  global_avgpool_bwd_kernel<<<pe_grid(n), kPeThreads, 0, stream>>>(dy, dx, N, HW, C);
  TGB_PE_CHECK("global_avgpool_bwd");
  return 0;
}

extern "C" int tgb_add_act_fwd(const float* a, const float* b, float* z, long long n, int relu,
                               cudaStream_t stream) {
  if (n % 4) return set_error("add_act: element count %% 4 != 0");
  add_act_fwd_kernel<<<pe_grid(n / 4), kPeThreads, 0, stream>>>(a, b, z, n / 4, relu);
  TGB_PE_CHECK("add_act_fwd");
  return 0;
}

extern "C" int tgb_relu_mask(const float* dz, const float* z, float* d, long long n,
                             cudaStream_t stream) {
  if (n % 4) return set_error("relu_mask: element count %% 4 != 0");
  relu_mask_kernel<<<pe_grid(n / 4), kPeThreads, 0, stream>>>(dz, z, d, n / 4);
  TGB_PE_CHECK("relu_mask");
  return 0;
}

extern "C" int tgb_bn_add_act_fwd(const float* y, const float* coef, const float* res, float* z,
                                  int P, int C, int relu, cudaStream_t stream) {
  if (C % 4) return set_error("bn_add_act_fwd: C %% 4 != 0");
  const size_t n4 = static_cast<size_t>(P) * (C / 4);
  bn_add_act_fwd_kernel<<<pe_grid(n4), kPeThreads, 0, stream>>>(y, coef, res, z, n4, C, relu);
  TGB_PE_CHECK("bn_add_act_fwd");
  return 0;
}

// sums [C,2] must be zero on entry; on return sums[c] = (sum dn, sum dn * xhat) = (dbeta, dgamma).
extern "C" int tgb_bn_add_act_bwd(const float* dz, const float* y, const float* z,
                                  const float* coef, const float* saved, float* sums, float* dy,
                                  float* dres, int P, int C, int relu, cudaStream_t stream) {
  if (C % 4) return set_error("bn_add_act_bwd: C %% 4 != 0");
  if (2 * C * sizeof(float) > 48 * 1024) return set_error("bn_add_act_bwd: C too large");
  const int slabs = bn_reduce_slabs(P, sm_count());
  bn_bwd_reduce_generic_kernel<true><<<slabs, kBnReduceThreads, 2 * C * sizeof(float), stream>>>(
      dz, y, z, coef, saved, sums, P, C, slabs, relu);
  TGB_PE_CHECK("bn_add_act_bwd_reduce");
  const size_t n4 = static_cast<size_t>(P) * (C / 4);
  bn_add_act_bwd_apply_kernel<<<pe_grid(n4), kPeThreads, 0, stream>>>(
      dz, y, z, coef, saved, sums, dy, dres, n4, C, 1.f / static_cast<float>(P), relu);
  TGB_PE_CHECK("bn_add_act_bwd_apply");
  return 0;
}
