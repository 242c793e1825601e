// Fused normalisation / activation kernels on NHWC fp32 tensors (memory-bound: 16-byte accesses,
// one read + one write per element, statistics taken from the convolution epilogue).
//
//  * per-(image, channel) or per-channel moment reduction (sum, sum of squares | sum dn, sum dn*u)
//  * InstanceNorm2d + Dropout2d + LeakyReLU forward / backward     (U-Net cell)
//  * BatchNorm2d (training) + optional residual add + ReLU forward / backward (ResNet /
//    AmoebaNet), with the DeferredBatchNorm accumulators fed from the same
//    statistics (reference torchgpipe/batchnorm.py:45-59 needs two extra full reductions).
//
// Layout convention: a tensor is [G][P][C] with G "groups" (images for InstanceNorm, 1 for
// BatchNorm with P = N*H*W), C % 4 == 0.
#include <cuda_runtime.h>
#include <cstdint>
#include "bn_reduce.cuh"
#include "tgb_internal.h"

namespace tgb {

constexpr int kEwThreads = 256;

__device__ __forceinline__ float4 ld4(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// ------------------------------------------------------------------ moments
// MODE 0: a = x,            out += (sum x, sum x^2)
// MODE 1: a = dz, b = z:    dn = dz * (z > 0 ? 1 : slope); u = z > 0 ? z : z / slope
//                           out += (sum dn, sum dn*u)            (norm+leaky backward reduction)
template <int MODE>
__global__ void __launch_bounds__(kEwThreads)
moments_kernel(const float* __restrict__ a, const float* __restrict__ b,
               float* __restrict__ out, int P, int C, int slabs, float slope) {
  // grid: (slabs, G)
  const int g = blockIdx.y;
  const int Q = C >> 2;
  const size_t base = static_cast<size_t>(g) * P * C;
  const int pix_per_slab = (P + slabs - 1) / slabs;
  const int p_begin = blockIdx.x * pix_per_slab;
  const int p_end = min(P, p_begin + pix_per_slab);
  extern __shared__ float sh[];      // [C][2] block accumulators
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();

  const float inv_slope = 1.f / slope;
  // No per-element integer division: a thread's channel quad is fixed. Q <= 256: quad = tid % Q
  // and floor(256/Q) pixel lanes walk the slab together; Q > 256: the thread owns quads
  // tid, tid+256, ... and walks the slab's pixels alone. Rows are read as coalesced float4.
  const int lanes = Q <= kEwThreads ? kEwThreads / Q : 1;
  const int lane = Q <= kEwThreads ? threadIdx.x / Q : 0;
  const int q_step = Q <= kEwThreads ? Q : kEwThreads;      // one pass when Q <= 256
  if (lane < lanes) {
    for (int q = Q <= kEwThreads ? threadIdx.x % Q : threadIdx.x; q < Q; q += q_step) {
      float4 s1 = make_float4(0, 0, 0, 0), s2 = make_float4(0, 0, 0, 0);
#pragma unroll 4
      for (int p = p_begin + lane; p < p_end; p += lanes) {
        const size_t off = base + static_cast<size_t>(p) * C + q * 4;
        const float4 va = ld4(a + off);
        if (MODE == 0) {
          s1.x += va.x; s1.y += va.y; s1.z += va.z; s1.w += va.w;
          s2.x = fmaf(va.x, va.x, s2.x); s2.y = fmaf(va.y, va.y, s2.y);
          s2.z = fmaf(va.z, va.z, s2.z); s2.w = fmaf(va.w, va.w, s2.w);
        } else if (MODE == 1) {
          const float4 vz = ld4(b + off);
#define TGB_ACC1(f)                                              \
  {                                                              \
    const float dn = vz.f > 0.f ? va.f : va.f * slope;           \
    const float u = vz.f > 0.f ? vz.f : vz.f * inv_slope;        \
    s1.f += dn;                                                  \
    s2.f = fmaf(dn, u, s2.f);                                    \
  }
          TGB_ACC1(x) TGB_ACC1(y) TGB_ACC1(z) TGB_ACC1(w)
#undef TGB_ACC1
        }
      }
      float* d = sh + q * 8;
      atomicAdd(d + 0, s1.x); atomicAdd(d + 1, s2.x);
      atomicAdd(d + 2, s1.y); atomicAdd(d + 3, s2.y);
      atomicAdd(d + 4, s1.z); atomicAdd(d + 5, s2.z);
      atomicAdd(d + 6, s1.w); atomicAdd(d + 7, s2.w);
      if (Q <= kEwThreads) break;
    }
  }
  __syncthreads();
  float* o = out + static_cast<size_t>(g) * C * 2;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) atomicAdd(o + i, sh[i]);
}

// ------------------------------------------------------------------ InstanceNorm cell
// coef[g][c] = ms / sqrt(ms^2 * var + eps), mean[g][c]; ms = dropout scale (0 or 1/(1-p)).
// drop_p > 0: `ms` holds uniform random numbers in [0,1) and the Dropout2d mask is made here
// (channel kept when u >= p, scaled by 1/(1-p)) instead of by separate compare / scale kernels.
__global__ void in_prepare_kernel(const float* __restrict__ stats, const float* __restrict__ ms,
                                  float* __restrict__ mean_coef, int GC, float inv_p, float eps,
                                  float drop_p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= GC) return;
  const float s1 = stats[2 * i], s2 = stats[2 * i + 1];
  const float mu = s1 * inv_p;
  const float var = fmaxf(s2 * inv_p - mu * mu, 0.f);
  float m = ms ? ms[i] : 1.f;
  if (ms && drop_p > 0.f) m = m >= drop_p ? 1.f / (1.f - drop_p) : 0.f;
  mean_coef[2 * i] = mu;
  mean_coef[2 * i + 1] = m * rsqrtf(m * m * var + eps);
}

// z = leaky(coef * (y - mean))
__global__ void __launch_bounds__(kEwThreads)
in_apply_kernel(const float* __restrict__ y, const float* __restrict__ mean_coef,
                float* __restrict__ z, int P, int C, float slope) {
  const int g = blockIdx.y;
  const int Q = C >> 2;
  const size_t n4 = static_cast<size_t>(P) * Q;
  const size_t base = static_cast<size_t>(g) * P * C;
  const float* mc = mean_coef + static_cast<size_t>(g) * C * 2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const float4 m01 = ld4(mc + q * 8);        // (mean0, coef0, mean1, coef1)
    const float4 m23 = ld4(mc + q * 8 + 4);
    float4 v = ld4(y + base + i * 4);
    v.x = (v.x - m01.x) * m01.y; v.y = (v.y - m01.z) * m01.w;
    v.z = (v.z - m23.x) * m23.y; v.w = (v.w - m23.z) * m23.w;
    v.x = v.x > 0.f ? v.x : v.x * slope; v.y = v.y > 0.f ? v.y : v.y * slope;
    v.z = v.z > 0.f ? v.z : v.z * slope; v.w = v.w > 0.f ? v.w : v.w * slope;
    st4(z + base + i * 4, v);
  }
}

// dy = coef * (dn - S1/P - u * S2/P), dn and u re-derived from (dz, z)
__global__ void __launch_bounds__(kEwThreads)
in_backward_kernel(const float* __restrict__ dz, const float* __restrict__ z,
                   const float* __restrict__ mean_coef, const float* __restrict__ sums,
                   float* __restrict__ dy, int P, int C, float slope) {
  const int g = blockIdx.y;
  const int Q = C >> 2;
  const size_t n4 = static_cast<size_t>(P) * Q;
  const size_t base = static_cast<size_t>(g) * P * C;
  const float* mc = mean_coef + static_cast<size_t>(g) * C * 2;
  const float* sm = sums + static_cast<size_t>(g) * C * 2;
  const float inv_p = 1.f / static_cast<float>(P);
  const float inv_slope = 1.f / slope;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const float4 m01 = ld4(mc + q * 8), m23 = ld4(mc + q * 8 + 4);
    const float4 s01 = ld4(sm + q * 8), s23 = ld4(sm + q * 8 + 4);
    const float4 g4 = ld4(dz + base + i * 4);
    const float4 z4 = ld4(z + base + i * 4);
    float4 o;
#define TGB_BWD(f, coef, S1, S2)                                 \
  {                                                              \
    const float dn = z4.f > 0.f ? g4.f : g4.f * slope;           \
    const float u = z4.f > 0.f ? z4.f : z4.f * inv_slope;        \
    o.f = (coef) * (dn - (S1) * inv_p - u * (S2) * inv_p);       \
  }
    TGB_BWD(x, m01.y, s01.x, s01.y)
    TGB_BWD(y, m01.w, s01.z, s01.w)
    TGB_BWD(z, m23.y, s23.x, s23.y)
    TGB_BWD(w, m23.w, s23.z, s23.w)
#undef TGB_BWD
    st4(dy + base + i * 4, o);
  }
}

// ------------------------------------------------------------------ BatchNorm (training)
// From per-channel (sum, sum of squares) over P elements: coef[c] = (a, b) with
// y_norm = y * a + b, a = gamma * rstd, b = beta - mean * a; saved[c] = (mean, rstd) for backward.
// Optionally folds the batch statistics into the running estimates exactly like nn.BatchNorm2d
// (unbiased variance for running_var, momentum < 0 selects the cumulative average 1/count).
// `tracked` (optional): nn.BatchNorm2d's num_batches_tracked. With it the kernel runs as ONE
// block: every thread reads the old count, then thread 0 adds one - the counter update costs no
// launch of its own, and momentum < 0 (cumulative average) uses 1/(count+1) without a host read.
__global__ void bn_prepare_kernel(const float* __restrict__ stats, const float* __restrict__ gamma,
                                  const float* __restrict__ beta, float* __restrict__ coef,
                                  float* __restrict__ saved, float* __restrict__ running_mean,
                                  float* __restrict__ running_var, int C, float inv_p,
                                  float unbias, float eps, float momentum,
                                  long long* __restrict__ tracked) {
  if (tracked != nullptr) {
    const long long count = *tracked + 1;
    if (momentum < 0.f) momentum = 1.f / static_cast<float>(count);
    __syncthreads();
    if (threadIdx.x == 0) *tracked = count;
  }
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < C; c += gridDim.x * blockDim.x) {
  const float mu = stats[2 * c] * inv_p;
  const float var = fmaxf(stats[2 * c + 1] * inv_p - mu * mu, 0.f);
  const float rstd = rsqrtf(var + eps);
  const float g = gamma ? gamma[c] : 1.f;
  const float a = g * rstd;
  coef[2 * c] = a;
  coef[2 * c + 1] = (beta ? beta[c] : 0.f) - mu * a;
  saved[2 * c] = mu;
  saved[2 * c + 1] = rstd;
  if (running_mean != nullptr) {
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mu;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * var * unbias;
  }
  }
}

// z = act(y * a + b), act = ReLU or identity
__global__ void __launch_bounds__(kEwThreads)
bn_apply_kernel(const float* __restrict__ y, const float* __restrict__ coef,
                float* __restrict__ z, size_t n4, int C, int relu) {
  const int Q = C >> 2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const float4 c01 = ld4(coef + q * 8), c23 = ld4(coef + q * 8 + 4);
    float4 v = ld4(y + i * 4);
    v.x = fmaf(v.x, c01.x, c01.y); v.y = fmaf(v.y, c01.z, c01.w);
    v.z = fmaf(v.z, c23.x, c23.y); v.w = fmaf(v.w, c23.z, c23.w);
    if (relu) {
      v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
    }
    st4(z + i * 4, v);
  }
}

// (the backward reduction lives in bn_reduce.cuh, shared with the residual variant)

// dy = a * (dn - S1/P - xhat * S2/P)
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_apply_kernel(const float* __restrict__ dz, const float* __restrict__ y,
                    const float* __restrict__ coef, const float* __restrict__ saved,
                    const float* __restrict__ sums, float* __restrict__ dy, size_t n4, int C,
                    float inv_p, int relu) {
  const int Q = C >> 2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(i % Q);
    const float4 g = ld4(dz + i * 4), v = ld4(y + i * 4);
    const float4 c01 = ld4(coef + q * 8), c23 = ld4(coef + q * 8 + 4);
    const float4 s01 = ld4(saved + q * 8), s23 = ld4(saved + q * 8 + 4);
    const float4 m01 = ld4(sums + q * 8), m23 = ld4(sums + q * 8 + 4);
    float4 o;
#define TGB_BN_BWD(f, ca, cb, mu, rs, S1, S2)                           \
  {                                                                     \
    const float dn = (!relu || fmaf(v.f, ca, cb) > 0.f) ? g.f : 0.f;    \
    const float xh = (v.f - (mu)) * (rs);                               \
    o.f = (ca) * (dn - (S1) * inv_p - xh * (S2) * inv_p);               \
  }
    TGB_BN_BWD(x, c01.x, c01.y, s01.x, s01.y, m01.x, m01.y)
    TGB_BN_BWD(y, c01.z, c01.w, s01.z, s01.w, m01.z, m01.w)
    TGB_BN_BWD(z, c23.x, c23.y, s23.x, s23.y, m23.x, m23.y)
    TGB_BN_BWD(w, c23.z, c23.w, s23.z, s23.w, m23.z, m23.w)
#undef TGB_BN_BWD
    st4(dy + i * 4, o);
  }
}

static int ew_grid(size_t n4) {
  size_t blocks = (n4 + kEwThreads - 1) / kEwThreads;
  const size_t cap = static_cast<size_t>(sm_count()) * 16;
  return static_cast<int>(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace tgb

using namespace tgb;

#define TGB_CHECK_LAUNCH(name)                                                       \
  do {                                                                               \
    cudaError_t e_ = cudaGetLastError();                                             \
    if (e_ != cudaSuccess) return set_error(name ": %s", cudaGetErrorString(e_));    \
    count_launch();                                                                  \
  } while (0)

extern "C" int tgb_moments(const float* x, float* out, int G, int P, int C,
                           cudaStream_t stream) {
  if (C % 4) return set_error("moments: C %% 4 != 0");
  int slabs = (sm_count() * 4 + G - 1) / G;
  if (slabs > (P + 63) / 64) slabs = (P + 63) / 64;
  if (slabs < 1) slabs = 1;
  dim3 grid(slabs, G);
  moments_kernel<0><<<grid, kEwThreads, 2 * C * sizeof(float), stream>>>(
      x, nullptr, out, P, C, slabs, 1.f);
  TGB_CHECK_LAUNCH("moments");
  return 0;
}

extern "C" int tgb_in_leaky_bwd_sums(const float* dz, const float* z, float* out, int G, int P,
                                     int C, float slope, cudaStream_t stream) {
  if (C % 4) return set_error("in_leaky_bwd_sums: C %% 4 != 0");
  int slabs = (sm_count() * 4 + G - 1) / G;
  if (slabs > (P + 63) / 64) slabs = (P + 63) / 64;
  if (slabs < 1) slabs = 1;
  dim3 grid(slabs, G);
  moments_kernel<1><<<grid, kEwThreads, 2 * C * sizeof(float), stream>>>(
      dz, z, out, P, C, slabs, slope);
  TGB_CHECK_LAUNCH("in_leaky_bwd_sums");
  return 0;
}

extern "C" int tgb_in_prepare(const float* stats, const float* drop_scale, float* mean_coef,
                              int G, int P, int C, float eps, cudaStream_t stream) {
  const int gc = G * C;
  in_prepare_kernel<<<(gc + 255) / 256, 256, 0, stream>>>(stats, drop_scale, mean_coef, gc,
                                                          1.f / static_cast<float>(P), eps, 0.f);
  TGB_CHECK_LAUNCH("in_prepare");
  return 0;
}

extern "C" int tgb_in_prepare_dropout(const float* stats, const float* uniform, float drop_p,
                                      float* mean_coef, int G, int P, int C, float eps,
                                      cudaStream_t stream) {
  if (!(drop_p >= 0.f && drop_p < 1.f)) return set_error("in_prepare_dropout: p must be in [0,1)");
  const int gc = G * C;
  in_prepare_kernel<<<(gc + 255) / 256, 256, 0, stream>>>(
      stats, drop_p > 0.f ? uniform : nullptr, mean_coef, gc, 1.f / static_cast<float>(P), eps,
      drop_p);
  TGB_CHECK_LAUNCH("in_prepare_dropout");
  return 0;
}

extern "C" int tgb_in_leaky_fwd(const float* y, const float* mean_coef, float* z, int G, int P,
                                int C, float slope, cudaStream_t stream) {
  if (C % 4) return set_error("in_leaky_fwd: C %% 4 != 0");
  const size_t n4 = static_cast<size_t>(P) * (C / 4);
  int gx = ew_grid(n4);
  if (G > 1) gx = (gx + G - 1) / G > 0 ? (gx + G - 1) / G : 1;
  dim3 grid(gx, G);
  in_apply_kernel<<<grid, kEwThreads, 0, stream>>>(y, mean_coef, z, P, C, slope);
  TGB_CHECK_LAUNCH("in_leaky_fwd");
  return 0;
}

extern "C" int tgb_in_leaky_bwd(const float* dz, const float* z, const float* mean_coef,
                                const float* sums, float* dy, int G, int P, int C, float slope,
                                cudaStream_t stream) {
  if (C % 4) return set_error("in_leaky_bwd: C %% 4 != 0");
  const size_t n4 = static_cast<size_t>(P) * (C / 4);
  int gx = ew_grid(n4);
  if (G > 1) gx = (gx + G - 1) / G > 0 ? (gx + G - 1) / G : 1;
  dim3 grid(gx, G);
  in_backward_kernel<<<grid, kEwThreads, 0, stream>>>(dz, z, mean_coef, sums, dy, P, C, slope);
  TGB_CHECK_LAUNCH("in_leaky_bwd");
  return 0;
}

extern "C" int tgb_bn_prepare(const float* stats, const float* gamma, const float* beta,
                              float* coef, float* saved, float* running_mean,
                              float* running_var, int P, int C, float eps, float momentum,
                              cudaStream_t stream) {
  const float unbias = P > 1 ? static_cast<float>(P) / static_cast<float>(P - 1) : 1.f;
  bn_prepare_kernel<<<(C + 255) / 256, 256, 0, stream>>>(stats, gamma, beta, coef, saved,
                                                        running_mean, running_var, C,
                                                        1.f / static_cast<float>(P), unbias, eps,
                                                        momentum, nullptr);
  TGB_CHECK_LAUNCH("bn_prepare");
  return 0;
}

extern "C" int tgb_bn_prepare_tracked(const float* stats, const float* gamma, const float* beta,
                                      float* coef, float* saved, float* running_mean,
                                      float* running_var, long long* num_batches_tracked, int P,
                                      int C, float eps, float momentum, cudaStream_t stream) {
  const float unbias = P > 1 ? static_cast<float>(P) / static_cast<float>(P - 1) : 1.f;
  if (num_batches_tracked == nullptr && momentum < 0.f)
    return set_error("bn_prepare_tracked: the cumulative average needs num_batches_tracked");
  bn_prepare_kernel<<<1, 1024, 0, stream>>>(stats, gamma, beta, coef, saved, running_mean,
                                            running_var, C, 1.f / static_cast<float>(P), unbias,
                                            eps, momentum, num_batches_tracked);
  TGB_CHECK_LAUNCH("bn_prepare_tracked");
  return 0;
}

extern "C" int tgb_bn_act_fwd(const float* y, const float* coef, float* z, int P, int C, int relu,
                              cudaStream_t stream) {
  if (C % 4) return set_error("bn_act_fwd: C %% 4 != 0");
  const size_t n4 = static_cast<size_t>(P) * (C / 4);
  bn_apply_kernel<<<ew_grid(n4), kEwThreads, 0, stream>>>(y, coef, z, n4, C, relu);
  TGB_CHECK_LAUNCH("bn_act_fwd");
  return 0;
}

extern "C" int tgb_bn_act_bwd(const float* dz, const float* y, const float* coef,
                              const float* saved, float* sums, float* dy, int P, int C, int relu,
                              cudaStream_t stream) {
  if (C % 4) return set_error("bn_act_bwd: C %% 4 != 0");
  const int slabs = bn_reduce_slabs(P, sm_count());
  bn_bwd_reduce_generic_kernel<false><<<slabs, kBnReduceThreads, 2 * C * sizeof(float), stream>>>(
      dz, y, nullptr, coef, saved, sums, P, C, slabs, relu);
  TGB_CHECK_LAUNCH("bn_bwd_reduce");
  const size_t n4 = static_cast<size_t>(P) * (C / 4);
  bn_bwd_apply_kernel<<<ew_grid(n4), kEwThreads, 0, stream>>>(
      dz, y, coef, saved, sums, dy, n4, C, 1.f / static_cast<float>(P), relu);
  TGB_CHECK_LAUNCH("bn_bwd_apply");
  return 0;
}
