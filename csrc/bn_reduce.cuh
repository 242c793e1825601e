// Per-channel backward sums of a training-mode BatchNorm(+ReLU)(+residual) layer over NHWC fp32:
//   sums[c] += (sum dn, sum dn * xhat),  dn = dz masked by the ReLU,  xhat = (y - mean) * rstd.
// MASK_FROM_Z: the mask is z > 0 (z = layer output; needed when a residual was added before the
// ReLU), otherwise it is re-derived as y * a + b > 0 from the coefficient table.
//
// One block per slab of pixels. No per-element integer division: a thread's channel quad is fixed
//   (a) Q <= 256 : quad = tid % Q, the floor(256/Q) pixel lanes of a block walk the slab together
//                  (threads beyond lanes*Q idle when Q does not divide 256);
//   (b) Q > 256  : quads tid, tid+256, ...; the thread walks the slab's pixels alone.
// Rows are read as 16-byte vectors, coalesced across the quads of a pixel. Partial sums meet in shared
// memory, then one global atomic per channel and block.
#pragma once
#include <cuda_runtime.h>

namespace tgb {

constexpr int kBnReduceThreads = 256;

__device__ __forceinline__ float4 bnr_ld(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}

template <bool MASK_FROM_Z>
struct BnReduceAcc {
  float mu[4], rs[4], ca[4], cb[4];
  float s1[4], s2[4];
  __device__ __forceinline__ void init(const float* __restrict__ coef,
                                       const float* __restrict__ saved, int q) {
    const float4 m01 = bnr_ld(saved + q * 8), m23 = bnr_ld(saved + q * 8 + 4);
    mu[0] = m01.x; rs[0] = m01.y; mu[1] = m01.z; rs[1] = m01.w;
    mu[2] = m23.x; rs[2] = m23.y; mu[3] = m23.z; rs[3] = m23.w;
    if (!MASK_FROM_Z) {
      const float4 c01 = bnr_ld(coef + q * 8), c23 = bnr_ld(coef + q * 8 + 4);
      ca[0] = c01.x; cb[0] = c01.y; ca[1] = c01.z; cb[1] = c01.w;
      ca[2] = c23.x; cb[2] = c23.y; ca[3] = c23.z; cb[3] = c23.w;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) s1[k] = s2[k] = 0.f;
  }
  __device__ __forceinline__ void add(const float* __restrict__ dz, const float* __restrict__ y,
                                      const float* __restrict__ z, size_t off, int relu) {
    const float4 g4 = bnr_ld(dz + off), v4 = bnr_ld(y + off);
    const float g[4] = {g4.x, g4.y, g4.z, g4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
    float m[4];
    if (MASK_FROM_Z) {
      const float4 z4 = bnr_ld(z + off);
      m[0] = z4.x; m[1] = z4.y; m[2] = z4.z; m[3] = z4.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) m[k] = fmaf(v[k], ca[k], cb[k]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float dn = (!relu || m[k] > 0.f) ? g[k] : 0.f;
      s1[k] += dn;
      s2[k] += dn * ((v[k] - mu[k]) * rs[k]);
    }
  }
};

template <bool MASK_FROM_Z>
__global__ void __launch_bounds__(kBnReduceThreads)
bn_bwd_reduce_generic_kernel(const float* __restrict__ dz, const float* __restrict__ y,
                             const float* __restrict__ z, const float* __restrict__ coef,
                             const float* __restrict__ saved, float* __restrict__ sums, int P,
                             int C, int slabs, int relu) {
  const int Q = C >> 2;
  const int per = (P + slabs - 1) / slabs;
  const int p0 = blockIdx.x * per, p1 = min(P, p0 + per);
  extern __shared__ float sh[];
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  BnReduceAcc<MASK_FROM_Z> acc;
  if (Q <= kBnReduceThreads) {                          // (a)
    const int lanes = kBnReduceThreads / Q;
    const int q = threadIdx.x % Q, lane = threadIdx.x / Q;
    if (lane < lanes) {
      acc.init(coef, saved, q);
#pragma unroll 4
      for (int p = p0 + lane; p < p1; p += lanes)
        acc.add(dz, y, z, static_cast<size_t>(p) * C + q * 4, relu);
      float* d = sh + q * 8;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        atomicAdd(d + 2 * k, acc.s1[k]);
        atomicAdd(d + 2 * k + 1, acc.s2[k]);
      }
    }
  } else {                                              // (b)
    for (int q = threadIdx.x; q < Q; q += kBnReduceThreads) {
      acc.init(coef, saved, q);
#pragma unroll 4
      for (int p = p0; p < p1; ++p) acc.add(dz, y, z, static_cast<size_t>(p) * C + q * 4, relu);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        sh[q * 8 + 2 * k] = acc.s1[k];
        sh[q * 8 + 2 * k + 1] = acc.s2[k];
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) atomicAdd(sums + i, sh[i]);
}

// Slab count: enough blocks to fill the machine (8 resident blocks of 256 threads per SM), at
// least 32 pixels per block so the per-block epilogue (2C global atomics) stays small.
inline int bn_reduce_slabs(int P, int sms) {
  int slabs = sms * 8;
  if (slabs > (P + 31) / 32) slabs = (P + 31) / 32;
  return slabs < 1 ? 1 : slabs;
}

}  // namespace tgb
