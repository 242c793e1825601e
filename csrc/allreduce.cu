// Gradient all-reduce across data-parallel replicas of a pipeline stage, fused with the gradient
// scale (and an optional bf16 cast), written directly against NVLink peer memory.
//
// Every replica keeps its flat fp32 gradient bucket in symmetric memory: each rank can address
// every peer's bucket (P2P over NVLink 5 / NVSwitch) and, where the fabric supports it, one
// multicast address that targets all replicas at once (NVLS). One kernel does everything:
//
//   barrier            all replicas have finished writing their gradients
//   reduce + scatter   rank r sums slice r of all replicas
//                        P2P     : ld.global from each peer's bucket
//                        NVLS    : multimem.ld_reduce.add (the switch adds, one load returns the sum)
//                      multiplies by `scale` (1 / replicas, loss scale, ...) - the fused part -
//                      and hands the finished values out in the same pass:
//                        P2P     : st.global into slice r of every replica's bucket (push)
//                        NVLS    : multimem.st (the switch replicates the store)
//                      and optionally writes a bf16 copy locally
//   barrier            every replica holds the full reduced bucket; buffers may be reused
//
// In place and race free without a barrier in the middle: element e of slice r is read - in any
// bucket - by rank r alone, and by exactly the thread that then overwrites it everywhere, after
// its loads have returned. Loads (inbound) and stores (outbound) keep both directions of the
// links busy at once, and posted stores do not pay the NVLink round trip. Bytes on the wire per
// direction per GPU: P2P 2N(R-1)/R; NVLS N(R+1)/R (the whole bucket goes up to the switch once
// for the in-switch reduction, every reduced slice comes down to every GPU once) - the roofline
// the benchmarks report against. NVLS therefore wins from three replicas up, P2P at two.
// (The earlier pull-only version - reduce, barrier, then every rank pulls the finished slices -
// is kept behind TGB_ALLREDUCE_PULL=1 for A/B timing.)
//
// Cross-GPU barriers use a per-rank signal pad (a second, zero-initialised symmetric allocation
// owned by the caller, tgb_allreduce_pad_bytes() bytes): block b of rank r
// stores a monotonically increasing ticket into slot [b][r] of every peer's pad and spins until
// its own slots [b][*] reach the ticket (st.release.sys / ld.acquire.sys). All blocks of all
// ranks are resident (grid <= SM count, 1 block/SM), so the spin cannot starve a peer.
//
// Reference: absent (SURVEY 2.3a last row; the reference has no data parallelism). Baseline for
// the numbers is NCCL all_reduce on the same bucket.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdlib>
#include "tgb_internal.h"

namespace tgb {

constexpr int kArMaxWorld = 8;
constexpr int kArThreads = 512;
constexpr int kArPadBase = 0;         // first uint32 slot of the signal pad used by this kernel

struct ArParams {
  float* bufs[kArMaxWorld];           // each replica's bucket (peer-mapped)
  uint32_t* pads[kArMaxWorld];        // each replica's signal pad (peer-mapped)
  float* mc;                          // multicast address of the bucket or nullptr
  __nv_bfloat16* cast_out;            // optional local bf16 copy of the result
  unsigned long long n;               // elements (multiple of 4 * world)
  float scale;
  int rank, world;
  uint32_t ticket;                    // 3 * launch index; the kernel uses ticket+1..ticket+3
  unsigned long long timeout_cycles;  // spin budget before giving up (sets *error)
  int* error;                         // device flag: 1 if a barrier timed out
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Block-wise barrier across replicas. Returns false on timeout.
__device__ __forceinline__ bool replica_barrier(const ArParams& p, uint32_t value) {
  __syncthreads();
  bool ok = true;
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(p.pads[peer] + kArPadBase + blockIdx.x * kArMaxWorld + p.rank, value);
    const uint32_t* mine = p.pads[p.rank] + kArPadBase + blockIdx.x * kArMaxWorld + peer;
    const long long start = clock64();
    // tickets only grow; compare with wrap-around-safe signed difference
    while (static_cast<int32_t>(ld_acquire_sys(mine) - value) < 0) {
      if (static_cast<unsigned long long>(clock64() - start) > p.timeout_cycles) {
        ok = false;
        *p.error = 1;
        break;
      }
      __nanosleep(64);
    }
  }
  ok = __syncthreads_and(ok);
  return ok;
}

// Plain (weak) vector accesses: the compiler may batch several loads per thread, which is what
// hides the ~2 us NVLink round trip. Ordering against the peers comes from the barriers
// (__syncthreads + fence + release/acquire), which the compiler does not move accesses across.
__device__ __forceinline__ float4 ld_f4(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
__device__ __forceinline__ void st_f4(float* p, float4 v) {
  *reinterpret_cast<float4*>(p) = v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

template <bool NVLS, bool PUSH>
__global__ void __launch_bounds__(kArThreads, 1) allreduce_scale_kernel(const ArParams p) {
  if (!replica_barrier(p, p.ticket + 1)) return;

  const unsigned long long slice = p.n / p.world;           // elements per rank, multiple of 4
  const unsigned long long begin = slice * p.rank;
  const unsigned long long vecs = slice / 4;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * blockDim.x;
  float* mine = p.bufs[p.rank];

  // ---- reduce-scatter (+ scale): this rank owns [begin, begin + slice)
#pragma unroll 4
  for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * blockDim.x + threadIdx.x;
       i < vecs; i += stride) {
    const unsigned long long e = begin + i * 4;
    float4 acc;
    if (NVLS) {
      acc = multimem_ld_reduce_f4(p.mc + e);
    } else {
      acc = ld_f4(mine + e);
#pragma unroll
      for (int k = 1; k < kArMaxWorld; ++k) {
        if (k >= p.world) break;
        const int peer = (p.rank + k) % p.world;             // spread the first reads over peers
        const float4 v = ld_f4(p.bufs[peer] + e);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    }
    acc.x *= p.scale; acc.y *= p.scale; acc.z *= p.scale; acc.w *= p.scale;
    if (NVLS) {
      // the finished values go straight to every replica through the switch; nobody reads
      // slice r of any bucket during the reduce phase except rank r itself
      multimem_st_f4(p.mc + e, acc);
    } else {
      // slice r of ANY bucket is read by this thread only (peers read their own slice index),
      // so it can be overwritten right away, here and - PUSH - in every peer's bucket
      st_f4(mine + e, acc);
      if (PUSH) {
#pragma unroll
        for (int k = 1; k < kArMaxWorld; ++k) {
          if (k >= p.world) break;
          st_f4(p.bufs[(p.rank + k) % p.world] + e, acc);
        }
      }
    }
    if (p.cast_out != nullptr) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y);
      __nv_bfloat162 hi = __floats2bfloat162_rn(acc.z, acc.w);
      *reinterpret_cast<__nv_bfloat162*>(p.cast_out + e) = lo;
      *reinterpret_cast<__nv_bfloat162*>(p.cast_out + e + 2) = hi;
    }
  }

  if (!replica_barrier(p, p.ticket + 2)) return;

  // ---- pull variant: fetch the other ranks' finished slices; NVLS / PUSH already delivered them
  if (!NVLS && !PUSH) {
    for (int k = 1; k < p.world; ++k) {
      const int peer = (p.rank + k) % p.world;
      const unsigned long long pb = slice * peer;
      const float* src = p.bufs[peer];
#pragma unroll 4
      for (unsigned long long i =
               static_cast<unsigned long long>(blockIdx.x) * blockDim.x + threadIdx.x;
           i < vecs; i += stride) {
        const float4 v = ld_f4(src + pb + i * 4);
        st_f4(mine + pb + i * 4, v);
        if (p.cast_out != nullptr) {
          *reinterpret_cast<__nv_bfloat162*>(p.cast_out + pb + i * 4) =
              __floats2bfloat162_rn(v.x, v.y);
          *reinterpret_cast<__nv_bfloat162*>(p.cast_out + pb + i * 4 + 2) =
              __floats2bfloat162_rn(v.z, v.w);
        }
      }
    }
  } else if (p.cast_out != nullptr) {
    // bf16 copy of the slices other ranks produced (they are in the local bucket now: the
    // barrier above was passed by every replica after its last store)
    for (int k = 1; k < p.world; ++k) {
      const unsigned long long pb = slice * ((p.rank + k) % p.world);
      for (unsigned long long i =
               static_cast<unsigned long long>(blockIdx.x) * blockDim.x + threadIdx.x;
           i < vecs; i += stride) {
        const float4 v = ld_f4(mine + pb + i * 4);
        *reinterpret_cast<__nv_bfloat162*>(p.cast_out + pb + i * 4) =
            __floats2bfloat162_rn(v.x, v.y);
        *reinterpret_cast<__nv_bfloat162*>(p.cast_out + pb + i * 4 + 2) =
            __floats2bfloat162_rn(v.z, v.w);
      }
    }
  }

  // pull variant only: peers are still reading this rank's slice until they pass this point
  if (!NVLS && !PUSH) replica_barrier(p, p.ticket + 3);
}

}  // namespace tgb

using namespace tgb;

// bufs / pads: host arrays of `world` device pointers (peer-mapped on the calling rank's GPU).
// n must be a multiple of 4 * world. `launch_index` must be the same on every rank and grow by
// one per call on this set of buffers (it derives the barrier tickets). `error_flag` is a
// device int that the kernel sets to 1 if a peer did not show up within `timeout_ms`.
extern "C" int tgb_allreduce_scale_f32(void* const* bufs, void* const* pads, void* multicast,
                                       void* cast_out_bf16, unsigned long long n, float scale,
                                       int rank, int world, unsigned int launch_index,
                                       int timeout_ms, int* error_flag, int max_blocks,
                                       cudaStream_t stream) {
  if (world < 1 || world > kArMaxWorld) return set_error("allreduce: world %d not in 1..8", world);
  if (rank < 0 || rank >= world) return set_error("allreduce: bad rank %d", rank);
  if (n % (4ull * world) != 0)
    return set_error("allreduce: n (%llu) must be a multiple of 4*world", n);
  ArParams p;
  for (int i = 0; i < kArMaxWorld; ++i) {
    p.bufs[i] = i < world ? static_cast<float*>(bufs[i]) : nullptr;
    p.pads[i] = i < world ? static_cast<uint32_t*>(pads[i]) : nullptr;
  }
  p.mc = static_cast<float*>(multicast);
  p.cast_out = static_cast<__nv_bfloat16*>(cast_out_bf16);
  p.n = n;
  p.scale = scale;
  p.rank = rank;
  p.world = world;
  p.ticket = launch_index * 3u;
  // cudaDevAttrClockRate is a slow query (~1 ms): look it up once per device
  static int khz_cache[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (khz_cache[dev] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrClockRate, dev);
    khz_cache[dev] = v > 0 ? v : 1500000;
  }
  const int khz = khz_cache[dev];
  p.timeout_cycles = static_cast<unsigned long long>(khz) * static_cast<unsigned long long>(
                         timeout_ms > 0 ? timeout_ms : 20000);
  p.error = error_flag;
  int blocks = sm_count();
  if (max_blocks > 0 && blocks > max_blocks) blocks = max_blocks;
  const unsigned long long vecs = n / world / 4;
  const unsigned long long want = (vecs + kArThreads - 1) / kArThreads;
  if (want < static_cast<unsigned long long>(blocks)) blocks = static_cast<int>(want ? want : 1);
  const char* pull_env = getenv("TGB_ALLREDUCE_PULL");         // read per call: A/B in one run
  const bool pull = pull_env != nullptr && pull_env[0] == '1';
  if (multicast != nullptr)
    allreduce_scale_kernel<true, false><<<blocks, kArThreads, 0, stream>>>(p);
  else if (pull)
    allreduce_scale_kernel<false, false><<<blocks, kArThreads, 0, stream>>>(p);
  else
    allreduce_scale_kernel<false, true><<<blocks, kArThreads, 0, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error("allreduce launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

// Slots of the signal pad this kernel uses, in bytes from the start of the pad (for callers that
// need to size / zero the pad): [kArPadBase, kArPadBase + max_blocks * 8) uint32 slots.
extern "C" unsigned long long tgb_allreduce_pad_bytes(int max_blocks) {
  return (static_cast<unsigned long long>(kArPadBase) +
          static_cast<unsigned long long>(max_blocks) * kArMaxWorld) * 4ull;
}
