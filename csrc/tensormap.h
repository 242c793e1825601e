// Host-side construction of TMA tensor maps without linking libcuda: the driver entry point
// cuTensorMapEncodeTiled is looked up through the runtime (cudaGetDriverEntryPoint).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

namespace tgb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) !=
            cudaSuccess ||
        p == nullptr) {
      return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// rank-`rank` fp32 (or 2-byte) tensor, dims[0] innermost. strides_bytes[i] is the stride of
// dim i+1 (dim 0 is contiguous). swizzle_mode: 0 none, 1 = 128B (16-byte atoms, for K-major
// operands and the epilogue), 2 = 128B with 32-byte atoms (the only layout tcgen05 accepts for
// MN-major 32-bit operands). Returns 0 on success.
inline int make_tensor_map(CUtensorMap* map, const void* base, int elem_bytes, int rank,
                           const uint64_t* dims, const uint64_t* strides_bytes,
                           const uint32_t* box, int swizzle_mode) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return -1;
  // cuTensorMapEncodeTiled is a driver call and wants a current context. A thread that has
  // only been handed work (an autograd worker, an executor thread) may not have touched the
  // runtime yet: bind the current device's primary context once per thread.
  static thread_local bool context_bound = false;
  if (!context_bound) {
    cudaFree(nullptr);
    context_bound = true;
  }
  cuuint64_t gdims[5];
  cuuint64_t gstrides[5];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i + 1 < rank) gstrides[i] = strides_bytes[i];
  }
  CUtensorMapDataType dt =
      elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUresult r = fn(map, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims,
                  gstrides, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle_mode == 1   ? CU_TENSOR_MAP_SWIZZLE_128B
                  : swizzle_mode == 2 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                      : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {       // the thread switched devices since: bind again
    cudaFree(nullptr);
    r = fn(map, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims, gstrides, gbox,
           estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           swizzle_mode == 1   ? CU_TENSOR_MAP_SWIZZLE_128B
           : swizzle_mode == 2 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                               : CU_TENSOR_MAP_SWIZZLE_NONE,
           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "cuTensorMapEncodeTiled failed: %d (rank %d)\n", static_cast<int>(r), rank);
    return static_cast<int>(r);
  }
  return 0;
}

}  // namespace tgb
