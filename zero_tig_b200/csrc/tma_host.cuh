// Host side of TMA: cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda), bf16 tensors,
// SWIZZLE_128B boxes, zero fill outside the tensor.  Encoded maps are cached by (pointer, dims, box).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <array>
#include <map>
#include <mutex>
#include "common.cuh"

namespace zt {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn tma_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
    (void)cudaGetLastError();
  }
  return fn;
}

// rank-2 or rank-3 bf16 tensor, innermost dimension first; strides in elements for dims 1..rank-1
inline int tma_map_bf16(const void* base, int rank, const int* dims, const int64_t* strides_el, const int* box,
                        CUtensorMap* out) {
  static std::mutex mu;
  static std::map<std::array<int64_t, 10>, CUtensorMap> cache;
  std::array<int64_t, 10> key = {(int64_t)(uintptr_t)base, rank, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < rank; ++i) { key[2 + i] = dims[i]; key[5 + i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) key[8 + i] = strides_el[i];
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return ZT_OK; }
  EncodeTiledFn fn = tma_encode_fn();
  if (!fn) { set_error("TMA: cuTensorMapEncodeTiled is unavailable (no CUDA driver)"); return ZT_ERR_CUDA; }
  cuuint64_t gd[3];
  cuuint64_t gs[2];
  cuuint32_t bx[3], es[3] = {1u, 1u, 1u};
  for (int i = 0; i < rank; ++i) { gd[i] = (cuuint64_t)dims[i]; bx[i] = (cuuint32_t)box[i]; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = (cuuint64_t)strides_el[i] * 2;
  CUtensorMap m;
  const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("TMA: cuTensorMapEncodeTiled failed (%d)", (int)r); return ZT_ERR_CUDA; }
  if (cache.size() > 512) cache.clear();
  cache[key] = m;
  *out = m;
  return ZT_OK;
}

}  // namespace zt
