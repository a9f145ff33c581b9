// Shared helpers for libzerotig (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/zerotig.h"

namespace zt {

void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;

inline int check_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return ZT_ERR_CUDA;
  }
  return ZT_OK;
}

#define ZT_REQUIRE(cond, ...)      \
  do {                             \
    if (!(cond)) {                 \
      zt::set_error(__VA_ARGS__);  \
      return ZT_ERR_ARG;           \
    }                              \
  } while (0)

#define ZT_CUDA(call)                                                     \
  do {                                                                    \
    cudaError_t e_ = (call);                                              \
    if (e_ != cudaSuccess) {                                              \
      zt::set_error("%s failed: %s", #call, cudaGetErrorString(e_));      \
      return ZT_ERR_CUDA;                                                 \
    }                                                                     \
  } while (0)

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case ZT_ACT_RELU: return fmaxf(v, 0.f);
    case ZT_ACT_LRELU02: return v > 0.f ? v : 0.2f * v;
    case ZT_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
    case ZT_ACT_TANH: return tanhf(v);
    default: return v;
  }
}

constexpr int kNumSMs = 148;  // B200

// ---- programmatic dependent launch (PDL) ----
// Kernels of the per-frame chain are launched with cudaLaunchAttributeProgrammaticStreamSerialization: the next
// kernel of the stream may start while this one is still running.  pdl_enter() = "my dependents may launch" +
// "wait until the kernel before me has completed and flushed"; everything after it sees the predecessor's
// writes.  Without the launch attribute both instructions are no-ops.  ZT_NO_PDL=1 disables the attribute.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace zt
