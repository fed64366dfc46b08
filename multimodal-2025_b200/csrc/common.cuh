// common.cuh — shared helpers for libmm25 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include "../../include/mm25.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libmm25 is written for sm_100a only"
#endif

namespace mm25 {

void set_last_error(cudaError_t e);

#define MM25_CUDA_CHECK(expr)                    \
  do {                                           \
    cudaError_t _e = (expr);                     \
    if (_e != cudaSuccess) {                     \
      ::mm25::set_last_error(_e);                \
      return MM25_ERR_CUDA;                      \
    }                                            \
  } while (0)

void count_launch();
#define MM25_LAUNCH_CHECK()                 \
  do {                                      \
    ::mm25::count_launch();                 \
    MM25_CUDA_CHECK(cudaGetLastError());    \
  } while (0)

// kernel-timing hook (prep.cu): token = timing_begin(slot, stream); <launch>; timing_end(token, stream)
bool timing_enabled();
void* timing_begin(int slot, cudaStream_t st);
void timing_end(void* tok, cudaStream_t st);

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

__device__ __forceinline__ float load_scale(const float* scale, int kind, float clamp_max) {
  float v = __ldg(scale);
  if (kind == MM25_SCALE_LOG) v = expf(fminf(v, clamp_max));
  return v;
}

// d(scale)/d(raw parameter): exp'd scale passes s; clamped log passes 0 above the clamp
// (torch.clamp passes gradient at equality).
__device__ __forceinline__ float scale_chain(const float* scale, int kind, float clamp_max) {
  if (kind != MM25_SCALE_LOG) return 1.f;
  float l = __ldg(scale);
  return (l <= clamp_max) ? expf(l) : 0.f;
}

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// (max, sum) pair algebra of the online log-sum-exp.
__device__ __forceinline__ void lse_merge(float& m, float& l, float m2, float l2) {
  float mn = fmaxf(m, m2);
  if (mn == -INFINITY) { m = mn; l = 0.f; return; }
  l = l * expf(m - mn) + l2 * expf(m2 - mn);
  m = mn;
}

}  // namespace mm25
