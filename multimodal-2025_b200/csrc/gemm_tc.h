// gemm_tc.h — host interface of the bf16 tcgen05 GEMM (gemm_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mm25 {
namespace tc {
bool gemm_eligible(int64_t Mg, int64_t Ng, int64_t K);
// C[Mg,Ng] fp32 (+)= alpha * alpha_dev[0] * Aop * Bop.  a_mn / b_mn: operand stored [K, MN].
int gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, float* C, int64_t ldc,
              int64_t Mg, int64_t Ng, int64_t K, float alpha, const float* alpha_dev,
              int accumulate, cudaStream_t st);
}  // namespace tc
}  // namespace mm25
