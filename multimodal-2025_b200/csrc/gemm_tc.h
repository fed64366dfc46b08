// gemm_tc.h — host interface of the bf16 tcgen05 GEMMs (gemm_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mm25 {
namespace tc {
bool gemm_eligible(int64_t Mg, int64_t Ng, int64_t K);

// One problem  C[Mg, Ng] fp32 (+)= alpha * Aop[Mg, K] * Bop[K, Ng]  (bf16 operands).
// a_mn / b_mn = 1: the operand is stored [K, Mg] / [K, Ng] instead of [Mg, K] / [Ng, K].
struct GemmDesc {
  const void* A;
  const void* B;
  float* C;
  int64_t ldc;
  int64_t Mg, K;
  int a_mn, b_mn;
  int accumulate;
  float alpha;
  // optional fused reduce-scatter: C is ignored and rows [q R, (q+1) R) of the result are
  // added (fp32 red.global over NVLink) into peers[q][row - q R, :]; npeers * R == Mg
  float* const* peers;
  int npeers;
  int64_t rows_per_rank;
  // scatter mode only: this rank's index.  The m-blocks are walked starting with the rows owned
  // by rank self + 1 and ending with the local ones, so at any moment the P ranks add into P
  // different owners (a ring order) instead of all hitting owner 0 first.
  int self_rank;
  // batched problems (B entries of identical shape stacked along the rows of every operand and
  // of C): the m-blocks run over batch * Mg / 128 (Mg, K = sizes of ONE entry).  For entry z the
  // K coordinate of an MN-major operand (its rows are K) is shifted by z * K, and the M coordinate
  // of an MN-major A stays local to the entry.  Needs b_mn = 1.  batch <= 1: plain problem.
  int batch;
};
constexpr int GEMM_MAX_PEERS = 16;

// Up to two problems sharing Ng in ONE persistent launch with a static balanced schedule
// (either pointer may be NULL).  alpha_dev: optional device scalar multiplied into both.
int gemm_grouped(const GemmDesc* d0, const GemmDesc* d1, int64_t Ng, const float* alpha_dev,
                 cudaStream_t st);

// EXPERIMENTAL CTA-pair (cta_group::2) variant of gemm_grouped (gemm_tc2.cu), taken only when
// the environment variable MM25_GEMM_2CTA is set.  Returns MM25_ERR_SHAPE for problems it does
// not cover (scatter, split-K, batched, odd 128-row block counts): gemm_grouped then runs the
// 1-CTA kernel.
int gemm_grouped2(const GemmDesc* d0, const GemmDesc* d1, int64_t Ng, const float* alpha_dev,
                  cudaStream_t st);

// ---- kept-E backward GEMMs with the E -> G' transform fused into the A-operand load (gemm_te.cu)
// One problem  C[batch * Mg, Ng] fp32 = T(E)op[Mg, K] * Bop[K, Ng]; E is the bf16 matrix the
// forward kept, [batch * Mg, K] (a_mn = 0) or [batch * K, Mg] (a_mn = 1) with row pitch ldE;
// B is MN-major [batch * K, Ng].  rowf / colf: per-row / per-column factors of E's storage.
struct GemmTeDesc {
  const void* E;
  int64_t ldE;
  const void* B;
  float* C;
  int64_t ldc;
  int64_t Mg, K;
  int a_mn;
  int batch;
  const float* rowf;
  const float* colf;
  const float* gdiag;         // exact fp32 G' of each E row's positive (NULL: none)
  int64_t diag_offset;        // positive of E's local row r: column r + diag_offset
  float* const* peers;        // fused reduce-scatter as in GemmDesc (rows_per_rank % 256 == 0)
  int npeers;
  int64_t rows_per_rank;
  int self_rank;
};
bool gemm_te_eligible(int64_t Mg, int64_t Ng, int64_t K);
int gemm_te_grouped(const GemmTeDesc* d0, const GemmTeDesc* d1, int64_t Ng, const int* e_flag,
                    const float* consts, int zero_c, cudaStream_t st);
// rowf [B M], colf [B N], consts [B][2] = {gs w_all, gs w_diag} of the static-shift gradient
// + gdiag [B M]: exact G' of the positives, from a_i . b_{i + diag_offset} in fp32
int te_coef(const void* a, const void* b, int64_t d, int64_t diag_offset, const float* row_lse,
            const float* col_lse, int64_t B, int64_t M, int64_t N,
            const float* scale, int scale_kind, float clamp_max, const float* gout, float w_row,
            float w_col, float w_diag, float w_all, float dot_bound, float* rowf, float* colf,
            float* consts, float* gdiag, cudaStream_t st);

// single problem (wrapper over gemm_grouped)
int gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, float* C, int64_t ldc,
              int64_t Mg, int64_t Ng, int64_t K, float alpha, const float* alpha_dev,
              int accumulate, cudaStream_t st);
}  // namespace tc
}  // namespace mm25
