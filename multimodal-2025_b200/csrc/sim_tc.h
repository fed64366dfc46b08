// sim_tc.h — host interface of the bf16 tcgen05 / TMEM / TMA similarity kernels (sim_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace mm25 {
namespace tc {

// `dot_bound`: caller-guaranteed bound on |a_i . b_j| (e.g. 1.01 for rows normalised by
// mm25_rownorm_fwd); <= 0 = unknown.  Enables the static-shift one-exp epilogues.
// PLAIN mode, bf16, d % 64 == 0 (64 <= d <= 1024), M % 128 == 0, N % 128 == 0.
bool eligible_fwd(int64_t B, int64_t M, int64_t N, int64_t d, int dtype, int mode);
bool eligible_bwd(int64_t B, int64_t M, int64_t N, int64_t d, int dtype, int mode);

size_t fwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d);
int launch_fwd(const void* a, const void* b, int64_t B, int64_t M, int64_t N, int64_t d,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               float* row_lse, float* col_max, float* col_sum, float* diag, void* ws,
               size_t ws_bytes, float dot_bound, void* e_out, int* e_flag, cudaStream_t st,
               int mask_diag = 0);

// `e_buf` (optional, [B, M, N] bf16 + flag written by launch_fwd): E = 2^(z log2e - m) kept by
// the forward; the backward then forms G' from it in place (no S-tile recompute) and needs no
// panel workspace.
size_t bwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d, bool with_e = false);
int64_t bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d);
int launch_bwd(const void* a, const void* b, int64_t B, int64_t M, int64_t N, int64_t d,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               const float* row_lse, const float* col_lse, float w_row, float w_col,
               float w_diag, float w_all, const float* gout, float* da, float* db,
               float* dscale_part, void* ws, size_t ws_bytes, float dot_bound,
               float* const* db_peers, int npeers, void* e_buf, const int* e_flag,
               cudaStream_t st, int mask_diag = 0);

// fp32-grade path on two-term bf16 splits (sim_tc_bwd.cu launch_bwd_split)
size_t bwd_split_workspace_bytes(int64_t M, int64_t N);
int64_t bwd_split_nparts(int64_t B, int64_t M, int64_t N);
int launch_bwd_split(const void* a4, const void* b4, const void* a_hi, const void* a_lo,
                     const void* b_hi, const void* b_lo, int64_t B, int64_t M, int64_t N,
                     int64_t d, const float* scale, int scale_kind, float clamp_max,
                     int64_t diag_offset, const float* row_lse, const float* col_lse,
                     float w_row, float w_col, float w_diag, float w_all, const float* gout,
                     float* da, float* db, float* dscale_part, void* ws, size_t ws_bytes,
                     float dot_bound, cudaStream_t st, int mask_diag = 0);

}  // namespace tc

// sim_tc_mod.cu — MODULATED mode on the tensor cores (two TMEM accumulators per tile), B = 1,
// row statistics only.  u, v are the fp32 normalised semantic features of the C-ABI (cast to
// bf16 into the workspace).
namespace tcm {
bool eligible(int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2, int dtype, bool bwd);
size_t workspace_bytes(int64_t M, int64_t N, int64_t d, int64_t d2, bool bwd);
int64_t bwd_nparts(int64_t M, int64_t N);
int launch_fwd(const void* a, const void* b, const float* u, const float* v,
               const float* diag_add, int64_t M, int64_t N, int64_t d, int64_t d2,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               float* row_lse, float* diag, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_bwd(const void* a, const void* b, const float* u, const float* v,
               const float* diag_add, int64_t M, int64_t N, int64_t d, int64_t d2,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               const float* row_lse, float w_row, float w_diag, float w_all, const float* gout,
               float* da, float* db, float* dscale_part, float* ddiag, void* ws,
               size_t ws_bytes, cudaStream_t st);
// fp32-grade Hopfield backward on two-term bf16 splits (sim_tc_mod.cu)
bool hopfield_split_eligible(int64_t M, int64_t N, int64_t d);
size_t hopfield_split_workspace_bytes(int64_t M, int64_t N, int64_t d);
int launch_hopfield_bwd_split(const void* q4, const void* k4, const void* q_hi, const void* q_lo,
                              const void* k_hi, const void* k_lo, const float* d_o, int64_t M,
                              int64_t N, int64_t d, const float* beta, const float* lse,
                              const float* delta, float* dq, float* dk, void* ws, size_t ws_bytes,
                              cudaStream_t st);
}  // namespace tcm
}  // namespace mm25
