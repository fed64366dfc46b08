// api_sim.cu — C-ABI entry points of the fused similarity forward / backward and the
// implementation choice between the fp32 SIMT kernels and the bf16 tcgen05 kernels.
#include "common.cuh"
#include "sim_tc.h"

namespace mm25 {
namespace simt {
size_t fwd_workspace_bytes(int64_t B, int64_t M, int64_t N);
int launch_fwd(const void* a, const void* b, int dtype, const float* u, const float* v,
               const float* diag_add, int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2,
               const float* scale, int scale_kind, float clamp_max, int mode,
               int64_t diag_offset, float* row_lse, float* col_max, float* col_sum, float* diag,
               void* ws, size_t ws_bytes, cudaStream_t st);
int64_t bwd_nparts(int64_t B, int64_t M);
int launch_bwd_pass(const void* a, const void* b, int dtype, const float* u, const float* v,
                    const float* diag_add, int diag_by_col, int64_t B, int64_t M, int64_t N,
                    int64_t d, int64_t d2, const float* scale, int scale_kind, float clamp_max,
                    int mode, int64_t diag_offset, const float* row_lse, const float* col_lse,
                    float w_row, float w_col, float w_diag, float w_all, const float* gout,
                    float* dA, float* dscale_part, float* ddiag, cudaStream_t st);
}  // namespace simt
}  // namespace mm25

using namespace mm25;

// which kernel family the last mm25_sim_fwd / mm25_sim_bwd call of this process took
// (0 = fp32 SIMT, 1 = tcgen05 PLAIN, 2 = tcgen05 MODULATED); tests use it to prove the path.
static int g_last_path = -1;
extern "C" int mm25_debug_last_sim_path(void) { return g_last_path; }

static bool use_tc(bool bwd, int impl, int64_t B, int64_t M, int64_t N, int64_t d, int dtype,
                   int mode, int* err) {
  *err = MM25_OK;
  const bool ok = bwd ? tc::eligible_bwd(B, M, N, d, dtype, mode)
                      : tc::eligible_fwd(B, M, N, d, dtype, mode);
  if (impl == 2 && !ok) *err = MM25_ERR_SHAPE;
  if (impl == 1) return false;
  return ok;
}

extern "C" size_t mm25_sim_fwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d,
                                               int dtype, int impl) {
  int err;
  size_t w = simt::fwd_workspace_bytes(B, M, N);
  if (use_tc(false, impl, B, M, N, d, dtype, MM25_SIM_PLAIN, &err)) {
    size_t t = tc::fwd_workspace_bytes(B, M, N, d);
    if (t > w) w = t;
  }
  return w;
}

extern "C" int mm25_sim_fwd(const void* a, const void* b, int dtype, const float* u,
                            const float* v, const float* diag_add, int64_t B, int64_t M,
                            int64_t N, int64_t d, int64_t d2, const float* scale, int scale_kind,
                            float clamp_max, int mode, int64_t diag_offset, float* row_lse,
                            float* col_max, float* col_sum, float* diag, void* workspace,
                            size_t workspace_bytes, int impl, float dot_bound, void* e_keep,
                            int* e_flag, void* stream) {
  if (B <= 0 || M <= 0 || N <= 0 || d <= 0) return MM25_ERR_SHAPE;
  if (!a || !b || !scale || !row_lse || !diag) return MM25_ERR_ARG;
  if ((e_keep == nullptr) != (e_flag == nullptr)) return MM25_ERR_ARG;
  if ((col_max == nullptr) != (col_sum == nullptr)) return MM25_ERR_ARG;
  if (mode < 0 || mode > 2) return MM25_ERR_ARG;
  if (mode == MM25_SIM_MODULATED && (!u || !v || d2 <= 0)) return MM25_ERR_ARG;
  if (dtype != MM25_F32 && dtype != MM25_BF16) return MM25_ERR_DTYPE;
  if (!workspace) return MM25_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == MM25_SIM_MODULATED && impl != 1 && !col_max &&
      tcm::eligible(B, M, N, d, d2, dtype, false) &&
      workspace_bytes >= tcm::workspace_bytes(M, N, d, d2, false)) {
    if (e_flag) MM25_CUDA_CHECK(cudaMemsetAsync(e_flag, 0, sizeof(int), st));
    g_last_path = 2;
    return tcm::launch_fwd(a, b, u, v, diag_add, M, N, d, d2, scale, scale_kind, clamp_max,
                           diag_offset, row_lse, diag, workspace, workspace_bytes, st);
  }
  int err;
  const bool t = use_tc(false, impl, B, M, N, d, dtype, mode, &err);
  if (err) return err;
  g_last_path = t ? 1 : 0;
  if (t)
    return tc::launch_fwd(a, b, B, M, N, d, scale, scale_kind, clamp_max, diag_offset, row_lse,
                          col_max, col_sum, diag, workspace, workspace_bytes, dot_bound,
                          mode == MM25_SIM_MASKDIAG ? nullptr : e_keep,
                          mode == MM25_SIM_MASKDIAG ? nullptr : e_flag, st,
                          mode == MM25_SIM_MASKDIAG);
  // the SIMT kernels keep nothing: tell the backward to recompute
  if (e_flag) MM25_CUDA_CHECK(cudaMemsetAsync(e_flag, 0, sizeof(int), st));
  return simt::launch_fwd(a, b, dtype, u, v, diag_add, B, M, N, d, d2, scale, scale_kind,
                          clamp_max, mode, diag_offset, row_lse, col_max, col_sum, diag,
                          workspace, workspace_bytes, st);
}

extern "C" size_t mm25_sim_bwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d,
                                               int dtype, int impl, int with_e) {
  int err;
  if (use_tc(true, impl, B, M, N, d, dtype, MM25_SIM_PLAIN, &err))
    return tc::bwd_workspace_bytes(B, M, N, d, with_e != 0);
  return 16;
}

extern "C" int64_t mm25_sim_bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d, int dtype,
                                       int impl) {
  int err;
  int64_t n = simt::bwd_nparts(B, M > N ? M : N);
  if (use_tc(true, impl, B, M, N, d, dtype, MM25_SIM_PLAIN, &err)) {
    int64_t t = tc::bwd_nparts(B, M, N, d);
    if (t > n) n = t;
  }
  return n;
}

// MODULATED mode on the tensor cores: workspace (bf16 copies of u, v + partials / G'' panel) and
// dscale partial count; 0 when the shape is not eligible (the SIMT kernels then run and the
// PLAIN-mode queries above apply).
extern "C" size_t mm25_sim_mod_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d,
                                               int64_t d2, int dtype, int bwd) {
  if (!tcm::eligible(B, M, N, d, d2, dtype, bwd != 0)) return 0;
  return tcm::workspace_bytes(M, N, d, d2, bwd != 0);
}

extern "C" int64_t mm25_sim_mod_bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d,
                                           int64_t d2, int dtype) {
  if (!tcm::eligible(B, M, N, d, d2, dtype, true)) return 0;
  return tcm::bwd_nparts(M, N);
}

static int sim_bwd_impl(const void* a, const void* b, int dtype, const float* u, const float* v,
                        const float* diag_add, int64_t B, int64_t M, int64_t N, int64_t d,
                        int64_t d2, const float* scale, int scale_kind, float clamp_max, int mode,
                        int64_t diag_offset, const float* row_lse, const float* col_lse,
                        float w_row, float w_col, float w_diag, float w_all, const float* gout,
                        float* da, float* db, float* dscale_part, float* ddiag, void* workspace,
                        size_t workspace_bytes, int impl, float dot_bound,
                        float* const* db_peers, int npeers, void* e_keep, const int* e_flag,
                        void* stream) {
  if (B <= 0 || M <= 0 || N <= 0 || d <= 0) return MM25_ERR_SHAPE;
  if (!a || !b || !scale || !row_lse) return MM25_ERR_ARG;
  if (!col_lse && w_col != 0.f) return MM25_ERR_ARG;
  if (mode < 0 || mode > 2) return MM25_ERR_ARG;
  if (mode == MM25_SIM_MODULATED && (!u || !v || d2 <= 0)) return MM25_ERR_ARG;
  if (dtype != MM25_F32 && dtype != MM25_BF16) return MM25_ERR_DTYPE;
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == MM25_SIM_MODULATED && impl != 1 && w_col == 0.f && !db_peers &&
      tcm::eligible(B, M, N, d, d2, dtype, true) && workspace &&
      workspace_bytes >= tcm::workspace_bytes(M, N, d, d2, true)) {
    g_last_path = 2;
    return tcm::launch_bwd(a, b, u, v, diag_add, M, N, d, d2, scale, scale_kind, clamp_max,
                           diag_offset, row_lse, w_row, w_diag, w_all, gout, da, db, dscale_part,
                           ddiag, workspace, workspace_bytes, st);
  }
  int err;
  const bool t = use_tc(true, impl, B, M, N, d, dtype, mode, &err);
  if (err) return err;
  g_last_path = t ? 1 : 0;
  if (t)
    return tc::launch_bwd(a, b, B, M, N, d, scale, scale_kind, clamp_max, diag_offset, row_lse,
                          col_lse, w_row, w_col, w_diag, w_all, gout, da, db, dscale_part,
                          workspace, workspace_bytes, dot_bound, db_peers, npeers,
                          mode == MM25_SIM_MASKDIAG ? nullptr : e_keep,
                          mode == MM25_SIM_MASKDIAG ? nullptr : e_flag, st,
                          mode == MM25_SIM_MASKDIAG);
  if (db_peers) return MM25_ERR_SHAPE;  // the fused reduce-scatter exists on the tcgen05 path only
  int rc = MM25_OK;
  if (da) {
    rc = simt::launch_bwd_pass(a, b, dtype, u, v, diag_add, 0, B, M, N, d, d2, scale, scale_kind,
                               clamp_max, mode, diag_offset, row_lse, col_lse, w_row, w_col,
                               w_diag, w_all, gout, da, dscale_part, ddiag, st);
    if (rc) return rc;
  }
  if (db) {
    // transposed roles: rows = b, columns = a; row statistics = col_lse
    rc = simt::launch_bwd_pass(b, a, dtype, v, u, diag_add, 1, B, N, M, d, d2, scale, scale_kind,
                               clamp_max, mode, -diag_offset, col_lse, row_lse, w_col, w_row,
                               w_diag, w_all, gout, db, da ? nullptr : dscale_part, nullptr, st);
  }
  return rc;
}

extern "C" int mm25_sim_bwd(const void* a, const void* b, int dtype, const float* u,
                            const float* v, const float* diag_add, int64_t B, int64_t M,
                            int64_t N, int64_t d, int64_t d2, const float* scale, int scale_kind,
                            float clamp_max, int mode, int64_t diag_offset, const float* row_lse,
                            const float* col_lse, float w_row, float w_col, float w_diag,
                            float w_all, const float* gout, float* da, float* db,
                            float* dscale_part, float* ddiag, void* workspace,
                            size_t workspace_bytes, int impl, float dot_bound, void* e_keep,
                            const int* e_flag, void* stream) {
  return sim_bwd_impl(a, b, dtype, u, v, diag_add, B, M, N, d, d2, scale, scale_kind, clamp_max,
                      mode, diag_offset, row_lse, col_lse, w_row, w_col, w_diag, w_all, gout, da,
                      db, dscale_part, ddiag, workspace, workspace_bytes, impl, dot_bound,
                      nullptr, 0, e_keep, e_flag, stream);
}

extern "C" int mm25_sim_bwd_rs(const void* a, const void* b, int dtype, int64_t M, int64_t N,
                               int64_t d, const float* scale, int scale_kind, float clamp_max,
                               int64_t diag_offset, const float* row_lse, const float* col_lse,
                               float w_row, float w_col, float w_diag, float w_all,
                               const float* gout, float* da, float* const* db_peers,
                               int64_t npeers, float* dscale_part, void* workspace,
                               size_t workspace_bytes, float dot_bound, void* e_keep,
                               const int* e_flag, void* stream) {
  if (!db_peers || npeers < 1 || npeers > 16) return MM25_ERR_ARG;
  for (int64_t i = 0; i < npeers; ++i)
    if (!db_peers[i]) return MM25_ERR_ARG;
  return sim_bwd_impl(a, b, dtype, nullptr, nullptr, nullptr, 1, M, N, d, 0, scale, scale_kind,
                      clamp_max, MM25_SIM_PLAIN, diag_offset, row_lse, col_lse, w_row, w_col,
                      w_diag, w_all, gout, da, nullptr, dscale_part, nullptr, workspace,
                      workspace_bytes, 2, dot_bound, db_peers, (int)npeers, e_keep, e_flag,
                      stream);
}

// fp32-grade backward on two-term bf16 splits (sim_tc_bwd.cu launch_bwd_split)
extern "C" size_t mm25_sim_bwd_split_workspace_bytes(int64_t M, int64_t N) {
  return tc::bwd_split_workspace_bytes(M, N);
}
extern "C" int64_t mm25_sim_bwd_split_nparts(int64_t B, int64_t M, int64_t N) {
  return tc::bwd_split_nparts(B, M, N);
}
extern "C" int mm25_sim_split_eligible(int64_t B, int64_t M, int64_t N, int64_t d) {
  return tc::eligible_fwd(B, M, N, MM25_SPLIT_K * d, MM25_BF16, MM25_SIM_PLAIN) && d % 256 == 0 &&
                 tc::eligible_bwd(B, M, N, d, MM25_BF16, MM25_SIM_PLAIN)
             ? 1 : 0;
}
extern "C" int mm25_sim_bwd_split(const void* a4, const void* b4, const void* a_hi,
                                  const void* a_lo, const void* b_hi, const void* b_lo,
                                  int64_t B, int64_t M, int64_t N, int64_t d, const float* scale,
                                  int scale_kind, float clamp_max, int64_t diag_offset,
                                  const float* row_lse, const float* col_lse, float w_row,
                                  float w_col, float w_diag, float w_all, const float* gout,
                                  float* da, float* db, float* dscale_part, void* workspace,
                                  size_t workspace_bytes, float dot_bound, int mode,
                                  void* stream) {
  if (mode != MM25_SIM_PLAIN && mode != MM25_SIM_MASKDIAG) return MM25_ERR_ARG;
  if (!a4 || !b4 || !a_hi || !a_lo || !b_hi || !b_lo || !scale || !row_lse || !workspace)
    return MM25_ERR_ARG;
  if (!col_lse && w_col != 0.f) return MM25_ERR_ARG;
  return tc::launch_bwd_split(a4, b4, a_hi, a_lo, b_hi, b_lo, B, M, N, d, scale, scale_kind,
                              clamp_max, diag_offset, row_lse, col_lse, w_row, w_col, w_diag,
                              w_all, gout, da, db, dscale_part, workspace, workspace_bytes,
                              dot_bound, (cudaStream_t)stream, mode == MM25_SIM_MASKDIAG);
}

extern "C" int mm25_hopfield_split_eligible(int64_t M, int64_t N, int64_t d) {
  return tcm::hopfield_split_eligible(M, N, d) ? 1 : 0;
}
extern "C" size_t mm25_hopfield_bwd_split_workspace_bytes(int64_t M, int64_t N, int64_t d) {
  return tcm::hopfield_split_eligible(M, N, d) ? tcm::hopfield_split_workspace_bytes(M, N, d) : 0;
}
extern "C" int mm25_hopfield_bwd_split(const void* q4, const void* k4, const void* q_hi,
                                       const void* q_lo, const void* k_hi, const void* k_lo,
                                       const float* d_o, int64_t M, int64_t N, int64_t d,
                                       const float* beta, const float* lse, const float* delta,
                                       float* dq, float* dk, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  if (!q4 || !k4 || !q_hi || !q_lo || !k_hi || !k_lo || !d_o || !beta || !lse || !delta ||
      !workspace)
    return MM25_ERR_ARG;
  return tcm::launch_hopfield_bwd_split(q4, k4, q_hi, q_lo, k_hi, k_lo, d_o, M, N, d, beta, lse,
                                        delta, dq, dk, workspace, workspace_bytes,
                                        (cudaStream_t)stream);
}
