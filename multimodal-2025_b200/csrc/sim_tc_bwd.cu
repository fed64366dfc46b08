// sim_tc_bwd.cu — tensor-core backward of the fused similarity loss (bf16, PLAIN mode).
//
// TMEM holds 128 x 512 fp32 per SM: exactly one 128-row block of dA at d = 512, leaving no room
// for the recomputed S tile, so a single fully-fused pass cannot keep both gradient accumulators
// on chip (DESIGN.md "backward").  The backward therefore runs per ROW PANEL of R rows:
//   1. G-panel kernel (this file): recompute the S tiles (same TMA/tcgen05 sweep as the
//      forward), form  G' = gout * s * (w_row e^{z-r_i} + w_col e^{z-c_j} - w_diag d_ij - w_all)
//      in registers and store it as bf16 into a bounded workspace [R, N]; also sum G.z/s.
//   2. dA[panel] = G' . Yhat        (gemm_tc.cu, A K-major, B MN-major)
//   3. dB       += G'^T . Xhat[panel] (gemm_tc.cu, both MN-major)
// R is chosen so the workspace stays <= 1 GiB; the N x N matrix never exists as a whole.
#include "sim_tc_sweep.cuh"
#include "sim_tc.h"
#include "gemm_tc.h"

#include <cstdlib>
#include <type_traits>

namespace mm25 {
namespace tc {

int tc_sm_count();

// Cc[j] = gout * s * w_col * 2^(m2 - c_j log2e): the column factor of the one-exp gradient
// (static-shift path, see sim_grad_tc_kernel).  One thread per column.
__global__ void __launch_bounds__(256)
grad_col_coef_kernel(const float* __restrict__ col_lse, int64_t N, const float* __restrict__ scale,
                     int scale_kind, float clamp_max, const float* __restrict__ gout, float w_col,
                     float dot_bound, float* __restrict__ Cc) {
  const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (j >= N) return;
  const int64_t bz = blockIdx.y;  // batch entry (grid.y = 1 for a single problem)
  const float s = load_scale(scale, scale_kind, clamp_max);
  const float m2 = dot_bound * fabsf(s) * LOG2E;
  const float go = gout ? __ldg(gout + bz) : 1.f;
  Cc[bz * N + j] = go * s * w_col * exp2f(m2 - col_lse[bz * N + j] * LOG2E);
}

// Epilogue paths:
//  * static shift (one exp per logit): when the caller bounds |a_i.b_j| <= dot_bound and
//    2 dot_bound |s| log2e <= SPREAD_LIMIT, every 2^(z log2e - m2) with m2 = dot_bound |s| log2e
//    lies in [2^-96, 1], so  G' = E (R_i + C_j) - K  with E = 2^(z log2e - m2),
//    R_i = gs w_row 2^(m2 - r_i log2e), C_j = gs w_col 2^(m2 - c_j log2e) (both <= 2^96).
//  * general (two exps per logit): any range (s up to the 100 clamp, un-normalised inputs).
__global__ void __launch_bounds__(FWD_THREADS, 1)
sim_grad_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                   const __grid_constant__ CUtensorMap tmap_b,
                   const __grid_constant__ CUtensorMap tmap_g, int64_t M, int64_t N, int d,
                   int first_rb, int num_rb, int nseg, int tps, const float* __restrict__ scale,
                   int scale_kind, float clamp_max, int64_t diag_offset,
                   const float* __restrict__ row_lse, const float* __restrict__ col_lse,
                   float w_row, float w_col, float w_diag, float w_all,
                   const float* __restrict__ gout, float* __restrict__ dscale_part,
                   const float* __restrict__ Cc, float dot_bound,
                   const int* __restrict__ skip_flag, int mask_diag) {
#define MM25_PAIR 0
#include "sim_tc_bwd_body.inc"
#undef MM25_PAIR
}

// EXPERIMENTAL (MM25_SWEEP_2CTA=1): clusters of two CTAs; `tmap_b` must have 128-row boxes;
// num_rb, first_rb and M / 128 even, grid even.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FWD_THREADS, 1)
sim_grad_tc2_kernel(const __grid_constant__ CUtensorMap tmap_a,
                   const __grid_constant__ CUtensorMap tmap_b,
                   const __grid_constant__ CUtensorMap tmap_g, int64_t M, int64_t N, int d,
                   int first_rb, int num_rb, int nseg, int tps, const float* __restrict__ scale,
                   int scale_kind, float clamp_max, int64_t diag_offset,
                   const float* __restrict__ row_lse, const float* __restrict__ col_lse,
                   float w_row, float w_col, float w_diag, float w_all,
                   const float* __restrict__ gout, float* __restrict__ dscale_part,
                   const float* __restrict__ Cc, float dot_bound,
                   const int* __restrict__ skip_flag, int mask_diag) {
#define MM25_PAIR 1
#include "sim_tc_bwd_body.inc"
#undef MM25_PAIR
}

// ------------------------------------------------------------------------------------------
// Kept-E path.  When the forward ran its static-shift epilogue it stored E = 2^(z log2e - m2) as
// bf16 [M, N] (sim_tc.cu) and raised *e_flag.  The backward then needs no S-tile recompute:
//   G'_ij = E_ij (R_i + C_j) - K  (- k_diag on the positives),   in place over E.
// HBM-bound: 2 B read + 2 B written per logit.  One CTA owns a 32-row x 2048-column block; each
// thread keeps its 8 column factors C_j in registers and walks the rows four at a time.
constexpr int GE_ROWS = 32, GE_COLS = 2048;

__global__ void __launch_bounds__(256)
grad_from_e_kernel(__nv_bfloat16* __restrict__ EG, const int* __restrict__ e_flag, int64_t M,
                   int64_t N, const float* __restrict__ scale, int scale_kind, float clamp_max,
                   int64_t diag_offset, const float* __restrict__ row_lse,
                   const float* __restrict__ Cc, float w_row, float w_diag, float w_all,
                   const float* __restrict__ gout, float dot_bound) {
  if (*e_flag == 0) return;
  const int64_t c0 = (int64_t)blockIdx.x * GE_COLS + threadIdx.x * 8;
  if (c0 >= N) return;
  const int64_t r0 = (int64_t)blockIdx.y * GE_ROWS;
  const float s = load_scale(scale, scale_kind, clamp_max);
  const float gs = (gout ? __ldg(gout) : 1.f) * s;
  const float m2 = dot_bound * fabsf(s) * LOG2E;
  const float k_all = gs * w_all, k_diag = gs * w_diag, gr = gs * w_row;
  float cj[8];
  {
    const float4 x = __ldg(reinterpret_cast<const float4*>(Cc + c0));
    const float4 y = __ldg(reinterpret_cast<const float4*>(Cc + c0 + 4));
    cj[0] = x.x; cj[1] = x.y; cj[2] = x.z; cj[3] = x.w;
    cj[4] = y.x; cj[5] = y.y; cj[6] = y.z; cj[7] = y.w;
  }
  const int64_t rend = (r0 + GE_ROWS < M) ? r0 + GE_ROWS : M;
  for (int64_t r = r0; r < rend; r += 4) {
    uint4 v[4];
    float Ri[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (r + u < rend) {
        v[u] = __ldcs(reinterpret_cast<const uint4*>(EG + (r + u) * N + c0));
        Ri[u] = gr * exp2f(m2 - __ldg(row_lse + r + u) * LOG2E);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (r + u >= rend) break;
      const int64_t rel = r + u + diag_offset - c0;  // position of the positive in this thread
      uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float e0 = __uint_as_float(w[k] << 16);
        const float e1 = __uint_as_float(w[k] & 0xffff0000u);
        float g0 = fmaf(e0, Ri[u] + cj[2 * k], -k_all);
        float g1 = fmaf(e1, Ri[u] + cj[2 * k + 1], -k_all);
        if (rel == 2 * k) g0 -= k_diag;
        if (rel == 2 * k + 1) g1 -= k_diag;
        __nv_bfloat162 h = __floats2bfloat162_rn(g0, g1);
        w[k] = *reinterpret_cast<uint32_t*>(&h);
      }
      __stcs(reinterpret_cast<uint4*>(EG + (r + u) * N + c0), make_uint4(w[0], w[1], w[2], w[3]));
    }
  }
}

// d(loss)/d(scale) of the kept-E path:  gout sum_ij G_ij v_ij = (1/s) sum_i a_i . dA_i  (dA = G' B
// already holds gout s G B), so the logits are not needed again.  `x` is the bf16 operand whose
// gradient `g` (fp32, same shape) the grouped GEMM just wrote; one partial per CTA.
__global__ void __launch_bounds__(1024)
dscale_from_grad_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ g,
                        int64_t n, const int* __restrict__ e_flag,
                        const float* __restrict__ scale, int scale_kind, float clamp_max,
                        float* __restrict__ part) {
  if (*e_flag == 0) return;
  float acc = 0.f;
  // 1024 threads per CTA (one CTA per partial slot): four times the loads in flight of the
  // 256-thread version, which was latency-bound at 2.5 TB/s
  for (int64_t i = ((int64_t)blockIdx.x * 1024 + threadIdx.x) * 4; i < n;
       i += (int64_t)gridDim.x * 1024 * 4) {
    const uint2 xv = __ldg(reinterpret_cast<const uint2*>(x + i));
    const float4 gv = __ldg(reinterpret_cast<const float4*>(g + i));
    acc = fmaf(__uint_as_float(xv.x << 16), gv.x, acc);
    acc = fmaf(__uint_as_float(xv.x & 0xffff0000u), gv.y, acc);
    acc = fmaf(__uint_as_float(xv.y << 16), gv.z, acc);
    acc = fmaf(__uint_as_float(xv.y & 0xffff0000u), gv.w, acc);
  }
  __shared__ float red[32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) tot += red[k];
    const float s = load_scale(scale, scale_kind, clamp_max);
    part[blockIdx.x] = s != 0.f ? tot / s : 0.f;
  }
}

// ------------------------------------------------------------------------------------------
static size_t g_workspace_cap = (size_t)1 << 30;   // bytes of the bf16 G panel
static size_t g_batched_cap = (size_t)6 << 30;   // bytes of all B panels of a batched launch
// test hook: a small cap forces the multi-panel (and per-entry) code paths
void set_g_cap(size_t bytes) {
  g_workspace_cap = bytes ? bytes : ((size_t)1 << 30);
  g_batched_cap = bytes ? bytes : ((size_t)6 << 30);
}

bool eligible_bwd(int64_t B, int64_t M, int64_t N, int64_t d, int dtype, int mode) {
  return eligible_fwd(B, M, N, d, dtype, mode) && d % 256 == 0 && gemm_eligible(M >= 128 ? 128 : M, d, N);
}

static int64_t panel_rows(int64_t M, int64_t N) {
  int64_t r = (int64_t)(g_workspace_cap / ((size_t)N * 2)) / BM * BM;
  if (r < BM) r = BM;
  if (r > M) r = M;
  return r;
}

// B > 1 (forward_sampled: S sample slices): when all B panels fit, ONE gradient-panel launch and
// ONE batched grouped GEMM cover every entry instead of 2 B launches of quarter-filled waves
static bool batched_fits(int64_t B, int64_t M, int64_t N) {
  return B > 1 && (size_t)B * (size_t)M * (size_t)N * 2 <= g_batched_cap &&
         B * M < (1ll << 31) && B * N < (1ll << 31);
}

// Kept-E backward with the transform fused into the GEMM's operand load (gemm_te.cu): default
// whenever the shapes allow; MM25_TE=0 restores the in-place pass + grouped GEMM (A/B switch).
static bool te_enabled() {
  static const bool on = !(getenv("MM25_TE") && atoi(getenv("MM25_TE")) == 0);
  return on;
}
static bool te_shapes_ok(int64_t M, int64_t N, int64_t d, int npeers) {
  return gemm_te_eligible(M, d, N) && gemm_te_eligible(N, d, M) &&
         (npeers <= 0 || (N % npeers == 0 && (N / npeers) % 256 == 0));
}
static size_t te_ws_bytes(int64_t B, int64_t M, int64_t N) {
  return 256 + sizeof(float) * (size_t)(B * N + 2 * B * M + 2 * B + 64) + 256;
}

size_t bwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d, bool with_e) {
  if (with_e && te_enabled() && te_shapes_ok(M, N, d, 0)) {
    const size_t legacy = 256 + (size_t)N * sizeof(float) + 256;
    const size_t te = te_ws_bytes(B, M, N);
    return te > legacy ? te : legacy;
  }
  // bf16 G panel (not needed when the forward kept E: G' is formed in place there) + the [N]
  // column factors of the static-shift epilogue
  if (!with_e && batched_fits(B, M, N))
    return (((size_t)B * M * N * 2 + 255) & ~(size_t)255) + 256 + (size_t)B * N * sizeof(float) +
           256;
  const size_t panel = with_e ? 0 : (size_t)panel_rows(M, N) * (size_t)N * 2;
  return panel + 256 + (size_t)N * sizeof(float) + 256;
}

int64_t bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d) {
  // one slot per CTA of every panel's G kernel + one set for the kept-E scale gradient
  const int64_t R = panel_rows(M, N);
  return B * (ceil_div(M, R) + 1) * tc_sm_count();
}

int launch_bwd(const void* a, const void* b, int64_t B, int64_t M, int64_t N, int64_t d,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               const float* row_lse, const float* col_lse, float w_row, float w_col,
               float w_diag, float w_all, const float* gout, float* da, float* db,
               float* dscale_part, void* ws, size_t ws_bytes, float dot_bound,
               float* const* db_peers, int npeers, void* e_buf, const int* e_flag,
               cudaStream_t st, int mask_diag) {
  if (!eligible_bwd(B, M, N, d, MM25_BF16, MM25_SIM_PLAIN)) return MM25_ERR_SHAPE;
  if (mask_diag && e_buf) return MM25_ERR_ARG;
  if (db_peers && (B != 1 || npeers < 1 || N % npeers != 0)) return MM25_ERR_SHAPE;
  if (!col_lse && w_col != 0.f) return MM25_ERR_ARG;
  if ((e_buf == nullptr) != (e_flag == nullptr)) return MM25_ERR_ARG;
  // kept-E path: the scale gradient comes from a . dA (or b . dB), so one of them must exist
  const bool use_e = e_buf && dot_bound > 0.f && (!dscale_part || da || (db && !db_peers));
  if (ws_bytes < bwd_workspace_bytes(B, M, N, d, use_e)) return MM25_ERR_WORKSPACE;
  const uintptr_t ws0 = (reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255);
  if (use_e && te_enabled() && te_shapes_ok(M, N, d, db_peers ? npeers : 0) &&
      ws_bytes >= te_ws_bytes(B, M, N)) {
    // ---- kept E, transform fused into the GEMM (gemm_te.cu) ---------------------------------
    float* colf = reinterpret_cast<float*>(ws0);
    float* rowf = colf + B * N;
    float* gdiag = rowf + B * M;
    float* consts = gdiag + B * M;
    int rc = te_coef(a, b, d, diag_offset, row_lse, col_lse, B, M, N, scale, scale_kind, clamp_max,
                     gout, w_row, w_col, w_diag, w_all, dot_bound, rowf, colf, consts, gdiag, st);
    if (rc) return rc;
    const int sms = tc_sm_count();
    const __nv_bfloat16* ab = (const __nv_bfloat16*)a;
    const __nv_bfloat16* bb = (const __nv_bfloat16*)b;
    __nv_bfloat16* E = reinterpret_cast<__nv_bfloat16*>(e_buf);
    if (dscale_part)
      MM25_CUDA_CHECK(cudaMemsetAsync(dscale_part, 0, sizeof(float) * 2 * B * sms, st));
    // The forward raises *e_flag only when its static-shift epilogue ran (decided on the device
    // from the scale).  Otherwise E holds nothing: the recompute kernel fills the same buffer
    // with G' (it returns at once when the flag is set) and the GEMM below copies instead of
    // transforming.
    {
      const size_t smem = sizeof(GradSmem) + 1024;
      MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc_kernel,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      const int num_rb = (int)(M / BM), ntiles = (int)(N / BN);
      int nseg = 1;
      {
        double best = 0.0;
        const int max_seg = ntiles >= 8 ? ntiles / 4 : 1;
        for (int ns = 1; ns <= max_seg && ns <= 64; ++ns) {
          const int tps = (ntiles + ns - 1) / ns;
          const int sg = (ntiles + tps - 1) / tps;
          const long items = (long)num_rb * sg;
          const double eff = (double)items / ((double)((items + sms - 1) / sms) * sms);
          if (eff > best + 0.01) { best = eff; nseg = sg; }
        }
      }
      const int tps = (ntiles + nseg - 1) / nseg;
      nseg = (ntiles + tps - 1) / tps;
      const long items = (long)num_rb * nseg;
      const int grid = (int)(items < sms ? items : sms);
      for (int64_t bz = 0; bz < B; ++bz) {
        CUtensorMap ta, tb, tg;
        if ((rc = make_tmap_bf16(&ta, ab + bz * M * d, (uint64_t)M, (uint64_t)d, BM, BK))) return rc;
        if ((rc = make_tmap_bf16(&tb, bb + bz * N * d, (uint64_t)N, (uint64_t)d, BN, BK))) return rc;
        if ((rc = make_tmap_bf16(&tg, E + bz * M * N, (uint64_t)M, (uint64_t)N, 32, 32, 64))) return rc;
        sim_grad_tc_kernel<<<grid, FWD_THREADS, smem, st>>>(
            ta, tb, tg, M, N, (int)d, 0, num_rb, nseg, tps, scale, scale_kind, clamp_max,
            diag_offset, row_lse + bz * M, col_lse ? col_lse + bz * N : colf + bz * N, w_row,
            w_col, w_diag, w_all, gout ? gout + bz : nullptr,
            dscale_part ? dscale_part + 2 * bz * sms : nullptr, colf + bz * N, dot_bound, e_flag,
            mask_diag);
        MM25_LAUNCH_CHECK();
      }
    }
    const bool have_db = db != nullptr || db_peers != nullptr;
    GemmTeDesc ga{E, N, b, da, d, M, N, 0, (int)B, rowf, colf, gdiag, diag_offset, nullptr, 0, 0, 0};
    GemmTeDesc gb{E, N, a, db, d, N, M, 1, (int)B, rowf, colf, gdiag, diag_offset, db_peers, npeers,
                  db_peers ? N / npeers : 0,
                  db_peers ? (int)((diag_offset / M) % npeers) : 0};
    rc = gemm_te_grouped(da ? &ga : nullptr, have_db ? &gb : nullptr, d, e_flag, consts, 1, st);
    if (rc) return rc;
    if (dscale_part) {
      for (int64_t bz = 0; bz < B; ++bz) {
        float* part = dscale_part + (2 * bz + 1) * sms;
        if (da)
          dscale_from_grad_kernel<<<sms, 1024, 0, st>>>(ab + bz * M * d, da + bz * M * d, M * d,
                                                       e_flag, scale, scale_kind, clamp_max, part);
        else
          dscale_from_grad_kernel<<<sms, 1024, 0, st>>>(bb + bz * N * d, db + bz * N * d, N * d,
                                                       e_flag, scale, scale_kind, clamp_max, part);
        MM25_LAUNCH_CHECK();
      }
    }
    return MM25_OK;
  }
  if (!use_e && !db_peers && batched_fits(B, M, N)) {
    __nv_bfloat16* G = reinterpret_cast<__nv_bfloat16*>(ws0);
    float* cbuf = reinterpret_cast<float*>(
        (ws0 + (size_t)B * M * N * 2 + 255) & ~uintptr_t(255));
    float* Cc = dot_bound > 0.f ? cbuf : nullptr;
    const int sms = tc_sm_count();
    const size_t smem = sizeof(GradSmem) + 1024;
    MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (!col_lse) MM25_CUDA_CHECK(cudaMemsetAsync(cbuf, 0, sizeof(float) * B * N, st));
    if (Cc && col_lse) {
      grad_col_coef_kernel<<<dim3((unsigned)ceil_div(N, 256), (unsigned)B), 256, 0, st>>>(
          col_lse, N, scale, scale_kind, clamp_max, gout, w_col, dot_bound, Cc);
      MM25_LAUNCH_CHECK();
    }
    CUtensorMap ta, tb, tg;
    int rc = make_tmap_bf16(&ta, a, (uint64_t)(B * M), (uint64_t)d, BM, BK);
    if (rc) return rc;
    if ((rc = make_tmap_bf16(&tb, b, (uint64_t)(B * N), (uint64_t)d, BN, BK))) return rc;
    if ((rc = make_tmap_bf16(&tg, G, (uint64_t)(B * M), (uint64_t)N, 32, 32, 64))) return rc;
    const int num_rb = (int)(B * M / BM), ntiles = (int)(N / BN);
    int nseg = 1;
    {
      double best = 0.0;
      const int max_seg = ntiles >= 8 ? ntiles / 4 : 1;
      for (int ns = 1; ns <= max_seg && ns <= 64; ++ns) {
        const int tps = (ntiles + ns - 1) / ns;
        const int sg = (ntiles + tps - 1) / tps;
        const long items = (long)num_rb * sg;
        const double eff = (double)items / ((double)((items + sms - 1) / sms) * sms);
        if (eff > best + 0.01) { best = eff; nseg = sg; }
      }
    }
    const int tps = (ntiles + nseg - 1) / nseg;
    nseg = (ntiles + tps - 1) / tps;
    const long items = (long)num_rb * nseg;
    const int grid = (int)(items < sms ? items : sms);
    if (dscale_part)
      MM25_CUDA_CHECK(cudaMemsetAsync(dscale_part, 0, sizeof(float) * bwd_nparts(B, M, N, d), st));
    // opt-in experiment (MM25_SWEEP_2CTA=1), as in the panel loop below; pairs of row blocks
    // must not straddle two batch entries
    static const bool sweep2b = getenv("MM25_SWEEP_2CTA") != nullptr;
    if (sweep2b && (M / BM) % 2 == 0 && sms >= 2) {
      CUtensorMap tb2;
      if ((rc = make_tmap_bf16(&tb2, b, (uint64_t)(B * N), (uint64_t)d, BN / 2, BK))) return rc;
      const long items2 = (long)(num_rb / 2) * nseg;
      const int pairs = (int)(items2 < sms / 2 ? items2 : sms / 2);
      MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc2_kernel,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem));
      sim_grad_tc2_kernel<<<2 * pairs, FWD_THREADS, smem, st>>>(
          ta, tb2, tg, M, N, (int)d, 0, num_rb, nseg, tps, scale, scale_kind, clamp_max,
          diag_offset, row_lse, col_lse ? col_lse : cbuf, w_row, w_col, w_diag, w_all, gout,
          dscale_part, Cc, dot_bound, nullptr, mask_diag);
    } else
    { void* tt_ = timing_begin(MM25_T_SIM_GRAD, st);
    sim_grad_tc_kernel<<<grid, FWD_THREADS, smem, st>>>(
        ta, tb, tg, M, N, (int)d, 0, num_rb, nseg, tps, scale, scale_kind, clamp_max, diag_offset,
        row_lse, col_lse ? col_lse : cbuf, w_row, w_col, w_diag, w_all, gout, dscale_part, Cc,
        dot_bound, nullptr, mask_diag);
    timing_end(tt_, st); }
    MM25_LAUNCH_CHECK();
    GemmDesc ga{G, b, da, d, M, N, 0, 1, 0, 1.f, nullptr, 0, 0, 0, (int)B};
    GemmDesc gb{G, a, db, d, N, M, 1, 1, 0, 1.f, nullptr, 0, 0, 0, (int)B};
    return gemm_grouped(da ? &ga : nullptr, db ? &gb : nullptr, d, nullptr, st);
  }
  const int64_t R = use_e ? M : panel_rows(M, N);
  // [N] floats after the panel: the column factors C_j of the static-shift path; for row-only
  // losses (col_lse = NULL, w_col = 0) also the placeholder the kernel's column loads read
  float* cbuf = reinterpret_cast<float*>(
      use_e ? ws0 : ((ws0 + (size_t)R * (size_t)N * 2 + 255) & ~uintptr_t(255)));
  float* Cc = dot_bound > 0.f ? cbuf : nullptr;
  if (!col_lse) MM25_CUDA_CHECK(cudaMemsetAsync(cbuf, 0, sizeof(float) * N, st));
  const int sms = tc_sm_count();
  const size_t smem = sizeof(GradSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const __nv_bfloat16* ab = (const __nv_bfloat16*)a;
  const __nv_bfloat16* bb = (const __nv_bfloat16*)b;
  int64_t part = 0;
  for (int64_t bz = 0; bz < B; ++bz) {
    const __nv_bfloat16* a_b = ab + bz * M * d;
    const __nv_bfloat16* b_b = bb + bz * N * d;
    __nv_bfloat16* G = use_e ? reinterpret_cast<__nv_bfloat16*>(e_buf) + bz * M * N
                             : reinterpret_cast<__nv_bfloat16*>(ws0);
    CUtensorMap ta, tb;
    int rc = make_tmap_bf16(&ta, a_b, (uint64_t)M, (uint64_t)d, BM, BK);
    if (rc) return rc;
    rc = make_tmap_bf16(&tb, b_b, (uint64_t)N, (uint64_t)d, BN, BK);
    if (rc) return rc;
    if (Cc && col_lse) {
      grad_col_coef_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(
          col_lse + bz * N, N, scale, scale_kind, clamp_max, gout ? gout + bz : nullptr, w_col,
          dot_bound, Cc);
      MM25_LAUNCH_CHECK();
    }
    if (use_e) {
      // E -> G' in place (no-op when the forward did not raise the flag; the recompute kernel
      // below then fills the same buffer)
      dim3 grid((unsigned)ceil_div(N, GE_COLS), (unsigned)ceil_div(M, GE_ROWS));
      grad_from_e_kernel<<<grid, 256, 0, st>>>(
          G, e_flag, M, N, scale, scale_kind, clamp_max, diag_offset, row_lse + bz * M, Cc, w_row,
          w_diag, w_all, gout ? gout + bz : nullptr, dot_bound);
      MM25_LAUNCH_CHECK();
    }
    for (int64_t p0 = 0; p0 < M; p0 += R) {
      const int64_t rows = (M - p0 < R) ? (M - p0) : R;
      const int num_rb = (int)(rows / BM), ntiles = (int)(N / BN);
      // column segments so that items fill the SMs (same heuristic as the forward)
      int nseg = 1;
      {
        double best = 0.0;
        const int max_seg = ntiles >= 8 ? ntiles / 4 : 1;
        for (int ns = 1; ns <= max_seg && ns <= 64; ++ns) {
          const int tps = (ntiles + ns - 1) / ns;
          const int sg = (ntiles + tps - 1) / tps;
          const long items = (long)num_rb * sg;
          const double eff = (double)items / ((double)((items + sms - 1) / sms) * sms);
          if (eff > best + 0.01) { best = eff; nseg = sg; }
        }
      }
      const int tps = (ntiles + nseg - 1) / nseg;
      nseg = (ntiles + tps - 1) / tps;
      const long items = (long)num_rb * nseg;
      const int grid = (int)(items < sms ? items : sms);
      if (dscale_part) {
        MM25_CUDA_CHECK(cudaMemsetAsync(dscale_part + part, 0, sizeof(float) * sms, st));
      }
      CUtensorMap tg;  // the [rows, N] bf16 panel, stored in 32 x 32 boxes (64B swizzle)
      rc = make_tmap_bf16(&tg, G, (uint64_t)rows, (uint64_t)N, 32, 32, 64);
      if (rc) return rc;
      // opt-in experiment (MM25_SWEEP_2CTA=1): CTA pairs sweep 256-row super-blocks with
      // `cta_group::2` MMAs (sim_tc_sweep.cuh); needs whole pairs of row blocks
      static const bool sweep2 = getenv("MM25_SWEEP_2CTA") != nullptr;
      if (sweep2 && !use_e && num_rb % 2 == 0 && (p0 / BM) % 2 == 0 && sms >= 2) {
        CUtensorMap tb2;  // B in 128-row boxes: each CTA of a pair stages half of the 256 columns
        rc = make_tmap_bf16(&tb2, b_b, (uint64_t)N, (uint64_t)d, BN / 2, BK);
        if (rc) return rc;
        const long items2 = (long)(num_rb / 2) * nseg;
        const int pairs = (int)(items2 < sms / 2 ? items2 : sms / 2);
        MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc2_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem));
        sim_grad_tc2_kernel<<<2 * pairs, FWD_THREADS, smem, st>>>(
            ta, tb2, tg, M, N, (int)d, (int)(p0 / BM), num_rb, nseg, tps, scale, scale_kind,
            clamp_max, diag_offset, row_lse + bz * M, col_lse ? col_lse + bz * N : cbuf, w_row,
            w_col, w_diag, w_all, gout ? gout + bz : nullptr,
            dscale_part ? dscale_part + part : nullptr, Cc, dot_bound, nullptr, mask_diag);
      } else
      { void* tt_ = timing_begin(MM25_T_SIM_GRAD, st);
      sim_grad_tc_kernel<<<grid, FWD_THREADS, smem, st>>>(
          ta, tb, tg, M, N, (int)d, (int)(p0 / BM), num_rb, nseg, tps, scale, scale_kind, clamp_max,
          diag_offset, row_lse + bz * M, col_lse ? col_lse + bz * N : cbuf, w_row, w_col, w_diag,
          w_all,
          gout ? gout + bz : nullptr, dscale_part ? dscale_part + part : nullptr, Cc,
          dot_bound, use_e ? e_flag : nullptr, mask_diag);
      timing_end(tt_, st); }
      MM25_LAUNCH_CHECK();
      part += sms;
      {
        // dA[panel] = G . Yhat and dB += G^T . Xhat[panel] in one grouped launch
        GemmDesc ga{G, b_b, da ? da + (bz * M + p0) * d : nullptr, d, rows, N, 0, 1, 0, 1.f,
                    nullptr, 0, 0};
        GemmDesc gb{G, a_b + p0 * d, db ? db + bz * N * d : nullptr, d, N, rows, 1, 1,
                    p0 > 0 ? 1 : 0, 1.f, db_peers, npeers, db_peers ? N / npeers : 0,
                    // row-sharded layout: diag_offset = rank * M (loss.py:94-95)
                    db_peers ? (int)((diag_offset / M) % npeers) : 0};
        const bool have_db = db != nullptr || db_peers != nullptr;
        static const bool split = getenv("MM25_GEMM_SPLIT") != nullptr;  // A/B switch
        if (split) {
          rc = gemm_grouped(da ? &ga : nullptr, nullptr, d, nullptr, st);
          if (rc) return rc;
          rc = gemm_grouped(nullptr, have_db ? &gb : nullptr, d, nullptr, st);
        } else {
          rc = gemm_grouped(da ? &ga : nullptr, have_db ? &gb : nullptr, d, nullptr, st);
        }
        if (rc) return rc;
      }
    }
    if (dscale_part) {
      MM25_CUDA_CHECK(cudaMemsetAsync(dscale_part + part, 0, sizeof(float) * sms, st));
      if (use_e) {
        if (da)
          dscale_from_grad_kernel<<<sms, 1024, 0, st>>>(a_b, da + bz * M * d, M * d, e_flag, scale,
                                                       scale_kind, clamp_max, dscale_part + part);
        else
          dscale_from_grad_kernel<<<sms, 1024, 0, st>>>(b_b, db + bz * N * d, N * d, e_flag, scale,
                                                       scale_kind, clamp_max, dscale_part + part);
        MM25_LAUNCH_CHECK();
      }
      part += sms;
    }
  }
  return MM25_OK;
}

// ------------------------------------------------------------------------------------------
// fp32-grade backward on the tensor cores (SURVEY §7 hard part 1; clipLoss.py:106 `.float()`).
// Operands are two-term bf16 splits x = x1 + x2 (16 significant bits, prep.cu split kernel):
//   a4 = [a1 a1 a2], b4 = [b1 b2 b1]  ([rows, 3 d] bf16)  =>  a4 . b4 = a1.b1 + a1.b2 + a2.b1
// (the a2.b2 product, 2^-18 of the dot product, is dropped) so the unchanged sweep recomputes
// fp32-grade logits with K = 3 d.  The gradient panel is split
// the same way, G' = G1 + G2: the sweep runs twice per panel (second pass emits g - bf16(g)), and
// every product G . Y becomes three accumulating grouped-GEMM launches
//   dA = G1 Y1 + G1 Y2 + G2 Y1,   dB likewise with X1, X2.
// Cost: 2 x 3 sweeps + 3 x GEMMs = 24 N^2 d tensor FLOPs against 6 on the bf16 path — and against
// the FFMA kernels' 6 N^2 d at ~10 TFLOP/s.  B entries are walked one by one (panel loop).
size_t bwd_split_workspace_bytes(int64_t M, int64_t N) {
  return (size_t)panel_rows(M, N) * (size_t)N * 2 + 256 + (size_t)N * sizeof(float) + 256;
}
int64_t bwd_split_nparts(int64_t B, int64_t M, int64_t N) {
  return B * ceil_div(M, panel_rows(M, N)) * tc_sm_count();
}

int launch_bwd_split(const void* a4, const void* b4, const void* a_hi, const void* a_lo,
                     const void* b_hi, const void* b_lo, int64_t B, int64_t M, int64_t N,
                     int64_t d, const float* scale, int scale_kind, float clamp_max,
                     int64_t diag_offset, const float* row_lse, const float* col_lse,
                     float w_row, float w_col, float w_diag, float w_all, const float* gout,
                     float* da, float* db, float* dscale_part, void* ws, size_t ws_bytes,
                     float dot_bound, cudaStream_t st, int mask_diag) {
  const int64_t d4 = MM25_SPLIT_K * d;   // width of the concatenated split operands
  if (!eligible_fwd(B, M, N, d4, MM25_BF16, MM25_SIM_PLAIN) || d % 256 != 0 ||
      !gemm_eligible(M >= 128 ? 128 : M, d, N))
    return MM25_ERR_SHAPE;
  if (ws_bytes < bwd_split_workspace_bytes(M, N)) return MM25_ERR_WORKSPACE;
  const uintptr_t ws0 = (reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255);
  const int64_t R = panel_rows(M, N);
  __nv_bfloat16* G = reinterpret_cast<__nv_bfloat16*>(ws0);
  float* cbuf = reinterpret_cast<float*>((ws0 + (size_t)R * (size_t)N * 2 + 255) & ~uintptr_t(255));
  float* Cc = dot_bound > 0.f ? cbuf : nullptr;
  if (!col_lse) MM25_CUDA_CHECK(cudaMemsetAsync(cbuf, 0, sizeof(float) * N, st));
  const int sms = tc_sm_count();
  const size_t smem = sizeof(GradSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const __nv_bfloat16* a4b = (const __nv_bfloat16*)a4;
  const __nv_bfloat16* b4b = (const __nv_bfloat16*)b4;
  const __nv_bfloat16* ah = (const __nv_bfloat16*)a_hi;
  const __nv_bfloat16* al = (const __nv_bfloat16*)a_lo;
  const __nv_bfloat16* bh = (const __nv_bfloat16*)b_hi;
  const __nv_bfloat16* bl = (const __nv_bfloat16*)b_lo;
  int64_t part = 0;
  for (int64_t bz = 0; bz < B; ++bz) {
    CUtensorMap ta, tb;
    int rc = make_tmap_bf16(&ta, a4b + bz * M * d4, (uint64_t)M, (uint64_t)d4, BM, BK);
    if (rc) return rc;
    if ((rc = make_tmap_bf16(&tb, b4b + bz * N * d4, (uint64_t)N, (uint64_t)d4, BN, BK))) return rc;
    if (Cc && col_lse) {
      grad_col_coef_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(
          col_lse + bz * N, N, scale, scale_kind, clamp_max, gout ? gout + bz : nullptr, w_col,
          dot_bound, Cc);
      MM25_LAUNCH_CHECK();
    }
    for (int64_t p0 = 0; p0 < M; p0 += R) {
      const int64_t rows = (M - p0 < R) ? (M - p0) : R;
      const int num_rb = (int)(rows / BM), ntiles = (int)(N / BN);
      int nseg = 1;
      {
        double best = 0.0;
        const int max_seg = ntiles >= 8 ? ntiles / 4 : 1;
        for (int ns = 1; ns <= max_seg && ns <= 64; ++ns) {
          const int tps = (ntiles + ns - 1) / ns;
          const int sg = (ntiles + tps - 1) / tps;
          const long items = (long)num_rb * sg;
          const double eff = (double)items / ((double)((items + sms - 1) / sms) * sms);
          if (eff > best + 0.01) { best = eff; nseg = sg; }
        }
      }
      const int tps = (ntiles + nseg - 1) / nseg;
      nseg = (ntiles + tps - 1) / tps;
      const long items = (long)num_rb * nseg;
      const int grid = (int)(items < sms ? items : sms);
      if (dscale_part) MM25_CUDA_CHECK(cudaMemsetAsync(dscale_part + part, 0, sizeof(float) * sms, st));
      CUtensorMap tg;
      if ((rc = make_tmap_bf16(&tg, G, (uint64_t)rows, (uint64_t)N, 32, 32, 64))) return rc;
      for (int lo = 0; lo < 2; ++lo) {
        { void* tt_ = timing_begin(MM25_T_SIM_GRAD, st);
        sim_grad_tc_kernel<<<grid, FWD_THREADS, smem, st>>>(
            ta, tb, tg, M, N, (int)d4, (int)(p0 / BM), num_rb, nseg, tps, scale, scale_kind,
            clamp_max, diag_offset, row_lse + bz * M, col_lse ? col_lse + bz * N : cbuf, w_row,
            w_col, w_diag, w_all, gout ? gout + bz : nullptr,
            (dscale_part && lo == 0) ? dscale_part + part : nullptr, Cc, dot_bound, nullptr,
            (lo ? 2 : 0) | (mask_diag ? 1 : 0));
        timing_end(tt_, st); }
        MM25_LAUNCH_CHECK();
        for (int t = 0; t < 2; ++t) {   // second operand: hi term, then lo term
          if (lo && t) continue;       // G_lo Y_lo: 2^-18 of the product, dropped
          const __nv_bfloat16* yb = (t ? bl : bh) + bz * N * d;
          const __nv_bfloat16* xb = (t ? al : ah) + (bz * M + p0) * d;
          const int first = (lo == 0 && t == 0);
          GemmDesc ga{G, yb, da ? da + (bz * M + p0) * d : nullptr, d, rows, N, 0, 1,
                      first ? 0 : 1, 1.f, nullptr, 0, 0};
          GemmDesc gb{G, xb, db ? db + bz * N * d : nullptr, d, N, rows, 1, 1,
                      (first && p0 == 0) ? 0 : 1, 1.f, nullptr, 0, 0};
          rc = gemm_grouped(da ? &ga : nullptr, db ? &gb : nullptr, d, nullptr, st);
          if (rc) return rc;
        }
      }
      part += sms;
    }
  }
  return MM25_OK;
}

}  // namespace tc
}  // namespace mm25

extern "C" void mm25_debug_set_g_panel_bytes(size_t bytes) { mm25::tc::set_g_cap(bytes); }
