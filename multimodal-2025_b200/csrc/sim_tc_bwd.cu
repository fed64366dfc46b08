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

namespace mm25 {
namespace tc {

int tc_sm_count();

__global__ void __launch_bounds__(FWD_THREADS, 1)
sim_grad_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                   const __grid_constant__ CUtensorMap tmap_b, int64_t M, int64_t N, int d,
                   int first_rb, int num_rb, int nseg, int tps, const float* __restrict__ scale,
                   int scale_kind, float clamp_max, int64_t diag_offset,
                   const float* __restrict__ row_lse, const float* __restrict__ col_lse,
                   float w_row, float w_col, float w_diag, float w_all,
                   const float* __restrict__ gout, __nv_bfloat16* __restrict__ G,
                   float* __restrict__ dscale_part) {
  extern __shared__ uint8_t smem_raw[];
  FwdSmem& sm = *reinterpret_cast<FwdSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Sweep w;
  w.KC = d / BK; w.ntiles = (int)(N / BN); w.nseg = nseg; w.tps = tps;
  w.nrb_per_batch = (int)(M / BM); w.first_rbg = first_rb; w.num_rbg = num_rb; w.N = N;
  sweep_setup(sm, &tmap_a, &tmap_b);
  const uint32_t tmem = sm.tmem_base;
  const int nitems = num_rb * nseg;

  if (warp == 0) {
    if (lane == 0) sweep_producer(sm, &tmap_a, &tmap_b, w);
  } else if (warp == 1) {
    if (lane == 0) sweep_mma(sm, tmem, w);
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const float s = load_scale(scale, scale_kind, clamp_max);
    const float c2 = s * LOG2E;
    const float go = gout ? __ldg(gout) : 1.f;
    const float gs = go * s;
    float ds = 0.f;
    uint32_t as = 0, accph = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const int rb = first_rb + item / nseg, seg = item % nseg;
      const int64_t row = (int64_t)rb * BM + q * 32 + lane;
      const int64_t dcol = row + diag_offset;
      const float r2 = __ldg(row_lse + row) * LOG2E;
      __nv_bfloat16* grow = G + (int64_t)(row - (int64_t)first_rb * BM) * N;
      const int t0 = seg * tps, t1 = min(w.ntiles, t0 + tps);
      for (int t = t0; t < t1; ++t) {
        const int64_t col0 = (int64_t)t * BN;
        mbar_wait(&sm.acc_full[as], accph);
        tc_fence_after();
#pragma unroll 1
        for (int ch = half * CH_PER_WARP; ch < (half + 1) * CH_PER_WARP; ++ch) {
          float v[32];
          tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + as * BN + ch * 32, v);
          const int64_t cbase = col0 + ch * 32;
          const int rel = (int)min((int64_t)64, max((int64_t)-1, dcol - cbase));
          const float4* cl4 = reinterpret_cast<const float4*>(col_lse + cbase);
          uint32_t packed[16];
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 cc = __ldg(cl4 + j4);
            const float cj[4] = {cc.x, cc.y, cc.z, cc.w};
            float gq[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int j = j4 * 4 + u;
              const float zl = v[j] * c2;
              float gg = w_row * ex2_approx(zl - r2) + w_col * ex2_approx(fmaf(-cj[u], LOG2E, zl));
              gg -= w_all;
              gg -= (j == rel) ? w_diag : 0.f;
              ds = fmaf(gg, v[j], ds);
              gq[u] = gg * gs;
            }
            __nv_bfloat162 lo = __floats2bfloat162_rn(gq[0], gq[1]);
            __nv_bfloat162 hi = __floats2bfloat162_rn(gq[2], gq[3]);
            packed[j4 * 2] = *reinterpret_cast<uint32_t*>(&lo);
            packed[j4 * 2 + 1] = *reinterpret_cast<uint32_t*>(&hi);
          }
          uint4* dst = reinterpret_cast<uint4*>(grow + cbase);
#pragma unroll
          for (int u = 0; u < 4; ++u)
            dst[u] = make_uint4(packed[4 * u], packed[4 * u + 1], packed[4 * u + 2],
                                packed[4 * u + 3]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
        if (++as == NACC) { as = 0; accph ^= 1; }
      }
    }
    // d(loss)/d(scale) partial of this CTA: sum G.z/s = sum G.v (times the upstream gradient)
    ds = warp_sum(ds) * go;
    if (lane == 0) sm.colm[0][0][warp - 2] = ds;
    named_bar_sync(1, 32 * NEPI);
    if (warp == 2 && lane == 0 && dscale_part) {
      float tot = 0.f;
#pragma unroll
      for (int k = 0; k < NEPI; ++k) tot += sm.colm[0][0][k];
      dscale_part[blockIdx.x] = tot;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------
static size_t g_workspace_cap = (size_t)1 << 30;   // bytes of the bf16 G panel
void set_g_cap(size_t bytes) { g_workspace_cap = bytes ? bytes : ((size_t)1 << 30); }

bool eligible_bwd(int64_t B, int64_t M, int64_t N, int64_t d, int dtype, int mode) {
  return eligible_fwd(B, M, N, d, dtype, mode) && d % 256 == 0 && gemm_eligible(M >= 128 ? 128 : M, d, N);
}

static int64_t panel_rows(int64_t M, int64_t N) {
  int64_t r = (int64_t)(g_workspace_cap / ((size_t)N * 2)) / BM * BM;
  if (r < BM) r = BM;
  if (r > M) r = M;
  return r;
}

size_t bwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d) {
  return (size_t)panel_rows(M, N) * (size_t)N * 2 + 256;
}

int64_t bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d) {
  const int64_t R = panel_rows(M, N);
  return B * ceil_div(M, R) * tc_sm_count();
}

int launch_bwd(const void* a, const void* b, int64_t B, int64_t M, int64_t N, int64_t d,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               const float* row_lse, const float* col_lse, float w_row, float w_col,
               float w_diag, float w_all, const float* gout, float* da, float* db,
               float* dscale_part, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (!eligible_bwd(B, M, N, d, MM25_BF16, MM25_SIM_PLAIN)) return MM25_ERR_SHAPE;
  if (!col_lse) return MM25_ERR_ARG;
  if (ws_bytes < bwd_workspace_bytes(B, M, N, d)) return MM25_ERR_WORKSPACE;
  __nv_bfloat16* G = reinterpret_cast<__nv_bfloat16*>(
      (reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255));
  const int64_t R = panel_rows(M, N);
  const int sms = tc_sm_count();
  const size_t smem = sizeof(FwdSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_grad_tc_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const __nv_bfloat16* ab = (const __nv_bfloat16*)a;
  const __nv_bfloat16* bb = (const __nv_bfloat16*)b;
  int64_t part = 0;
  for (int64_t bz = 0; bz < B; ++bz) {
    const __nv_bfloat16* a_b = ab + bz * M * d;
    const __nv_bfloat16* b_b = bb + bz * N * d;
    CUtensorMap ta, tb;
    int rc = make_tmap_bf16(&ta, a_b, (uint64_t)M, (uint64_t)d, BM, BK);
    if (rc) return rc;
    rc = make_tmap_bf16(&tb, b_b, (uint64_t)N, (uint64_t)d, BN, BK);
    if (rc) return rc;
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
      sim_grad_tc_kernel<<<grid, FWD_THREADS, smem, st>>>(
          ta, tb, M, N, (int)d, (int)(p0 / BM), num_rb, nseg, tps, scale, scale_kind, clamp_max,
          diag_offset, row_lse + bz * M, col_lse + bz * N, w_row, w_col, w_diag, w_all,
          gout ? gout + bz : nullptr, G, dscale_part ? dscale_part + part : nullptr);
      MM25_LAUNCH_CHECK();
      part += sms;
      if (da) {
        rc = gemm_bf16(G, 0, b_b, 1, da + (bz * M + p0) * d, d, rows, d, N, 1.f, nullptr, 0, st);
        if (rc) return rc;
      }
      if (db) {
        rc = gemm_bf16(G, 1, a_b + p0 * d, 1, db + bz * N * d, d, N, d, rows, 1.f, nullptr,
                       p0 > 0 ? 1 : 0, st);
        if (rc) return rc;
      }
    }
  }
  return MM25_OK;
}

}  // namespace tc
}  // namespace mm25

extern "C" void mm25_debug_set_g_panel_bytes(size_t bytes) { mm25::tc::set_g_cap(bytes); }
