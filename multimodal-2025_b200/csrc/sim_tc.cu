// sim_tc.cu — bf16 tensor-core path of the fused similarity kernels (sm_100a):
// TMA (128B-swizzled tiles) -> shared memory ring -> tcgen05.mma (fp32 accumulators in TMEM)
// -> tcgen05.ld -> softmax-statistics / gradient epilogues in registers.
//
// Forward (this file, part 1): persistent CTAs; a work item is (128-row block, column segment).
// The A row block stays resident in shared memory for the whole item, B tiles stream through a
// 4-stage ring, 4 TMEM accumulator stages decouple the MMA warp from the 4 epilogue warps.
#include "sim_tc_sweep.cuh"
#include "sim_tc.h"

#include <cstdlib>
#include <mutex>

namespace mm25 {
namespace tc {

// sum over the 32 lanes of e[j] for every j: afterwards lane j holds column j in e[0]
__device__ __forceinline__ float butterfly_colsum(float (&e)[32], int lane) {
#pragma unroll
  for (int w = 16; w >= 1; w >>= 1) {
    const bool up = (lane & w) != 0;
#pragma unroll
    for (int k = 0; k < w; ++k) {
      const float send = up ? e[k] : e[k + w];
      const float keep = up ? e[k + w] : e[k];
      e[k] = keep + __shfl_xor_sync(0xffffffffu, send, w);
    }
  }
  return e[0];
}

__global__ void __launch_bounds__(FWD_THREADS, 1)
sim_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                  const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_e, int store_e,
                  int* __restrict__ e_flag, int64_t M, int64_t N, int d,
                  int nrb_per_batch, int total_rb, int nseg, int tps,
                  const float* __restrict__ scale, int scale_kind, float clamp_max,
                  int64_t diag_offset, float* __restrict__ row_pm, float* __restrict__ row_pl,
                  float* __restrict__ col_pm, float* __restrict__ col_pl,
                  float* __restrict__ diag, int force_slow, float dot_bound) {
#define MM25_PAIR 0
#include "sim_tc_fwd_body.inc"
#undef MM25_PAIR
}

// EXPERIMENTAL (MM25_SWEEP_2CTA=1): clusters of two CTAs; `tmap_b` must have 128-row boxes;
// total_rb and nrb_per_batch even, grid even.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FWD_THREADS, 1)
sim_fwd_tc2_kernel(const __grid_constant__ CUtensorMap tmap_a,
                  const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_e, int store_e,
                  int* __restrict__ e_flag, int64_t M, int64_t N, int d,
                  int nrb_per_batch, int total_rb, int nseg, int tps,
                  const float* __restrict__ scale, int scale_kind, float clamp_max,
                  int64_t diag_offset, float* __restrict__ row_pm, float* __restrict__ row_pl,
                  float* __restrict__ col_pm, float* __restrict__ col_pl,
                  float* __restrict__ diag, int force_slow, float dot_bound) {
#define MM25_PAIR 1
#include "sim_tc_fwd_body.inc"
#undef MM25_PAIR
}

// partials are in log2-scaled units: value = 2^m * l.  Block = 32 elements x 8 partial lanes;
// lane k merges partials k, k+8, ... in order, then the 8 lanes are merged in order.
__device__ __forceinline__ void lse2_merge(float& m, float& l, float m2, float l2) {
  const float mn = fmaxf(m, m2);
  if (mn == -INFINITY) return;
  l = l * exp2f(m - mn) + l2 * exp2f(m2 - mn);
  m = mn;
}

// block = 32 elements x 32 partial lanes: lane ty merges partials ty, ty+32, ... in order, then
// the 32 lanes are merged in order (fixed order -> deterministic).  1024 threads per 32 columns
// keep ~64 MB of partials streaming at HBM speed.
__global__ void __launch_bounds__(1024) fwd_tc_merge_kernel(
    const float* __restrict__ row_pm, const float* __restrict__ row_pl, int nseg, int64_t M,
    float* __restrict__ row_lse, const float* __restrict__ col_pm,
    const float* __restrict__ col_pl, int nrb, int64_t N, float* __restrict__ col_max,
    float* __restrict__ col_sum, const float* __restrict__ scale, int scale_kind,
    float clamp_max, float dot_bound, int force_slow) {
  // same predicate as the forward kernel: on the static-shift path the column partials carry no
  // maxima (all equal m2s)
  const float m2s = dot_bound * fabsf(load_scale(scale, scale_kind, clamp_max) * LOG2E);
  const bool fast_static = !(force_slow & 1) && !(force_slow & 4) && dot_bound > 0.f &&
                           2.f * m2s <= SPREAD_LIMIT;
  __shared__ float sm_m[32][33], sm_l[32][33];
  const int bz = blockIdx.y;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t t = (int64_t)blockIdx.x * 32 + tx;
  {
    float m = -INFINITY, l = 0.f;
    if (t < M)
      for (int p = ty; p < nseg; p += 32)
        lse2_merge(m, l, row_pm[((int64_t)bz * nseg + p) * M + t],
                   row_pl[((int64_t)bz * nseg + p) * M + t]);
    sm_m[ty][tx] = m; sm_l[ty][tx] = l;
    __syncthreads();
    if (ty == 0 && t < M) {
      for (int k = 1; k < 32; ++k) lse2_merge(m, l, sm_m[k][tx], sm_l[k][tx]);
      row_lse[(int64_t)bz * M + t] = (m + log2f(l)) * LN2;
    }
    __syncthreads();
  }
  if (col_max && !fast_static) {  // the static-shift case runs fwd_tc_colsum_static_kernel
    float m = -INFINITY, l = 0.f;
    if (t < N)
      for (int p = ty; p < nrb; p += 32)
        lse2_merge(m, l, fast_static ? m2s : col_pm[((int64_t)bz * nrb + p) * N + t],
                   col_pl[((int64_t)bz * nrb + p) * N + t]);
    sm_m[ty][tx] = m; sm_l[ty][tx] = l;
    __syncthreads();
    if (ty == 0 && t < N) {
      for (int k = 1; k < 32; ++k) lse2_merge(m, l, sm_m[k][tx], sm_l[k][tx]);
      col_max[(int64_t)bz * N + t] = m * LN2;
      col_sum[(int64_t)bz * N + t] = l;
    }
  }
}

// Column sums of the static-shift path: the partials are plain sums under the common maximum
// m2s, so they are added with independent 16-byte loads (512 B contiguous per warp and partial
// row) in a fixed order.  Returns at once on the other paths (same predicate as the forward).
__global__ void __launch_bounds__(1024) fwd_tc_colsum_static_kernel(
    const float* __restrict__ col_pl, int nrb, int64_t N, float* __restrict__ col_max,
    float* __restrict__ col_sum, const float* __restrict__ scale, int scale_kind,
    float clamp_max, float dot_bound, int force_slow) {
  const float m2s = dot_bound * fabsf(load_scale(scale, scale_kind, clamp_max) * LOG2E);
  const bool fast_static = !(force_slow & 1) && !(force_slow & 4) && dot_bound > 0.f &&
                           2.f * m2s <= SPREAD_LIMIT;
  if (!fast_static) return;
  __shared__ float4 sm[32][33];
  const int bz = blockIdx.y;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t c4 = (int64_t)blockIdx.x * 32 + tx, n4 = N >> 2;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c4 < n4) {
    const float4* src = reinterpret_cast<const float4*>(col_pl + (int64_t)bz * nrb * N) + c4;
    int p = ty;
    for (; p + 96 < nrb; p += 128) {
      const float4 v0 = src[(int64_t)p * n4], v1 = src[(int64_t)(p + 32) * n4];
      const float4 v2 = src[(int64_t)(p + 64) * n4], v3 = src[(int64_t)(p + 96) * n4];
      acc.x += (v0.x + v1.x) + (v2.x + v3.x); acc.y += (v0.y + v1.y) + (v2.y + v3.y);
      acc.z += (v0.z + v1.z) + (v2.z + v3.z); acc.w += (v0.w + v1.w) + (v2.w + v3.w);
    }
    for (; p < nrb; p += 32) {
      const float4 v = src[(int64_t)p * n4];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  sm[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && c4 < n4) {
    float4 t = sm[0][tx];
    for (int k = 1; k < 32; ++k) {
      const float4 v = sm[k][tx];
      t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
    }
    const float m = m2s * LN2;
    reinterpret_cast<float4*>(col_max + (int64_t)bz * N)[c4] = make_float4(m, m, m, m);
    reinterpret_cast<float4*>(col_sum + (int64_t)bz * N)[c4] = t;
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                   uint32_t box_rows, uint32_t box_cols, int swizzle_bytes) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return MM25_ERR_CUDA;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (cols * 2) % 16) return MM25_ERR_ALIGN;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {cols * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim,
                   gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? MM25_OK : MM25_ERR_CUDA;
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

bool eligible_fwd(int64_t B, int64_t M, int64_t N, int64_t d, int dtype, int mode) {
  return dtype == MM25_BF16 && (mode == MM25_SIM_PLAIN || mode == MM25_SIM_MASKDIAG) && d % BK == 0 && d >= BK &&
         d <= MAX_KC * BK && M % BM == 0 && N % BN == 0 && NEPI * 32 == BN && B * M < (1ll << 31) &&
         B * N < (1ll << 31);
}

struct FwdGeom { int nrb, total_rb, ntiles, nseg, tps, grid; };

static FwdGeom fwd_geom(int64_t B, int64_t M, int64_t N) {
  FwdGeom g;
  g.nrb = (int)(M / BM);
  g.total_rb = (int)(B * g.nrb);
  g.ntiles = (int)(N / BN);
  const int sms = sm_count();
  int best = 1;
  double best_eff = 0.0;
  const int max_seg = g.ntiles >= 8 ? g.ntiles / 4 : 1;
  for (int ns = 1; ns <= max_seg && ns <= 64; ++ns) {
    const int tps = (g.ntiles + ns - 1) / ns;
    const int nseg = (g.ntiles + tps - 1) / tps;
    const long items = (long)g.total_rb * nseg;
    const double waves = (double)((items + sms - 1) / sms);
    const double eff = (double)items / (waves * sms);
    if (eff > best_eff + 0.01) { best_eff = eff; best = nseg; }
  }
  g.nseg = best;
  g.tps = (g.ntiles + best - 1) / best;
  g.nseg = (g.ntiles + g.tps - 1) / g.tps;
  const long items = (long)g.total_rb * g.nseg;
  g.grid = (int)(items < sms ? items : sms);
  return g;
}

size_t fwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d) {
  FwdGeom g = fwd_geom(B, M, N);
  return sizeof(float) * 2 * (size_t)B * ((size_t)g.nseg * 2 * M + (size_t)g.nrb * N);
}

static int g_force_slow = 0;  // test hook: exercise the robust epilogue path
void set_force_slow(int v) { g_force_slow = v; }

int launch_fwd(const void* a, const void* b, int64_t B, int64_t M, int64_t N, int64_t d,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               float* row_lse, float* col_max, float* col_sum, float* diag, void* ws,
               size_t ws_bytes, float dot_bound, void* e_out, int* e_flag, cudaStream_t st,
               int mask_diag) {
  if (mask_diag && e_out) return MM25_ERR_ARG;
  if (!eligible_fwd(B, M, N, d, MM25_BF16, MM25_SIM_PLAIN)) return MM25_ERR_SHAPE;
  if ((e_out == nullptr) != (e_flag == nullptr)) return MM25_ERR_ARG;
  if (ws_bytes < fwd_workspace_bytes(B, M, N, d)) return MM25_ERR_WORKSPACE;
  FwdGeom g = fwd_geom(B, M, N);
  static const int legacy_partials = getenv("MM25_COL_PARTIAL_MAX") ? 4 : 0;  // A/B switch
  const int flags = g_force_slow | (mask_diag ? 2 : 0) | legacy_partials;
  CUtensorMap ta, tb;
  int rc = make_tmap_bf16(&ta, a, (uint64_t)(B * M), (uint64_t)d, BM, BK);
  if (rc) return rc;
  rc = make_tmap_bf16(&tb, b, (uint64_t)(B * N), (uint64_t)d, BN, BK);
  if (rc) return rc;
  CUtensorMap te = ta;  // placeholder when E is not kept
  if (e_out) {
    rc = make_tmap_bf16(&te, e_out, (uint64_t)(B * M), (uint64_t)N, 32, 32, 64);
    if (rc) return rc;
  }
  float* row_pm = (float*)ws;
  float* row_pl = row_pm + (size_t)B * g.nseg * 2 * M;
  float* col_pm = row_pl + (size_t)B * g.nseg * 2 * M;
  float* col_pl = col_pm + (size_t)B * g.nrb * N;
  const size_t smem = sizeof(FwdSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_fwd_tc_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // opt-in experiment (MM25_SWEEP_2CTA=1): CTA pairs sweep 256-row super-blocks with
  // `cta_group::2` MMAs (sim_tc_sweep.cuh); needs whole pairs of row blocks per batch entry
  static const bool sweep2 = getenv("MM25_SWEEP_2CTA") != nullptr;
  if (sweep2 && !e_out && g.nrb % 2 == 0 && sm_count() >= 2) {
    CUtensorMap tb2;  // B in 128-row boxes: each CTA of a pair stages half of the 256 columns
    rc = make_tmap_bf16(&tb2, b, (uint64_t)(B * N), (uint64_t)d, BN / 2, BK);
    if (rc) return rc;
    const long items2 = (long)(g.total_rb / 2) * g.nseg;
    const int half = sm_count() / 2;
    const int pairs = (int)(items2 < half ? items2 : half);
    MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_fwd_tc2_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    sim_fwd_tc2_kernel<<<2 * pairs, FWD_THREADS, smem, st>>>(
        ta, tb2, te, 0, e_flag, M, N, (int)d, g.nrb, g.total_rb, g.nseg, g.tps, scale, scale_kind,
        clamp_max, diag_offset, row_pm, row_pl, col_pm, col_pl, diag, flags, dot_bound);
  } else
  { void* tt_ = timing_begin(MM25_T_SIM_FWD, st);
  sim_fwd_tc_kernel<<<g.grid, FWD_THREADS, smem, st>>>(
      ta, tb, te, e_out ? 1 : 0, e_flag, M, N, (int)d, g.nrb, g.total_rb, g.nseg, g.tps, scale, scale_kind, clamp_max,
      diag_offset, row_pm, row_pl, col_pm, col_pl, diag, flags, dot_bound);
  timing_end(tt_, st); }
  MM25_LAUNCH_CHECK();
  const int64_t mx = M > N ? M : N;
  fwd_tc_merge_kernel<<<dim3((unsigned)ceil_div(mx, 32), (unsigned)B), 1024, 0, st>>>(
      row_pm, row_pl, g.nseg * 2, M, row_lse, col_pm, col_pl, g.nrb, N, col_max, col_sum, scale,
      scale_kind, clamp_max, dot_bound, flags);
  MM25_LAUNCH_CHECK();
  if (col_max && dot_bound > 0.f) {
    fwd_tc_colsum_static_kernel<<<dim3((unsigned)ceil_div(N, 128), (unsigned)B), 1024, 0, st>>>(
        col_pl, g.nrb, N, col_max, col_sum, scale, scale_kind, clamp_max, dot_bound, flags);
    MM25_LAUNCH_CHECK();
  }
  return MM25_OK;
}

// shared with sim_tc_bwd.cu
int tc_sm_count() { return sm_count(); }

}  // namespace tc
}  // namespace mm25

extern "C" void mm25_debug_force_slow_epilogue(int v) { mm25::tc::set_force_slow(v); }
