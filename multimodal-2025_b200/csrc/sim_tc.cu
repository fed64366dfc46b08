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
  extern __shared__ uint8_t smem_raw[];
  FwdSmem& sm = *reinterpret_cast<FwdSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = (int)(N / BN);
  const int nitems = total_rb * nseg;
  Sweep w;
  w.KC = d / BK; w.ntiles = ntiles; w.nseg = nseg; w.tps = tps;
  w.nrb_per_batch = nrb_per_batch; w.first_rbg = 0; w.num_rbg = total_rb; w.N = N;
  sweep_setup(sm, &tmap_a, &tmap_b);
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) sweep_producer(sm, &tmap_a, &tmap_b, w);
  } else if (warp == 1) {
    if (lane == 0) sweep_mma(sm, tmem, w);
  } else {
    // ===================================================================== epilogue warps
    const int q = warp & 3;        // TMEM lane quadrant this warp may read
    const int ew = warp - 2;       // 0..7
    const int half = ew >> 2;      // which half of the tile's columns this warp handles
    const float s = load_scale(scale, scale_kind, clamp_max);
    const float c2 = s * LOG2E;
    const float m2s = dot_bound * fabsf(c2);
    // force_slow carries two flags: bit 0 = test hook (robust epilogue), bit 1 = MASKDIAG mode
    // (InfoLOOB, cloobLoss.py:28-31: the positive enters the LSE as the constant -10000)
    const bool mask_diag = (force_slow & 2) != 0;
    const bool sums_only_ok = (force_slow & 4) == 0;  // A/B switch MM25_COL_PARTIAL_MAX
    force_slow &= 1;
    const float vmask = s != 0.f ? -10000.f / s : 0.f;
    const bool fast_static = !force_slow && dot_bound > 0.f && 2.f * m2s <= SPREAD_LIMIT;
    // E = 2^(z log2e - m2s) in bf16 is kept for the backward (static shift only): it replaces
    // the S-tile recompute there.  e_flag tells the backward whether E is valid.
    const bool keep_e = store_e && fast_static;
    if (store_e && blockIdx.x == 0 && threadIdx.x == 64) *e_flag = keep_e ? 1 : 0;
    uint8_t* stage = sm.stage[ew];
    const uint32_t st_row = smem_u32(stage) + lane * 64;
    const uint32_t st_xor = (lane >> 1) & 3;
    if (keep_e && lane == 0) tma_prefetch_desc(&tmap_e);
    uint32_t as = 0, accph = 0, buf = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const int rbg = item / nseg, seg = item % nseg;
      const int bz = rbg / nrb_per_batch, rb = rbg % nrb_per_batch;
      const int64_t row = (int64_t)rb * BM + q * 32 + lane;  // row inside this batch entry
      const int64_t dcol = row + diag_offset;                // its positive column
      float rm = fast_static ? m2s : -INFINITY, rl = 0.f;
      const int t0 = seg * tps, t1 = min(ntiles, t0 + tps);
      for (int t = t0; t < t1; ++t) {
        const int64_t col0 = (int64_t)t * BN;
        mbar_wait(&sm.acc_full[as], accph);
        tc_fence_after();
        const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + as * BN +
                               half * CH_PER_WARP * 32;

        auto process = [&](uint32_t (&r)[32], int chl) {
          const int ch = half * CH_PER_WARP + chl;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          const int64_t cbase = col0 + ch * 32;
          // positive logit (warp-uniform test: this chunk crosses the warp's diagonal range)
          {
            const int64_t d0 = (int64_t)rb * BM + q * 32 + diag_offset;
            if (d0 + 31 >= cbase && d0 < cbase + 32) {
              const int rel = (int)(dcol - cbase);
              float dv = 0.f;
#pragma unroll
              for (int j = 0; j < 32; ++j) dv = (j == rel) ? v[j] : dv;
              if (rel >= 0 && rel < 32) diag[(int64_t)bz * M + row] = s * dv;
              if (mask_diag) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = (j == rel) ? vmask : v[j];
              }
            }
          }
          float e[32];
          float cm;
          if (fast_static) {
            // static shift m2s >= every z log2e: one exp per logit, no running maximum
            float rs = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) { e[j] = ex2_approx(fmaf(v[j], c2, -m2s)); rs += e[j]; }
            rl += rs;
            cm = m2s;
            if (keep_e) {
              uint32_t pk[16];
#pragma unroll
              for (int j2 = 0; j2 < 16; ++j2) {
                __nv_bfloat162 h = __floats2bfloat162_rn(e[2 * j2], e[2 * j2 + 1]);
                pk[j2] = *reinterpret_cast<uint32_t*>(&h);
              }
              if (lane == 0) tma_store_wait_read<0>();
              __syncwarp();
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const uint32_t addr = st_row + ((u ^ st_xor) << 4);
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr),
                             "r"(pk[4 * u]), "r"(pk[4 * u + 1]), "r"(pk[4 * u + 2]),
                             "r"(pk[4 * u + 3])
                             : "memory");
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_2d(&tmap_e, stage, (int32_t)cbase,
                             (int32_t)((int64_t)bz * M + (int64_t)rb * BM + q * 32));
                tma_store_commit();
              }
            }
          } else {
            float tmx = v[0], tmn = v[0];
#pragma unroll
            for (int j = 1; j < 32; ++j) { tmx = fmaxf(tmx, v[j]); tmn = fminf(tmn, v[j]); }
            const float bmx = warp_max(tmx), bmn = -warp_max(-tmn);
            const float hi2 = fmaxf(bmx * c2, bmn * c2), lo2 = fminf(bmx * c2, bmn * c2);
            if (!force_slow && hi2 - lo2 <= SPREAD_LIMIT) {
              // fast path: one exp per logit, shared block shift
              const float m2 = hi2;
              float rs = 0.f;
#pragma unroll
              for (int j = 0; j < 32; ++j) { e[j] = ex2_approx(fmaf(v[j], c2, -m2)); rs += e[j]; }
              const float mn = fmaxf(rm, m2);
              rl = rl * ex2_approx(rm - mn) + rs * ex2_approx(m2 - mn);
              rm = mn;
              cm = m2;
            } else {
              // robust path: per-row and per-column shifts (two exps per logit)
              float tz = v[0] * c2;
#pragma unroll
              for (int j = 1; j < 32; ++j) tz = fmaxf(tz, v[j] * c2);
              const float mn = fmaxf(rm, tz);
              float rs = 0.f;
#pragma unroll
              for (int j = 0; j < 32; ++j) rs += ex2_approx(fmaf(v[j], c2, -mn));
              rl = rl * ex2_approx(rm - mn) + rs;
              rm = mn;
              cm = 0.f;
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float cj = warp_max(v[j] * c2);
                e[j] = ex2_approx(fmaf(v[j], c2, -cj));
                cm = (j == lane) ? cj : cm;
              }
            }
          }
          const float cs = butterfly_colsum(e, lane);
          sm.colm[buf][q][ch * 32 + lane] = cm;
          sm.cols[buf][q][ch * 32 + lane] = cs;
        };

        static_assert(CH_PER_WARP % 2 == 0, "pipeline is unrolled by two");
        uint32_t ra[32], rb_[32];
        tmem_ld_32x32_issue(tbase, ra);
        tmem_ld_wait(ra);
#pragma unroll
        for (int c = 0; c < CH_PER_WARP; c += 2) {
          tmem_ld_32x32_issue(tbase + (c + 1) * 32, rb_);
          process(ra, c);
          tmem_ld_wait(rb_);
          if (c + 2 < CH_PER_WARP) tmem_ld_32x32_issue(tbase + (c + 2) * 32, ra);
          process(rb_, c + 1);
          if (c + 2 < CH_PER_WARP) tmem_ld_wait(ra);
        }
        // all TMEM reads of this stage are complete (wait::ld inside the load helper)
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
        if (++as == NACC) { as = 0; accph ^= 1; }
        // merge the four 32-row groups of this tile; warp ew owns columns [32 ew, 32 ew + 32)
        named_bar_sync(1, 32 * NEPI);
        {
          const int c = ew * 32 + lane;   // 8 warps x 32 lanes = the tile's 256 columns
          const int64_t o = ((int64_t)bz * nrb_per_batch + rb) * N + col0 + c;
          if (fast_static && sums_only_ok) {
            // every partial shares the shift m2s: plain sums, and the merge kernel knows the
            // maximum without reading it (half the partial traffic)
            col_pl[o] = (sm.cols[buf][0][c] + sm.cols[buf][1][c]) +
                        (sm.cols[buf][2][c] + sm.cols[buf][3][c]);
          } else {
            float m = sm.colm[buf][0][c], l = sm.cols[buf][0][c];
#pragma unroll
            for (int g = 1; g < 4; ++g) {
              const float m2 = sm.colm[buf][g][c], l2 = sm.cols[buf][g][c];
              const float mn = fmaxf(m, m2);
              l = l * ex2_approx(m - mn) + l2 * ex2_approx(m2 - mn);
              m = mn;
            }
            col_pm[o] = m;
            col_pl[o] = l;
          }
        }
        buf ^= 1;
      }
      const int64_t o = ((int64_t)bz * nseg * 2 + seg * 2 + half) * M + row;
      row_pm[o] = rm;
      row_pl[o] = rl;
    }
    if (keep_e && lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
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
  sim_fwd_tc_kernel<<<g.grid, FWD_THREADS, smem, st>>>(
      ta, tb, te, e_out ? 1 : 0, e_flag, M, N, (int)d, g.nrb, g.total_rb, g.nseg, g.tps, scale, scale_kind, clamp_max,
      diag_offset, row_pm, row_pl, col_pm, col_pl, diag, flags, dot_bound);
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
