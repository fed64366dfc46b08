// sim_tc_mod.cu — tensor-core path of the MODULATED similarity loss (AlignCLIP semantic term,
// reference AlignCLIP/align_clip/loss.py:229-242):
//     z_ij = s [ (a_i . b_j) (1 - u_i . v_j) + delta_ij D_i ]
// Every 128 x 128 logit tile needs TWO contractions (features over d, normalised semantic
// features over d2), so a tile owns two fp32 accumulators in TMEM (2 x 128 columns; two such
// stages fill the 512 columns) and the K-loop runs the (a, b) chunks into the first and the
// (u, v) chunks into the second.  Same warp roles as sim_tc.cu: warp 0 TMA producer, warp 1 MMA
// issuer, 8 epilogue warps (32 rows x 64 columns each).
//   forward : row log-sum-exp (online max, one exp per logit) + the positive logit
//   backward: G''_ij = gout s (1 - w_ij) G_ij as a bf16 panel (TMA store), d/dD_i, d/ds partials;
//             dA = G'' B and dB += G''^T A then run on the grouped GEMM of gemm_tc.cu.
// Only row statistics exist in this mode (w_col = 0), which is all the reference's CE needs.
#include "tc_common.cuh"
#include "sim_tc.h"
#include "gemm_tc.h"

namespace mm25 {
namespace tc {
int tc_sm_count();
}
namespace tcm {

using namespace mm25::tc;

constexpr int MBM = 128, MBN = 128, MBK = 64;
constexpr int MNST = 6;                        // ring stages (A 16 KB + B 16 KB each)
constexpr int MNACC = 2;                       // accumulator stages of 2 x 128 TMEM columns
constexpr int MT_BYTES = MBM * MBK * 2;        // 16384, both operands
constexpr int MNEPI = 8;
constexpr int MTHREADS = 64 + 32 * MNEPI;
constexpr int MCH = MBN / 32 / 2;              // 32-column chunks per epilogue warp (2)
constexpr float K_LOG2E = 1.4426950408889634f;
constexpr float K_LN2 = 0.6931471805599453f;

struct ModSmem {
  uint8_t a[MNST][MT_BYTES];
  uint8_t b[MNST][MT_BYTES];
  uint8_t stage[MNEPI][2048];  // bf16 G'' chunk of each warp (32 rows x 64 B, 64B swizzle)
  uint8_t stage2[MNEPI][2048];  // second panel (Hopfield: P)
  float red[MNEPI];
  uint64_t full[MNST], empty[MNST];
  uint64_t acc_full[MNACC], acc_empty[MNACC];
  uint32_t tmem_base;
};

// items are ordered segment-major (item = seg * num_rb + row block): the CTAs of one wave then
// sweep the same few column segments, whose B tiles (N (d + d2) 2 B in total, more than L2 at
// cfg5) stay L2-resident while every row block passes over them.
struct ModGeom {
  int KC1, KC2, ntiles, nseg, tps, first_rb, num_rb;
};

// EPI 0: MODULATED forward; 1: MODULATED gradient panel; 2: Hopfield-retrieval gradient panels
// (cloobLoss.py:43-47 backward: a = Q, b = K, u = dO, v = K, so the second accumulator is
// dP = dO K^T; row_lse = LSE_j beta q.k_j, diag_add = delta_i = dO_i . O_i;
// panel 1 = dZ = beta P (dP - delta), panel 2 = P).
constexpr int EPI_FWD = 0, EPI_GRAD = 1, EPI_HOPF = 2;

template <int EPI>
__global__ void __launch_bounds__(MTHREADS, 1)
sim_mod_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                  const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_u,
                  const __grid_constant__ CUtensorMap tmap_v,
                  const __grid_constant__ CUtensorMap tmap_g,
                  const __grid_constant__ CUtensorMap tmap_p, ModGeom g, int64_t M, int64_t N,
                  const float* __restrict__ scale, int scale_kind, float clamp_max,
                  int64_t diag_offset, const float* __restrict__ diag_add,
                  // forward outputs: partial (max, sum) per (segment, half) and the positives
                  float* __restrict__ row_pm, float* __restrict__ row_pl,
                  float* __restrict__ diag,
                  // backward inputs / outputs
                  const float* __restrict__ row_lse, float w_row, float w_diag, float w_all,
                  const float* __restrict__ gout, float* __restrict__ dscale_part,
                  float* __restrict__ ddiag) {
  constexpr bool GRAD = EPI != EPI_FWD;
  extern __shared__ uint8_t smem_raw[];
  ModSmem& sm = *reinterpret_cast<ModSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < MNST; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < MNACC; ++i) {
      mbar_init(&sm.acc_full[i], 1);
      mbar_init(&sm.acc_empty[i], MNEPI);
    }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&sm.tmem_base, 512);
    if (lane == 0) {
      tma_prefetch_desc(&tmap_a); tma_prefetch_desc(&tmap_b);
      tma_prefetch_desc(&tmap_u); tma_prefetch_desc(&tmap_v);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const int nitems = g.num_rb * g.nseg;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int rb = g.first_rb + item % g.num_rb, seg = item / g.num_rb;
        const int32_t arow = rb * MBM;
        const int t0 = seg * g.tps, t1 = min(g.ntiles, t0 + g.tps);
        for (int t = t0; t < t1; ++t) {
          const int32_t brow = t * MBN;
          for (int kc = 0; kc < g.KC1 + g.KC2; ++kc) {
            mbar_wait(&sm.empty[st], ph ^ 1);
            mbar_arrive_expect_tx(&sm.full[st], 2 * MT_BYTES);
            if (kc < g.KC1) {
              tma_load_2d(sm.a[st], &tmap_a, &sm.full[st], kc * MBK, arow);
              tma_load_2d(sm.b[st], &tmap_b, &sm.full[st], kc * MBK, brow);
            } else {
              tma_load_2d(sm.a[st], &tmap_u, &sm.full[st], (kc - g.KC1) * MBK, arow);
              tma_load_2d(sm.b[st], &tmap_v, &sm.full[st], (kc - g.KC1) * MBK, brow);
            }
            if (++st == MNST) { st = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(MBM, MBN, 0, 0);
      uint32_t st = 0, ph = 0, as = 0, accph = 0;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int seg = item / g.num_rb;
        const int t0 = seg * g.tps, t1 = min(g.ntiles, t0 + g.tps);
        for (int t = t0; t < t1; ++t) {
          mbar_wait(&sm.acc_empty[as], accph ^ 1);
          tc_fence_after();
          for (int kc = 0; kc < g.KC1 + g.KC2; ++kc) {
            // first accumulator: a . b over d; second: u . v over d2
            const uint32_t dcol = tmem + as * (2 * MBN) + (kc < g.KC1 ? 0 : MBN);
            const bool first = (kc == 0) || (kc == g.KC1);
            mbar_wait(&sm.full[st], ph);
            tc_fence_after();
            const uint32_t abase = smem_u32(sm.a[st]);
            const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
            for (int k = 0; k < MBK / 16; ++k) {
              const uint64_t ad = make_smem_desc(abase + k * 32, 16, 1024);
              const uint64_t bd = make_smem_desc(bbase + k * 32, 16, 1024);
              umma_bf16(dcol, ad, bd, idesc, !(first && k == 0));
            }
            umma_commit(&sm.empty[st]);
            if (++st == MNST) { st = 0; ph ^= 1; }
          }
          umma_commit(&sm.acc_full[as]);
          if (++as == MNACC) { as = 0; accph ^= 1; }
        }
      }
    }
  } else {
    const int q = warp & 3;
    const int ew = warp - 2;
    const int half = ew >> 2;
    const float s = load_scale(scale, scale_kind, clamp_max);
    const float c2 = s * K_LOG2E;
    const float go = (GRAD && gout) ? __ldg(gout) : 1.f;
    const float gs = go * s;
    uint8_t* stage = sm.stage[ew];
    const uint32_t st_row = smem_u32(stage) + lane * 64;
    const uint32_t st_xor = (lane >> 1) & 3;
    if (GRAD && lane == 0) tma_prefetch_desc(&tmap_g);
    if (EPI == EPI_HOPF && lane == 0) tma_prefetch_desc(&tmap_p);
    const uint32_t st2_row = smem_u32(sm.stage2[ew]) + lane * 64;
    float ds = 0.f;
    uint32_t as = 0, accph = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const int rb = g.first_rb + item % g.num_rb, seg = item / g.num_rb;
      const int64_t row = (int64_t)rb * MBM + q * 32 + lane;
      const int64_t dcol = row + diag_offset;
      const float Di = diag_add ? __ldg(diag_add + row) : 0.f;
      const float r2 = GRAD ? __ldg(row_lse + row) * K_LOG2E : 0.f;
      const int32_t grow0 = (int32_t)((int64_t)(rb - g.first_rb) * MBM + q * 32);
      float rm = -INFINITY, rl = 0.f;
      const int t0 = seg * g.tps, t1 = min(g.ntiles, t0 + g.tps);
      for (int t = t0; t < t1; ++t) {
        const int64_t col0 = (int64_t)t * MBN;
        mbar_wait(&sm.acc_full[as], accph);
        tc_fence_after();
        const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + as * (2 * MBN) + half * 64;
#pragma unroll
        for (int c = 0; c < MCH; ++c) {
          uint32_t rv[32], rw[32];
          tmem_ld_32x32_issue(tbase + c * 32, rv);
          tmem_ld_32x32_issue(tbase + MBN + c * 32, rw);
          tmem_ld_wait(rv);
          tmem_ld_wait(rw);
          const int64_t cbase = col0 + half * 64 + c * 32;
          if (EPI == EPI_HOPF) {
            uint32_t pz[16], pp[16];
#pragma unroll
            for (int j2 = 0; j2 < 16; ++j2) {
              float zq[2], pq[2];
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                const int j = 2 * j2 + u;
                const float p = ex2_approx(fmaf(__uint_as_float(rv[j]), c2, -r2));
                pq[u] = p;
                zq[u] = s * p * (__uint_as_float(rw[j]) - Di);
              }
              if (w_all != 0.f) {   // EPI_HOPF reuses w_all as the "emit the low split half" flag
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                  zq[u] -= __bfloat162float(__float2bfloat16_rn(zq[u]));
                  pq[u] -= __bfloat162float(__float2bfloat16_rn(pq[u]));
                }
              }
              __nv_bfloat162 hz = __floats2bfloat162_rn(zq[0], zq[1]);
              __nv_bfloat162 hp = __floats2bfloat162_rn(pq[0], pq[1]);
              pz[j2] = *reinterpret_cast<uint32_t*>(&hz);
              pp[j2] = *reinterpret_cast<uint32_t*>(&hp);
            }
            if (lane == 0) tma_store_wait_read<0>();
            __syncwarp();
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t off = (u ^ st_xor) << 4;
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(st_row + off),
                           "r"(pz[4 * u]), "r"(pz[4 * u + 1]), "r"(pz[4 * u + 2]),
                           "r"(pz[4 * u + 3])
                           : "memory");
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(st2_row + off),
                           "r"(pp[4 * u]), "r"(pp[4 * u + 1]), "r"(pp[4 * u + 2]),
                           "r"(pp[4 * u + 3])
                           : "memory");
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmap_g, stage, (int32_t)cbase, grow0);
              tma_store_2d(&tmap_p, sm.stage2[ew], (int32_t)cbase, grow0);
              tma_store_commit();
            }
            continue;
          }
          const int rel = (int)min((int64_t)64, max((int64_t)-1, dcol - cbase));
          float base[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float wv = __uint_as_float(rw[j]);
            base[j] = __uint_as_float(rv[j]) * (1.f - wv);
            if (j == rel) base[j] += Di;
          }
          if (EPI == EPI_FWD) {
            if (rel >= 0 && rel < 32) {
              float dv = 0.f;
#pragma unroll
              for (int j = 0; j < 32; ++j) dv = (j == rel) ? base[j] : dv;
              diag[row] = s * dv;
            }
            float tm = base[0] * c2;
#pragma unroll
            for (int j = 1; j < 32; ++j) tm = fmaxf(tm, base[j] * c2);
            const float mn = fmaxf(rm, tm);
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) acc += ex2_approx(fmaf(base[j], c2, -mn));
            rl = rl * ex2_approx(rm - mn) + acc;
            rm = mn;
          } else {
            uint32_t packed[16];
#pragma unroll
            for (int j2 = 0; j2 < 16; ++j2) {
              float gq[2];
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                const int j = 2 * j2 + u;
                const float p = ex2_approx(fmaf(base[j], c2, -r2));
                float gg = fmaf(w_row, p, -w_all);
                if (j == rel) {
                  gg -= w_diag;
                  if (ddiag) ddiag[row] = gs * gg;
                }
                ds = fmaf(gg, base[j], ds);
                gq[u] = gs * (1.f - __uint_as_float(rw[j])) * gg;
              }
              __nv_bfloat162 h = __floats2bfloat162_rn(gq[0], gq[1]);
              packed[j2] = *reinterpret_cast<uint32_t*>(&h);
            }
            if (lane == 0) tma_store_wait_read<0>();
            __syncwarp();
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t addr = st_row + ((u ^ st_xor) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr),
                           "r"(packed[4 * u]), "r"(packed[4 * u + 1]), "r"(packed[4 * u + 2]),
                           "r"(packed[4 * u + 3])
                           : "memory");
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmap_g, stage, (int32_t)cbase, grow0);
              tma_store_commit();
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
        if (++as == MNACC) { as = 0; accph ^= 1; }
      }
      if (EPI == EPI_FWD) {
        const int64_t o = ((int64_t)seg * 2 + half) * M + row;
        row_pm[o] = rm;
        row_pl[o] = rl;
      }
    }
    if (GRAD) {
      if (lane == 0) tma_store_wait_all<0>();
    }
    if (EPI == EPI_GRAD) {
      ds = warp_sum(ds) * go;
      if (lane == 0) sm.red[ew] = ds;
      named_bar_sync(1, 32 * MNEPI);
      if (warp == 2 && lane == 0 && dscale_part) {
        float tot = 0.f;
#pragma unroll
        for (int k = 0; k < MNEPI; ++k) tot += sm.red[k];
        dscale_part[blockIdx.x] = tot;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// row_lse[i] = ln2 (m + log2 sum_p l_p 2^(m_p - m)) over the nparts = 2 nseg partials, fixed order
__global__ void __launch_bounds__(256)
mod_row_merge_kernel(const float* __restrict__ row_pm, const float* __restrict__ row_pl,
                     int nparts, int64_t M, float* __restrict__ row_lse) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= M) return;
  float m = -INFINITY;
  for (int p = 0; p < nparts; ++p) m = fmaxf(m, row_pm[(int64_t)p * M + i]);
  float l = 0.f;
  for (int p = 0; p < nparts; ++p)
    l += row_pl[(int64_t)p * M + i] * exp2f(row_pm[(int64_t)p * M + i] - m);
  row_lse[i] = (m + log2f(l)) * K_LN2;
}

__global__ void __launch_bounds__(256)
cast_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i + 3 < n) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(x + i));
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo);
    o.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(y + i) = o;
  } else {
    for (int64_t k = i; k < n; ++k) y[k] = __float2bfloat16(x[k]);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static size_t g_panel_cap = (size_t)1 << 30;

bool eligible(int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2, int dtype, bool bwd) {
  if (dtype != MM25_BF16 || B != 1) return false;
  if (M % MBM || N % MBN || d % MBK || d2 % MBK || d < MBK || d2 < MBK) return false;
  if (M > (int64_t)1 << 30 || N > (int64_t)1 << 30) return false;
  if (bwd && !(d % 256 == 0 && d <= 1024 && N % 256 == 0 && gemm_eligible(128, d, N)))
    return false;
  return true;
}

static int64_t panel_rows(int64_t M, int64_t N) {
  int64_t r = (int64_t)(g_panel_cap / ((size_t)N * 2)) / MBM * MBM;
  if (r < MBM) r = MBM;
  return r > M ? M : r;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static void pick_segments(int num_rb, int ntiles, int sms, int* nseg_out, int* tps_out) {
  int nseg = 1;
  double best = 0.0;
  const int max_seg = ntiles >= 8 ? ntiles / 4 : 1;
  for (int ns = 1; ns <= max_seg && ns <= 64; ++ns) {
    const int tps = (ntiles + ns - 1) / ns;
    const int sg = (ntiles + tps - 1) / tps;
    const long items = (long)num_rb * sg;
    const double eff = (double)items / ((double)((items + sms - 1) / sms) * sms);
    if (eff > best + 0.01) { best = eff; nseg = sg; }
  }
  const int tps = (ntiles + nseg - 1) / nseg;
  *nseg_out = (ntiles + tps - 1) / tps;
  *tps_out = tps;
}

// workspace: bf16 copies of u and v, then (fwd) the row partials or (bwd) the G'' panel
size_t workspace_bytes(int64_t M, int64_t N, int64_t d, int64_t d2, bool bwd) {
  size_t w = 256 + align256((size_t)M * d2 * 2) + align256((size_t)N * d2 * 2);
  if (bwd) return w + align256((size_t)panel_rows(M, N) * (size_t)N * 2);
  return w + 2 * align256((size_t)2 * 64 * M * sizeof(float));  // nseg <= 64, two halves
}

int64_t bwd_nparts(int64_t M, int64_t N) {
  return ceil_div(M, panel_rows(M, N)) * tc::tc_sm_count();
}

struct Prep {
  __nv_bfloat16 *ub, *vb;
  uint8_t* rest;
};

static int prepare(const float* u, const float* v, int64_t M, int64_t N, int64_t d2, void* ws,
                   Prep* p, cudaStream_t st) {
  uintptr_t w0 = (reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255);
  p->ub = reinterpret_cast<__nv_bfloat16*>(w0);
  w0 += align256((size_t)M * d2 * 2);
  p->vb = reinterpret_cast<__nv_bfloat16*>(w0);
  w0 += align256((size_t)N * d2 * 2);
  p->rest = reinterpret_cast<uint8_t*>(w0);
  const int64_t nu = M * d2, nv = N * d2;
  cast_bf16_kernel<<<(unsigned)ceil_div(nu, 1024), 256, 0, st>>>(u, p->ub, nu);
  MM25_LAUNCH_CHECK();
  if (v == u && M == N) {
    p->vb = p->ub;
  } else {
    cast_bf16_kernel<<<(unsigned)ceil_div(nv, 1024), 256, 0, st>>>(v, p->vb, nv);
    MM25_LAUNCH_CHECK();
  }
  return MM25_OK;
}

int launch_fwd(const void* a, const void* b, const float* u, const float* v,
               const float* diag_add, int64_t M, int64_t N, int64_t d, int64_t d2,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               float* row_lse, float* diag, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (!eligible(1, M, N, d, d2, MM25_BF16, false)) return MM25_ERR_SHAPE;
  if (ws_bytes < workspace_bytes(M, N, d, d2, false)) return MM25_ERR_WORKSPACE;
  Prep p;
  int rc = prepare(u, v, M, N, d2, ws, &p, st);
  if (rc) return rc;
  const int sms = tc::tc_sm_count();
  ModGeom g;
  g.KC1 = (int)(d / MBK); g.KC2 = (int)(d2 / MBK); g.ntiles = (int)(N / MBN);
  g.first_rb = 0; g.num_rb = (int)(M / MBM);
  pick_segments(g.num_rb, g.ntiles, sms, &g.nseg, &g.tps);
  float* row_pm = reinterpret_cast<float*>(p.rest);
  float* row_pl = reinterpret_cast<float*>(p.rest + align256((size_t)2 * 64 * M * sizeof(float)));
  CUtensorMap ta, tb, tu, tv;
  if ((rc = make_tmap_bf16(&ta, a, (uint64_t)M, (uint64_t)d, MBM, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tb, b, (uint64_t)N, (uint64_t)d, MBN, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tu, p.ub, (uint64_t)M, (uint64_t)d2, MBM, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tv, p.vb, (uint64_t)N, (uint64_t)d2, MBN, MBK))) return rc;
  const size_t smem = sizeof(ModSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_mod_tc_kernel<EPI_FWD>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long items = (long)g.num_rb * g.nseg;
  const int grid = (int)(items < sms ? items : sms);
  // rows whose positive lies outside [0, N) keep diag = 0
  MM25_CUDA_CHECK(cudaMemsetAsync(diag, 0, sizeof(float) * M, st));
{ void* tt_ = timing_begin(MM25_T_SIM_MOD, st);
  sim_mod_tc_kernel<EPI_FWD><<<grid, MTHREADS, smem, st>>>(
      ta, tb, tu, tv, ta, ta, g, M, N, scale, scale_kind, clamp_max, diag_offset, diag_add, row_pm,
      row_pl, diag, nullptr, 0.f, 0.f, 0.f, nullptr, nullptr, nullptr);
  timing_end(tt_, st); }
  MM25_LAUNCH_CHECK();
  mod_row_merge_kernel<<<(unsigned)ceil_div(M, 256), 256, 0, st>>>(row_pm, row_pl, 2 * g.nseg, M,
                                                                   row_lse);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

int launch_bwd(const void* a, const void* b, const float* u, const float* v,
               const float* diag_add, int64_t M, int64_t N, int64_t d, int64_t d2,
               const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
               const float* row_lse, float w_row, float w_diag, float w_all, const float* gout,
               float* da, float* db, float* dscale_part, float* ddiag, void* ws,
               size_t ws_bytes, cudaStream_t st) {
  if (!eligible(1, M, N, d, d2, MM25_BF16, true)) return MM25_ERR_SHAPE;
  if (ws_bytes < workspace_bytes(M, N, d, d2, true)) return MM25_ERR_WORKSPACE;
  Prep p;
  int rc = prepare(u, v, M, N, d2, ws, &p, st);
  if (rc) return rc;
  __nv_bfloat16* G = reinterpret_cast<__nv_bfloat16*>(p.rest);
  const int sms = tc::tc_sm_count();
  const int64_t R = panel_rows(M, N);
  CUtensorMap ta, tb, tu, tv;
  if ((rc = make_tmap_bf16(&ta, a, (uint64_t)M, (uint64_t)d, MBM, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tb, b, (uint64_t)N, (uint64_t)d, MBN, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tu, p.ub, (uint64_t)M, (uint64_t)d2, MBM, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tv, p.vb, (uint64_t)N, (uint64_t)d2, MBN, MBK))) return rc;
  const size_t smem = sizeof(ModSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_mod_tc_kernel<EPI_GRAD>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (ddiag) MM25_CUDA_CHECK(cudaMemsetAsync(ddiag, 0, sizeof(float) * M, st));
  const __nv_bfloat16* ab = (const __nv_bfloat16*)a;
  const __nv_bfloat16* bb = (const __nv_bfloat16*)b;
  int64_t part = 0;
  for (int64_t p0 = 0; p0 < M; p0 += R) {
    const int64_t rows = (M - p0 < R) ? (M - p0) : R;
    ModGeom g;
    g.KC1 = (int)(d / MBK); g.KC2 = (int)(d2 / MBK); g.ntiles = (int)(N / MBN);
    g.first_rb = (int)(p0 / MBM); g.num_rb = (int)(rows / MBM);
    pick_segments(g.num_rb, g.ntiles, sms, &g.nseg, &g.tps);
    const long items = (long)g.num_rb * g.nseg;
    const int grid = (int)(items < sms ? items : sms);
    if (dscale_part) MM25_CUDA_CHECK(cudaMemsetAsync(dscale_part + part, 0, sizeof(float) * sms, st));
    CUtensorMap tg;
    if ((rc = make_tmap_bf16(&tg, G, (uint64_t)rows, (uint64_t)N, 32, 32, 64))) return rc;
{ void* tt_ = timing_begin(MM25_T_SIM_MOD, st);
    sim_mod_tc_kernel<EPI_GRAD><<<grid, MTHREADS, smem, st>>>(
        ta, tb, tu, tv, tg, tg, g, M, N, scale, scale_kind, clamp_max, diag_offset, diag_add, nullptr,
        nullptr, nullptr, row_lse, w_row, w_diag, w_all, gout,
        dscale_part ? dscale_part + part : nullptr, ddiag);
  timing_end(tt_, st); }
    MM25_LAUNCH_CHECK();
    part += sms;
    GemmDesc ga{G, bb, da ? da + p0 * d : nullptr, d, rows, N, 0, 1, 0, 1.f, nullptr, 0, 0};
    GemmDesc gb{G, ab + p0 * d, db, d, N, rows, 1, 1, p0 > 0 ? 1 : 0, 1.f, nullptr, 0, 0};
    rc = gemm_grouped(da ? &ga : nullptr, db ? &gb : nullptr, d, nullptr, st);
    if (rc) return rc;
  }
  return MM25_OK;
}

// ------------------------------------------------------------------------------------------
// Hopfield retrieval backward on the tensor cores (cloobLoss.py:43-47, O = softmax(beta Q K^T) K):
//   dP = dO K^T,  dZ = beta P o (dP - delta_i),  dQ = dZ K,  dK = dZ^T Q + P^T dO.
// One dual-accumulator sweep writes the bf16 panels dZ and P; three GEMMs follow (two grouped).
// ------------------------------------------------------------------------------------------
static int64_t hopf_panel_rows(int64_t M, int64_t N) {
  int64_t r = (int64_t)(g_panel_cap / ((size_t)N * 4)) / MBM * MBM;  // two bf16 panels
  if (r < MBM) r = MBM;
  return r > M ? M : r;
}

bool hopfield_eligible(int64_t M, int64_t N, int64_t d) {
  return eligible(1, M, N, d, d, MM25_BF16, true);
}

size_t hopfield_workspace_bytes(int64_t M, int64_t N, int64_t d) {
  return 256 + align256((size_t)M * d * 2) +
         2 * align256((size_t)hopf_panel_rows(M, N) * (size_t)N * 2);
}

int launch_hopfield_bwd(const void* q, const void* k, const float* d_o, int64_t M, int64_t N,
                        int64_t d, const float* beta, const float* lse, const float* delta,
                        float* dq, float* dk, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (!hopfield_eligible(M, N, d)) return MM25_ERR_SHAPE;
  if (ws_bytes < hopfield_workspace_bytes(M, N, d)) return MM25_ERR_WORKSPACE;
  uintptr_t w0 = (reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255);
  __nv_bfloat16* dob = reinterpret_cast<__nv_bfloat16*>(w0);
  w0 += align256((size_t)M * d * 2);
  const int64_t R = hopf_panel_rows(M, N);
  __nv_bfloat16* GZ = reinterpret_cast<__nv_bfloat16*>(w0);
  __nv_bfloat16* GP = reinterpret_cast<__nv_bfloat16*>(w0 + align256((size_t)R * (size_t)N * 2));
  cast_bf16_kernel<<<(unsigned)ceil_div(M * d, 1024), 256, 0, st>>>(d_o, dob, M * d);
  MM25_LAUNCH_CHECK();
  const int sms = tc::tc_sm_count();
  CUtensorMap ta, tb, tu;
  int rc;
  if ((rc = make_tmap_bf16(&ta, q, (uint64_t)M, (uint64_t)d, MBM, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tb, k, (uint64_t)N, (uint64_t)d, MBN, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tu, dob, (uint64_t)M, (uint64_t)d, MBM, MBK))) return rc;
  const size_t smem = sizeof(ModSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_mod_tc_kernel<EPI_HOPF>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const __nv_bfloat16* qb = (const __nv_bfloat16*)q;
  const __nv_bfloat16* kb = (const __nv_bfloat16*)k;
  for (int64_t p0 = 0; p0 < M; p0 += R) {
    const int64_t rows = (M - p0 < R) ? (M - p0) : R;
    ModGeom g;
    g.KC1 = g.KC2 = (int)(d / MBK); g.ntiles = (int)(N / MBN);
    g.first_rb = (int)(p0 / MBM); g.num_rb = (int)(rows / MBM);
    pick_segments(g.num_rb, g.ntiles, sms, &g.nseg, &g.tps);
    const long items = (long)g.num_rb * g.nseg;
    const int grid = (int)(items < sms ? items : sms);
    CUtensorMap tg, tp;
    if ((rc = make_tmap_bf16(&tg, GZ, (uint64_t)rows, (uint64_t)N, 32, 32, 64))) return rc;
    if ((rc = make_tmap_bf16(&tp, GP, (uint64_t)rows, (uint64_t)N, 32, 32, 64))) return rc;
{ void* tt_ = timing_begin(MM25_T_SIM_MOD, st);
    sim_mod_tc_kernel<EPI_HOPF><<<grid, MTHREADS, smem, st>>>(
        ta, tb, tu, tb, tg, tp, g, M, N, beta, MM25_SCALE_MUL, 0.f, 0, delta, nullptr, nullptr,
        nullptr, lse, 0.f, 0.f, 0.f, nullptr, nullptr, nullptr);
  timing_end(tt_, st); }
    MM25_LAUNCH_CHECK();
    // dQ[panel] = dZ K  and  dK (+)= dZ^T Q[panel]  in one grouped launch, then dK += P^T dO
    GemmDesc ga{GZ, kb, dq ? dq + p0 * d : nullptr, d, rows, N, 0, 1, 0, 1.f, nullptr, 0, 0};
    GemmDesc gb{GZ, qb + p0 * d, dk, d, N, rows, 1, 1, p0 > 0 ? 1 : 0, 1.f, nullptr, 0, 0};
    rc = gemm_grouped(dq ? &ga : nullptr, dk ? &gb : nullptr, d, nullptr, st);
    if (rc) return rc;
    if (dk) {
      GemmDesc gc{GP, dob + p0 * d, dk, d, N, rows, 1, 1, 1, 1.f, nullptr, 0, 0};
      rc = gemm_grouped(nullptr, &gc, d, nullptr, st);
      if (rc) return rc;
    }
  }
  return MM25_OK;
}

// ------------------------------------------------------------------------------------------
// fp32-grade Hopfield backward (cloobLoss.py:43-47 in fp32; contract 1e-5 / 1e-4): the same
// dual-accumulator sweep on two-term bf16 splits.  q4 / do4 = [hi hi lo], k4 = [hi lo hi]
// ([rows, 3 d], prep.cu split kernel), so S = Q K^T and T = dO K^T accumulate with 16 significant
// bits per operand; the panels dZ and P are split too (second pass emits x - bf16(x)) and every
// product with them becomes accumulating grouped-GEMM launches over the (hi, lo) operand terms.
bool hopfield_split_eligible(int64_t M, int64_t N, int64_t d) {
  return eligible(1, M, N, MM25_SPLIT_K * d, MM25_SPLIT_K * d, MM25_BF16, false) &&
         d % 256 == 0 && d <= 1024 &&
         N % 256 == 0 && gemm_eligible(128, d, N);
}

size_t hopfield_split_workspace_bytes(int64_t M, int64_t N, int64_t d) {
  return 256 + align256((size_t)M * MM25_SPLIT_K * d * 2) + 2 * align256((size_t)M * d * 2) +
         2 * align256((size_t)hopf_panel_rows(M, N) * (size_t)N * 2);
}

extern "C" int mm25_split_bf16x2(const float* x, int64_t rows, int64_t d, int role, void* out4,
                                 void* hi, void* lo, void* stream);

int launch_hopfield_bwd_split(const void* q4, const void* k4, const void* q_hi, const void* q_lo,
                              const void* k_hi, const void* k_lo, const float* d_o, int64_t M,
                              int64_t N, int64_t d, const float* beta, const float* lse,
                              const float* delta, float* dq, float* dk, void* ws, size_t ws_bytes,
                              cudaStream_t st) {
  if (!hopfield_split_eligible(M, N, d)) return MM25_ERR_SHAPE;
  if (ws_bytes < hopfield_split_workspace_bytes(M, N, d)) return MM25_ERR_WORKSPACE;
  const int64_t d4 = MM25_SPLIT_K * d;
  uintptr_t w0 = (reinterpret_cast<uintptr_t>(ws) + 255) & ~uintptr_t(255);
  __nv_bfloat16* do4 = reinterpret_cast<__nv_bfloat16*>(w0);
  w0 += align256((size_t)M * d4 * 2);
  __nv_bfloat16* doh = reinterpret_cast<__nv_bfloat16*>(w0);
  w0 += align256((size_t)M * d * 2);
  __nv_bfloat16* dol = reinterpret_cast<__nv_bfloat16*>(w0);
  w0 += align256((size_t)M * d * 2);
  const int64_t R = hopf_panel_rows(M, N);
  __nv_bfloat16* GZ = reinterpret_cast<__nv_bfloat16*>(w0);
  __nv_bfloat16* GP = reinterpret_cast<__nv_bfloat16*>(w0 + align256((size_t)R * (size_t)N * 2));
  int rc = mm25_split_bf16x2(d_o, M, d, 0, do4, doh, dol, st);
  if (rc) return rc;
  const int sms = tc::tc_sm_count();
  CUtensorMap ta, tb, tu;
  if ((rc = make_tmap_bf16(&ta, q4, (uint64_t)M, (uint64_t)d4, MBM, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tb, k4, (uint64_t)N, (uint64_t)d4, MBN, MBK))) return rc;
  if ((rc = make_tmap_bf16(&tu, do4, (uint64_t)M, (uint64_t)d4, MBM, MBK))) return rc;
  const size_t smem = sizeof(ModSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_mod_tc_kernel<EPI_HOPF>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const __nv_bfloat16* qs[2] = {(const __nv_bfloat16*)q_hi, (const __nv_bfloat16*)q_lo};
  const __nv_bfloat16* ks[2] = {(const __nv_bfloat16*)k_hi, (const __nv_bfloat16*)k_lo};
  const __nv_bfloat16* dos[2] = {doh, dol};
  for (int64_t p0 = 0; p0 < M; p0 += R) {
    const int64_t rows = (M - p0 < R) ? (M - p0) : R;
    ModGeom g;
    g.KC1 = g.KC2 = (int)(d4 / MBK); g.ntiles = (int)(N / MBN);
    g.first_rb = (int)(p0 / MBM); g.num_rb = (int)(rows / MBM);
    pick_segments(g.num_rb, g.ntiles, sms, &g.nseg, &g.tps);
    const long items = (long)g.num_rb * g.nseg;
    const int grid = (int)(items < sms ? items : sms);
    CUtensorMap tg, tp;
    if ((rc = make_tmap_bf16(&tg, GZ, (uint64_t)rows, (uint64_t)N, 32, 32, 64))) return rc;
    if ((rc = make_tmap_bf16(&tp, GP, (uint64_t)rows, (uint64_t)N, 32, 32, 64))) return rc;
    for (int lo = 0; lo < 2; ++lo) {
      { void* tt_ = timing_begin(MM25_T_SIM_MOD, st);
      sim_mod_tc_kernel<EPI_HOPF><<<grid, MTHREADS, smem, st>>>(
          ta, tb, tu, tb, tg, tp, g, M, N, beta, MM25_SCALE_MUL, 0.f, 0, delta, nullptr, nullptr,
          nullptr, lse, 0.f, 0.f, lo ? 1.f : 0.f, nullptr, nullptr, nullptr);
      timing_end(tt_, st); }
      MM25_LAUNCH_CHECK();
      for (int t = 0; t < 2; ++t) {
        if (lo && t) continue;   // lo x lo products: 2^-18, dropped
        const int first = (lo == 0 && t == 0);
        // dQ[panel] (+)= dZ K_t  and  dK (+)= dZ^T Q_t[panel]
        GemmDesc ga{GZ, ks[t], dq ? dq + p0 * d : nullptr, d, rows, N, 0, 1, first ? 0 : 1, 1.f,
                    nullptr, 0, 0};
        GemmDesc gb{GZ, qs[t] + p0 * d, dk, d, N, rows, 1, 1, (first && p0 == 0) ? 0 : 1, 1.f,
                    nullptr, 0, 0};
        rc = gemm_grouped(dq ? &ga : nullptr, dk ? &gb : nullptr, d, nullptr, st);
        if (rc) return rc;
        if (dk) {   // dK += P^T dO_t[panel]
          GemmDesc gc{GP, dos[t] + p0 * d, dk, d, N, rows, 1, 1, 1, 1.f, nullptr, 0, 0};
          rc = gemm_grouped(nullptr, &gc, d, nullptr, st);
          if (rc) return rc;
        }
      }
    }
  }
  return MM25_OK;
}

}  // namespace tcm
}  // namespace mm25

extern "C" size_t mm25_hopfield_bwd_tc_workspace_bytes(int64_t M, int64_t N, int64_t d) {
  if (!mm25::tcm::hopfield_eligible(M, N, d)) return 0;
  return mm25::tcm::hopfield_workspace_bytes(M, N, d);
}

extern "C" int mm25_hopfield_bwd_tc(const void* q, const void* k, const float* d_o, int64_t M,
                                    int64_t N, int64_t d, const float* beta, const float* lse,
                                    const float* delta, float* dq, float* dk, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  if (!q || !k || !d_o || !beta || !lse || !delta || !workspace) return MM25_ERR_ARG;
  if (!dq && !dk) return MM25_OK;
  return mm25::tcm::launch_hopfield_bwd(q, k, d_o, M, N, d, beta, lse, delta, dq, dk, workspace,
                                        workspace_bytes, (cudaStream_t)stream);
}
