// sim_simt.cu — fp32 SIMT fused similarity + online LSE forward, statistic merge, and the
// tile-recomputing backward.  General shapes, three logit modes (mm25.h MM25_SIM_*).
#include "sim_simt.cuh"

namespace mm25 {
namespace simt {

// -----------------------------------------------------------------------------------------
// forward: grid (nsplit, row blocks, batch).  Each CTA sweeps its share of column blocks,
// keeps running row (max,sum) in registers and emits per-tile column (max,sum) partials.
// -----------------------------------------------------------------------------------------
template <typename T, int MODE>
__global__ void __launch_bounds__(NT) sim_fwd_kernel(
    const T* __restrict__ a, const T* __restrict__ b, const float* __restrict__ u,
    const float* __restrict__ v, const float* __restrict__ diag_add, int64_t M, int64_t N, int d,
    int d2, const float* __restrict__ scale, int scale_kind, float clamp_max,
    int64_t diag_offset, float* __restrict__ row_pm, float* __restrict__ row_pl,
    float* __restrict__ col_pm, float* __restrict__ col_pl, float* __restrict__ diag,
    int nsplit, int jb_per_split) {
  __shared__ __align__(16) float As[BK][BM];
  __shared__ __align__(16) float Bs[BK][BN];
  __shared__ float red_m[16][64];
  __shared__ float red_l[16][64];

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int split = blockIdx.x, rb = blockIdx.y, bz = blockIdx.z;
  const int nrb = gridDim.y;
  a += (int64_t)bz * M * d;
  b += (int64_t)bz * N * d;
  if (MODE == MM25_SIM_MODULATED) {
    u += (int64_t)bz * M * d2;
    v += (int64_t)bz * N * d2;
    if (diag_add) diag_add += (int64_t)bz * M;
  }
  const float s = load_scale(scale, scale_kind, clamp_max);
  const int64_t row0 = (int64_t)rb * BM;
  const int njb = (int)ceil((double)N / BN);

  float rm[4], rl[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { rm[i] = -INFINITY; rl[i] = 0.f; }

  const int jb0 = split * jb_per_split;
  const int jb1 = min(njb, jb0 + jb_per_split);
  for (int jb = jb0; jb < jb1; ++jb) {
    const int64_t col0 = (int64_t)jb * BN;
    float acc[4][4] = {};
    tile_dot<T>(acc, As, Bs, a, b, row0, M, col0, N, d);
    float z[4][4];
    if (MODE == MM25_SIM_MODULATED) {
      float acc2[4][4] = {};
      tile_dot<float>(acc2, As, Bs, u, v, row0, M, col0, N, d2);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] *= (1.f - acc2[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t row = row0 + ty * 4 + i;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t col = col0 + tx * 4 + j;
        const bool valid = row < M && col < N;
        const bool isdiag = (col == row + diag_offset);
        float base = acc[i][j];
        if (MODE == MM25_SIM_MODULATED && isdiag && valid && diag_add) base += diag_add[row];
        float zz = s * base;
        if (isdiag && valid) diag[(int64_t)bz * M + row] = zz;
        if (MODE == MM25_SIM_MASKDIAG && isdiag) zz = -10000.f;
        z[i][j] = valid ? zz : -INFINITY;
      }
    }
    // running row statistics
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float tmax = fmaxf(fmaxf(z[i][0], z[i][1]), fmaxf(z[i][2], z[i][3]));
      float mn = fmaxf(rm[i], tmax);
      if (mn > -INFINITY) {
        float l = rl[i] * expf(rm[i] - mn);
#pragma unroll
        for (int j = 0; j < 4; ++j) l += expf(z[i][j] - mn);
        rl[i] = l;
        rm[i] = mn;
      }
    }
    // per-tile column partials
    if (col_pm) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float cm = fmaxf(fmaxf(z[0][j], z[1][j]), fmaxf(z[2][j], z[3][j]));
        float cl = 0.f;
        if (cm > -INFINITY) {
#pragma unroll
          for (int i = 0; i < 4; ++i) cl += expf(z[i][j] - cm);
        }
        red_m[ty][tx * 4 + j] = cm;
        red_l[ty][tx * 4 + j] = cl;
      }
      __syncthreads();
      if (threadIdx.x < 64) {
        float m = -INFINITY, l = 0.f;
#pragma unroll
        for (int t = 0; t < 16; ++t) lse_merge(m, l, red_m[t][threadIdx.x], red_l[t][threadIdx.x]);
        const int64_t col = col0 + threadIdx.x;
        if (col < N) {
          col_pm[((int64_t)bz * nrb + rb) * N + col] = m;
          col_pl[((int64_t)bz * nrb + rb) * N + col] = l;
        }
      }
      __syncthreads();
    }
  }
  // merge row statistics across the 16 tx threads
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    red_m[tx][ty * 4 + i] = rm[i];
    red_l[tx][ty * 4 + i] = rl[i];
  }
  __syncthreads();
  if (threadIdx.x < 64) {
    float m = -INFINITY, l = 0.f;
#pragma unroll
    for (int t = 0; t < 16; ++t) lse_merge(m, l, red_m[t][threadIdx.x], red_l[t][threadIdx.x]);
    const int64_t row = row0 + threadIdx.x;
    if (row < M) {
      row_pm[((int64_t)bz * nsplit + split) * M + row] = m;
      row_pl[((int64_t)bz * nsplit + split) * M + row] = l;
    }
  }
}

// merge partials in fixed order: row_lse = LSE over splits; column (max,sum) over row blocks
__global__ void __launch_bounds__(256) stats_merge_kernel(
    const float* __restrict__ row_pm, const float* __restrict__ row_pl, int nsplit, int64_t M,
    float* __restrict__ row_lse, const float* __restrict__ col_pm,
    const float* __restrict__ col_pl, int nrb, int64_t N, float* __restrict__ col_max,
    float* __restrict__ col_sum) {
  const int bz = blockIdx.y;
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t < M) {
    float m = -INFINITY, l = 0.f;
    for (int p = 0; p < nsplit; ++p)
      lse_merge(m, l, row_pm[((int64_t)bz * nsplit + p) * M + t],
                row_pl[((int64_t)bz * nsplit + p) * M + t]);
    row_lse[(int64_t)bz * M + t] = m + logf(l);
  }
  if (col_max && t < N) {
    float m = -INFINITY, l = 0.f;
    for (int p = 0; p < nrb; ++p)
      lse_merge(m, l, col_pm[((int64_t)bz * nrb + p) * N + t],
                col_pl[((int64_t)bz * nrb + p) * N + t]);
    col_max[(int64_t)bz * N + t] = m;
    col_sum[(int64_t)bz * N + t] = l;
  }
}

// -----------------------------------------------------------------------------------------
// backward, row-stationary pass: dA[rows, d] = sum_j G'_ij b_j with the S tile recomputed.
// grid (d chunks of DC, row blocks, batch); the dA chunk [64 x DC] accumulates in smem.
// -----------------------------------------------------------------------------------------
constexpr int DC = 512;

template <typename T, int MODE>
__global__ void __launch_bounds__(NT) sim_bwd_kernel(
    const T* __restrict__ a, const T* __restrict__ b, const float* __restrict__ u,
    const float* __restrict__ v, const float* __restrict__ diag_add, int diag_by_col, int64_t M,
    int64_t N, int d, int d2, const float* __restrict__ scale, int scale_kind, float clamp_max,
    int64_t diag_offset, const float* __restrict__ row_lse, const float* __restrict__ col_lse,
    float w_row, float w_col, float w_diag, float w_all, const float* __restrict__ gout,
    float* __restrict__ dA, float* __restrict__ dscale_part, float* __restrict__ ddiag) {
  extern __shared__ __align__(16) float smem[];
  float(*As)[BM] = reinterpret_cast<float(*)[BM]>(smem);                 // [BK][64]
  float(*Bs)[BN] = reinterpret_cast<float(*)[BN]>(smem + BK * BM);       // [BK][64]
  float(*Gt)[BM] = reinterpret_cast<float(*)[BM]>(smem + 2 * BK * BM);   // [64 j][64 i]
  float(*Bc)[64] = reinterpret_cast<float(*)[64]>(smem + 2 * BK * BM + 64 * BM);  // [64 j][64 c]
  float* dAs = smem + 2 * BK * BM + 2 * 64 * 64;                         // [64][DC]
  __shared__ float red[NT];

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int dchunk = blockIdx.x, rb = blockIdx.y, bz = blockIdx.z;
  a += (int64_t)bz * M * d;
  b += (int64_t)bz * N * d;
  if (row_lse) row_lse += (int64_t)bz * M;
  if (col_lse) col_lse += (int64_t)bz * N;
  if (MODE == MM25_SIM_MODULATED) {
    u += (int64_t)bz * M * d2;
    v += (int64_t)bz * N * d2;
    if (diag_add) diag_add += (int64_t)bz * (diag_by_col ? N : M);
  }
  const float s = load_scale(scale, scale_kind, clamp_max);
  const float go = gout ? __ldg(gout + bz) : 1.f;
  const int64_t row0 = (int64_t)rb * BM;
  const int dc0 = dchunk * DC;
  const int dcw = min(DC, d - dc0);  // width of this chunk
  const int njb = (int)ceil((double)N / BN);

  for (int t = threadIdx.x; t < 64 * DC; t += NT) dAs[t] = 0.f;
  float ds_acc = 0.f;

  float rl[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t row = row0 + ty * 4 + i;
    rl[i] = (row_lse && row < M) ? row_lse[row] : 0.f;
  }

  for (int jb = 0; jb < njb; ++jb) {
    const int64_t col0 = (int64_t)jb * BN;
    float acc[4][4] = {};
    tile_dot<T>(acc, As, Bs, a, b, row0, M, col0, N, d);
    float chain[4][4];
    if (MODE == MM25_SIM_MODULATED) {
      float acc2[4][4] = {};
      tile_dot<float>(acc2, As, Bs, u, v, row0, M, col0, N, d2);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) chain[i][j] = 1.f - acc2[i][j];
    }
    float cl[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t col = col0 + tx * 4 + j;
      cl[j] = (col_lse && col < N) ? col_lse[col] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t row = row0 + ty * 4 + i;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t col = col0 + tx * 4 + j;
        const bool valid = row < M && col < N;
        const bool isdiag = (col == row + diag_offset);
        float base = acc[i][j];
        float ch = 1.f;
        if (MODE == MM25_SIM_MODULATED) {
          ch = chain[i][j];
          base *= ch;
          if (isdiag && valid && diag_add) base += diag_add[diag_by_col ? col : row];
        }
        float zz = s * base;
        if (MODE == MM25_SIM_MASKDIAG && isdiag) zz = -10000.f;
        float g = 0.f;
        if (valid) {
          if (w_row != 0.f) g += w_row * expf(zz - rl[i]);
          if (w_col != 0.f) g += w_col * expf(zz - cl[j]);
          // d(loss)/d(scale) uses dz/ds = base, except the masked constant
          float glse = g;
          if (MODE == MM25_SIM_MASKDIAG && isdiag) { glse = 0.f; base = acc[i][j]; }
          g = glse - (isdiag ? w_diag : 0.f) - w_all;
          ds_acc = fmaf(g, base, ds_acc);
          if (MODE == MM25_SIM_MODULATED && isdiag && ddiag && dchunk == 0)
            ddiag[(int64_t)bz * (diag_by_col ? N : M) + (diag_by_col ? col : row)] = go * s * g;
        }
        Gt[tx * 4 + j][ty * 4 + i] = go * s * ch * g;
      }
    }
    // dA chunk += G . b_j[:, dc0 : dc0 + dcw]
    for (int c0 = 0; c0 < dcw; c0 += 64) {
      __syncthreads();  // Gt complete / previous Bc consumed
      for (int t = threadIdx.x; t < 64 * 64; t += NT) {
        const int jj = t >> 6, cc = t & 63;
        const int64_t col = col0 + jj;
        const int dcol = dc0 + c0 + cc;
        Bc[jj][cc] = (col < N && dcol < d) ? to_f32<T>(b[col * d + dcol]) : 0.f;
      }
      __syncthreads();
      float o[4][4] = {};
#pragma unroll 8
      for (int kk = 0; kk < 64; ++kk) {
        float4 gv = *reinterpret_cast<const float4*>(&Gt[kk][ty * 4]);
        float4 bv = *reinterpret_cast<const float4*>(&Bc[kk][tx * 4]);
        const float gr[4] = {gv.x, gv.y, gv.z, gv.w};
        const float br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) o[i][j] = fmaf(gr[i], br[j], o[i][j]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float4* p = reinterpret_cast<float4*>(&dAs[(ty * 4 + i) * DC + c0 + tx * 4]);
        float4 cur = *p;
        cur.x += o[i][0]; cur.y += o[i][1]; cur.z += o[i][2]; cur.w += o[i][3];
        *p = cur;
      }
    }
    __syncthreads();
  }
  // write the chunk
  for (int t = threadIdx.x; t < 64 * DC; t += NT) {
    const int r = t / DC, c = t % DC;
    const int64_t row = row0 + r;
    if (row < M && c < dcw) dA[((int64_t)bz * M + row) * d + dc0 + c] = dAs[t];
  }
  if (dscale_part && dchunk == 0) {
    red[threadIdx.x] = ds_acc * go;
    __syncthreads();
    for (int st = NT / 2; st > 0; st >>= 1) {
      if (threadIdx.x < st) red[threadIdx.x] += red[threadIdx.x + st];
      __syncthreads();
    }
    if (threadIdx.x == 0) dscale_part[(int64_t)bz * gridDim.y + rb] = red[0];
  }
}

constexpr size_t bwd_smem_bytes() { return sizeof(float) * (2 * BK * BM + 2 * 64 * 64 + 64 * DC); }

// ---------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------
static void fwd_geometry(int64_t B, int64_t M, int64_t N, int* nrb, int* nsplit, int* jbps) {
  *nrb = (int)ceil_div(M, BM);
  const int njb = (int)ceil_div(N, BN);
  // aim for >= 2 waves of 148 SMs x 2 resident CTAs without splitting below 4 column blocks
  int want = (int)ceil_div(592, (int64_t)(*nrb) * B);
  int ns = want < 1 ? 1 : want;
  if (ns > njb / 4) ns = njb / 4;
  if (ns < 1) ns = 1;
  *jbps = (int)ceil_div(njb, ns);
  *nsplit = (int)ceil_div(njb, *jbps);
}

size_t fwd_workspace_bytes(int64_t B, int64_t M, int64_t N) {
  int nrb, nsplit, jbps;
  fwd_geometry(B, M, N, &nrb, &nsplit, &jbps);
  return sizeof(float) * 2 * (size_t)B * ((size_t)nsplit * M + (size_t)nrb * N);
}

template <typename T>
static int launch_fwd_t(const T* a, const T* b, const float* u, const float* v,
                        const float* diag_add, int64_t B, int64_t M, int64_t N, int64_t d,
                        int64_t d2, const float* scale, int scale_kind, float clamp_max, int mode,
                        int64_t diag_offset, float* row_lse, float* col_max, float* col_sum,
                        float* diag, void* ws, size_t ws_bytes, cudaStream_t st) {
  int nrb, nsplit, jbps;
  fwd_geometry(B, M, N, &nrb, &nsplit, &jbps);
  if (ws_bytes < fwd_workspace_bytes(B, M, N)) return MM25_ERR_WORKSPACE;
  float* row_pm = (float*)ws;
  float* row_pl = row_pm + (size_t)B * nsplit * M;
  float* col_pm = row_pl + (size_t)B * nsplit * M;
  float* col_pl = col_pm + (size_t)B * nrb * N;
  if (!col_max) col_pm = col_pl = nullptr;
  dim3 grid(nsplit, nrb, (unsigned)B);
#define LAUNCH(MODE)                                                                          \
  sim_fwd_kernel<T, MODE><<<grid, NT, 0, st>>>(a, b, u, v, diag_add, M, N, (int)d, (int)d2,   \
                                               scale, scale_kind, clamp_max, diag_offset,     \
                                               row_pm, row_pl, col_pm, col_pl, diag, nsplit,  \
                                               jbps)
  { void* tt_ = timing_begin(MM25_T_SIMT_FWD, st);
  if (mode == MM25_SIM_PLAIN) LAUNCH(MM25_SIM_PLAIN);
  else if (mode == MM25_SIM_MASKDIAG) LAUNCH(MM25_SIM_MASKDIAG);
  else LAUNCH(MM25_SIM_MODULATED);
  timing_end(tt_, st); }
#undef LAUNCH
  MM25_LAUNCH_CHECK();
  const int64_t mx = M > N ? M : N;
  stats_merge_kernel<<<dim3((unsigned)ceil_div(mx, 256), (unsigned)B), 256, 0, st>>>(
      row_pm, row_pl, nsplit, M, row_lse, col_pm, col_pl, nrb, N, col_max, col_sum);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

int launch_fwd(const void* a, const void* b, int dtype, const float* u, const float* v,
               const float* diag_add, int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2,
               const float* scale, int scale_kind, float clamp_max, int mode,
               int64_t diag_offset, float* row_lse, float* col_max, float* col_sum, float* diag,
               void* ws, size_t ws_bytes, cudaStream_t st) {
  if (dtype == MM25_F32)
    return launch_fwd_t<float>((const float*)a, (const float*)b, u, v, diag_add, B, M, N, d, d2,
                               scale, scale_kind, clamp_max, mode, diag_offset, row_lse, col_max,
                               col_sum, diag, ws, ws_bytes, st);
  if (dtype == MM25_BF16)
    return launch_fwd_t<__nv_bfloat16>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, u, v,
                                       diag_add, B, M, N, d, d2, scale, scale_kind, clamp_max,
                                       mode, diag_offset, row_lse, col_max, col_sum, diag, ws,
                                       ws_bytes, st);
  return MM25_ERR_DTYPE;
}

int64_t bwd_nparts(int64_t B, int64_t M) { return B * ceil_div(M, BM); }

// one row-stationary pass: rows = `a`, columns = `b`
template <typename T>
static int launch_bwd_pass_t(const T* a, const T* b, const float* u, const float* v,
                             const float* diag_add, int diag_by_col, int64_t B, int64_t M,
                             int64_t N, int64_t d, int64_t d2, const float* scale, int scale_kind,
                             float clamp_max, int mode, int64_t diag_offset, const float* row_lse,
                             const float* col_lse, float w_row, float w_col, float w_diag,
                             float w_all, const float* gout, float* dA, float* dscale_part,
                             float* ddiag, cudaStream_t st) {
  dim3 grid((unsigned)ceil_div(d, DC), (unsigned)ceil_div(M, BM), (unsigned)B);
  const size_t smem = bwd_smem_bytes();
#define LAUNCH(MODE)                                                                           \
  do {                                                                                         \
    MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_bwd_kernel<T, MODE>,                              \
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                         (int)smem));                                          \
    sim_bwd_kernel<T, MODE><<<grid, NT, smem, st>>>(                                           \
        a, b, u, v, diag_add, diag_by_col, M, N, (int)d, (int)d2, scale, scale_kind,           \
        clamp_max, diag_offset, row_lse, col_lse, w_row, w_col, w_diag, w_all, gout, dA,       \
        dscale_part, ddiag);                                                                   \
  } while (0)
  void* tt_ = timing_begin(MM25_T_SIMT_BWD, st);
  if (mode == MM25_SIM_PLAIN) LAUNCH(MM25_SIM_PLAIN);
  else if (mode == MM25_SIM_MASKDIAG) LAUNCH(MM25_SIM_MASKDIAG);
  else LAUNCH(MM25_SIM_MODULATED);
  timing_end(tt_, st);
#undef LAUNCH
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

int launch_bwd_pass(const void* a, const void* b, int dtype, const float* u, const float* v,
                    const float* diag_add, int diag_by_col, int64_t B, int64_t M, int64_t N,
                    int64_t d, int64_t d2, const float* scale, int scale_kind, float clamp_max,
                    int mode, int64_t diag_offset, const float* row_lse, const float* col_lse,
                    float w_row, float w_col, float w_diag, float w_all, const float* gout,
                    float* dA, float* dscale_part, float* ddiag, cudaStream_t st) {
  if (dtype == MM25_F32)
    return launch_bwd_pass_t<float>((const float*)a, (const float*)b, u, v, diag_add, diag_by_col,
                                    B, M, N, d, d2, scale, scale_kind, clamp_max, mode,
                                    diag_offset, row_lse, col_lse, w_row, w_col, w_diag, w_all,
                                    gout, dA, dscale_part, ddiag, st);
  if (dtype == MM25_BF16)
    return launch_bwd_pass_t<__nv_bfloat16>(
        (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, u, v, diag_add, diag_by_col, B, M, N, d,
        d2, scale, scale_kind, clamp_max, mode, diag_offset, row_lse, col_lse, w_row, w_col,
        w_diag, w_all, gout, dA, dscale_part, ddiag, st);
  return MM25_ERR_DTYPE;
}

}  // namespace simt
}  // namespace mm25
