// hopfield.cu — backward of the CLOOB Hopfield retrieval (reference losses/cloobLoss.py:43-47),
//   U = rownorm(O),  O_q = sum_k p_qk K_k,  p_qk = softmax_k(beta q.k)      (attention, K == V)
// The forward re-uses the fused similarity kernels (row LSE, then the softmax-weighted sum is the
// row pass of mm25_sim_bwd with w_row = 1).  The backward needs two contractions per tile
// (S = Q K^T and dP = dO K^T), so it has its own fp32 SIMT kernel:
//   dS_qk = p_qk (dP_qk - D_q),  D_q = dO_q . O_q
//   dQ_q  = beta sum_k dS_qk K_k
//   dK_k  = beta sum_q dS_qk Q_q + sum_q p_qk dO_q
// One row-stationary pass computes  out[rows] = sum_cols G1 . bm1[cols] (+ G2 . bm2[cols]); it is
// run once with rows = queries (dQ) and once with rows = stored patterns (dK).
#include "sim_simt.cuh"

namespace mm25 {
namespace simt {

constexpr int HDC = 512;  // output columns accumulated per CTA

// out chunk += Gt^T-tile . Bmat tile, Bmat rows = columns of the logit tile
__device__ __forceinline__ void accum_chunk(float* dAs, float (*Gt)[BM], float (*Bc)[64],
                                            const float* __restrict__ bm, int64_t col0,
                                            int64_t N, int d, int dc0, int dcw) {
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int c0 = 0; c0 < dcw; c0 += 64) {
    __syncthreads();
    for (int t = threadIdx.x; t < 64 * 64; t += NT) {
      const int jj = t >> 6, cc = t & 63;
      const int64_t col = col0 + jj;
      const int dcol = dc0 + c0 + cc;
      Bc[jj][cc] = (col < N && dcol < d) ? bm[col * d + dcol] : 0.f;
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
      float4* p = reinterpret_cast<float4*>(&dAs[(ty * 4 + i) * HDC + c0 + tx * 4]);
      float4 cur = *p;
      cur.x += o[i][0]; cur.y += o[i][1]; cur.z += o[i][2]; cur.w += o[i][3];
      *p = cur;
    }
  }
}

// a,b: S operands; a2,b2: dP operands (rows / cols of the tile); lse, D indexed by the QUERY
// index, which is the row (stat_by_col = 0) or the column (stat_by_col = 1) of the tile.
__global__ void __launch_bounds__(NT) hopfield_bwd_kernel(
    const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ a2,
    const float* __restrict__ b2, const float* __restrict__ bm1, const float* __restrict__ bm2,
    int64_t M, int64_t N, int d, float beta, const float* __restrict__ lse,
    const float* __restrict__ D, int stat_by_col, float* __restrict__ out) {
  extern __shared__ __align__(16) float smem[];
  float(*As)[BM] = reinterpret_cast<float(*)[BM]>(smem);
  float(*Bs)[BN] = reinterpret_cast<float(*)[BN]>(smem + BK * BM);
  float(*G1)[BM] = reinterpret_cast<float(*)[BM]>(smem + 2 * BK * BM);
  float(*G2)[BM] = reinterpret_cast<float(*)[BM]>(smem + 2 * BK * BM + 64 * 64);
  float(*Bc)[64] = reinterpret_cast<float(*)[64]>(smem + 2 * BK * BM + 2 * 64 * 64);
  float* dAs = smem + 2 * BK * BM + 3 * 64 * 64;

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t row0 = (int64_t)blockIdx.y * BM;
  const int dc0 = blockIdx.x * HDC;
  const int dcw = min(HDC, d - dc0);
  const int njb = (int)((N + BN - 1) / BN);
  for (int t = threadIdx.x; t < 64 * HDC; t += NT) dAs[t] = 0.f;

  for (int jb = 0; jb < njb; ++jb) {
    const int64_t col0 = (int64_t)jb * BN;
    float s1[4][4] = {}, s2[4][4] = {};
    tile_dot<float>(s1, As, Bs, a, b, row0, M, col0, N, d);
    tile_dot<float>(s2, As, Bs, a2, b2, row0, M, col0, N, d);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t row = row0 + ty * 4 + i;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t col = col0 + tx * 4 + j;
        float g1 = 0.f, g2 = 0.f;
        if (row < M && col < N) {
          const int64_t qi = stat_by_col ? col : row;
          const float p = expf(beta * s1[i][j] - lse[qi]);
          g1 = beta * p * (s2[i][j] - D[qi]);
          g2 = p;
        }
        G1[tx * 4 + j][ty * 4 + i] = g1;
        G2[tx * 4 + j][ty * 4 + i] = g2;
      }
    }
    accum_chunk(dAs, G1, Bc, bm1, col0, N, d, dc0, dcw);
    if (bm2) accum_chunk(dAs, G2, Bc, bm2, col0, N, d, dc0, dcw);
    __syncthreads();
  }
  for (int t = threadIdx.x; t < 64 * HDC; t += NT) {
    const int r = t / HDC, c = t % HDC;
    const int64_t row = row0 + r;
    if (row < M && c < dcw) out[row * d + dc0 + c] = dAs[t];
  }
}

__global__ void __launch_bounds__(256) rowdot_kernel(const float* __restrict__ x,
                                                     const float* __restrict__ y,
                                                     float* __restrict__ out, int64_t rows,
                                                     int d) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  float acc = 0.f;
  for (int k = lane; k < d; k += 32) acc = fmaf(x[row * d + k], y[row * d + k], acc);
  acc = warp_sum(acc);
  if (lane == 0) out[row] = acc;
}

}  // namespace simt
}  // namespace mm25

using namespace mm25;

extern "C" int mm25_rowdot(const float* x, const float* y, float* out, int64_t rows, int64_t d,
                           void* stream) {
  if (!x || !y || !out || rows <= 0 || d <= 0) return MM25_ERR_ARG;
  simt::rowdot_kernel<<<(unsigned)ceil_div(rows, 8), 256, 0, (cudaStream_t)stream>>>(
      x, y, out, rows, (int)d);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_hopfield_bwd_pass(const float* a, const float* b, const float* a2,
                                      const float* b2, const float* bm1, const float* bm2,
                                      int64_t M, int64_t N, int64_t d, float beta,
                                      const float* lse, const float* D, int stat_by_col,
                                      float* out, void* stream) {
  if (!a || !b || !a2 || !b2 || !bm1 || !lse || !D || !out) return MM25_ERR_ARG;
  if (M <= 0 || N <= 0 || d <= 0) return MM25_ERR_SHAPE;
  const size_t smem =
      sizeof(float) * (2 * simt::BK * simt::BM + 3 * 64 * 64 + 64 * simt::HDC);
  MM25_CUDA_CHECK(cudaFuncSetAttribute(simt::hopfield_bwd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)ceil_div(d, simt::HDC), (unsigned)ceil_div(M, simt::BM));
  void* tt_ = timing_begin(MM25_T_SIMT_BWD, (cudaStream_t)stream);
  simt::hopfield_bwd_kernel<<<grid, simt::NT, smem, (cudaStream_t)stream>>>(
      a, b, a2, b2, bm1, bm2, M, N, (int)d, beta, lse, D, stat_by_col, out);
  timing_end(tt_, (cudaStream_t)stream);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
