// rank.cu — position of a target entry in the descending order of its row AND of its column of
// an fp32 score slab, without sorting (AlignCLIP/main/train.py:338-356 `get_clip_metrics`:
// `argsort(logits, descending=True)` of the N x N matrix on the CPU, then
// `where(ranking == ground_truth)`; main/retrieval.py:55-102 does the same argsort twice).
//
// The rank of the target t in row i is  #{j : s_ij > s_it} + #{j < t : s_ij == s_it}  (the
// position a stable descending sort gives it), so one pass over the scores replaces the
// O(N^2 log N) sort: each element is read ONCE and feeds the count of its row and of its column.
// Entries and targets are folded into 64-bit composites (rank_keys.h: order-preserving key in
// the high word, inverted position in the low word), so "sorts ahead" is one unsigned 64-bit
// compare per direction.  Integer work: bit-exact, independent of launch geometry.
//
// HBM-bound: algorithmic bytes = Q * N * 4 (one read of the slab); outputs are Q + N ints.
// A CTA owns RK_ROWS rows x (RK_THREADS * 4) columns; a thread owns 4 adjacent columns (one
// 16-byte load per row), keeps their 4 column counts in registers across the rows of the
// block, and the row counts are reduced with `redux.sync` + one shared atomic per warp.
#include "common.cuh"
#include "rank_keys.h"

namespace mm25 {

constexpr int RK_THREADS = 256;
constexpr int RK_ROWS = 32;
constexpr int RK_COLS = RK_THREADS * 4;

// One row of the thread's 4 columns.  NV = how many of them are inside the matrix (4 on the
// fast path: no per-element bounds test).
template <int NV, bool DO_ROWS, bool DO_COLS>
__device__ __forceinline__ void rank_row4(const float (&v)[4], uint32_t cpos0, uint32_t rpos,
                                          uint64_t row_target, const uint64_t (&col_target)[4],
                                          int& row_cnt, int (&col_cnt)[4]) {
#pragma unroll
  for (int q = 0; q < NV; ++q) {
    const uint32_t key = mm25_rank_key_bits(__float_as_uint(v[q]));
    if (DO_ROWS) row_cnt += mm25_rank_composite(key, cpos0 + q) > row_target ? 1 : 0;
    if (DO_COLS) col_cnt[q] += mm25_rank_composite(key, rpos) > col_target[q] ? 1 : 0;
  }
}

template <bool VEC, bool DO_ROWS, bool DO_COLS>
__global__ void __launch_bounds__(RK_THREADS)
rank_both_kernel(const float* __restrict__ scores, int64_t Q, int64_t N, int64_t ld, int64_t row0,
                 const float* __restrict__ row_tval, const int64_t* __restrict__ row_tcol,
                 const float* __restrict__ col_tval, const int64_t* __restrict__ col_trow,
                 int* __restrict__ row_rank, int* __restrict__ col_rank) {
  __shared__ int s_row[RK_ROWS];
  __shared__ unsigned long long s_rt[RK_ROWS];   // row targets (composites)
  const int tid = threadIdx.x;
  const int64_t r_base = (int64_t)blockIdx.y * RK_ROWS;
  const int64_t c_base = (int64_t)blockIdx.x * RK_COLS + (int64_t)tid * 4;
  const int rows = (int)((Q - r_base) < (int64_t)RK_ROWS ? (Q - r_base) : (int64_t)RK_ROWS);
  // columns of this thread inside the matrix: 4 (fast path), 1..3 (right edge) or 0
  const int nv = c_base + 4 <= N ? 4 : (c_base < N ? (int)(N - c_base) : 0);

  if (tid < RK_ROWS) {
    s_row[tid] = 0;
    if (DO_ROWS && tid < rows)
      s_rt[tid] = mm25_rank_composite(mm25_rank_key_bits(__float_as_uint(row_tval[r_base + tid])),
                                      (uint32_t)row_tcol[r_base + tid]);
  }
  uint64_t ct[4];
  int ccnt[4] = {0, 0, 0, 0};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    ct[q] = mm25_rank_composite_never();
    if (DO_COLS && q < nv)
      ct[q] = mm25_rank_composite(mm25_rank_key_bits(__float_as_uint(col_tval[c_base + q])),
                                  (uint32_t)col_trow[c_base + q]);
  }
  __syncthreads();

  const uint32_t cpos0 = (uint32_t)c_base;
#pragma unroll 4
  for (int r = 0; r < rows; ++r) {
    const uint32_t rpos = (uint32_t)(row0 + r_base + r);  // global row (column-rank tie order)
    const float* p = scores + (r_base + r) * ld + c_base;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (VEC && nv == 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(p));
      v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (q < nv) v[q] = __ldg(p + q);
    }
    const uint64_t rt = DO_ROWS ? (uint64_t)s_rt[r] : 0ull;
    int rc = 0;
    if (nv == 4) rank_row4<4, DO_ROWS, DO_COLS>(v, cpos0, rpos, rt, ct, rc, ccnt);
    else if (nv == 3) rank_row4<3, DO_ROWS, DO_COLS>(v, cpos0, rpos, rt, ct, rc, ccnt);
    else if (nv == 2) rank_row4<2, DO_ROWS, DO_COLS>(v, cpos0, rpos, rt, ct, rc, ccnt);
    else if (nv == 1) rank_row4<1, DO_ROWS, DO_COLS>(v, cpos0, rpos, rt, ct, rc, ccnt);
    if (DO_ROWS) {
      rc = __reduce_add_sync(0xffffffffu, rc);
      if ((tid & 31) == 0 && rc) atomicAdd(&s_row[r], rc);
    }
  }
  __syncthreads();
  if (DO_ROWS && tid < rows && s_row[tid]) atomicAdd(&row_rank[r_base + tid], s_row[tid]);
  if (DO_COLS) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (q < nv && ccnt[q]) atomicAdd(&col_rank[c_base + q], ccnt[q]);
  }
}

template <bool VEC>
static void launch_rank(dim3 grid, cudaStream_t st, const float* scores, int64_t Q, int64_t N,
                        int64_t ld, int64_t row0, const float* row_tval, const int64_t* row_tcol,
                        const float* col_tval, const int64_t* col_trow, int* row_rank,
                        int* col_rank) {
#define MM25_RK(DR, DC)                                                                        \
  rank_both_kernel<VEC, DR, DC><<<grid, RK_THREADS, 0, st>>>(scores, Q, N, ld, row0, row_tval, \
                                                             row_tcol, col_tval, col_trow,    \
                                                             row_rank, col_rank)
  if (row_rank && col_rank) MM25_RK(true, true);
  else if (row_rank) MM25_RK(true, false);
  else MM25_RK(false, true);
#undef MM25_RK
}

}  // namespace mm25

extern "C" int mm25_rank_both(const float* scores, int64_t Q, int64_t N, int64_t ld, int64_t row0,
                              const float* row_tval, const int64_t* row_tcol,
                              const float* col_tval, const int64_t* col_trow, int* row_rank,
                              int* col_rank, void* stream) {
  if (Q <= 0 || N <= 0 || ld < N || row0 < 0 || N >= (1ll << 31) || row0 + Q >= (1ll << 31))
    return MM25_ERR_SHAPE;
  if (!scores || (!row_rank && !col_rank)) return MM25_ERR_ARG;
  if (row_rank && (!row_tval || !row_tcol)) return MM25_ERR_ARG;
  if (col_rank && (!col_tval || !col_trow)) return MM25_ERR_ARG;
  const int64_t gx = mm25::ceil_div(N, mm25::RK_COLS), gy = mm25::ceil_div(Q, mm25::RK_ROWS);
  if (gy > 65535) return MM25_ERR_SHAPE;  // callers slab the rows (2M rows per call)
  const dim3 grid((unsigned)gx, (unsigned)gy);
  const bool vec = (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(scores) & 15u) == 0);
  if (vec)
    mm25::launch_rank<true>(grid, (cudaStream_t)stream, scores, Q, N, ld, row0, row_tval,
                            row_tcol, col_tval, col_trow, row_rank, col_rank);
  else
    mm25::launch_rank<false>(grid, (cudaStream_t)stream, scores, Q, N, ld, row0, row_tval,
                             row_tcol, col_tval, col_trow, row_rank, col_rank);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
