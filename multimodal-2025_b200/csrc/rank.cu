// rank.cu — position of a target entry in the descending order of its row AND of its column of
// an fp32 score slab, without sorting (AlignCLIP/main/train.py:338-356 `get_clip_metrics`:
// `argsort(logits, descending=True)` of the N x N matrix on the CPU, then
// `where(ranking == ground_truth)`; main/retrieval.py:55-102 does the same argsort twice).
//
// The rank of the target t in row i is  #{j : s_ij > s_it} + #{j < t : s_ij == s_it}  (the
// position a stable descending sort gives it), so one pass over the scores replaces the
// O(N^2 log N) sort: each element is read ONCE and feeds the count of its row and of its column.
// Integer work on order-preserving keys: bit-exact, independent of launch geometry.
//
// HBM-bound: algorithmic bytes = Q * N * 4 (one read of the slab); outputs are Q + N ints.
// A CTA owns RK_ROWS rows x (RK_THREADS * 4) columns; a thread owns 4 adjacent columns (one
// 16-byte load per row), keeps their 4 column counts in registers across the rows of the
// block, and the row counts are reduced with `redux.sync` + one shared atomic per warp.
#include "common.cuh"

namespace mm25 {

constexpr int RK_THREADS = 256;
constexpr int RK_ROWS = 32;
constexpr int RK_COLS = RK_THREADS * 4;

__device__ __forceinline__ uint32_t rk_key(float f) {
  if (f != f) return 0xFFFFFFFFu;  // NaN ranks above +inf (torch sort order)
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// before(key, pos | tkey, tpos): does the element sort ahead of the target?
__device__ __forceinline__ int rk_before(uint32_t key, int64_t pos, uint32_t tkey, int64_t tpos) {
  return (key > tkey) || (key == tkey && pos < tpos);
}

template <bool VEC>
__global__ void __launch_bounds__(RK_THREADS)
rank_both_kernel(const float* __restrict__ scores, int64_t Q, int64_t N, int64_t ld, int64_t row0,
                 const float* __restrict__ row_tval, const int64_t* __restrict__ row_tcol,
                 const float* __restrict__ col_tval, const int64_t* __restrict__ col_trow,
                 int* __restrict__ row_rank, int* __restrict__ col_rank) {
  __shared__ int s_row[RK_ROWS];
  __shared__ uint32_t s_rkey[RK_ROWS];
  __shared__ long long s_rcol[RK_ROWS];
  const int tid = threadIdx.x;
  const int64_t r_base = (int64_t)blockIdx.y * RK_ROWS;
  const int64_t c_base = (int64_t)blockIdx.x * RK_COLS + (int64_t)tid * 4;
  const int rows = (int)((Q - r_base) < (int64_t)RK_ROWS ? (Q - r_base) : (int64_t)RK_ROWS);
  const bool do_rows = row_rank != nullptr;
  const bool do_cols = col_rank != nullptr;

  if (tid < RK_ROWS) {
    s_row[tid] = 0;
    if (do_rows && tid < rows) {
      s_rkey[tid] = rk_key(row_tval[r_base + tid]);
      s_rcol[tid] = (long long)row_tcol[r_base + tid];
    }
  }
  // column targets of this thread's 4 columns
  uint32_t ckey[4];
  long long crow[4];
  int ccnt[4] = {0, 0, 0, 0};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int64_t c = c_base + q;
    const bool ok = do_cols && c < N;
    ckey[q] = ok ? rk_key(col_tval[c]) : 0xFFFFFFFFu;
    crow[q] = ok ? (long long)col_trow[c] : -1;  // (max key, row -1): nothing sorts ahead
  }
  __syncthreads();

#pragma unroll 4
  for (int r = 0; r < rows; ++r) {
    const int64_t gr = row0 + r_base + r;  // global row index (tie order of the column ranks)
    const float* p = scores + (r_base + r) * ld + c_base;
    float v[4];
    if (VEC) {
      if (c_base + 3 < N) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) v[q] = (c_base + q < N) ? __ldg(p + q) : 0.f;
      }
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = (c_base + q < N) ? __ldg(p + q) : 0.f;
    }
    int rc = 0;
    const uint32_t tkey = do_rows ? s_rkey[r] : 0u;
    const long long tcol = do_rows ? s_rcol[r] : 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const bool in = c_base + q < N;
      const uint32_t key = rk_key(v[q]);
      if (do_rows) rc += (in && rk_before(key, c_base + q, tkey, tcol)) ? 1 : 0;
      ccnt[q] += (in && rk_before(key, gr, ckey[q], crow[q])) ? 1 : 0;
    }
    if (do_rows) {
      rc = __reduce_add_sync(0xffffffffu, rc);
      if ((tid & 31) == 0 && rc) atomicAdd(&s_row[r], rc);
    }
  }
  __syncthreads();
  if (do_rows && tid < rows && s_row[tid]) atomicAdd(&row_rank[r_base + tid], s_row[tid]);
  if (do_cols) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (c_base + q < N && ccnt[q]) atomicAdd(&col_rank[c_base + q], ccnt[q]);
  }
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
    mm25::rank_both_kernel<true><<<grid, mm25::RK_THREADS, 0, (cudaStream_t)stream>>>(
        scores, Q, N, ld, row0, row_tval, row_tcol, col_tval, col_trow, row_rank, col_rank);
  else
    mm25::rank_both_kernel<false><<<grid, mm25::RK_THREADS, 0, (cudaStream_t)stream>>>(
        scores, Q, N, ld, row0, row_tval, row_tcol, col_tval, col_trow, row_rank, col_rank);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
