// topk.cu — exact top-k indices per row of an fp32 score matrix (metrics/topk.py:43-48:
// `similarities.topk(k).indices`).  One CTA per row:
//   1. radix select (4 passes of 8 bits over order-preserving uint32 keys, shared-memory
//      histograms) finds the k-th largest key T and how many keys are strictly larger;
//   2. every element > T and the lowest-index elements == T (ordered block compaction) are
//      gathered: exactly k candidates;
//   3. a bitonic sort on (key desc, index asc) emits them in torch.topk's sorted order.
// Ties break lowest-index-first (torch leaves tie order unspecified; parity is defined on
// tie-free rows).  NaN ranks above +inf as in torch.  Integer work: bit-exact.
#include "common.cuh"

namespace mm25 {

constexpr int TK_THREADS = 1024;
constexpr int TK_MAXK = 2048;

__device__ __forceinline__ uint32_t order_key(float f) {
  if (f != f) return 0xFFFFFFFFu;  // NaN: largest
  uint32_t u = __float_as_uint(f);
  if ((u << 1) == 0u) u = 0u;  // -0.0 ties with +0.0 (torch.topk compares values)
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
  return __uint_as_float(u);
}

__global__ void __launch_bounds__(TK_THREADS)
topk_rows_kernel(const float* __restrict__ scores, int64_t N, int64_t ld, int k,
                 int64_t* __restrict__ out_idx, float* __restrict__ out_val) {
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long cand[TK_MAXK];
  __shared__ unsigned int warp_cnt[TK_THREADS / 32];
  __shared__ unsigned int s_prefix, s_remaining, s_gt_count, s_eq_base;
  const float* row = scores + (int64_t)blockIdx.x * ld;
  const int tid = threadIdx.x;

  // ---- 1. radix select of the k-th largest key ------------------------------------------
  if (tid == 0) { s_prefix = 0; s_remaining = (unsigned)k; }
  __syncthreads();
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    const uint32_t mask_hi = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    for (int64_t i = tid; i < N; i += TK_THREADS) {
      const uint32_t key = order_key(row[i]);
      if ((key & mask_hi) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      unsigned rem = s_remaining;
      int b = 255;
      for (; b > 0; --b) {
        if (hist[b] >= rem) break;
        rem -= hist[b];
      }
      s_prefix = prefix | ((uint32_t)b << shift);
      s_remaining = rem;  // rank of the target inside bucket b (1-based, from the top)
    }
    __syncthreads();
  }
  const uint32_t T = s_prefix;
  const unsigned need_eq = s_remaining;  // how many keys == T belong to the top-k
  if (tid == 0) { s_gt_count = 0; s_eq_base = 0; }
  __syncthreads();

  // ---- 2. gather: all keys > T (any order), then the first need_eq keys == T by index -----
  const unsigned n_gt = (unsigned)k - need_eq;
  const int64_t nchunks = (N + TK_THREADS - 1) / TK_THREADS;
  for (int64_t c = 0; c < nchunks; ++c) {
    const int64_t i = c * TK_THREADS + tid;
    uint32_t key = 0;
    bool gt = false, eq = false;
    if (i < N) {
      key = order_key(row[i]);
      gt = key > T;
      eq = key == T;
    }
    if (gt) {
      const unsigned p = atomicAdd(&s_gt_count, 1u);
      cand[p] = ((unsigned long long)key << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)i);
    }
    // ordered compaction of the equal keys
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0) warp_cnt[warp] = __popc(bal);
    __syncthreads();
    if (eq) {
      unsigned before = s_eq_base;
      for (int w = 0; w < warp; ++w) before += warp_cnt[w];
      before += __popc(bal & ((1u << lane) - 1u));
      if (before < need_eq)
        cand[n_gt + before] =
            ((unsigned long long)key << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)i);
    }
    __syncthreads();
    if (tid == 0) {
      unsigned tot = 0;
      for (int w = 0; w < TK_THREADS / 32; ++w) tot += warp_cnt[w];
      s_eq_base += tot;
    }
    __syncthreads();
  }

  // ---- 3. bitonic sort, descending on the composite (key, ~index) -------------------------
  int P = 1;
  while (P < k) P <<= 1;
  for (int i = k + tid; i < P; i += TK_THREADS) cand[i] = 0ull;  // pads sort last
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = tid; t < P / 2; t += TK_THREADS) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const unsigned long long a = cand[lo], b = cand[hi];
        if ((a < b) == desc) { cand[lo] = b; cand[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (int j = tid; j < k; j += TK_THREADS) {
    const unsigned long long c = cand[j];
    const uint32_t idx = 0xFFFFFFFFu - (uint32_t)(c & 0xFFFFFFFFull);
    out_idx[(int64_t)blockIdx.x * k + j] = (int64_t)idx;
    if (out_val) out_val[(int64_t)blockIdx.x * k + j] = row[idx];
  }
}

}  // namespace mm25

extern "C" int mm25_topk_rows(const float* scores, int64_t Q, int64_t N, int64_t ld, int64_t k,
                              int64_t* out_idx, float* out_val, void* stream) {
  if (Q <= 0 || N <= 0 || k <= 0 || k > N || k > mm25::TK_MAXK || ld < N ||
      N >= (1ll << 32))
    return MM25_ERR_SHAPE;
  if (!scores || !out_idx) return MM25_ERR_ARG;
  mm25::topk_rows_kernel<<<(unsigned)Q, mm25::TK_THREADS, 0, (cudaStream_t)stream>>>(
      scores, N, ld, (int)k, out_idx, out_val);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
