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

#include <cstdlib>

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
                 int64_t* __restrict__ out_idx, float* __restrict__ out_val,
                 const int* __restrict__ only_if = nullptr) {
  // only_if: per-row flags of the compaction kernel below — rows it finished are skipped
  if (only_if && only_if[blockIdx.x] == 0) return;
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


// ------------------------------------------------------------------------------------------
// Sliced path (round 2): a row is cut into C slices of at most TK_SLICE elements; one CTA per
// (slice, row) loads its slice ONCE (16-byte loads when aligned) into shared memory as keys, so
// the four radix passes, the gather and the sort never go back to HBM, and small-Q calls
// (the reference's shape: one query against all texts, metrics/topk.py:43-48) spread over
// every SM instead of one.  The top-k of the row is contained in the union of the slices'
// top-k (the composite (key, ~index) order is total), so a merge kernel sorts the C k
// candidates.  Same result as topk_rows_kernel, bit for bit.
// ------------------------------------------------------------------------------------------
constexpr int TK_SLICE = 16384;       // 64 KB of keys: two CTAs per SM
constexpr int TK_HCOPIES = 8;
constexpr int TK_MERGE_MAX = 8192;    // candidates the merge kernel sorts in shared memory

struct SliceSmem {
  uint32_t keys[TK_SLICE];
  unsigned long long cand[TK_MAXK];
  unsigned int hist[TK_HCOPIES][256];
  unsigned int warp_cnt[TK_THREADS / 32];
  unsigned int prefix, remaining, gt_count, eq_base;
};

__device__ __forceinline__ void bitonic_desc(unsigned long long* a, int P, int tid) {
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = tid; t < P / 2; t += TK_THREADS) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const unsigned long long x = a[lo], y = a[hi];
        if ((x < y) == desc) { a[lo] = y; a[hi] = x; }
      }
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(TK_THREADS, 2)
topk_slice_kernel(const float* __restrict__ scores, int64_t N, int64_t ld, int k, int C,
                  int64_t slice_len, int64_t* __restrict__ out_idx, float* __restrict__ out_val,
                  unsigned long long* __restrict__ ws) {
  extern __shared__ __align__(16) uint8_t tk_smem[];
  SliceSmem& sm = *reinterpret_cast<SliceSmem*>(tk_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t rowi = blockIdx.y;
  const float* row = scores + rowi * ld;
  const int64_t s0 = (int64_t)blockIdx.x * slice_len;
  const int n = (int)min(slice_len, N - s0);
  const int kk = min(k, n);
  const float* src = row + s0;

  // ---- the one HBM pass: slice -> order-preserving keys in shared memory ----------------
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    const int n4 = n >> 2;
    for (int i = tid; i < n4; i += TK_THREADS) {
      const float4 v = __ldg(s4 + i);
      uint4 kq = make_uint4(order_key(v.x), order_key(v.y), order_key(v.z), order_key(v.w));
      *reinterpret_cast<uint4*>(&sm.keys[4 * i]) = kq;
    }
    for (int i = 4 * n4 + tid; i < n; i += TK_THREADS) sm.keys[i] = order_key(src[i]);
  } else {
    for (int i = tid; i < n; i += TK_THREADS) sm.keys[i] = order_key(src[i]);
  }
  if (tid == 0) { sm.prefix = 0; sm.remaining = (unsigned)kk; }
  __syncthreads();

  // ---- radix select of the kk-th largest key of the slice (shared memory only) -----------
  unsigned int* myhist = sm.hist[warp & (TK_HCOPIES - 1)];
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    const uint32_t mask_hi = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
    for (int i = tid; i < TK_HCOPIES * 256; i += TK_THREADS) (&sm.hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t prefix = sm.prefix;
    for (int i0 = 0; i0 < n; i0 += TK_THREADS) {
      const int i = i0 + tid;
      const bool in = i < n && ((sm.keys[i] & mask_hi) == prefix);
      const unsigned act = __ballot_sync(0xffffffffu, in);
      if (in) {
        const uint32_t bin = (sm.keys[i] >> shift) & 255u;
        const unsigned peers = __match_any_sync(act, bin);   // one atomic per distinct bin
        if (lane == __ffs(peers) - 1) atomicAdd(&myhist[bin], (unsigned)__popc(peers));
      }
    }
    __syncthreads();
    // bin b = the highest bin whose count, added to everything above it, reaches `remaining`:
    // a parallel suffix scan over the 256 bins (thread t holds bin 255 - t) instead of a
    // 256-step walk by one thread
    unsigned cnt = 0, incl = 0;
    const unsigned rem = sm.remaining;   // read by everyone before the barrier below; written after it
    if (tid < 256) {
#pragma unroll
      for (int c = 0; c < TK_HCOPIES; ++c) cnt += sm.hist[c][255 - tid];
      incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
      }
      if (lane == 31) sm.warp_cnt[warp] = incl;
    }
    __syncthreads();
    if (tid < 256) {
      unsigned base = 0;
      for (int w8 = 0; w8 < warp; ++w8) base += sm.warp_cnt[w8];
      incl += base;
      // bin 0 (tid 255) takes whatever is left, as the serial walk did
      if ((incl >= rem && incl - cnt < rem) || (tid == 255 && incl < rem)) {
        sm.prefix = prefix | ((uint32_t)(255 - tid) << shift);
        sm.remaining = rem - (incl - cnt);
      }
    }
    __syncthreads();
  }
  const uint32_t T = sm.prefix;
  const unsigned need_eq = sm.remaining;
  if (tid == 0) { sm.gt_count = 0; sm.eq_base = 0; }
  __syncthreads();

  // ---- gather: keys > T in any order, then the first need_eq keys == T by index ----------
  const unsigned n_gt = (unsigned)kk - need_eq;
  for (int i0 = 0; i0 < n; i0 += TK_THREADS) {
    const int i = i0 + tid;
    uint32_t key = 0;
    bool gt = false, eq = false;
    if (i < n) { key = sm.keys[i]; gt = key > T; eq = key == T; }
    const unsigned long long comp =
        ((unsigned long long)key << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)(s0 + i));
    if (gt) sm.cand[atomicAdd(&sm.gt_count, 1u)] = comp;
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    if (sm.eq_base < need_eq) {            // uniform: read before the barrier below
      if (lane == 0) sm.warp_cnt[warp] = __popc(bal);
      __syncthreads();
      if (eq) {
        unsigned before = sm.eq_base;
        for (int w = 0; w < warp; ++w) before += sm.warp_cnt[w];
        before += __popc(bal & ((1u << lane) - 1u));
        if (before < need_eq) sm.cand[n_gt + before] = comp;
      }
      __syncthreads();
      if (tid == 0) {
        unsigned tot = 0;
        for (int w = 0; w < TK_THREADS / 32; ++w) tot += sm.warp_cnt[w];
        sm.eq_base += tot;
      }
      __syncthreads();
    }
  }
  __syncthreads();

  // ---- sort the slice's candidates, descending on (key, ~index) --------------------------
  int P = 1;
  while (P < kk) P <<= 1;
  for (int i = kk + tid; i < P; i += TK_THREADS) sm.cand[i] = 0ull;
  __syncthreads();
  bitonic_desc(sm.cand, P, tid);
  if (C == 1) {
    for (int j = tid; j < k; j += TK_THREADS) {
      const uint32_t idx = 0xFFFFFFFFu - (uint32_t)(sm.cand[j] & 0xFFFFFFFFull);
      out_idx[rowi * k + j] = (int64_t)idx;
      if (out_val) out_val[rowi * k + j] = row[idx];
    }
  } else {
    unsigned long long* dst = ws + (rowi * C + blockIdx.x) * (int64_t)k;
    for (int j = tid; j < k; j += TK_THREADS) dst[j] = j < kk ? sm.cand[j] : 0ull;  // 0 = pad
  }
}

__global__ void __launch_bounds__(TK_THREADS)
topk_merge_kernel(const unsigned long long* __restrict__ ws, int total, int k,
                  const float* __restrict__ scores, int64_t ld, int64_t* __restrict__ out_idx,
                  float* __restrict__ out_val) {
  extern __shared__ __align__(16) uint8_t tk_smem[];
  unsigned long long* a = reinterpret_cast<unsigned long long*>(tk_smem);
  const int tid = threadIdx.x;
  const int64_t rowi = blockIdx.x;
  int P = 1;
  while (P < total) P <<= 1;
  for (int i = tid; i < P; i += TK_THREADS) a[i] = i < total ? ws[rowi * total + i] : 0ull;
  __syncthreads();
  bitonic_desc(a, P, tid);
  const float* row = scores + rowi * ld;
  for (int j = tid; j < k; j += TK_THREADS) {
    const uint32_t idx = 0xFFFFFFFFu - (uint32_t)(a[j] & 0xFFFFFFFFull);
    out_idx[rowi * k + j] = (int64_t)idx;
    if (out_val) out_val[rowi * k + j] = row[idx];
  }
}

// ------------------------------------------------------------------------------------------
// Many long rows (Q >= 2 SMs-worth, N > one slice): one CTA per row, two passes over the row
// instead of five.  Pass 1 histograms the top 12 key bits (sign, exponent, 3 mantissa bits:
// 4096 bins); a parallel suffix scan finds the bin b0 holding the k-th largest key.  Pass 2
// (L2-resident for rows up to a few MB) sends every key above b0 straight to the candidates and
// COMPACTS bin b0, in index order, into shared memory; the remaining 20 key bits are resolved
// there (radix passes, gather of the lowest-index equals, bitonic sort) exactly as the slice
// kernel does on its keys.  A bin that does not fit (heavy ties) raises the row's flag in
// `redo` and the five-pass kernel repeats that row.  Same result, bit for bit.
// ------------------------------------------------------------------------------------------
constexpr int TK_BCAP = 8192;   // compacted keys per row held in shared memory

struct BigSmem {
  uint32_t keys[TK_BCAP];
  uint32_t idxs[TK_BCAP];
  unsigned long long cand[TK_MAXK];
  unsigned int hist[4096];
  unsigned int wcnt[2][TK_THREADS / 32];
  unsigned int b0, rem, nb, gt_count, prefix, remaining, eq_base;
};

__global__ void __launch_bounds__(TK_THREADS, 2)
topk_rows_compact_kernel(const float* __restrict__ scores, int64_t N, int64_t ld, int k,
                         int64_t* __restrict__ out_idx, float* __restrict__ out_val,
                         int* __restrict__ redo) {
  extern __shared__ __align__(16) uint8_t tk_smem[];
  BigSmem& sm = *reinterpret_cast<BigSmem*>(tk_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t rowi = blockIdx.x;
  const float* row = scores + rowi * ld;
  const bool vec = (reinterpret_cast<uintptr_t>(row) & 15) == 0;
  const int64_t n4 = vec ? (N >> 2) : 0;

  // ---- pass 1: 4096-bin histogram of the top 12 key bits ---------------------------------
  for (int i = tid; i < 4096; i += TK_THREADS) sm.hist[i] = 0;
  __syncthreads();
  for (int64_t i = tid; i < n4; i += 2 * TK_THREADS) {
    // two 16-byte loads in flight per thread
    const float4 v = __ldg(reinterpret_cast<const float4*>(row) + i);
    const bool two = i + TK_THREADS < n4;
    const float4 u = two ? __ldg(reinterpret_cast<const float4*>(row) + i + TK_THREADS) : v;
    atomicAdd(&sm.hist[order_key(v.x) >> 20], 1u);
    atomicAdd(&sm.hist[order_key(v.y) >> 20], 1u);
    atomicAdd(&sm.hist[order_key(v.z) >> 20], 1u);
    atomicAdd(&sm.hist[order_key(v.w) >> 20], 1u);
    if (two) {
      atomicAdd(&sm.hist[order_key(u.x) >> 20], 1u);
      atomicAdd(&sm.hist[order_key(u.y) >> 20], 1u);
      atomicAdd(&sm.hist[order_key(u.z) >> 20], 1u);
      atomicAdd(&sm.hist[order_key(u.w) >> 20], 1u);
    }
  }
  for (int64_t i = 4 * n4 + tid; i < N; i += TK_THREADS)
    atomicAdd(&sm.hist[order_key(row[i]) >> 20], 1u);
  __syncthreads();
  // suffix scan from the top: thread t owns bins 4095 - 4t .. 4092 - 4t
  {
    unsigned c[4], tot = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) { c[j] = sm.hist[4095 - 4 * tid - j]; tot += c[j]; }
    unsigned incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    if (lane == 31) sm.wcnt[0][warp] = incl;
    __syncthreads();
    unsigned base = 0;
    for (int w8 = 0; w8 < warp; ++w8) base += sm.wcnt[0][w8];
    unsigned run = base + incl - tot;      // keys strictly above this thread's first bin
    const unsigned kk = (unsigned)k;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (run < kk && run + c[j] >= kk) {  // the k-th largest key lies in this bin
        sm.b0 = (unsigned)(4095 - 4 * tid - j);
        sm.rem = kk - run;                 // how many keys of the bin belong to the top-k
        sm.nb = c[j];
      }
      run += c[j];
    }
    if (tid == 0) { sm.gt_count = 0; }
    __syncthreads();
  }
  const unsigned b0 = sm.b0, nb = sm.nb;
  if (nb > (unsigned)TK_BCAP) {            // heavy ties: the five-pass kernel repeats this row
    if (tid == 0) redo[rowi] = 1;
    return;
  }
  if (tid == 0) redo[rowi] = 0;

  // ---- pass 2: keys above b0 -> candidates; bin b0 -> shared memory, in index order --------
  // Warp w owns the contiguous elements [w L, (w + 1) L) and appends its bin-b0 keys to its own
  // region of TK_BCAP / 32 slots with a warp-private counter: no block barrier inside the pass.
  // Slot order (warp, position) is index order.  A region that overflows hands the row back.
  constexpr int WCAP = TK_BCAP / (TK_THREADS / 32);   // 256 slots per warp
  const int64_t L = (N + (TK_THREADS / 32) - 1) / (TK_THREADS / 32);
  const int64_t w0 = (int64_t)warp * L, w1 = min(N, w0 + L);
  unsigned wn = 0;   // this warp's count (same value in every lane)
  for (int64_t i0 = w0; i0 < w1; i0 += 128) {
    // four independent loads in flight per lane (the pass is L2-latency bound otherwise); the
    // four 32-element chunks are then appended in order
    float v[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int64_t i = i0 + 32 * c + lane;
      v[c] = i < w1 ? __ldg(row + i) : 0.f;
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int64_t i = i0 + 32 * c + lane;
      uint32_t key = 0;
      bool in_bin = false;
      if (i < w1) {
        key = order_key(v[c]);
        const uint32_t bin = key >> 20;
        if (bin > b0)
          sm.cand[atomicAdd(&sm.gt_count, 1u)] =
              ((unsigned long long)key << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)i);
        in_bin = bin == b0;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, in_bin);
      if (in_bin) {
        const unsigned pos = wn + __popc(bal & ((1u << lane) - 1u));
        if (pos < (unsigned)WCAP) {
          sm.keys[warp * WCAP + pos] = key;
          sm.idxs[warp * WCAP + pos] = (uint32_t)i;
        }
      }
      wn += __popc(bal);
    }
  }
  if (lane == 0) sm.wcnt[0][warp] = wn;
  __syncthreads();
  bool over = false;
#pragma unroll
  for (int w8 = 0; w8 < TK_THREADS / 32; ++w8) over |= sm.wcnt[0][w8] > (unsigned)WCAP;
  if (over) {              // uniform: every thread read the same counters
    if (tid == 0) redo[rowi] = 1;
    return;
  }
  const unsigned n_above = (unsigned)k - sm.rem;
  if (tid == 0) { sm.prefix = b0 << 20; sm.remaining = sm.rem; }
  __syncthreads();
  // slot s = (warp s / WCAP, position s % WCAP) is live when position < that warp's count
  auto live = [&](int slot) { return (unsigned)(slot & (WCAP - 1)) < sm.wcnt[0][slot / WCAP]; };

  // ---- radix select of the rem-th largest key of the list on the low 20 bits ---------------
  for (int pass = 0; pass < 3; ++pass) {
    const int shift = pass == 0 ? 12 : (pass == 1 ? 4 : 0);
    const int bits = pass == 2 ? 4 : 8;
    const uint32_t mask_hi = 0xFFFFFFFFu << (shift + bits);
    for (int i = tid; i < 256; i += TK_THREADS) sm.hist[i] = 0;
    __syncthreads();
    const uint32_t prefix = sm.prefix;
    for (int slot = tid; slot < TK_BCAP; slot += TK_THREADS) {
      if (!live(slot)) continue;
      const uint32_t key = sm.keys[slot];
      if ((key & mask_hi) == (prefix & mask_hi))
        atomicAdd(&sm.hist[(key >> shift) & ((1u << bits) - 1u)], 1u);
    }
    __syncthreads();
    if (warp == 0) {
      // warp-parallel walk from the top bin: lane l owns m consecutive bins
      const int nbins = 1 << bits, m = nbins >= 32 ? nbins / 32 : 1;
      const unsigned rem = sm.remaining;
      unsigned c[8], tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int b = nbins - 1 - lane * m - j;
        c[j] = (j < m && b >= 0) ? sm.hist[b] : 0u;
        tot += c[j];
      }
      unsigned incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
      }
      unsigned run = incl - tot;
      const unsigned all = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int b = nbins - 1 - lane * m - j;
        if (j < m && b >= 0) {
          // bin 0 takes whatever is left, as the serial walk did
          if ((run < rem && run + c[j] >= rem) || (b == 0 && all < rem)) {
            sm.prefix = prefix | ((uint32_t)b << shift);
            sm.remaining = rem - run;
          }
          run += c[j];
        }
      }
    }
    __syncthreads();
  }
  const uint32_t T = sm.prefix;
  const unsigned need_eq = sm.remaining;
  __syncthreads();
  // ---- gather from the list: keys > T, then the first need_eq keys == T in slot (= index) order
  const unsigned n_gt_list = sm.rem - need_eq;      // keys of the bin strictly above T
  unsigned eqn = 0;
  for (int s0 = 0; s0 < TK_BCAP; s0 += TK_THREADS) {
    const int slot = s0 + tid;
    uint32_t key = 0, idx = 0;
    bool gt = false, eq = false;
    if (live(slot)) { key = sm.keys[slot]; idx = sm.idxs[slot]; gt = key > T; eq = key == T; }
    const unsigned long long comp =
        ((unsigned long long)key << 32) | (unsigned long long)(0xFFFFFFFFu - idx);
    if (gt) sm.cand[atomicAdd(&sm.gt_count, 1u)] = comp;
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    if (lane == 0) sm.wcnt[1][warp] = __popc(bal);   // wcnt[1]: wcnt[0] holds the list counts
    __syncthreads();
    unsigned before = 0, total = 0;
#pragma unroll
    for (int w8 = 0; w8 < TK_THREADS / 32; ++w8) {
      const unsigned c = sm.wcnt[1][w8];
      before += w8 < warp ? c : 0u;
      total += c;
    }
    if (eq) {
      const unsigned pos = eqn + before + __popc(bal & ((1u << lane) - 1u));
      if (pos < need_eq) sm.cand[n_above + n_gt_list + pos] = comp;
    }
    eqn += total;
    __syncthreads();       // wcnt[1] is rewritten by the next iteration
  }
  __syncthreads();

  // ---- sort the k candidates, descending on (key, ~index) ----------------------------------
  int P = 1;
  while (P < k) P <<= 1;
  for (int i = k + tid; i < P; i += TK_THREADS) sm.cand[i] = 0ull;
  __syncthreads();
  bitonic_desc(sm.cand, P, tid);
  for (int j = tid; j < k; j += TK_THREADS) {
    const uint32_t idx = 0xFFFFFFFFu - (uint32_t)(sm.cand[j] & 0xFFFFFFFFull);
    out_idx[rowi * k + j] = (int64_t)idx;
    if (out_val) out_val[rowi * k + j] = row[idx];
  }
}

// slices per row for the sliced path; 0 = not applicable (the one-CTA-per-row kernel runs)
static int topk_slices(int64_t Q, int64_t N, int64_t k) {
  if (Q > 65535) return 0;
  const int64_t c_min = (N + TK_SLICE - 1) / TK_SLICE;
  const int64_t c_fill = (2 * 148 + Q - 1) / Q;          // fill the SMs when Q is small
  int64_t C = c_fill < N / 8192 ? c_fill : N / 8192;     // ... with slices of >= 8192 elements
  if (C < c_min) C = c_min;
  if (C < 1) C = 1;
  if (c_min > 1 && Q >= 2 * 148) return 0;  // many long rows: the SMs are full already and the
                                            // per-slice select + merge costs more than it saves
  if (C > 1 && C * k > TK_MERGE_MAX) C = TK_MERGE_MAX / k;
  if (C < c_min) return 0;                               // merge would not fit: plain kernel
  return (int)C;
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

extern "C" size_t mm25_topk_workspace_bytes(int64_t Q, int64_t N, int64_t k) {
  if (Q <= 0 || N <= 0 || k <= 0 || k > mm25::TK_MAXK) return 0;
  const int C = mm25::topk_slices(Q, N, k);
  if (C == 0) return Q <= 65535ll * 32768 ? sizeof(int) * (size_t)Q : 0;  // per-row redo flags
  return C > 1 ? sizeof(unsigned long long) * (size_t)Q * C * k : 0;
}

extern "C" int mm25_topk_rows_ws(const float* scores, int64_t Q, int64_t N, int64_t ld, int64_t k,
                                 int64_t* out_idx, float* out_val, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  if (Q <= 0 || N <= 0 || k <= 0 || k > N || k > mm25::TK_MAXK || ld < N ||
      N >= (1ll << 32))
    return MM25_ERR_SHAPE;
  if (!scores || !out_idx) return MM25_ERR_ARG;
  const int C = mm25::topk_slices(Q, N, k);
  cudaStream_t st = (cudaStream_t)stream;
  if (C == 0) {
    // many long rows: two-pass compaction kernel, five-pass kernel for the rows it hands back
    static const bool plain_env = getenv("MM25_TOPK_NO_COMPACT") != nullptr;
    if (plain_env || !workspace || workspace_bytes < sizeof(int) * (size_t)Q || Q >= (1ll << 31))
      return mm25_topk_rows(scores, Q, N, ld, k, out_idx, out_val, stream);
    int* redo = (int*)workspace;
    const size_t bsm = sizeof(mm25::BigSmem);
    MM25_CUDA_CHECK(cudaFuncSetAttribute(mm25::topk_rows_compact_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bsm));
    mm25::topk_rows_compact_kernel<<<(unsigned)Q, mm25::TK_THREADS, bsm, st>>>(
        scores, N, ld, (int)k, out_idx, out_val, redo);
    MM25_LAUNCH_CHECK();
    mm25::topk_rows_kernel<<<(unsigned)Q, mm25::TK_THREADS, 0, st>>>(scores, N, ld, (int)k,
                                                                      out_idx, out_val, redo);
    MM25_LAUNCH_CHECK();
    return MM25_OK;
  }
  if (C > 1 && (!workspace || workspace_bytes < mm25_topk_workspace_bytes(Q, N, k)))
    return MM25_ERR_WORKSPACE;
  const int64_t slice_len = ((N + C - 1) / C + 3) & ~int64_t(3);  // keeps 16-byte alignment
  const int Cn = (int)((N + slice_len - 1) / slice_len);
  const size_t smem = sizeof(mm25::SliceSmem);
  MM25_CUDA_CHECK(cudaFuncSetAttribute(mm25::topk_slice_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // the slice count of the launch may be smaller than C after rounding slice_len up; the
  // workspace layout uses the launched count
  mm25::topk_slice_kernel<<<dim3((unsigned)Cn, (unsigned)Q), mm25::TK_THREADS, smem, st>>>(
      scores, N, ld, (int)k, Cn, slice_len, out_idx, out_val, (unsigned long long*)workspace);
  MM25_LAUNCH_CHECK();
  if (Cn > 1) {
    const int total = Cn * (int)k;
    int P = 1;
    while (P < total) P <<= 1;
    const size_t msmem = sizeof(unsigned long long) * (size_t)P;
    MM25_CUDA_CHECK(cudaFuncSetAttribute(mm25::topk_merge_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
    mm25::topk_merge_kernel<<<(unsigned)Q, mm25::TK_THREADS, msmem, st>>>(
        (const unsigned long long*)workspace, total, (int)k, scores, ld, out_idx, out_val);
    MM25_LAUNCH_CHECK();
  }
  return MM25_OK;
}
