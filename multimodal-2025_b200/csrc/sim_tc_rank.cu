// sim_tc_rank.cu — retrieval ranks fused into the similarity GEMM:
//     rank[q] += #{ j != tcol[q] : a_q . b_j sorts ahead of the row's target (tval[q], tcol[q]) }
// for Q query rows against N gallery rows, on the S-tile sweep of the forward statistics kernel
// (sim_tc_sweep.cuh: TMA ring -> tcgen05.mma -> TMEM) with a counting epilogue, so the Q x N
// scores never leave TMEM / registers (reference: AlignCLIP/main/train.py:338-356 materialises
// `logits_per_image` N x N, copies it to the CPU and argsorts it once per direction;
// main/retrieval.py:62-71 the same for recall@k; metrics/topk.py:64-82 one query at a time).
// "Ahead" is the stable descending order of mm25_rank_both: larger score, or equal score and
// smaller column.  The target's own column is excluded by POSITION, not by value: `tval` comes
// from the caller (an fp32 dot product of the same bf16 rows) and may differ from the
// accumulator's value of that entry in the last bits.
// One launch counts the row direction; the column direction of a paired batch is the same
// launch with the operands swapped (the row-only epilogue is ~2 instructions per score, so each
// sweep stays MMA-bound; a ballot-per-column epilogue for both directions at once would cost
// more than the second sweep).
// Ragged shapes need no padding: the tensor maps carry the real row counts, TMA zero-fills the
// out-of-range rows of the last boxes, and the epilogue masks rows >= Q and columns >= N.
#include "sim_tc_sweep.cuh"
#include "sim_tc.h"

namespace mm25 {
namespace tc {

int tc_sm_count();

struct RankSmem {
  uint8_t a[NST][A_BYTES];
  uint8_t b[NST][B_BYTES];
  uint64_t full[NST], empty[NST];
  uint64_t acc_full[NACC], acc_empty[NACC];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(FWD_THREADS, 1)
sim_rank_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                   const __grid_constant__ CUtensorMap tmap_b, int64_t Q, int64_t N, int d,
                   int nrb, int nseg, int tps, const float* __restrict__ tval,
                   const int64_t* __restrict__ tcol, int* __restrict__ rank) {
  extern __shared__ uint8_t smem_raw[];
  RankSmem& sm = *reinterpret_cast<RankSmem*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = (int)((N + BN - 1) / BN);
  const int nitems = nrb * nseg;
  Sweep w;
  w.KC = d / BK; w.ntiles = ntiles; w.nseg = nseg; w.tps = tps;
  w.nrb_per_batch = nrb; w.first_rbg = 0; w.num_rbg = nrb; w.N = N;
  sweep_setup(sm, &tmap_a, &tmap_b);
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) sweep_producer(sm, &tmap_a, &tmap_b, w);
  } else if (warp == 1) {
    if (lane == 0) sweep_mma(sm, tmem, w);
  } else {
    const int q = warp & 3;        // TMEM lane quadrant this warp may read
    const int ew = warp - 2;
    const int half = ew >> 2;      // which half of the tile's columns this warp handles
    uint32_t as = 0, accph = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const int rb = item / nseg, seg = item % nseg;
      const int64_t row = (int64_t)rb * BM + q * 32 + lane;
      const bool live = row < Q;
      const float t = live ? tval[row] : 0.f;
      const int64_t tc_ = live ? tcol[row] : -1;
      int cnt = 0;
      const int t0 = seg * tps, t1 = min(ntiles, t0 + tps);
      for (int tl = t0; tl < t1; ++tl) {
        const int64_t col0 = (int64_t)tl * BN;
        mbar_wait(&sm.acc_full[as], accph);
        tc_fence_after();
        const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + as * BN +
                               half * CH_PER_WARP * 32;
        auto process = [&](uint32_t (&r)[32], int chl) {
          const int64_t cbase = col0 + (half * CH_PER_WARP + chl) * 32;
          if (cbase + 32 <= N && (tc_ >= cbase + 32 || tc_ < cbase)) {
            if (tc_ >= cbase + 32) {        // every column precedes the target: ties count
#pragma unroll
              for (int j = 0; j < 32; ++j) cnt += __uint_as_float(r[j]) >= t ? 1 : 0;
            } else {                        // every column follows the target: ties do not
#pragma unroll
              for (int j = 0; j < 32; ++j) cnt += __uint_as_float(r[j]) > t ? 1 : 0;
            }
          } else {                          // the chunk holds the target or the ragged edge
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int64_t c = cbase + j;
              const float v = __uint_as_float(r[j]);
              const bool ahead = c < tc_ ? v >= t : v > t;
              cnt += (c < N && c != tc_ && ahead) ? 1 : 0;
            }
          }
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
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
        if (++as == NACC) { as = 0; accph ^= 1; }
      }
      if (live && cnt) atomicAdd(&rank[row], cnt);   // integer adds: exact, order-free
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

bool eligible_rank(int64_t Q, int64_t N, int64_t d) {
  return Q > 0 && N > 0 && d % BK == 0 && d >= BK && d <= MAX_KC * BK && Q < (1ll << 31) &&
         N < (1ll << 31);
}

int launch_rank(const void* a, const void* b, int64_t Q, int64_t N, int64_t d, const float* tval,
                const int64_t* tcol, int* rank, cudaStream_t st) {
  if (!eligible_rank(Q, N, d)) return MM25_ERR_SHAPE;
  const int nrb = (int)((Q + BM - 1) / BM);
  const int ntiles = (int)((N + BN - 1) / BN);
  const int sms = tc_sm_count();
  // column segments: enough items to fill the SMs, as few row-count atomics as possible
  int nseg = 1;
  while ((long)nrb * nseg < 2L * sms && nseg < ntiles) ++nseg;
  const int tps = (ntiles + nseg - 1) / nseg;
  nseg = (ntiles + tps - 1) / tps;
  CUtensorMap ta, tb;
  int rc = make_tmap_bf16(&ta, a, (uint64_t)Q, (uint64_t)d, BM, BK);
  if (rc) return rc;
  rc = make_tmap_bf16(&tb, b, (uint64_t)N, (uint64_t)d, BN, BK);
  if (rc) return rc;
  const size_t smem = sizeof(RankSmem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sim_rank_tc_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long items = (long)nrb * nseg;
  const int grid = (int)(items < sms ? items : sms);
  { void* tt_ = timing_begin(MM25_T_SIM_RANK, st);
  sim_rank_tc_kernel<<<grid, FWD_THREADS, smem, st>>>(ta, tb, Q, N, (int)d, nrb, nseg, tps, tval,
                                                      tcol, rank);
  timing_end(tt_, st); }
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

}  // namespace tc
}  // namespace mm25

extern "C" int mm25_sim_rank(const void* a, const void* b, int dtype, int64_t Q, int64_t N,
                             int64_t d, const float* tval, const int64_t* tcol, int* rank,
                             void* stream) {
  if (!a || !b || !tval || !tcol || !rank) return MM25_ERR_ARG;
  if (dtype != MM25_BF16) return MM25_ERR_DTYPE;
  return mm25::tc::launch_rank(a, b, Q, N, d, tval, tcol, rank, (cudaStream_t)stream);
}
