// sim_tc_sweep.cuh — shared pieces of the "row block resident, column tiles streaming" tcgen05
// kernels (forward statistics kernel and backward G-panel kernel): tile constants, shared-memory
// layout, the TMA producer loop and the MMA issuer loop.
#pragma once
#include "tc_common.cuh"

namespace mm25 {
namespace tc {

constexpr int BM = 128, BN = 128, BK = 64;
constexpr int MAX_KC = 8;            // d <= 512 keeps the A block resident (128 KB)
constexpr int NST = 4;               // B ring stages (16 KB each)
constexpr int NACC = 4;              // TMEM accumulator stages (128 columns each)
constexpr int CHUNK_BYTES = BM * BK * 2;  // 16384
constexpr int NEPI = 8;              // epilogue warps: 2 per SM sub-partition
constexpr int FWD_THREADS = 64 + 32 * NEPI;  // warp0 TMA, warp1 MMA, warps 2.. epilogue
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float SPREAD_LIMIT = 96.f;  // log2 units: larger in-block spread takes the slow path

struct FwdSmem {
  uint8_t a[MAX_KC * CHUNK_BYTES];
  uint8_t b[NST * CHUNK_BYTES];
  float colm[2][4][BN];
  float cols[2][4][BN];
  uint64_t a_full, a_empty;
  uint64_t b_full[NST], b_empty[NST];
  uint64_t acc_full[NACC], acc_empty[NACC];
  uint32_t tmem_base;
};


// work decomposition: items = (row block, column segment); row blocks [first_rbg, first_rbg+num_rbg)
struct Sweep {
  int KC, ntiles, nseg, tps, nrb_per_batch, first_rbg, num_rbg;
  int64_t N;
};

__device__ __forceinline__ void sweep_setup(FwdSmem& sm, const CUtensorMap* ta,
                                            const CUtensorMap* tb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&sm.a_full, 1);
    mbar_init(&sm.a_empty, 1);
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < NACC; ++i) { mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.acc_empty[i], NEPI); }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&sm.tmem_base, 512);
    if (lane == 0) { tma_prefetch_desc(ta); tma_prefetch_desc(tb); }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
}

// one elected thread of warp 0
__device__ __forceinline__ void sweep_producer(FwdSmem& sm, const CUtensorMap* ta,
                                               const CUtensorMap* tb, const Sweep& w) {
  uint32_t bst = 0, bph = 0, aph = 0;
  const int nitems = w.num_rbg * w.nseg;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const int rbg = w.first_rbg + item / w.nseg, seg = item % w.nseg;
    const int bz = rbg / w.nrb_per_batch;
    const int64_t arow = (int64_t)rbg * BM;  // row in the flattened [B*M, d] matrix
    mbar_wait(&sm.a_empty, aph ^ 1);
    aph ^= 1;
    mbar_arrive_expect_tx(&sm.a_full, (uint32_t)(w.KC * CHUNK_BYTES));
    for (int kc = 0; kc < w.KC; ++kc)
      tma_load_2d(sm.a + kc * CHUNK_BYTES, ta, &sm.a_full, kc * BK, (int32_t)arow);
    const int t0 = seg * w.tps, t1 = min(w.ntiles, t0 + w.tps);
    for (int t = t0; t < t1; ++t) {
      const int64_t brow = (int64_t)bz * w.N + (int64_t)t * BN;
      for (int kc = 0; kc < w.KC; ++kc) {
        mbar_wait(&sm.b_empty[bst], bph ^ 1);
        mbar_arrive_expect_tx(&sm.b_full[bst], CHUNK_BYTES);
        tma_load_2d(sm.b + bst * CHUNK_BYTES, tb, &sm.b_full[bst], kc * BK, (int32_t)brow);
        if (++bst == NST) { bst = 0; bph ^= 1; }
      }
    }
  }
}

// one elected thread of warp 1
__device__ __forceinline__ void sweep_mma(FwdSmem& sm, uint32_t tmem, const Sweep& w) {
  constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 0, 0);
  uint32_t bst = 0, bph = 0, aph = 0, as = 0, accph = 0;
  const int nitems = w.num_rbg * w.nseg;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const int seg = item % w.nseg;
    mbar_wait(&sm.a_full, aph);
    aph ^= 1;
    tc_fence_after();
    const int t0 = seg * w.tps, t1 = min(w.ntiles, t0 + w.tps);
    for (int t = t0; t < t1; ++t) {
      mbar_wait(&sm.acc_empty[as], accph ^ 1);
      tc_fence_after();
      const uint32_t dcol = tmem + as * BN;
      for (int kc = 0; kc < w.KC; ++kc) {
        mbar_wait(&sm.b_full[bst], bph);
        tc_fence_after();
        const uint32_t abase = smem_u32(sm.a + kc * CHUNK_BYTES);
        const uint32_t bbase = smem_u32(sm.b + bst * CHUNK_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t ad = make_smem_desc(abase + k * 32, 16, 1024);
          const uint64_t bd = make_smem_desc(bbase + k * 32, 16, 1024);
          umma_bf16(dcol, ad, bd, idesc, (kc | k) != 0);
        }
        umma_commit(&sm.b_empty[bst]);
        if (++bst == NST) { bst = 0; bph ^= 1; }
      }
      umma_commit(&sm.acc_full[as]);
      if (++as == NACC) { as = 0; accph ^= 1; }
    }
    umma_commit(&sm.a_empty);
  }
}

}  // namespace tc
}  // namespace mm25
