// sim_tc_sweep.cuh — shared pieces of the S-tile sweep kernels (forward statistics kernel and
// backward G-panel kernel): tile constants, shared-memory layout, the TMA producer loop and the
// MMA issuer loop.  A work item is (128-row block, column segment); every 128 x 256 S tile is a
// K-loop over d in 64-wide chunks, A and B chunks streamed together through a 4-stage ring
// (192 KB in flight: the v1 kernel with a resident A block and 16 KB B stages was bound by
// prefetch depth and by 128-wide UMMA operand bandwidth, profiles/r1_v1_ncu_summary.md).
#pragma once
#include "tc_common.cuh"

namespace mm25 {
namespace tc {

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int MAX_KC = 64;           // K <= 4096 (d <= 1024 features, x4 for the split operands)
constexpr int NST = 4;               // ring stages (A 16 KB + B 32 KB each)
constexpr int NACC = 2;              // TMEM accumulator stages (256 columns each)
constexpr int A_BYTES = BM * BK * 2;  // 16384
constexpr int B_BYTES = BN * BK * 2;  // 32768
constexpr int NEPI = 8;              // epilogue warps: 2 per SM sub-partition
constexpr int FWD_THREADS = 64 + 32 * NEPI;  // warp0 TMA, warp1 MMA, warps 2.. epilogue
constexpr int CH_PER_WARP = BN / 32 / 2;     // 32-column chunks each epilogue warp handles
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float SPREAD_LIMIT = 96.f;  // log2 units: larger in-block spread takes the slow path

struct FwdSmem {
  uint8_t a[NST][A_BYTES];
  uint8_t b[NST][B_BYTES];
  float colm[2][4][BN];
  float cols[2][4][BN];
  uint8_t stage[NEPI][2048];  // per-warp staging of the bf16 E chunk (TMA store, 64B swizzle)
  uint64_t full[NST], empty[NST];
  uint64_t acc_full[NACC], acc_empty[NACC];
  uint32_t tmem_base;
};

// gradient-panel kernel: same rings; per-warp staging of the bf16 G chunk for the TMA store
// (32 rows x 64 B, 64B-swizzled) and of the column factors of the warp's 128 columns.
struct GradSmem {
  uint8_t a[NST][A_BYTES];
  uint8_t b[NST][B_BYTES];
  uint8_t stage[NEPI][2048];
  float ccw[NEPI][BN / 2];
  float red[NEPI];
  uint64_t full[NST], empty[NST];
  uint64_t acc_full[NACC], acc_empty[NACC];
  uint32_t tmem_base;
};

// work decomposition: items = (row block, column segment); row blocks [first_rbg, first_rbg+num_rbg)
struct Sweep {
  int KC, ntiles, nseg, tps, nrb_per_batch, first_rbg, num_rbg;
  int64_t N;
};

template <class SM>
__device__ __forceinline__ void sweep_setup(SM& sm, const CUtensorMap* ta,
                                            const CUtensorMap* tb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
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
template <class SM>
__device__ __forceinline__ void sweep_producer(SM& sm, const CUtensorMap* ta,
                                               const CUtensorMap* tb, const Sweep& w) {
  uint32_t st = 0, ph = 0;
  const int nitems = w.num_rbg * w.nseg;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const int rbg = w.first_rbg + item / w.nseg, seg = item % w.nseg;
    const int bz = rbg / w.nrb_per_batch;
    const int32_t arow = (int32_t)((int64_t)rbg * BM);  // row in the flattened [B*M, d] matrix
    const int t0 = seg * w.tps, t1 = min(w.ntiles, t0 + w.tps);
    for (int t = t0; t < t1; ++t) {
      const int32_t brow = (int32_t)((int64_t)bz * w.N + (int64_t)t * BN);
      for (int kc = 0; kc < w.KC; ++kc) {
        mbar_wait(&sm.empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&sm.full[st], A_BYTES + B_BYTES);
        tma_load_2d(sm.a[st], ta, &sm.full[st], kc * BK, arow);
        tma_load_2d(sm.b[st], tb, &sm.full[st], kc * BK, brow);
        if (++st == NST) { st = 0; ph ^= 1; }
      }
    }
  }
}

// one elected thread of warp 1
template <class SM>
__device__ __forceinline__ void sweep_mma(SM& sm, uint32_t tmem, const Sweep& w) {
  constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 0, 0);
  uint32_t st = 0, ph = 0, as = 0, accph = 0;
  const int nitems = w.num_rbg * w.nseg;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const int seg = item % w.nseg;
    const int t0 = seg * w.tps, t1 = min(w.ntiles, t0 + w.tps);
    for (int t = t0; t < t1; ++t) {
      mbar_wait(&sm.acc_empty[as], accph ^ 1);
      tc_fence_after();
      const uint32_t dcol = tmem + as * BN;
      for (int kc = 0; kc < w.KC; ++kc) {
        mbar_wait(&sm.full[st], ph);
        tc_fence_after();
        const uint32_t abase = smem_u32(sm.a[st]);
        const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t ad = make_smem_desc(abase + k * 32, 16, 1024);
          const uint64_t bd = make_smem_desc(bbase + k * 32, 16, 1024);
          umma_bf16(dcol, ad, bd, idesc, (kc | k) != 0);
        }
        umma_commit(&sm.empty[st]);
        if (++st == NST) { st = 0; ph ^= 1; }
      }
      umma_commit(&sm.acc_full[as]);
      if (++as == NACC) { as = 0; accph ^= 1; }
    }
  }
}

// ------------------------------------------------------------------------------------------
// EXPERIMENTAL CTA-pair variants (opt-in MM25_SWEEP_2CTA=1, never taken by default; protocol of
// gemm_tc2.cu).  A cluster of two CTAs sweeps a 256-row super-block: CTA r owns row block
// 2 g + r, stages its own A chunk (16 KB) and HALF of the B chunk (columns r 128 .., 16 KB, in
// the first half of the 32 KB slot), the leader (rank 0) issues `tcgen05.mma.cta_group::2`
// with M = 256, and each CTA's tensor memory receives its 128 rows x all 256 columns — so the
// epilogues are unchanged.  Per SM the operands staged and read per k-block drop from 48 KB to
// 32 KB (the gradient-panel kernel is bound by shared-memory bandwidth).
// Items are numbered over PAIRS: item = (super-block, column segment).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pair_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void pair_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t pair_mapa(const void* p, uint32_t rank) {
  uint32_t out;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(smem_u32(p)), "r"(rank));
  return out;
}
__device__ __forceinline__ void pair_mbar_arrive_leader(uint64_t* bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(
                   pair_mapa(bar, 0))
               : "memory");
}
__device__ __forceinline__ void pair_tma_load_2d(void* smem_dst, const CUtensorMap* m,
                                                 uint32_t bar_cluster, int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void pair_umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void pair_umma_commit(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(smem_u32(bar)),
      "h"(mask)
      : "memory");
}

template <class SM>
__device__ __forceinline__ void sweep_setup_pair(SM& sm, const CUtensorMap* ta,
                                                 const CUtensorMap* tb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < NACC; ++i) {
      mbar_init(&sm.acc_full[i], 1);
      mbar_init(&sm.acc_empty[i], 2 * NEPI);  // the epilogue warps of BOTH CTAs (leader's copy)
    }
    fence_barrier_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&sm.tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    if (lane == 0) { tma_prefetch_desc(ta); tma_prefetch_desc(tb); }
  }
  tc_fence_before();
  pair_sync_all();  // both CTAs' barriers exist before any remote arrive / peer-completing TMA
  tc_fence_after();
}

template <class SM>
__device__ __forceinline__ void sweep_teardown_pair(SM& sm, uint32_t tmem) {
  tc_fence_before();
  __syncwarp();
  pair_sync_all();  // the peer may still read my shared memory / arrive on my barriers
  if ((threadIdx.x >> 5) == 0)
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u)
                 : "memory");
}

// one elected thread of warp 0 of EACH CTA; `tb_half` has 128-row boxes
template <class SM>
__device__ __forceinline__ void sweep_producer_pair(SM& sm, const CUtensorMap* ta,
                                                    const CUtensorMap* tb_half, const Sweep& w) {
  const uint32_t rank = pair_rank();
  uint32_t st = 0, ph = 0;
  const int nitems = (w.num_rbg >> 1) * w.nseg;
  for (int item = blockIdx.x >> 1; item < nitems; item += gridDim.x >> 1) {
    const int rbg = w.first_rbg + 2 * (item / w.nseg) + (int)rank, seg = item % w.nseg;
    const int bz = rbg / w.nrb_per_batch;
    const int32_t arow = (int32_t)((int64_t)rbg * BM);
    const int t0 = seg * w.tps, t1 = min(w.ntiles, t0 + w.tps);
    for (int t = t0; t < t1; ++t) {
      const int32_t brow =
          (int32_t)((int64_t)bz * w.N + (int64_t)t * BN + (int64_t)rank * (BN / 2));
      for (int kc = 0; kc < w.KC; ++kc) {
        mbar_wait(&sm.empty[st], ph ^ 1);
        const uint32_t fbar = pair_mapa(&sm.full[st], 0);
        if (rank == 0) mbar_arrive_expect_tx(&sm.full[st], 2 * (A_BYTES + B_BYTES / 2));
        pair_tma_load_2d(sm.a[st], ta, fbar, kc * BK, arow);
        pair_tma_load_2d(sm.b[st], tb_half, fbar, kc * BK, brow);
        if (++st == NST) { st = 0; ph ^= 1; }
      }
    }
  }
}

// one elected thread of warp 1 of the LEADER CTA
template <class SM>
__device__ __forceinline__ void sweep_mma_pair(SM& sm, uint32_t tmem, const Sweep& w) {
  constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, 0, 0);
  uint32_t st = 0, ph = 0, as = 0, accph = 0;
  const int nitems = (w.num_rbg >> 1) * w.nseg;
  for (int item = blockIdx.x >> 1; item < nitems; item += gridDim.x >> 1) {
    const int seg = item % w.nseg;
    const int t0 = seg * w.tps, t1 = min(w.ntiles, t0 + w.tps);
    for (int t = t0; t < t1; ++t) {
      mbar_wait(&sm.acc_empty[as], accph ^ 1);
      tc_fence_after();
      const uint32_t dcol = tmem + as * BN;
      for (int kc = 0; kc < w.KC; ++kc) {
        mbar_wait(&sm.full[st], ph);
        tc_fence_after();
        const uint32_t abase = smem_u32(sm.a[st]);
        const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t ad = make_smem_desc(abase + k * 32, 16, 1024);
          const uint64_t bd = make_smem_desc(bbase + k * 32, 16, 1024);
          pair_umma_bf16(dcol, ad, bd, idesc, (kc | k) != 0);
        }
        pair_umma_commit(&sm.empty[st]);
        if (++st == NST) { st = 0; ph ^= 1; }
      }
      pair_umma_commit(&sm.acc_full[as]);
      if (++as == NACC) { as = 0; accph ^= 1; }
    }
  }
}

}  // namespace tc
}  // namespace mm25
