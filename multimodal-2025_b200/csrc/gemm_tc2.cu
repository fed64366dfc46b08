// gemm_tc2.cu — EXPERIMENTAL (opt-in: MM25_GEMM_2CTA=1; never taken by default) CTA-pair variant
// of the grouped bf16 GEMM of the tensor-core backward (gemm_tc.cu):
//     C[Mg, Ng] (fp32) (+)= alpha * Aop[Mg, K] * Bop[K, Ng]      for up to two problems
// Two CTAs of a cluster (one TPC) own a 256 x 256 output tile.  `tcgen05.mma.cta_group::2`
// with M = 256: CTA r of the pair holds A rows [r 128, r 128 + 128) and B columns
// [r 128, r 128 + 128) of the tile in ITS shared memory (same offsets in both CTAs) and the
// accumulator rows r 128 .. of all 256 columns in ITS tensor memory.  Against the 1-CTA kernel
// (A 16 KB + B 32 KB per k-block and SM) every SM stages and reads A 16 KB + B 16 KB: a third
// less shared-memory and L2 operand traffic per FLOP — the 1-CTA kernel sits at the pool's
// cuBLAS rate (1.37 PF/s, power-capped), so the point of the experiment is energy per FLOP.
//
// Protocol (differences from gemm_tc.cu):
//   * both CTAs run a TMA producer for their own slices; every load completes on the LEADER's
//     (cluster rank 0) `full` barrier (`cp.async.bulk.tensor ... .cta_group::2`, barrier address
//     mapped with `mapa`), on which the leader expects the bytes of both CTAs;
//   * only the leader issues MMAs; `tcgen05.commit ... multicast::cluster` (mask 0b11) arrives
//     on the `empty` / `acc_full` barriers of BOTH CTAs;
//   * the epilogue warps of both CTAs arrive on the leader's `acc_empty` (count 8);
//   * TMEM is allocated / freed with cta_group::2 by warp 0 of each CTA, a cluster barrier
//     before the first remote arrive and before the free.
// Not supported here (the caller falls back to the 1-CTA kernel): scatter (fused
// reduce-scatter), split-K, batched problems, an odd number of 128-row blocks.
// The instruction forms and who-executes-what were cross-checked against the CUTLASS headers
// vendored under site-packages/flashinfer/data/cutlass/include (behaviour only, no code taken):
// `SM100_TMA_2SM_LOAD_2D` (both CTAs issue, bytes complete on CTA 0's barrier),
// `Allocator2Sm` (the same warp of BOTH CTAs allocates / frees, same smem slot),
// `umma_arrive_multicast_2x1SM`, and the 2-SM pipeline of sm100_gemm_tma_warpspecialized.hpp
// (leader expects the bytes of both CTAs; both epilogues release the leader's accumulator).
// Written without GPU time left in the round: compiled (ptxas) but NOT yet run; every mbarrier
// wait is the bounded spin of tc_common.cuh, so a protocol error traps instead of hanging.
#include "tc_common.cuh"
#include "gemm_tc.h"

#include <cstdlib>
#include <cstring>

namespace mm25 {
namespace tc {

void gemm_schedule(int64_t mb0, int64_t nk0, int64_t mb1, int64_t nk1, int ngroups, int* r,
                   int* nbA, int* nbB, int* nsA, int* nsB);

namespace g2 {
constexpr int BM = 128;    // rows per CTA (the pair: 256)
constexpr int BN = 256;    // tile columns (each CTA stages BN / 2 of B)
constexpr int BNH = 128;
constexpr int BK = 64;
constexpr int NST = 6;
constexpr int NACC = 2;
constexpr int A_BYTES = BM * BK * 2;    // 16 KB
constexpr int B_BYTES = BNH * BK * 2;   // 16 KB
constexpr int THREADS = 192;

struct Smem {
  uint8_t a[NST][A_BYTES];
  uint8_t b[NST][B_BYTES];
  uint64_t full[NST], empty[NST];
  uint64_t acc_full[NACC], acc_empty[NACC];
  uint32_t tmem_base;
};

struct Problem {
  CUtensorMap ta, tb;
  float* C;
  int64_t ldc;
  int nk;        // k-blocks of BK
  int sblocks;   // Mg / 256
  int a_mn, b_mn;
  int accumulate;
  float alpha;
};

struct Params {
  Problem p[2];
  int group, ngroups, r, nbA, nbB, nsA, nsB;
  const float* alpha_dev;
};

// job k of group `grp` (same static schedule as gemm_tc.cu, on 256-row super-blocks)
__device__ __forceinline__ bool job_at(const Params& P, int grp, int k, int* prob, int* sb) {
  const bool A = grp < P.r;
  const int nb = A ? P.nbA : P.nbB;
  const int ns = A ? P.nsA : P.nsB;
  int job;
  if (k < nb) {
    *prob = 0;
    job = grp + k * P.ngroups;
  } else {
    const int ks = k - nb;
    if (ks >= ns) return false;
    *prob = 1;
    job = (A ? grp * P.nsA : P.r * P.nsA + (grp - P.r) * P.nsB) + ks;
  }
  if (job >= P.p[*prob].sblocks) return false;
  *sb = job;
  return true;
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (an address in MY shared memory) inside CTA `rank`
__device__ __forceinline__ uint32_t mapa_u32(const void* p, uint32_t rank) {
  uint32_t out;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(smem_u32(p)), "r"(rank));
  return out;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr)
               : "memory");
}
// 2-D tile load into MY shared memory whose bytes complete on a barrier that may live in the
// peer CTA of the pair (shared::cluster address `bar_cluster`)
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m,
                                                 uint32_t bar_cluster, int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(smem_u32(bar)),
      "h"(mask)
      : "memory");
}
}  // namespace g2

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(g2::THREADS, 1)
gemm_grouped2_kernel(const __grid_constant__ g2::Params P) {
  using namespace g2;
  extern __shared__ uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                      ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int grp = pair / P.group, nt = pair % P.group;
  const int n0 = nt * BN;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < NACC; ++i) { mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.acc_empty[i], 8); }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc_pair(&sm.tmem_base, 512);
    if (lane == 0) {
      tma_prefetch_desc(&P.p[0].ta); tma_prefetch_desc(&P.p[0].tb);
      tma_prefetch_desc(&P.p[1].ta); tma_prefetch_desc(&P.p[1].tb);
    }
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / TMA
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      int prob, sb;
      for (int k = 0; job_at(P, grp, k, &prob, &sb); ++k) {
        const Problem& pr = P.p[prob];
        const int m0 = (sb * 2 + (int)rank) * BM;      // my 128 rows of the 256-row super-block
        const int nb0 = n0 + (int)rank * BNH;          // my 128 columns of B
        for (int kb = 0; kb < pr.nk; ++kb) {
          const int k0 = kb * BK;
          mbar_wait(&sm.empty[st], ph ^ 1);            // my slot is free (multicast commit)
          const uint32_t fbar = mapa_u32(&sm.full[st], 0);
          if (leader) mbar_arrive_expect_tx(&sm.full[st], 2 * (A_BYTES + B_BYTES));
          if (pr.a_mn) {
#pragma unroll
            for (int at = 0; at < BM / 64; ++at)
              tma_load_2d_pair(sm.a[st] + at * 8192, &pr.ta, fbar, m0 + at * 64, k0);
          } else {
            tma_load_2d_pair(sm.a[st], &pr.ta, fbar, k0, m0);
          }
          if (pr.b_mn) {
#pragma unroll
            for (int at = 0; at < BNH / 64; ++at)
              tma_load_2d_pair(sm.b[st] + at * 8192, &pr.tb, fbar, nb0 + at * 64, k0);
          } else {
            tma_load_2d_pair(sm.b[st], &pr.tb, fbar, k0, nb0);
          }
          if (++st == NST) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && lane == 0) {
      uint32_t st = 0, ph = 0, as = 0, accph = 0;
      int prob, sb;
      for (int k = 0; job_at(P, grp, k, &prob, &sb); ++k) {
        const Problem& pr = P.p[prob];
        const int a_mn = pr.a_mn, b_mn = pr.b_mn, nk = pr.nk;
        const uint32_t idesc = make_idesc_bf16(2 * BM, BN, a_mn, b_mn);
        const uint32_t a_step = a_mn ? 2048 : 32, a_lbo = a_mn ? 8192 : 16;
        const uint32_t b_step = b_mn ? 2048 : 32, b_lbo = b_mn ? 8192 : 16;
        mbar_wait(&sm.acc_empty[as], accph ^ 1);   // the epilogues of BOTH CTAs drained it
        tc_fence_after();
        const uint32_t dcol = tmem + as * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&sm.full[st], ph);             // bytes of both CTAs have landed
          tc_fence_after();
          const uint32_t abase = smem_u32(sm.a[st]);
          const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t ad = make_smem_desc(abase + kk * a_step, a_lbo, 1024);
            const uint64_t bd = make_smem_desc(bbase + kk * b_step, b_lbo, 1024);
            umma_bf16_pair(dcol, ad, bd, idesc, (kb | kk) != 0);
          }
          umma_commit_pair(&sm.empty[st]);
          if (++st == NST) { st = 0; ph ^= 1; }
        }
        umma_commit_pair(&sm.acc_full[as]);
        if (++as == NACC) { as = 0; accph ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;
    const float ad = P.alpha_dev ? __ldg(P.alpha_dev) : 1.f;
    uint32_t as = 0, accph = 0;
    int prob, sb;
    for (int k = 0; job_at(P, grp, k, &prob, &sb); ++k) {
      const Problem& pr = P.p[prob];
      const float al = pr.alpha * ad;
      const int accumulate = pr.accumulate;
      const int grow = (sb * 2 + (int)rank) * BM + q * 32 + lane;
      float* crow = pr.C + (int64_t)grow * pr.ldc + n0;
      mbar_wait(&sm.acc_full[as], accph);
      tc_fence_after();
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        float v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + as * BN + ch * 32, v);
        float4* dst = reinterpret_cast<float4*>(crow + ch * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float4 o = make_float4(v[4 * j] * al, v[4 * j + 1] * al, v[4 * j + 2] * al,
                                 v[4 * j + 3] * al);
          if (accumulate) {
            const float4 c = dst[j];
            o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
          }
          dst[j] = o;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(&sm.acc_empty[as], 0));
      if (++as == NACC) { as = 0; accph ^= 1; }
    }
  }
  tc_fence_before();
  __syncwarp();        // lanes of the single-thread roles reconverge before the aligned barrier
  cluster_sync_all();  // nobody frees tensor memory / exits while the peer still uses the pair
  if (warp == 0) tmem_dealloc_pair(tmem, 512);
}

static int sm_count_g2() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

static int fill_problem2(g2::Problem* p, const GemmDesc& d, int64_t Ng) {
  int rc;
  if (d.a_mn) rc = make_tmap_bf16(&p->ta, d.A, (uint64_t)d.K, (uint64_t)d.Mg, 64, 64);
  else rc = make_tmap_bf16(&p->ta, d.A, (uint64_t)d.Mg, (uint64_t)d.K, g2::BM, 64);
  if (rc) return rc;
  if (d.b_mn) rc = make_tmap_bf16(&p->tb, d.B, (uint64_t)d.K, (uint64_t)Ng, 64, 64);
  else rc = make_tmap_bf16(&p->tb, d.B, (uint64_t)Ng, (uint64_t)d.K, g2::BNH, 64);
  if (rc) return rc;
  p->C = d.C;
  p->ldc = d.ldc;
  p->nk = (int)(d.K / g2::BK);
  p->sblocks = (int)(d.Mg / (2 * g2::BM));
  p->a_mn = d.a_mn;
  p->b_mn = d.b_mn;
  p->accumulate = d.accumulate;
  p->alpha = d.alpha;
  return MM25_OK;
}

static bool plain2(const GemmDesc* d) {
  return d->peers == nullptr && d->batch <= 1 && d->Mg % (2 * g2::BM) == 0;
}

// MM25_ERR_SHAPE = "not for this kernel": the caller runs the 1-CTA kernel instead
int gemm_grouped2(const GemmDesc* d0, const GemmDesc* d1, int64_t Ng, const float* alpha_dev,
                  cudaStream_t st) {
  if (!d0 && !d1) return MM25_OK;
  if (!d0) { d0 = d1; d1 = nullptr; }
  if (d1 && d1->K > d0->K) { const GemmDesc* t = d0; d0 = d1; d1 = t; }
  if (!gemm_eligible(d0->Mg, Ng, d0->K) || !plain2(d0)) return MM25_ERR_SHAPE;
  if (d1 && (!gemm_eligible(d1->Mg, Ng, d1->K) || !plain2(d1))) return MM25_ERR_SHAPE;
  g2::Params P;
  memset(&P, 0, sizeof(P));
  int rc = fill_problem2(&P.p[0], *d0, Ng);
  if (rc) return rc;
  if (d1) {
    rc = fill_problem2(&P.p[1], *d1, Ng);
    if (rc) return rc;
  } else {
    P.p[1] = P.p[0];
    P.p[1].sblocks = 0;
  }
  P.alpha_dev = alpha_dev;
  const int group = (int)(Ng / g2::BN);             // CTA pairs per group
  int ngroups = (sm_count_g2() / 2) / group;
  const int64_t total = P.p[0].sblocks + (int64_t)P.p[1].sblocks;
  if (ngroups > total) ngroups = (int)total;
  if (ngroups < 1) ngroups = 1;
  // an under-filled long-K problem wants split-K: leave it to the 1-CTA kernel
  if (P.p[0].sblocks < ngroups && P.p[0].nk >= 128) return MM25_ERR_SHAPE;
  P.group = group;
  P.ngroups = ngroups;
  gemm_schedule(P.p[0].sblocks, P.p[0].nk, P.p[1].sblocks, P.p[1].nk, ngroups, &P.r, &P.nbA,
                &P.nbB, &P.nsA, &P.nsB);
  const size_t smem = sizeof(g2::Smem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(gemm_grouped2_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  { void* tt_ = timing_begin(MM25_T_GEMM_GROUPED, st);
  gemm_grouped2_kernel<<<ngroups * group * 2, g2::THREADS, smem, st>>>(P);
  timing_end(tt_, st); }
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

}  // namespace tc
}  // namespace mm25
