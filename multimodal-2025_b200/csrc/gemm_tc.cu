// gemm_tc.cu — bf16 tcgen05 GEMMs of the tensor-core backward:
//     C[Mg, Ng] (fp32)  (+)=  alpha * Aop[Mg, K] * Bop[K, Ng]
// Operands are fed by TMA straight from their natural row-major storage, so each of them may
// be "K-major" (memory [rows = M/N index, cols = K]) or "MN-major" (memory [rows = K,
// cols = M/N index]); the UMMA descriptors carry the difference (no transposes in HBM).
//   dA = G  . Yhat   : A = G   K-major  [R, N],   B = Yhat MN-major [N, d]
//   dB = G^T. Xhat   : A = G   MN-major [R, N],   B = Xhat MN-major [R, d]
// Tile 128 x 256 x 64, 4-stage TMA ring, 2 TMEM accumulator stages, persistent CTAs.
//
// One launch runs up to TWO problems that share Ng (the "grouped" kernel): with separate
// launches each problem pays its own wave quantisation (256 and 512 tiles on 148 SMs: 86 %),
// together 1024 cost units level out at 7 per SM (98.8 %).  Work is assigned by a static
// host-computed schedule (deterministic, no atomics): CTAs form groups of Ng/256 neighbours
// that walk the same list of m-blocks, CTA g of a group owning n-tile g, so the two CTAs that
// read one G tile run side by side and share it in L2.  Every group first takes its m-blocks
// of the long-K problem round-robin, then a contiguous run of m-blocks of the short-K problem
// sized by water-filling.
#include "tc_common.cuh"
#include "gemm_tc.h"

#include <cstdlib>

namespace mm25 {
namespace tc {

namespace g {
constexpr int BM = 128, BN = 256, BK = 64;
constexpr int NST = 4;
constexpr int NACC = 2;
constexpr int A_BYTES = BM * BK * 2;   // 16 KB
constexpr int B_BYTES = BN * BK * 2;   // 32 KB
constexpr int THREADS = 192;

struct Smem {
  uint8_t a[NST][A_BYTES];
  uint8_t b[NST][B_BYTES];
  uint64_t full[NST], empty[NST];
  uint64_t acc_full[NACC], acc_empty[NACC];
  uint32_t tmem_base;
};

// Opt-in staged scatter (MM25_RS_MODE, default 0 = per-lane 16-byte red.global): the plain
// scatter epilogue has every lane add 16 bytes of its own row, i.e. 32 separate 16-byte remote
// atomics per warp instruction over NVLink (measured: the 56 MB a rank pushes at P = 8 take
// ~180 us on top of the math, ~0.3 TB/s of a 0.9 TB/s link).  The staged modes pass each
// 32 x 32 chunk through a padded per-warp shared-memory buffer (row stride 144 B: 16-byte
// stores of a quarter-warp land in distinct banks) and send whole 128-byte row segments:
//   mode 1: one `cp.reduce.async.bulk ... add.f32` (TMA reduce) of 128 B per lane and chunk;
//   mode 2: `red.global.add.v4.f32` re-read row-contiguously (8 lanes cover one 128-byte row).
// The buffer lives after Smem and is only allocated when a staged mode is selected.
constexpr int STG_LD = 36;                         // floats per staged row (128 B + 16 B pad)
constexpr int STG_BYTES = 4 * 32 * STG_LD * 4;     // 4 epilogue warps x 32 rows

struct Problem {
  CUtensorMap ta, tb;
  float* C;
  int64_t ldc;
  int nk;          // k-blocks (of BK) per JOB = (K / BK) / ksplit
  int mblocks;     // Mg / BM
  int ksplit;      // split-K factor: jobs = mblocks * ksplit, job j -> (m-block j % mblocks,
                   // k-part j / mblocks); partial tiles are added with red.global (C zeroed)
  int a_mn, b_mn;
  int accumulate;
  float alpha;
  // scatter mode (fused reduce-scatter): m-block rows [q R, (q+1) R) are ADDED (red.global) into
  // peer[q] + (row - q R) * ldc — peer-mapped fp32 buffers of the other ranks over NVLink
  int scatter;
  int rows_per_rank;
  int rot;         // scatter: m-block rotation (ring order over the owners, see GemmDesc)
  int mb_per_batch;     // batched: m-blocks per batch entry (0 = not batched)
  int a_k_rows, b_k_rows;  // batched: K-coordinate shift per entry of A (MN-major) and B
  float* peer[GEMM_MAX_PEERS];
};

// groups [0, r) are class A, the rest class B; problem 0 is walked round-robin over groups
// (nb0 blocks each), problem 1 in contiguous runs (ns blocks each, class A runs first).
struct Schedule {
  int group;       // CTAs per group = Ng / BN
  int ngroups;
  int r;
  int nbA, nbB;
  int nsA, nsB;
  int p1_first;    // walk problem 1 before problem 0 (problem 1 scatters over NVLink: its
                   // fire-and-forget adds then drain under problem 0's tiles)
};

struct Params {
  Problem p[2];
  Schedule s;
  const float* alpha_dev;
};

// job k of group `grp`: which problem and which m-block; false when the list is exhausted
__device__ __forceinline__ bool job_at(const Params& P, int grp, int k, int* prob, int* mb,
                                       int* kb0) {
  const Schedule& s = P.s;
  const bool A = grp < s.r;
  const int nb = A ? s.nbA : s.nbB;
  const int ns = A ? s.nsA : s.nsB;
  const int base = A ? grp * s.nsA : s.r * s.nsA + (grp - s.r) * s.nsB;
  int job;
  if (s.p1_first) {
    // this group's run of problem 1 may be cut short by the end of the problem
    const int tot1 = P.p[1].mblocks * P.p[1].ksplit;
    const int n1 = min(ns, max(0, tot1 - base));
    if (k < n1) {
      *prob = 1;
      job = base + k;
    } else {
      if (k - n1 >= nb) return false;
      *prob = 0;
      job = grp + (k - n1) * s.ngroups;
    }
  } else if (k < nb) {
    *prob = 0;
    job = grp + k * s.ngroups;
  } else {
    const int ks = k - nb;
    if (ks >= ns) return false;
    *prob = 1;
    job = base + ks;
  }
  const Problem& pr = P.p[*prob];
  if (job >= pr.mblocks * pr.ksplit) return false;
  *mb = job % pr.mblocks;
  if (pr.scatter) *mb = (*mb + pr.rot) % pr.mblocks;
  *kb0 = (job / pr.mblocks) * pr.nk;
  return true;
}
}  // namespace g

template <int RS>
__global__ void __launch_bounds__(g::THREADS, 1)
gemm_grouped_kernel(const __grid_constant__ g::Params P) {
  using namespace g;
  extern __shared__ uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                      ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = blockIdx.x / P.s.group, nt = blockIdx.x % P.s.group;
  const int n0 = nt * BN;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < NACC; ++i) { mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.acc_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&sm.tmem_base, 512);
    if (lane == 0) {
      tma_prefetch_desc(&P.p[0].ta); tma_prefetch_desc(&P.p[0].tb);
      tma_prefetch_desc(&P.p[1].ta); tma_prefetch_desc(&P.p[1].tb);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      int prob, mb, kb0;
      for (int k = 0; job_at(P, grp, k, &prob, &mb, &kb0); ++k) {
        const Problem& pr = P.p[prob];
        const int bz = pr.mb_per_batch ? mb / pr.mb_per_batch : 0;
        // MN-major A of a batched problem: M coordinate local to the entry, K shifted
        const int m0 = ((pr.mb_per_batch && pr.a_mn) ? mb % pr.mb_per_batch : mb) * BM;
        const int ka_off = pr.a_mn ? bz * pr.a_k_rows : 0, kb_off = bz * pr.b_k_rows;
        for (int kb = 0; kb < pr.nk; ++kb) {
          const int k0 = (kb0 + kb) * BK;
          mbar_wait(&sm.empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&sm.full[st], A_BYTES + B_BYTES);
          if (pr.a_mn) {
#pragma unroll
            for (int at = 0; at < BM / 64; ++at)
              tma_load_2d(sm.a[st] + at * 8192, &pr.ta, &sm.full[st], m0 + at * 64, k0 + ka_off);
          } else {
            tma_load_2d(sm.a[st], &pr.ta, &sm.full[st], k0, m0);
          }
          if (pr.b_mn) {
#pragma unroll
            for (int at = 0; at < BN / 64; ++at)
              tma_load_2d(sm.b[st] + at * 8192, &pr.tb, &sm.full[st], n0 + at * 64, k0 + kb_off);
          } else {
            tma_load_2d(sm.b[st], &pr.tb, &sm.full[st], k0 + kb_off, n0);
          }
          if (++st == NST) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0, as = 0, accph = 0;
      int prob, mb, kb0;
      for (int k = 0; job_at(P, grp, k, &prob, &mb, &kb0); ++k) {
        const Problem& pr = P.p[prob];
        const int a_mn = pr.a_mn, b_mn = pr.b_mn, nk = pr.nk;
        const uint32_t idesc = make_idesc_bf16(BM, BN, a_mn, b_mn);
        // K-major: +32 B per 16 k inside the 128 B swizzle row; SBO = 8 rows.
        // MN-major: +16 k-rows (2048 B); LBO = next 64-wide MN atom, SBO = 8 k-rows.
        const uint32_t a_step = a_mn ? 2048 : 32, a_lbo = a_mn ? 8192 : 16;
        const uint32_t b_step = b_mn ? 2048 : 32, b_lbo = b_mn ? 8192 : 16;
        mbar_wait(&sm.acc_empty[as], accph ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + as * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&sm.full[st], ph);
          tc_fence_after();
          const uint32_t abase = smem_u32(sm.a[st]);
          const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t ad = make_smem_desc(abase + kk * a_step, a_lbo, 1024);
            const uint64_t bd = make_smem_desc(bbase + kk * b_step, b_lbo, 1024);
            umma_bf16(dcol, ad, bd, idesc, (kb | kk) != 0);
          }
          umma_commit(&sm.empty[st]);
          if (++st == NST) { st = 0; ph ^= 1; }
        }
        umma_commit(&sm.acc_full[as]);
        if (++as == NACC) { as = 0; accph ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;
    const float ad = P.alpha_dev ? __ldg(P.alpha_dev) : 1.f;
    uint32_t as = 0, accph = 0;
    int prob, mb, kb0;
    for (int k = 0; job_at(P, grp, k, &prob, &mb, &kb0); ++k) {
      const Problem& pr = P.p[prob];
      const float al = pr.alpha * ad;
      const int accumulate = pr.accumulate;
      const int grow = mb * BM + q * 32 + lane;
      float* crow;
      if (pr.scatter) {
        const int owner = (mb * BM) / pr.rows_per_rank;  // BM divides rows_per_rank
        crow = pr.peer[owner] + (int64_t)(grow - owner * pr.rows_per_rank) * pr.ldc + n0;
      } else {
        crow = pr.C + (int64_t)grow * pr.ldc + n0;
      }
      const int scatter = pr.scatter | (pr.ksplit > 1);  // both add with red.global
      mbar_wait(&sm.acc_full[as], accph);
      tc_fence_after();
      if (RS != 0 && pr.scatter) {
        // staged scatter: 128-byte row segments instead of per-lane 16-byte atomics
        float* stg = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(&sm) +
                                              ((sizeof(Smem) + 127) & ~size_t(127))) +
                     q * 32 * STG_LD;
        float* mine = stg + lane * STG_LD;
        // row 0 of this warp's 32 rows in the owner's buffer (all 32 rows share the owner)
        float* crow0 = crow - (int64_t)lane * pr.ldc;
#pragma unroll 1
        for (int ch = 0; ch < BN / 32; ++ch) {
          float v[32];
          tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + as * BN + ch * 32, v);
          if (RS == 1) tma_store_wait_read<0>();  // my previous bulk op has read my row
          else __syncwarp();                      // everyone has re-read the previous chunk
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(mine + 4 * j) = make_float4(
                v[4 * j] * al, v[4 * j + 1] * al, v[4 * j + 2] * al, v[4 * j + 3] * al);
          if (RS == 1) {
            fence_proxy_async_smem();
            asm volatile(
                "cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 128;" ::
                    "l"(crow + ch * 32), "r"(smem_u32(mine))
                : "memory");
            tma_store_commit();
          } else {
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int r = it * 4 + (lane >> 3), u = lane & 7;
              const float4 o = *reinterpret_cast<const float4*>(stg + r * STG_LD + 4 * u);
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(
                               crow0 + (int64_t)r * pr.ldc + ch * 32 + 4 * u),
                           "f"(o.x), "f"(o.y), "f"(o.z), "f"(o.w)
                           : "memory");
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
        if (++as == NACC) { as = 0; accph ^= 1; }
        continue;
      }
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        float v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + as * BN + ch * 32, v);
        float4* dst = reinterpret_cast<float4*>(crow + ch * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float4 o = make_float4(v[4 * j] * al, v[4 * j + 1] * al, v[4 * j + 2] * al,
                                 v[4 * j + 3] * al);
          if (scatter) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j),
                         "f"(o.x), "f"(o.y), "f"(o.z), "f"(o.w)
                         : "memory");
            continue;
          }
          if (accumulate) {
            const float4 c = dst[j];
            o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
          }
          dst[j] = o;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
      if (++as == NACC) { as = 0; accph ^= 1; }
    }
    if (RS == 1) {  // every bulk reduction of this thread has completed before the CTA exits
      tma_store_wait_all<0>();
      __threadfence_system();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

static int sm_count_g() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

bool gemm_eligible(int64_t Mg, int64_t Ng, int64_t K) {
  return Mg > 0 && Ng > 0 && K > 0 && Mg % g::BM == 0 && Ng % g::BN == 0 && K % g::BK == 0 &&
         Ng / g::BN <= 16 && Mg < (1ll << 31) && Ng < (1ll << 31) && K < (1ll << 31);
}

static int fill_problem(g::Problem* p, const GemmDesc& d, int64_t Ng) {
  int rc;
  const int64_t nb = d.batch > 1 ? d.batch : 1;
  if (nb > 1 && (!d.b_mn || d.peers)) return MM25_ERR_SHAPE;
  if (nb * d.Mg >= (1ll << 31) || nb * d.K >= (1ll << 31)) return MM25_ERR_SHAPE;
  if (d.a_mn) rc = make_tmap_bf16(&p->ta, d.A, (uint64_t)(nb * d.K), (uint64_t)d.Mg, 64, 64);
  else rc = make_tmap_bf16(&p->ta, d.A, (uint64_t)(nb * d.Mg), (uint64_t)d.K, g::BM, 64);
  if (rc) return rc;
  if (d.b_mn) rc = make_tmap_bf16(&p->tb, d.B, (uint64_t)(nb * d.K), (uint64_t)Ng, 64, 64);
  else rc = make_tmap_bf16(&p->tb, d.B, (uint64_t)Ng, (uint64_t)d.K, g::BN, 64);
  if (rc) return rc;
  p->C = d.C;
  p->ldc = d.ldc;
  p->nk = (int)(d.K / g::BK);
  p->mblocks = (int)(nb * d.Mg / g::BM);
  p->mb_per_batch = nb > 1 ? (int)(d.Mg / g::BM) : 0;
  p->a_k_rows = p->b_k_rows = nb > 1 ? (int)d.K : 0;
  p->ksplit = 1;
  p->a_mn = d.a_mn;
  p->b_mn = d.b_mn;
  p->accumulate = d.accumulate;
  p->alpha = d.alpha;
  p->scatter = 0;
  if (d.peers) {
    if (d.npeers < 1 || d.npeers > GEMM_MAX_PEERS || d.rows_per_rank <= 0 ||
        d.rows_per_rank % g::BM != 0 || d.rows_per_rank * d.npeers != d.Mg)
      return MM25_ERR_SHAPE;
    p->scatter = 1;
    p->rows_per_rank = (int)d.rows_per_rank;
    p->rot = (int)(((d.self_rank + 1) % d.npeers) * (d.rows_per_rank / g::BM));
    for (int i = 0; i < d.npeers; ++i) p->peer[i] = d.peers[i];
  }
  return MM25_OK;
}

// static schedule: problem 0 = the one with the longer K (costlier blocks), see file header
void gemm_schedule(int64_t mb0, int64_t nk0, int64_t mb1, int64_t nk1, int ngroups, int* r,
                   int* nbA, int* nbB, int* nsA, int* nsB) {
  const int64_t q = mb0 / ngroups, rem = mb0 % ngroups;
  *r = (int)rem;
  *nbA = (int)(q + (rem ? 1 : 0));
  *nbB = (int)q;
  if (rem == 0) { *r = ngroups; *nbA = (int)q; }
  *nsA = *nsB = 0;
  if (mb1 <= 0) return;
  // water-filling over the finishing level T (in k-blocks)
  const int64_t la = (int64_t)(*nbA) * nk0, lb = (int64_t)(*nbB) * nk0;
  const int64_t na = *r, nb = ngroups - *r;
  int64_t T = (mb0 * nk0 + mb1 * nk1 + ngroups - 1) / ngroups;
  if (T < la) T = la;
  for (;;) {
    const int64_t xa = T > la ? (T - la) / nk1 : 0, xb = T > lb ? (T - lb) / nk1 : 0;
    if (na * xa + nb * xb >= mb1) { *nsA = (int)xa; *nsB = (int)xb; return; }
    // next level at which some group can take one more block
    const int64_t ta = la + (xa + 1) * nk1, tb = lb + (xb + 1) * nk1;
    T = (nb > 0 && tb < ta) ? tb : ta;
  }
}

int gemm_grouped(const GemmDesc* d0, const GemmDesc* d1, int64_t Ng, const float* alpha_dev,
                 cudaStream_t st) {
  if (!d0 && !d1) return MM25_OK;
  // opt-in experiment: CTA-pair kernel (gemm_tc2.cu) for the problems it covers
  static const bool two_cta = getenv("MM25_GEMM_2CTA") != nullptr;
  if (two_cta) {
    const int rc2 = gemm_grouped2(d0, d1, Ng, alpha_dev, st);
    if (rc2 != MM25_ERR_SHAPE) return rc2;
  }
  if (!d0) { d0 = d1; d1 = nullptr; }
  if (d1 && d1->K > d0->K) { const GemmDesc* t = d0; d0 = d1; d1 = t; }
  if (!gemm_eligible(d0->Mg, Ng, d0->K)) return MM25_ERR_SHAPE;
  if (d1 && !gemm_eligible(d1->Mg, Ng, d1->K)) return MM25_ERR_SHAPE;
  g::Params P;
  memset(&P, 0, sizeof(P));
  int rc = fill_problem(&P.p[0], *d0, Ng);
  if (rc) return rc;
  if (d1) {
    rc = fill_problem(&P.p[1], *d1, Ng);
    if (rc) return rc;
  } else {
    P.p[1] = P.p[0];
    P.p[1].mblocks = 0;
  }
  P.alpha_dev = alpha_dev;
  const int group = (int)(Ng / g::BN);
  int ngroups = sm_count_g() / group;
  const int64_t total_mb = P.p[0].mblocks + (int64_t)P.p[1].mblocks;
  if (ngroups > total_mb) ngroups = (int)total_mb;
  if (ngroups < 1) ngroups = 1;
  P.s.group = group;
  P.s.ngroups = ngroups;
  // split-K for an under-filled long-K problem (e.g. the dA GEMM of a row shard: 32 m-blocks
  // of 512 k-blocks on 74 CTA groups): its jobs become (m-block, k-part) and the partial tiles
  // are added into a zeroed C with red.global — otherwise one long job bounds the makespan.
  if (!P.p[0].scatter && P.p[0].mblocks < ngroups && P.p[0].nk >= 128 && P.p[0].ldc == Ng) {
    int ks = 1;
    while (P.p[0].mblocks * ks < ngroups && ks < 8 && P.p[0].nk % (2 * ks) == 0 &&
           P.p[0].nk / (2 * ks) >= 32)
      ks *= 2;
    if (ks > 1) {
      P.p[0].ksplit = ks;
      P.p[0].nk /= ks;
      if (!P.p[0].accumulate)
        MM25_CUDA_CHECK(cudaMemsetAsync(P.p[0].C, 0,
                                        sizeof(float) * (size_t)P.p[0].mblocks * g::BM *
                                            (size_t)P.p[0].ldc, st));
    }
  }
  gemm_schedule((int64_t)P.p[0].mblocks * P.p[0].ksplit, P.p[0].nk,
                (int64_t)P.p[1].mblocks * P.p[1].ksplit, P.p[1].nk, ngroups, &P.s.r, &P.s.nbA,
                &P.s.nbB, &P.s.nsA, &P.s.nsB);
  // A/B switch: walking the scatter problem first measured no better at 8 GPUs (0.871 vs
  // 0.860 ms per step), so the default keeps problem 0 first
  static const bool p1_first = getenv("MM25_GEMM_P1_FIRST") != nullptr;
  P.s.p1_first = (P.p[1].mblocks > 0 && P.p[1].scatter && p1_first) ? 1 : 0;
  // staged scatter modes: opt-in A/B switch (see g::STG_LD); only when a problem scatters
  static const int rs_env = getenv("MM25_RS_MODE") ? atoi(getenv("MM25_RS_MODE")) : 0;
  const int rs = (P.p[0].scatter || (P.p[1].mblocks > 0 && P.p[1].scatter)) ? rs_env : 0;
  if (rs == 1 || rs == 2) {
    const size_t smem = ((sizeof(g::Smem) + 127) & ~size_t(127)) + g::STG_BYTES + 1024;
    auto kern = rs == 1 ? gemm_grouped_kernel<1> : gemm_grouped_kernel<2>;
    MM25_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem));
    kern<<<ngroups * group, g::THREADS, smem, st>>>(P);
    MM25_LAUNCH_CHECK();
    return MM25_OK;
  }
  const size_t smem = sizeof(g::Smem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(gemm_grouped_kernel<0>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  { void* tt_ = timing_begin(MM25_T_GEMM_GROUPED, st);
  gemm_grouped_kernel<0><<<ngroups * group, g::THREADS, smem, st>>>(P);
  timing_end(tt_, st); }
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

int gemm_bf16(const void* A, int a_mn, const void* Bm, int b_mn, float* C, int64_t ldc,
              int64_t Mg, int64_t Ng, int64_t K, float alpha, const float* alpha_dev,
              int accumulate, cudaStream_t st) {
  GemmDesc d{A, Bm, C, ldc, Mg, K, a_mn, b_mn, accumulate, alpha, nullptr, 0, 0};
  return gemm_grouped(&d, nullptr, Ng, alpha_dev, st);
}

}  // namespace tc
}  // namespace mm25

// debug / test entry: C = alpha * op(A) op(B) on bf16 device matrices (tests/test_tc_gpu.py)
extern "C" int mm25_debug_gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, float* C,
                                    int64_t Mg, int64_t Ng, int64_t K, float alpha,
                                    int accumulate, void* stream) {
  return mm25::tc::gemm_bf16(A, a_mn, B, b_mn, C, Ng, Mg, Ng, K, alpha, nullptr, accumulate,
                             (cudaStream_t)stream);
}

// debug / test entry: two problems sharing Ng in one launch (C0 = A0 op B0, C1 = A1 op B1)
extern "C" int mm25_debug_gemm_grouped(const void* A0, int a0_mn, const void* B0, int b0_mn,
                                       float* C0, int64_t M0, int64_t K0, const void* A1,
                                       int a1_mn, const void* B1, int b1_mn, float* C1,
                                       int64_t M1, int64_t K1, int64_t Ng, void* stream) {
  mm25::tc::GemmDesc d0{A0, B0, C0, Ng, M0, K0, a0_mn, b0_mn, 0, 1.f, nullptr, 0, 0};
  mm25::tc::GemmDesc d1{A1, B1, C1, Ng, M1, K1, a1_mn, b1_mn, 0, 1.f, nullptr, 0, 0};
  return mm25::tc::gemm_grouped(&d0, &d1, Ng, nullptr, (cudaStream_t)stream);
}

// debug / test entry: the scatter-mode problem (fused reduce-scatter epilogue) with the P owner
// buffers given as a HOST array of device pointers — on one GPU they are P local [R, Ng] fp32
// buffers, which exercises the same epilogue code (plain, TMA-reduce or staged per
// MM25_RS_MODE) without NVLink.  Adds op(A) op(B) rows [q R, (q+1) R) into owners[q].
extern "C" int mm25_debug_gemm_scatter(const void* A, int a_mn, const void* B, int b_mn,
                                       float* const* owners, int nowners, int64_t rows_per_rank,
                                       int self_rank, int64_t Mg, int64_t Ng, int64_t K,
                                       void* stream) {
  if (!owners || nowners < 1 || nowners > mm25::tc::GEMM_MAX_PEERS) return MM25_ERR_ARG;
  mm25::tc::GemmDesc d{A, B, nullptr, Ng, Mg, K, a_mn, b_mn, 0, 1.f, owners, nowners,
                       rows_per_rank, self_rank, 0};
  return mm25::tc::gemm_grouped(&d, nullptr, Ng, nullptr, (cudaStream_t)stream);
}

// host-only test hook: the static schedule of the grouped GEMM (no device work).
// out = {r, nbA, nbB, nsA, nsB}: groups [0, r) are class A (nbA round-robin blocks of problem 0,
// then a run of nsA blocks of problem 1), the others class B.
extern "C" int mm25_debug_gemm_schedule(int64_t mb0, int64_t nk0, int64_t mb1, int64_t nk1,
                                        int ngroups, int* out) {
  if (!out || ngroups < 1 || mb0 < 0 || mb1 < 0 || nk0 < 1 || (mb1 > 0 && nk1 < 1))
    return MM25_ERR_ARG;
  mm25::tc::gemm_schedule(mb0, nk0, mb1, nk1, ngroups, &out[0], &out[1], &out[2], &out[3],
                          &out[4]);
  return MM25_OK;
}
