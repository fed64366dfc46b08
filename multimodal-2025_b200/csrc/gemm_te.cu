// gemm_te.cu — the backward GEMMs of the kept-E path, with the E -> G' transform fused into the
// A-operand load ("transform-on-load"):
//     C[Mg, Ng] (fp32) = T(E)op[Mg, K] * Bop[K, Ng],   T(E)_rc = E_rc (rowf_r + colf_c) - k_all,
//                                                     T(E)_rp = gdiag_r on the positive p of row r
// E [rows, cols] bf16 is what the forward kept (E = 2^(z log2e - m), sim_tc.cu); rowf / colf are
// the per-row / per-column factors gs w_row 2^(m - r_i log2e), gs w_col 2^(m - c_j log2e)
// (te_coef_kernel).  The two problems of the backward read E in its natural storage:
//     dA = T(E)   . Yhat  : A K-major  (memory rows = M index), B = Yhat MN-major [N, d]
//     dB = T(E)^T . Xhat  : A MN-major (memory rows = K index), B = Xhat MN-major [M, d]
// so neither the 2 N^2 d S-tile recompute nor the separate HBM pass E -> G' (4 B per logit)
// of the earlier kept-E path exists: E is read once per problem, by the kernel that consumes it.
//
// Structure: clusters of two CTAs (one TPC), `tcgen05.mma.cta_group::2` with M = 256.  A pair
// owns 256 rows x 512 columns of C: two N = 256 MMAs per k-step share one A tile, so with
// Ng = d = 512 every E element is loaded, transformed and staged exactly once per problem, and
// the 512 TMEM columns hold the one accumulator (the K loop is thousands of k-blocks long; the
// un-overlapped epilogue is < 1 % of a job).  Ng = 256 (mod 512) runs one MMA per k-step with
// two accumulator stages.  Warp roles per CTA (320 threads, 168 registers):
//   warp 0      B producer: TMA loads of this CTA's half of the B tile (MN-major 64 x 64 boxes),
//               bytes complete on the LEADER's `full` barrier (as gemm_tc2.cu);
//   warp 1      MMA issuer (leader CTA only);
//   warps 2-9   A producers, four groups of two warps, group g fills ring stage g: 16-byte global
//               loads of E one whole tile ahead (register ring), transform in fp32, bf16 pack,
//               st.shared into the SWIZZLE_128B image a TMA load of the same box would have
//               produced, fence.proxy.async, one arrive per warp on the leader's `full` barrier
//               (count 1 + 4);
//   warps 2-5   after the last k-block of a piece also run its epilogue (the one accumulator
//               cannot overlap the next piece anyway): TMEM -> registers -> fp32 C (plain
//               store, or red.global.add for a tile whose K range is split between two pairs /
//               for the fused reduce-scatter).
// Work: the (problem, batch entry, 256-row super-block, 512-column tile) jobs are flattened
// into one sequence of k-blocks that is cut into equal contiguous ranges, one per pair
// (stream-K): every pair executes the same number of k-blocks whatever the shapes, a job is
// split between at most two pairs when its K is not longer than a range, and split tiles are
// added with red.global into a zeroed C (0 + a + b: order-independent, deterministic).
#include "tc_common.cuh"
#include "gemm_tc.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace mm25 {
namespace tc {

namespace ge {
constexpr int BM = 128;        // rows per CTA (pair: 256)
constexpr int BK = 64;
constexpr int NST = 4;
constexpr int A_BYTES = BM * BK * 2;        // 16 KB
constexpr int BH_BYTES = 128 * BK * 2;      // 16 KB: this CTA's 128 columns of one 256-wide MMA
constexpr int B_BYTES = 2 * BH_BYTES;       // up to two MMAs per k-step
constexpr int NPROD = 8;                    // A-producer warps
constexpr int PROD0 = 2;                    // first A-producer warp; warps 2..5 also run the epilogue
constexpr int THREADS = 32 * (PROD0 + NPROD);   // 320 (168 registers per thread)
constexpr int STG_LD = 36;                  // floats per staged row (128 B + 16 B pad)
constexpr int STG_BYTES = 4 * 32 * STG_LD * 4;   // 4 epilogue warps x 32 rows
constexpr int MAX_PAIRS = 96;               // >= SMs / 2
constexpr int NFULL = 1 + 2 * (NPROD / NST);  // arrivals on the leader's `full` barrier: the TMA
                                            // thread + the two warps of the stage's group per CTA

struct Smem {
  uint8_t a[NST][A_BYTES];
  uint8_t b[NST][B_BYTES];
  uint64_t full[NST], empty[NST];
  uint64_t acc_full[2], acc_empty[2];
  uint32_t tmem_base;
};

struct Problem {
  CUtensorMap tb;          // B operand, MN-major: [batch * K rows, Ng cols], 64 x 64 boxes
  CUtensorMap te;          // E itself, only for L2 prefetches of the A tiles (128 x 64 / 64 x 64 boxes)
  const __nv_bfloat16* E;  // A source
  int64_t ldE;
  const float* rowf;       // [batch * e_rows]   factor of E's memory row
  const float* colf;       // [batch * e_cols]   factor of E's memory column
  const float* gdiag;      // [batch * e_rows]   exact fp32 G' of the positive of each E row (or NULL)
  float* C;
  int64_t ldc;
  int64_t e_rows, e_cols;  // one batch entry of E
  int64_t diag_offset;     // positive of local E row r sits at column r + diag_offset
  int a_mn;
  int nk;                  // k-blocks (of BK) per job
  int sblocks;             // 256-row super-blocks of C per batch entry (Mg / 256)
  int njobs;               // batch * sblocks * ntiles
  int kb_total;            // njobs * nk
  // fused reduce-scatter: super-block rows [q R, (q+1) R) are added into peer[q]
  int scatter, rows_per_rank, rot;
  float* peer[GEMM_MAX_PEERS];
};

struct Params {
  Problem p[2];
  int ntiles;              // column tiles of C (Ng / (256 NH))
  int nh;                  // 256-wide MMAs per k-step (1 or 2)
  int share;               // k-blocks per pair (uniform split; start[] is what the kernel reads)
  int start[MAX_PAIRS + 1];  // pair i executes k-blocks [start[i], start[i + 1]) of the sequence
  int total;               // p[0].kb_total + p[1].kb_total
  const int* e_flag;       // device flag: 0 = E already holds G' (recompute path ran) -> copy
  const float* consts;     // device {k_all, k_diag} per batch entry
  int debug_skip_a;        // measurement only (MM25_TE_DEBUG_SKIP_A): producers arrive without data
  int rs_bulk;             // scatter epilogue: 128-byte TMA reductions through a staging row
  int pf_dist;             // L2 prefetch distance of the A tiles in k-blocks (0 = off)
  long long* stats;        // measurement only: cycle sums of one producer thread (or NULL)
};

// One contiguous piece of a pair's range: k-blocks [kb0, kb1) of job `job` of problem `prob`.
struct Piece {
  int prob, bz, sb, nt, kb0, kb1, partial;
};

__device__ __forceinline__ bool next_piece(const Params& P, int& g, int g1, Piece* pc) {
  if (g >= g1) return false;
  const int prob = g < P.p[0].kb_total ? 0 : 1;
  const Problem& pr = P.p[prob];
  const int local = g - (prob ? P.p[0].kb_total : 0);
  const int job = local / pr.nk;
  pc->prob = prob;
  pc->kb0 = local - job * pr.nk;
  pc->kb1 = min(pr.nk, pc->kb0 + (g1 - g));
  pc->partial = !(pc->kb0 == 0 && pc->kb1 == pr.nk);
  pc->nt = job % P.ntiles;
  const int rest = job / P.ntiles;
  int sb = rest % pr.sblocks;
  if (pr.scatter) sb = (sb + pr.rot) % pr.sblocks;
  pc->sb = sb;
  pc->bz = rest / pr.sblocks;
  g += pc->kb1 - pc->kb0;
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
__device__ __forceinline__ uint32_t mapa_u32(const void* p, uint32_t rank) {
  uint32_t out;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(smem_u32(p)), "r"(rank));
  return out;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default semantics (.release.cta): `.release.cluster` compiles to MEMBAR.ALL.GPU + ERRBAR per
  // arrive (8 % of the kernel's stall samples, profiles/r2_gemm_te_notes.md); the data the
  // barrier guards is made visible to the tensor core by fence.proxy.async before the arrive
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait with cluster-scope acquire: the data guarded by the barrier was written by BOTH CTAs
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return;
    if (++spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m,
                                                 uint32_t bar_cluster, int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
// bring one box of a tensor into L2 (no shared-memory destination, no completion to wait for)
__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* m, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(c0), "r"(c1)
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
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// streaming 16-byte load of E: read once by this SM, never re-used from L1
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// G' of one 16-byte chunk (8 consecutive columns of one E row)
__device__ __forceinline__ uint4 transform8(const uint4 v, const float rf, const float (&cf)[8],
                                            const float k_all) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float e0 = __uint_as_float(w[k] << 16);
    const float e1 = __uint_as_float(w[k] & 0xffff0000u);
    o[k] = pack_bf16x2(fmaf(e0, rf + cf[2 * k], -k_all), fmaf(e1, rf + cf[2 * k + 1], -k_all));
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}
}  // namespace ge

// A producers of one piece.  The 8 producer warps form NST = 4 groups of two warps; group g owns
// ring stage g, i.e. the tiles t = g (mod 4) of the pair's running tile count, and fills each of
// them ALONE: thread `gtid` (0..63) owns the 16-byte chunks
//   (smem row sr = i * 8 + gtid / 8, chunk cc = gtid % 8),  i = 0..15
// of the 128 x 128-byte tile image (8 neighbouring threads read 128 contiguous bytes of one E
// row).  All 16 loads of a thread's NEXT tile are issued as soon as the current one is staged, so
// they have the rest of the group period (4 k-blocks of MMA time, ~2.5 us) to land, and the
// proxy fence + barrier arrive are paid once per tile and warp.
template <bool A_MN>
__device__ __forceinline__ void produce_piece(ge::Smem& sm, const ge::Problem& pr,
                                              const ge::Piece& pc, const int rank, const int grp,
                                              const int gtid, const bool xform, const float k_all,
                                              const uint32_t t0, const uint32_t full_leader0,
                                              long long* stats) {
  using namespace ge;
  const int r8 = gtid >> 3, cc = gtid & 7;
  const uint32_t dst0 = smem_u32(sm.a[grp]) + (uint32_t)(r8 * 128 + ((cc ^ r8) << 4));  // + i * 1024
  const int64_t erow0 = (int64_t)pc.bz * pr.e_rows;   // first E row of the batch entry
  const int64_t ccol0 = (int64_t)pc.bz * pr.e_cols;   // colf offset of the batch entry
  const int m0 = pc.sb * 256 + rank * BM;             // first C row (= M index) of this CTA
  const int64_t ldE = pr.ldE;
  const int nkb = pc.kb1 - pc.kb0;
  // first tile of this piece that belongs to the group: running tile t0 + j with (t0 + j) % 4 == grp
  int j = (int)((grp - t0) & 3u);
  if (j >= nkb) return;
  // E address of chunk (i, kb):
  //   K-major:  row m0 + i * 8 + r8,                     col kb * 64 + cc * 8
  //   MN-major: row kb * 64 + (i & 7) * 8 + r8,          col m0 + (i >> 3) * 64 + cc * 8
  const __nv_bfloat16* src0;
  int64_t step;                // elements per k-block
  const int64_t i_stride = 8 * ldE;
  // position of the positive: element `rel` of chunk i holds it when
  //   K-major:  rel = dg0 + i * 8 - kb * 64               is in [0, 8)
  //   MN-major: rel = dg0 + kb * 64 + (i & 7) * 8 - (i >> 3) * 64
  int dg0;
  const bool patch = xform && pr.gdiag != nullptr;
  if (!A_MN) {
    src0 = pr.E + (erow0 + m0 + r8) * ldE + cc * 8;
    step = BK;
    const int64_t t = (int64_t)m0 + r8 + pr.diag_offset - cc * 8;
    dg0 = (int)(t < -(1ll << 30) ? -(1ll << 30) : (t > (1ll << 30) ? (1ll << 30) : t));
  } else {
    src0 = pr.E + (erow0 + r8) * ldE + m0 + cc * 8;
    step = (int64_t)BK * ldE;
    const int64_t t = (int64_t)r8 + pr.diag_offset - m0 - cc * 8;
    dg0 = (int)(t < -(1ll << 30) ? -(1ll << 30) : (t > (1ll << 30) ? (1ll << 30) : t));
  }
  // factors that are fixed for the piece: K-major rows / MN-major columns
  float fix[16];
  if (xform) {
    if (!A_MN) {
#pragma unroll
      for (int i = 0; i < 16; ++i) fix[i] = __ldg(pr.rowf + erow0 + m0 + i * 8 + r8);
    } else {
#pragma unroll
      for (int at = 0; at < 2; ++at) {
        const float* cp = pr.colf + ccol0 + m0 + at * 64 + cc * 8;
        const float4 x = __ldg(reinterpret_cast<const float4*>(cp));
        const float4 y = __ldg(reinterpret_cast<const float4*>(cp + 4));
        fix[at * 8 + 0] = x.x; fix[at * 8 + 1] = x.y; fix[at * 8 + 2] = x.z; fix[at * 8 + 3] = x.w;
        fix[at * 8 + 4] = y.x; fix[at * 8 + 5] = y.y; fix[at * 8 + 6] = y.z; fix[at * 8 + 7] = y.w;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) fix[i] = 0.f;
  }
  auto src_of = [&](int i, int kb) -> const __nv_bfloat16* {
    const __nv_bfloat16* p = src0 + (int64_t)kb * step;
    if (!A_MN) return p + i * i_stride;
    return p + (i & 7) * i_stride + (i >> 3) * 64;
  };
  // Two phases per tile, never mixed: (1) transform + st.shared of the 16 chunks whose loads were
  // issued a whole group period earlier, (2) issue ALL loads of the group's next tile (16 chunks
  // + the 8 per-tile factors).  A rolling prefetch inside the transform loop serialises on the
  // load latency: the 6 scoreboard slots are counters, so the first use of any loaded register
  // also waits for the prefetches just issued on the same slot (measured: 3x slower).
  uint4 buf[16];
  float var[8];   // per-tile factors: K-major columns / MN-major rows (i and i + 8 share a row)
  auto issue = [&](int kb) {
#pragma unroll
    for (int i = 0; i < 16; ++i) buf[i] = ld_stream16(src_of(i, kb));
    if (xform) {
      if (!A_MN) {
        const float* cp = pr.colf + ccol0 + (int64_t)kb * BK + cc * 8;
        const float4 x = __ldg(reinterpret_cast<const float4*>(cp));
        const float4 y = __ldg(reinterpret_cast<const float4*>(cp + 4));
        var[0] = x.x; var[1] = x.y; var[2] = x.z; var[3] = x.w;
        var[4] = y.x; var[5] = y.y; var[6] = y.z; var[7] = y.w;
      } else {
        const float* rp = pr.rowf + erow0 + (int64_t)kb * BK + r8;
#pragma unroll
        for (int u = 0; u < 8; ++u) var[u] = __ldg(rp + u * 8);
      }
    }
  };
  issue(pc.kb0 + j);
  long long t_issue = stats ? clock64() : 0;
  for (; j < nkb; j += NST) {
    const int kb = pc.kb0 + j;
    if (stats) {  // [0] cycles from the end of the loads' issue until the tile's data is usable
      unsigned acc = 0;
#pragma unroll
      for (int i = 0; i < 16; ++i) acc |= buf[i].x ^ buf[i].w;
      if (acc == 0x9e3779b9u) stats[7] = 1;   // consumes every load
      const long long t = clock64();
      stats[0] += t - t_issue;
      t_issue = t;
    }
    // The positive of an E row is NOT formed from the bf16 E (on trained embeddings
    // E_ii (R_i + C_i) nearly cancels against k_diag, so E's rounding would dominate): its exact
    // fp32 G' comes from te_diag_kernel and is patched into the staged tile.  A thread holds at
    // most one positive per tile.
    float gd = 0.f;
    uint32_t gd_addr = 0;
    if (patch) {
      int i_hit = -1, rel = 0;
      int64_t row = 0;
      if (!A_MN) {
        const int t = kb * BK - dg0;           // need i * 8 - t in [0, 8)
        const int i = (t + 7) >> 3;
        if (i >= 0 && i < 16) { i_hit = i; rel = i * 8 - t; row = (int64_t)m0 + i * 8 + r8; }
      } else {
#pragma unroll
        for (int at = 0; at < 2; ++at) {
          const int v = dg0 + kb * BK - at * 64;   // need v + 8 b in [0, 8)
          const int bq = (7 - v) >> 3;
          if (bq >= 0 && bq < 8) { i_hit = at * 8 + bq; rel = v + 8 * bq; row = (int64_t)kb * BK + bq * 8 + r8; }
        }
      }
      if (i_hit >= 0) {
        gd = __ldg(pr.gdiag + erow0 + row);
        gd_addr = dst0 + (uint32_t)i_hit * 1024 + (uint32_t)rel * 2;
      }
    }
    // the MMAs that read this stage (tile t - 4) have completed
    const uint32_t use = (t0 + (uint32_t)j) >> 2;       // how many times the group used its stage
    mbar_wait(&sm.empty[grp], (use & 1u) ^ 1u);
    if (stats) {  // [1] further wait for the stage to be free
      const long long t = clock64();
      stats[1] += t - t_issue;
      t_issue = t;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      uint4 o = buf[i];
      if (xform) {
        if (!A_MN) {
          o = transform8(o, fix[i], var, k_all);
        } else {
          float cf[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) cf[u] = fix[(i >> 3) * 8 + u];
          o = transform8(o, var[i & 7], cf, k_all);
        }
      }
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst0 + i * 1024),
                   "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w)
                   : "memory");
    }
    if (j + NST < nkb) issue(kb + NST);   // registers are free: next tile's loads go out now
    if (gd_addr != 0) {
      const __nv_bfloat16 h = __float2bfloat16_rn(gd);
      asm volatile("st.shared.b16 [%0], %1;" ::"r"(gd_addr), "h"(*reinterpret_cast<const uint16_t*>(&h))
                   : "memory");
    }
    fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core (async proxy)
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive_cluster(full_leader0 + grp * 8);
    if (stats) {  // [2] transform + stores + next issue + fence + arrive; [3] tiles
      const long long t = clock64();
      stats[2] += t - t_issue;
      stats[3] += 1;
      t_issue = t;
    }
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(ge::THREADS, 1)
gemm_te_kernel(const __grid_constant__ ge::Params P) {
  using namespace ge;
  extern __shared__ uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                      ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int nh = P.nh, nacc = nh == 2 ? 1 : 2;
  const int g_begin = P.start[pair], g_end = P.start[pair + 1];

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.full[i], NFULL); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.acc_empty[i], 8); }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc_pair(&sm.tmem_base, 512);
    if (lane == 0) { tma_prefetch_desc(&P.p[0].tb); tma_prefetch_desc(&P.p[1].tb); }
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------------ B producer (TMA)
    // The same thread runs an L2 prefetch cursor PF_DIST tiles ahead of the loads over this
    // CTA's A tiles (E is streamed from HBM exactly once; the producer warps' 16-byte loads then
    // hit L2 instead of paying the HBM latency inside their short register ring).
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      Piece pc, pq;
      int gq = g_begin, kbq = 0;
      bool have_q = next_piece(P, gq, g_end, &pq);
      if (have_q) kbq = pq.kb0;
      auto prefetch_one = [&]() {
        if (!have_q) return;
        const Problem& prq = P.p[pq.prob];
        const int mq = pq.sb * 256 + (int)rank * BM;
        if (prq.a_mn) {
          const int kr = pq.bz * prq.nk * BK + kbq * BK;
          tma_prefetch_l2_2d(&prq.te, mq, kr);
          tma_prefetch_l2_2d(&prq.te, mq + 64, kr);
        } else {
          tma_prefetch_l2_2d(&prq.te, kbq * BK, pq.bz * prq.sblocks * 256 + mq);
        }
        if (++kbq == pq.kb1) {
          have_q = next_piece(P, gq, g_end, &pq);
          if (have_q) kbq = pq.kb0;
        }
      };
      for (int i = 0; i < P.pf_dist; ++i) prefetch_one();
      for (int g = g_begin; next_piece(P, g, g_end, &pc);) {
        const Problem& pr = P.p[pc.prob];
        const int ncol0 = pc.nt * 256 * nh + (int)rank * 128;
        const int krow0 = pc.bz * pr.nk * BK;
        for (int kb = pc.kb0; kb < pc.kb1; ++kb) {
          if (P.pf_dist > 0) prefetch_one();
          mbar_wait(&sm.empty[st], ph ^ 1);
          const uint32_t fbar = mapa_u32(&sm.full[st], 0);
          if (leader) mbar_arrive_expect_tx(&sm.full[st], 2 * nh * BH_BYTES);
          for (int h = 0; h < nh; ++h) {
#pragma unroll
            for (int at = 0; at < 2; ++at)
              tma_load_2d_pair(sm.b[st] + h * BH_BYTES + at * 8192, &pr.tb, fbar,
                               ncol0 + h * 256 + at * 64, krow0 + kb * BK);
          }
          if (++st == NST) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (leader && lane == 0) {
      uint32_t st = 0, ph = 0, as = 0, accph = 0;
      Piece pc;
      for (int g = g_begin; next_piece(P, g, g_end, &pc);) {
        const Problem& pr = P.p[pc.prob];
        const int a_mn = pr.a_mn;
        const uint32_t idesc = make_idesc_bf16(2 * BM, 256, a_mn, 1);
        const uint32_t a_step = a_mn ? 2048 : 32, a_lbo = a_mn ? 8192 : 16;
        mbar_wait(&sm.acc_empty[as], accph ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + as * 256;
        for (int kb = pc.kb0; kb < pc.kb1; ++kb) {
          // CTA-scope acquire: a cluster-scope one costs a CCTL.IVALL (L1 invalidate) per k-block;
          // the operands were published to the async proxy by fence.proxy.async / TMA
          mbar_wait(&sm.full[st], ph);   // TMA bytes + the stage's producer warps of the pair
          tc_fence_after();
          const uint32_t abase = smem_u32(sm.a[st]);
          const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t ad = make_smem_desc(abase + kk * a_step, a_lbo, 1024);
            for (int h = 0; h < nh; ++h) {
              const uint64_t bd = make_smem_desc(bbase + h * BH_BYTES + kk * 2048, 8192, 1024);
              umma_bf16_pair(dcol + h * 256, ad, bd, idesc, (kb > pc.kb0 || kk > 0) ? 1u : 0u);
            }
          }
          umma_commit_pair(&sm.empty[st]);
          if (++st == NST) { st = 0; ph ^= 1; }
        }
        umma_commit_pair(&sm.acc_full[as]);
        if (++as == (uint32_t)nacc) { as = 0; accph ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------ A producers (+ epilogue, warps 2-5)
    const int grp = (warp - PROD0) >> 1, gtid = ((warp - PROD0) & 1) * 32 + lane;
    const bool epi = warp < PROD0 + 4;
    const int q = warp & 3;
    const bool xform = P.e_flag ? (__ldg(P.e_flag) != 0) : true;
    const uint32_t full_leader0 = mapa_u32(&sm.full[0], 0);
    long long* st_ptr = (P.stats && blockIdx.x == 20 && threadIdx.x == PROD0 * 32) ? P.stats : nullptr;
    long long clk0 = 0;
    unsigned long long ns0 = 0;
    if (st_ptr) {
      clk0 = clock64();
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns0));
    }
    uint32_t t0 = 0, as = 0, accph = 0;   // t0: tiles (k-blocks) of the pieces before this one
    Piece pc;
    for (int g = g_begin; next_piece(P, g, g_end, &pc);) {
      const Problem& pr = P.p[pc.prob];
      const float k_all = (xform && P.consts) ? __ldg(P.consts + 2 * pc.bz) : 0.f;
      if (P.debug_skip_a) {
        const int nkb = pc.kb1 - pc.kb0;
        for (int j = (int)((grp - t0) & 3u); j < nkb; j += NST) {
          const uint32_t use = (t0 + (uint32_t)j) >> 2;
          mbar_wait(&sm.empty[grp], (use & 1u) ^ 1u);
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(full_leader0 + grp * 8);
        }
      } else if (pr.a_mn)
        produce_piece<true>(sm, pr, pc, (int)rank, grp, gtid, xform, k_all, t0, full_leader0, st_ptr);
      else
        produce_piece<false>(sm, pr, pc, (int)rank, grp, gtid, xform, k_all, t0, full_leader0, st_ptr);
      t0 += (uint32_t)(pc.kb1 - pc.kb0);
      if (!epi) continue;
      const int lrow = pc.sb * 256 + (int)rank * BM + q * 32 + lane;   // row inside the entry
      const int n0 = pc.nt * 256 * nh;
      float* crow;
      bool red = pc.partial != 0;
      if (pr.scatter) {
        const int owner = (pc.sb * 256) / pr.rows_per_rank;
        crow = pr.peer[owner] + (int64_t)(lrow - owner * pr.rows_per_rank) * pr.ldc + n0;
        red = true;
      } else {
        crow = pr.C + ((int64_t)pc.bz * pr.sblocks * 256 + lrow) * pr.ldc + n0;
      }
      mbar_wait(&sm.acc_full[as], accph);
      tc_fence_after();
      if (pr.scatter && P.rs_bulk) {
        // Fused reduce-scatter, staged: each lane parks the 32 floats of its row chunk in its
        // own padded shared-memory row and sends them as ONE 128-byte TMA reduction
        // (cp.reduce.async.bulk add.f32) into the owner's buffer over NVLink, instead of eight
        // separate 16-byte red.global per lane (measured at 8 ranks on the grouped GEMM:
        // 0.842 -> 0.693 ms per step, profiles/r2_ab_n8_summary.txt).
        float* mine = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(&sm) +
                                               ((sizeof(Smem) + 127) & ~size_t(127))) +
                      (q * 32 + lane) * STG_LD;
#pragma unroll 1
        for (int ch = 0; ch < 8 * nh; ++ch) {
          float v[32];
          tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + as * 256 + ch * 32, v);
          tma_store_wait_read<0>();   // my previous bulk reduction has read my row
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(mine + 4 * j) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          fence_proxy_async_smem();
          asm volatile(
              "cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 128;" ::
                  "l"(crow + ch * 32), "r"(smem_u32(mine))
              : "memory");
          tma_store_commit();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_u32(&sm.acc_empty[as], 0));
        if (++as == (uint32_t)nacc) { as = 0; accph ^= 1; }
        continue;
      }
#pragma unroll 1
      for (int ch = 0; ch < 8 * nh; ++ch) {
        float v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + as * 256 + ch * 32, v);
        float4* dst = reinterpret_cast<float4*>(crow + ch * 32);
        if (red) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j),
                         "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3])
                         : "memory");
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(&sm.acc_empty[as], 0));
      if (++as == (uint32_t)nacc) { as = 0; accph ^= 1; }
    }
    if (epi && P.rs_bulk) {  // every bulk reduction of this thread has completed before exit
      tma_store_wait_all<0>();
      __threadfence_system();
    }
    if (st_ptr) {  // [4] SM cycles and [5] nanoseconds of this thread's whole run -> effective clock
      unsigned long long ns1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
      st_ptr[4] += clock64() - clk0;
      st_ptr[5] += (long long)(ns1 - ns0);
    }
  }
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc_pair(tmem, 512);
}

// rowf[i] = gs w_row 2^(m2 - r_i log2e), colf[j] = gs w_col 2^(m2 - c_j log2e), consts = {gs w_all,
// gs w_diag} per entry; gs = gout s, m2 = dot_bound |s| log2e (the static shift of the forward).  One launch
// covers the rows and columns of every batch entry (grid.y = batch entry).
__global__ void __launch_bounds__(256)
te_coef_kernel(const float* __restrict__ row_lse, const float* __restrict__ col_lse, int64_t M,
               int64_t N, const float* __restrict__ scale, int scale_kind, float clamp_max,
               const float* __restrict__ gout, float w_row, float w_col, float w_diag,
               float w_all, float dot_bound, float* __restrict__ rowf, float* __restrict__ colf,
               float* __restrict__ consts) {
  constexpr float LOG2E = 1.4426950408889634f;
  const int64_t bz = blockIdx.y;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const float s = load_scale(scale, scale_kind, clamp_max);
  const float m2 = dot_bound * fabsf(s) * LOG2E;
  const float gs = (gout ? __ldg(gout + bz) : 1.f) * s;
  if (i < M) rowf[bz * M + i] = gs * w_row * exp2f(m2 - row_lse[bz * M + i] * LOG2E);
  if (i < N)
    colf[bz * N + i] = col_lse ? gs * w_col * exp2f(m2 - col_lse[bz * N + i] * LOG2E) : 0.f;
  if (i == 0) {
    consts[2 * bz] = gs * w_all;
    consts[2 * bz + 1] = gs * w_diag;
  }
}

// gdiag[i] = exact fp32 G' of the positive of row i:
//   gs (w_row e^{z_ii - r_i} + w_col e^{z_ii - c_p} - w_diag - w_all),  p = i + diag_offset,
// with z_ii = s a_i . b_p recomputed from the bf16 operands in fp32 (one warp per row).  Rows
// whose positive lies outside [0, N) get 0 and are never patched.
__global__ void __launch_bounds__(256)
te_diag_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
               int64_t M, int64_t N, int d, int64_t diag_offset,
               const float* __restrict__ row_lse, const float* __restrict__ col_lse,
               const float* __restrict__ scale, int scale_kind, float clamp_max,
               const float* __restrict__ gout, float w_row, float w_col, float w_diag,
               float w_all, float* __restrict__ gdiag) {
  const int64_t bz = blockIdx.y;
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= M) return;
  const int64_t p = i + diag_offset;
  if (p < 0 || p >= N) {
    if (lane == 0) gdiag[bz * M + i] = 0.f;
    return;
  }
  const __nv_bfloat16* ar = a + (bz * M + i) * d;
  const __nv_bfloat16* br = b + (bz * N + p) * d;
  float acc = 0.f;
  for (int k = lane * 8; k < d; k += 256) {
    const uint4 x = __ldg(reinterpret_cast<const uint4*>(ar + k));
    const uint4 y = __ldg(reinterpret_cast<const uint4*>(br + k));
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w}, yw[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      acc = fmaf(__uint_as_float(xw[u] << 16), __uint_as_float(yw[u] << 16), acc);
      acc = fmaf(__uint_as_float(xw[u] & 0xffff0000u), __uint_as_float(yw[u] & 0xffff0000u), acc);
    }
  }
  acc = warp_sum(acc);
  if (lane == 0) {
    const float s = load_scale(scale, scale_kind, clamp_max);
    const float gs = (gout ? __ldg(gout + bz) : 1.f) * s;
    const float z = s * acc;
    float g = w_row * expf(z - row_lse[bz * M + i]) - w_diag - w_all;
    if (col_lse) g += w_col * expf(z - col_lse[bz * N + p]);
    gdiag[bz * M + i] = gs * g;
  }
}

// measurement only: device buffer of 8 counters filled by one producer thread
static long long* g_te_stats = nullptr;

static int sm_count_te() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

bool gemm_te_eligible(int64_t Mg, int64_t Ng, int64_t K) {
  return Mg > 0 && Ng > 0 && K > 0 && Mg % 256 == 0 && Ng % 256 == 0 && K % ge::BK == 0 &&
         Mg < (1ll << 31) && Ng < (1ll << 31) && K < (1ll << 31);
}

static int fill_te(ge::Problem* p, const GemmTeDesc& d, int64_t Ng, int ntiles) {
  const int64_t nb = d.batch > 1 ? d.batch : 1;
  if (nb * d.K >= (1ll << 31) || nb * d.Mg >= (1ll << 31)) return MM25_ERR_SHAPE;
  int rc = make_tmap_bf16(&p->tb, d.B, (uint64_t)(nb * d.K), (uint64_t)Ng, 64, 64);
  if (rc) return rc;
  if (d.ldE != (d.a_mn ? d.Mg : d.K)) return MM25_ERR_SHAPE;   // E is dense
  if (d.a_mn) rc = make_tmap_bf16(&p->te, d.E, (uint64_t)(nb * d.K), (uint64_t)d.Mg, 64, 64);
  else rc = make_tmap_bf16(&p->te, d.E, (uint64_t)(nb * d.Mg), (uint64_t)d.K, 128, 64);
  if (rc) return rc;
  p->E = reinterpret_cast<const __nv_bfloat16*>(d.E);
  p->ldE = d.ldE;
  p->rowf = d.rowf;
  p->colf = d.colf;
  p->gdiag = d.gdiag;
  p->C = d.C;
  p->ldc = d.ldc;
  p->e_rows = d.a_mn ? d.K : d.Mg;
  p->e_cols = d.a_mn ? d.Mg : d.K;
  p->diag_offset = d.diag_offset;
  p->a_mn = d.a_mn;
  p->nk = (int)(d.K / ge::BK);
  p->sblocks = (int)(d.Mg / 256);
  const int64_t njobs = nb * p->sblocks * ntiles;
  if (njobs * p->nk >= (1ll << 30)) return MM25_ERR_SHAPE;
  p->njobs = (int)njobs;
  p->kb_total = (int)(njobs * p->nk);
  p->scatter = 0;
  p->rot = 0;
  p->rows_per_rank = 0;
  if (d.peers) {
    if (nb != 1 || d.npeers < 1 || d.npeers > GEMM_MAX_PEERS || d.rows_per_rank <= 0 ||
        d.rows_per_rank % 256 != 0 || d.rows_per_rank * d.npeers != d.Mg)
      return MM25_ERR_SHAPE;
    p->scatter = 1;
    p->rows_per_rank = (int)d.rows_per_rank;
    p->rot = (int)(((d.self_rank + 1) % d.npeers) * (d.rows_per_rank / 256));
    for (int i = 0; i < d.npeers; ++i) p->peer[i] = d.peers[i];
  }
  return MM25_OK;
}

// Up to two problems over the same kept E (either may be NULL).  C of a non-scatter problem must
// be zeroed by the caller when `*needs_zero` comes back 1 (split tiles are added with red.global);
// gemm_te_grouped zeroes it itself with cudaMemsetAsync when `zero_c` is set.
int gemm_te_grouped(const GemmTeDesc* d0, const GemmTeDesc* d1, int64_t Ng, const int* e_flag,
                    const float* consts, int zero_c, cudaStream_t st) {
  if (!d0 && !d1) return MM25_OK;
  if (!d0) { d0 = d1; d1 = nullptr; }
  if (!gemm_te_eligible(d0->Mg, Ng, d0->K)) return MM25_ERR_SHAPE;
  if (d1 && !gemm_te_eligible(d1->Mg, Ng, d1->K)) return MM25_ERR_SHAPE;
  ge::Params P;
  memset(&P, 0, sizeof(P));
  P.nh = (Ng % 512 == 0) ? 2 : 1;
  P.ntiles = (int)(Ng / (256 * P.nh));
  int rc = fill_te(&P.p[0], *d0, Ng, P.ntiles);
  if (rc) return rc;
  if (d1) {
    rc = fill_te(&P.p[1], *d1, Ng, P.ntiles);
    if (rc) return rc;
  } else {
    P.p[1] = P.p[0];
    P.p[1].njobs = 0;
    P.p[1].kb_total = 0;
  }
  P.e_flag = e_flag;
  P.consts = consts;
  static const bool skip_a = getenv("MM25_TE_DEBUG_SKIP_A") != nullptr;
  P.debug_skip_a = skip_a ? 1 : 0;
  static const int pf = getenv("MM25_TE_PF") ? atoi(getenv("MM25_TE_PF")) : 0;
  P.pf_dist = pf;
  P.stats = g_te_stats;
  static const bool rs_bulk_env = !(getenv("MM25_TE_RS") && atoi(getenv("MM25_TE_RS")) == 0);
  P.rs_bulk = (rs_bulk_env && (P.p[0].scatter || P.p[1].scatter)) ? 1 : 0;
  P.total = P.p[0].kb_total + P.p[1].kb_total;
  int maxpairs = sm_count_te() / 2;
  if (maxpairs > ge::MAX_PAIRS) maxpairs = ge::MAX_PAIRS;
  // equal contiguous ranges, not shorter than 8 k-blocks
  int share = (P.total + maxpairs - 1) / maxpairs;
  if (share < 8) share = 8;
  P.share = share;
  const int pairs = (P.total + share - 1) / share;
  for (int i = 0; i <= pairs; ++i) P.start[i] = (int)std::min<int64_t>((int64_t)i * share, P.total);
  // Fused reduce-scatter: the epilogue of a scatter job pushes its 256 x Ng fp32 tile over
  // NVLink with red.global and is NOT overlapped with MMAs (one accumulator), so a job of the
  // scatter problem costs nk + epi k-block times, not nk.  The cut points are placed at equal
  // COST instead of equal k-block counts (epi in k-block units: MM25_TE_EPI_KB, default 24 —
  // measured at 8 ranks, profiles/r2_n8_timeline.txt: ~41 us per epilogue, ~1 us per k-block).
  {
    static const int epi_env = getenv("MM25_TE_EPI_KB") ? atoi(getenv("MM25_TE_EPI_KB")) : 24;
    const int ep[2] = {P.p[0].scatter ? epi_env : 0, P.p[1].scatter ? epi_env : 0};
    if ((ep[0] > 0 || ep[1] > 0) && pairs > 1) {
      const int64_t c0 = (int64_t)P.p[0].njobs * (P.p[0].nk + ep[0]);
      const int64_t c1 = (int64_t)P.p[1].njobs * (P.p[1].nk + ep[1]);
      const int64_t ctot = c0 + c1;
      for (int i = 1; i < pairs; ++i) {
        int64_t c = ctot * i / pairs;
        const int pb = c < c0 ? 0 : 1;
        if (pb) c -= c0;
        const int64_t per = P.p[pb].nk + ep[pb];
        const int64_t job = c / per;
        int64_t kb = c - job * per;
        if (kb > P.p[pb].nk) kb = P.p[pb].nk;   // a cut inside the epilogue share: after the job
        int64_t g = (pb ? P.p[0].kb_total : 0) + job * P.p[pb].nk + kb;
        if (g < P.start[i - 1]) g = P.start[i - 1];
        P.start[i] = (int)std::min<int64_t>(g, P.total);
      }
      P.start[pairs] = P.total;
    }
  }
  if (zero_c) {
    const GemmTeDesc* ds[2] = {d0, d1};
    for (int i = 0; i < 2; ++i) {
      if (!ds[i] || ds[i]->peers) continue;
      const int64_t nb = ds[i]->batch > 1 ? ds[i]->batch : 1;
      if (ds[i]->ldc == Ng) {
        MM25_CUDA_CHECK(cudaMemsetAsync(ds[i]->C, 0, sizeof(float) * (size_t)(nb * ds[i]->Mg * Ng), st));
      } else {
        MM25_CUDA_CHECK(cudaMemset2DAsync(ds[i]->C, sizeof(float) * ds[i]->ldc, 0, sizeof(float) * Ng,
                                          (size_t)(nb * ds[i]->Mg), st));
      }
    }
  }
  const size_t smem = sizeof(ge::Smem) + 1024 + (P.rs_bulk ? ge::STG_BYTES + 128 : 0);
  MM25_CUDA_CHECK(cudaFuncSetAttribute(gemm_te_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
  { void* tt_ = timing_begin(MM25_T_GEMM_TE, st);
  gemm_te_kernel<<<2 * pairs, ge::THREADS, smem, st>>>(P);
  timing_end(tt_, st); }
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

int te_coef(const void* a, const void* b, int64_t d, int64_t diag_offset, const float* row_lse,
            const float* col_lse, int64_t B, int64_t M, int64_t N,
            const float* scale, int scale_kind, float clamp_max, const float* gout, float w_row,
            float w_col, float w_diag, float w_all, float dot_bound, float* rowf, float* colf,
            float* consts, float* gdiag, cudaStream_t st) {
  if (d % 8 != 0) return MM25_ERR_SHAPE;
  te_diag_kernel<<<dim3((unsigned)((M + 7) / 8), (unsigned)B), 256, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b), M, N,
      (int)d, diag_offset, row_lse, col_lse, scale, scale_kind, clamp_max, gout, w_row, w_col,
      w_diag, w_all, gdiag);
  MM25_LAUNCH_CHECK();
  const int64_t mx = M > N ? M : N;
  te_coef_kernel<<<dim3((unsigned)((mx + 255) / 256), (unsigned)B), 256, 0, st>>>(
      row_lse, col_lse, M, N, scale, scale_kind, clamp_max, gout, w_row, w_col, w_diag, w_all,
      dot_bound, rowf, colf, consts);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

}  // namespace tc
}  // namespace mm25

// measurement hook (tools/bench_te.py): allocate / read / clear the producer cycle counters
extern "C" int mm25_debug_te_stats(long long* out8, int enable) {
  using mm25::tc::g_te_stats;
  if (enable && !g_te_stats) {
    if (cudaMalloc(&g_te_stats, 8 * sizeof(long long)) != cudaSuccess) return MM25_ERR_CUDA;
    cudaMemset(g_te_stats, 0, 8 * sizeof(long long));
  }
  if (g_te_stats && out8) {
    cudaDeviceSynchronize();
    cudaMemcpy(out8, g_te_stats, 8 * sizeof(long long), cudaMemcpyDeviceToHost);
    cudaMemset(g_te_stats, 0, 8 * sizeof(long long));
  }
  if (!enable && g_te_stats) { cudaFree(g_te_stats); g_te_stats = nullptr; }
  return MM25_OK;
}

// debug / test entry (tests/test_tc_gpu.py): C0 = T(E) B0 and C1 = T(E)^T B1 for a bf16
// E [Me, Ne], with T(E)_rc = E_rc (rowf_r + colf_c) - consts[2 z], and T(E)_rc = gdiag_r where
// c == r + diag_offset (gdiag may be NULL: no special positive), when *e_flag != 0;
// T = identity otherwise.  B0 [Ne, Ng], B1 [Me, Ng]
// bf16 row-major; C0 [Me, Ng], C1 [Ne, Ng] fp32.  Either C may be NULL.
extern "C" int mm25_debug_gemm_te(const void* E, int64_t Me, int64_t Ne, const float* rowf,
                                  const float* colf, const float* consts, const float* gdiag,
                                  const int* e_flag, int64_t diag_offset, const void* B0,
                                  float* C0, const void* B1, float* C1, int64_t Ng, int64_t batch,
                                  void* stream) {
  if (!E || Me <= 0 || Ne <= 0 || Ng <= 0) return MM25_ERR_ARG;
  const int nb = batch > 1 ? (int)batch : 1;
  mm25::tc::GemmTeDesc a{E, Ne, B0, C0, Ng, Me, Ne, 0, nb, rowf, colf, gdiag, diag_offset, nullptr, 0, 0, 0};
  mm25::tc::GemmTeDesc b{E, Ne, B1, C1, Ng, Ne, Me, 1, nb, rowf, colf, gdiag, diag_offset, nullptr, 0, 0, 0};
  return mm25::tc::gemm_te_grouped(C0 ? &a : nullptr, C1 ? &b : nullptr, Ng, e_flag, consts, 1,
                                   (cudaStream_t)stream);
}
