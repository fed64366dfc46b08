// gemm_tc.cu — bf16 tcgen05 GEMM used by the tensor-core backward:
//     C[Mg, Ng] (fp32)  (+)=  alpha * Aop[Mg, K] * Bop[K, Ng]
// Operands are fed by TMA straight from their natural row-major storage, so each of them may
// be "K-major" (memory [rows = M/N index, cols = K]) or "MN-major" (memory [rows = K,
// cols = M/N index]); the UMMA descriptors carry the difference (no transposes in HBM).
//   dA = G  . Yhat   : A = G   K-major  [R, N],   B = Yhat MN-major [N, d]
//   dB = G^T. Xhat   : A = G   MN-major [R, N],   B = Xhat MN-major [R, d]
// Tile 128 x 256 x 64, 4-stage TMA ring, 2 TMEM accumulator stages, persistent CTAs.
#include "tc_common.cuh"
#include "gemm_tc.h"

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
}  // namespace g

template <int A_MN, int B_MN>
__global__ void __launch_bounds__(g::THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
               const __grid_constant__ CUtensorMap tmap_b, float* __restrict__ C, int64_t ldc,
               int Mg, int Ng, int K, float alpha, const float* __restrict__ alpha_dev,
               int accumulate) {
  using namespace g;
  extern __shared__ uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                      ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = Mg / BM, tiles_n = Ng / BN;
  const int ntiles = tiles_m * tiles_n;
  const int nk = K / BK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    for (int i = 0; i < NACC; ++i) { mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.acc_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&sm.tmem_base, 512);
    if (lane == 0) { tma_prefetch_desc(&tmap_a); tma_prefetch_desc(&tmap_b); }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        // n fastest: consecutive CTAs share the A rows (G is the big operand)
        const int m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
        for (int kb = 0; kb < nk; ++kb) {
          const int k0 = kb * BK;
          mbar_wait(&sm.empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&sm.full[st], A_BYTES + B_BYTES);
          if (A_MN) {
#pragma unroll
            for (int at = 0; at < BM / 64; ++at)
              tma_load_2d(sm.a[st] + at * 8192, &tmap_a, &sm.full[st], m0 + at * 64, k0);
          } else {
            tma_load_2d(sm.a[st], &tmap_a, &sm.full[st], k0, m0);
          }
          if (B_MN) {
#pragma unroll
            for (int at = 0; at < BN / 64; ++at)
              tma_load_2d(sm.b[st] + at * 8192, &tmap_b, &sm.full[st], n0 + at * 64, k0);
          } else {
            tma_load_2d(sm.b[st], &tmap_b, &sm.full[st], k0, n0);
          }
          if (++st == NST) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      uint32_t st = 0, ph = 0, as = 0, accph = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        mbar_wait(&sm.acc_empty[as], accph ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + as * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&sm.full[st], ph);
          tc_fence_after();
          const uint32_t abase = smem_u32(sm.a[st]);
          const uint32_t bbase = smem_u32(sm.b[st]);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: +32 B per 16 k inside the 128 B swizzle row; SBO = 8 rows.
            // MN-major: +16 k-rows (2048 B); LBO = next 64-wide MN atom, SBO = 8 k-rows.
            const uint64_t ad = A_MN ? make_smem_desc(abase + k * 2048, 8192, 1024)
                                     : make_smem_desc(abase + k * 32, 16, 1024);
            const uint64_t bd = B_MN ? make_smem_desc(bbase + k * 2048, 8192, 1024)
                                     : make_smem_desc(bbase + k * 32, 16, 1024);
            umma_bf16(dcol, ad, bd, idesc, (kb | k) != 0);
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
    const float al = alpha * (alpha_dev ? __ldg(alpha_dev) : 1.f);
    uint32_t as = 0, accph = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
      float* crow = C + (int64_t)(m0 + q * 32 + lane) * ldc + n0;
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
      if (lane == 0) mbar_arrive(&sm.acc_empty[as]);
      if (++as == NACC) { as = 0; accph ^= 1; }
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
         Mg < (1ll << 31) && Ng < (1ll << 31) && K < (1ll << 31);
}

template <int A_MN, int B_MN>
static int launch_t(const CUtensorMap& ta, const CUtensorMap& tb, float* C, int64_t ldc,
                    int64_t Mg, int64_t Ng, int64_t K, float alpha, const float* alpha_dev,
                    int accumulate, cudaStream_t st) {
  const size_t smem = sizeof(g::Smem) + 1024;
  MM25_CUDA_CHECK(cudaFuncSetAttribute(gemm_tc_kernel<A_MN, B_MN>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t ntiles = (Mg / g::BM) * (Ng / g::BN);
  const int grid = (int)(ntiles < sm_count_g() ? ntiles : sm_count_g());
  gemm_tc_kernel<A_MN, B_MN><<<grid, g::THREADS, smem, st>>>(
      ta, tb, C, ldc, (int)Mg, (int)Ng, (int)K, alpha, alpha_dev, accumulate);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

// a_rows/a_cols (b_rows/b_cols): the row-major bf16 matrix as stored in memory.
int gemm_bf16(const void* A, int a_mn, const void* Bm, int b_mn, float* C, int64_t ldc,
              int64_t Mg, int64_t Ng, int64_t K, float alpha, const float* alpha_dev,
              int accumulate, cudaStream_t st) {
  if (!gemm_eligible(Mg, Ng, K)) return MM25_ERR_SHAPE;
  CUtensorMap ta, tb;
  int rc;
  if (a_mn) rc = make_tmap_bf16(&ta, A, (uint64_t)K, (uint64_t)Mg, 64, 64);
  else rc = make_tmap_bf16(&ta, A, (uint64_t)Mg, (uint64_t)K, g::BM, 64);
  if (rc) return rc;
  if (b_mn) rc = make_tmap_bf16(&tb, Bm, (uint64_t)K, (uint64_t)Ng, 64, 64);
  else rc = make_tmap_bf16(&tb, Bm, (uint64_t)Ng, (uint64_t)K, g::BN, 64);
  if (rc) return rc;
  if (a_mn && b_mn) return launch_t<1, 1>(ta, tb, C, ldc, Mg, Ng, K, alpha, alpha_dev, accumulate, st);
  if (a_mn) return launch_t<1, 0>(ta, tb, C, ldc, Mg, Ng, K, alpha, alpha_dev, accumulate, st);
  if (b_mn) return launch_t<0, 1>(ta, tb, C, ldc, Mg, Ng, K, alpha, alpha_dev, accumulate, st);
  return launch_t<0, 0>(ta, tb, C, ldc, Mg, Ng, K, alpha, alpha_dev, accumulate, st);
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
