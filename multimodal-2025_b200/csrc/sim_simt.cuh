// sim_simt.cuh — fp32 (FFMA) tile machinery of the fused similarity kernels.
// This is the precision path (fp32 contract: rel 1e-5 on the loss, SURVEY §7 hard part 1) and
// the general-shape path; the bf16 tcgen05 kernels live in sim_tc.cu.
#pragma once
#include "common.cuh"

namespace mm25 {
namespace simt {

constexpr int BM = 64;   // rows of a per CTA tile
constexpr int BN = 64;   // rows of b (= logit columns) per tile
constexpr int BK = 16;   // contraction chunk
constexpr int NT = 256;  // threads: 16 x 16, each a 4 x 4 micro tile

// load a [ROWS x BK] chunk of a row-major [nrows, d] matrix into k-major smem dst[BK][ROWS]
template <typename T, int ROWS>
__device__ __forceinline__ void load_chunk(float (*dst)[ROWS], const T* __restrict__ src,
                                           int64_t row0, int64_t nrows, int k0, int d) {
  static_assert(ROWS * BK == NT * 4, "one 4-vector per thread");
  const int m = threadIdx.x % ROWS;
  const int kq = (threadIdx.x / ROWS) * 4;
  const int64_t r = row0 + m;
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (r < nrows) {
    const T* p = src + r * d + k0 + kq;
    if (k0 + kq + 3 < d && ((d & 3) == 0)) {
      if constexpr (sizeof(T) == 4) {
        float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
      } else {
        uint2 t = *reinterpret_cast<const uint2*>(p);
        __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&t.x);
        __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&t.y);
        v[0] = __low2float(lo); v[1] = __high2float(lo);
        v[2] = __low2float(hi); v[3] = __high2float(hi);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (k0 + kq + e < d) v[e] = to_f32<T>(p[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < 4; ++e) dst[kq + e][m] = v[e];
}

// acc[4][4] += A_tile . B_tile^T over the full contraction dimension d
template <typename T>
__device__ __forceinline__ void tile_dot(float (&acc)[4][4], float (*As)[BM], float (*Bs)[BN],
                                         const T* __restrict__ a, const T* __restrict__ b,
                                         int64_t row0, int64_t M, int64_t col0, int64_t N,
                                         int d) {
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int k0 = 0; k0 < d; k0 += BK) {
    __syncthreads();
    load_chunk<T, BM>(As, a, row0, M, k0, d);
    load_chunk<T, BN>(Bs, b, col0, N, k0, d);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float ar[4] = {av.x, av.y, av.z, av.w};
      const float br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
    }
  }
}

// logits of one micro tile: z (for the LSE) and zraw (un-masked diagonal value), plus the
// chain factor dz/d(a.b) / s  (1 for PLAIN / MASKDIAG, 1 - u.v for MODULATED).
struct TileCtx {
  float s;
  int mode;
  int64_t diag_offset;
};

}  // namespace simt
}  // namespace mm25
