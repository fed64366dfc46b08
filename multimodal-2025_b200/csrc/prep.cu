// prep.cu — HBM-bound prologue / epilogue kernels of the loss path:
//   row L2-normalise (+ deterministic column sums), its backward, statistic merges and the
//   closed-form CLIP loss value.  One warp per embedding row, 16-byte vector accesses.
#include "common.cuh"

#include <atomic>

namespace mm25 {

static thread_local cudaError_t g_last_err = cudaSuccess;
void set_last_error(cudaError_t e) { g_last_err = e; }
static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---------------------------------------------------------------------------------------
// row normalise: x [rows, d] -> xhat (TO), inv_norm.  Replaces F.normalize (clipLoss.py:95-96)
// ---------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) rownorm_fwd_kernel(const TI* __restrict__ x,
                                                          TO* __restrict__ xhat,
                                                          float* __restrict__ inv_norm,
                                                          int64_t rows, int d, int normalize,
                                                          float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const TI* xr = x + row * d;
  float inv = 1.f;
  if (normalize) {
    float ss = 0.f;
    for (int k = lane; k < d; k += 32) {
      float v = to_f32<TI>(xr[k]);
      ss = fmaf(v, v, ss);
    }
    ss = warp_sum(ss);
    inv = 1.f / fmaxf(sqrtf(ss), eps);
  }
  TO* o = xhat + row * d;
  for (int k = lane; k < d; k += 32) o[k] = from_f32<TO>(to_f32<TI>(xr[k]) * inv);
  if (lane == 0 && inv_norm) inv_norm[row] = inv;
}

// deterministic column sums: block = 32 columns x 8 row lanes, fixed-order reduction
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x,
                                                     float* __restrict__ out, int64_t M, int d) {
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rl = threadIdx.x >> 5;
  const T* xb = x + (int64_t)blockIdx.y * M * d;
  float acc = 0.f;
  if (c < d)
    for (int64_t r = rl; r < M; r += 8) acc += to_f32<T>(xb[r * d + c]);
  red[rl][threadIdx.x & 31] = acc;
  __syncthreads();
  if (rl == 0 && c < d) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][threadIdx.x];
    out[(int64_t)blockIdx.y * d + c] = s;
  }
}

// dx = inv * (g - xhat (xhat . g))   (backward of F.normalize away from the eps clamp)
template <typename TH, typename TO>
__global__ void __launch_bounds__(256) rownorm_bwd_kernel(const float* __restrict__ g,
                                                          const TH* __restrict__ xhat,
                                                          const float* __restrict__ inv_norm,
                                                          TO* __restrict__ dx, int64_t rows,
                                                          int d, int normalize,
                                                          const float* __restrict__ gscale,
                                                          int accumulate) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float gs = gscale ? __ldg(gscale) : 1.f;
  const float* gr = g + row * d;
  const TH* hr = xhat + row * d;
  TO* o = dx + row * d;
  float dot = 0.f, inv = gs;
  if (normalize) {
    for (int k = lane; k < d; k += 32) dot = fmaf(gr[k], to_f32<TH>(hr[k]), dot);
    dot = warp_sum(dot);
    inv = inv_norm[row] * gs;
  }
  for (int k = lane; k < d; k += 32) {
    float v = normalize ? (gr[k] - to_f32<TH>(hr[k]) * dot) * inv : gr[k] * inv;
    if (accumulate) v += to_f32<TO>(o[k]);
    o[k] = from_f32<TO>(v);
  }
}

__global__ void __launch_bounds__(256) sum_parts_kernel(const float* __restrict__ p, int64_t n,
                                                        float alpha, float* __restrict__ out) {
  __shared__ float red[256];
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += 256) acc += p[i];
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = alpha * red[0];
}

__global__ void __launch_bounds__(256) col_merge_kernel(const float* __restrict__ cmax,
                                                        const float* __restrict__ csum,
                                                        int64_t P, int64_t N,
                                                        float* __restrict__ col_lse) {
  int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (j >= N) return;
  float m = -INFINITY, l = 0.f;
  for (int64_t p = 0; p < P; ++p) lse_merge(m, l, cmax[p * N + j], csum[p * N + j]);
  col_lse[j] = m + logf(l);
}

// one CTA per batch entry; fixed-order tree reduction -> deterministic
__global__ void __launch_bounds__(256) clip_finalize_kernel(
    const float* __restrict__ row_lse, const float* __restrict__ diag,
    const float* __restrict__ col_lse, const float* __restrict__ sx,
    const float* __restrict__ sy, int64_t M, int64_t N, int d, int64_t n_norm,
    const float* __restrict__ scale, int scale_kind, float clamp_max, float eps_ls,
    float* __restrict__ loss) {
  __shared__ double red[256];
  const int b = blockIdx.x;
  row_lse += (int64_t)b * M;
  diag += (int64_t)b * M;
  double acc = 0.0;
  const double nn = (double)n_norm;
  for (int64_t i = threadIdx.x; i < M; i += 256)
    acc += (double)row_lse[i] / (2.0 * nn) - (1.0 - (double)eps_ls) / nn * (double)diag[i];
  if (col_lse) {
    col_lse += (int64_t)b * N;
    for (int64_t j = threadIdx.x; j < N; j += 256) acc += (double)col_lse[j] / (2.0 * nn);
  }
  if (eps_ls != 0.f && sx && sy) {
    const float s = load_scale(scale, scale_kind, clamp_max);
    double dot = 0.0;
    for (int k = threadIdx.x; k < d; k += 256)
      dot += (double)sx[(int64_t)b * d + k] * (double)sy[(int64_t)b * d + k];
    acc -= (double)eps_ls / (nn * nn) * (double)s * dot;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[b] = (float)red[0];
}

}  // namespace mm25

using namespace mm25;

extern "C" const char* mm25_version(void) { return "mm25 0.1 (sm_100a)"; }
extern "C" const char* mm25_last_cuda_error(void) { return cudaGetErrorString(g_last_err); }
extern "C" long long mm25_launch_count(void) { return g_launches.load(); }
extern "C" int mm25_device_ok(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess)
    return 0;
  return major == 10;
}

extern "C" int mm25_rownorm_fwd(const void* x, int dtype, void* xhat, int out_dtype,
                                float* inv_norm, float* colsum, int64_t B, int64_t M, int64_t d,
                                int normalize, float eps, void* stream) {
  if (B <= 0 || M <= 0 || d <= 0 || d > (1 << 20)) return MM25_ERR_SHAPE;
  if (!x || !xhat) return MM25_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = B * M;
  dim3 grid((unsigned)ceil_div(rows, 8));
#define RN(TI, TO)                                                                         \
  rownorm_fwd_kernel<TI, TO><<<grid, 256, 0, st>>>((const TI*)x, (TO*)xhat, inv_norm, rows, \
                                                   (int)d, normalize, eps)
  if (dtype == MM25_F32 && out_dtype == MM25_F32) RN(float, float);
  else if (dtype == MM25_F32 && out_dtype == MM25_BF16) RN(float, __nv_bfloat16);
  else if (dtype == MM25_BF16 && out_dtype == MM25_BF16) RN(__nv_bfloat16, __nv_bfloat16);
  else if (dtype == MM25_BF16 && out_dtype == MM25_F32) RN(__nv_bfloat16, float);
  else return MM25_ERR_DTYPE;
#undef RN
  MM25_LAUNCH_CHECK();
  if (colsum) {
    dim3 g2((unsigned)ceil_div(d, 32), (unsigned)B);
    if (out_dtype == MM25_F32)
      colsum_kernel<float><<<g2, 256, 0, st>>>((const float*)xhat, colsum, M, (int)d);
    else
      colsum_kernel<__nv_bfloat16><<<g2, 256, 0, st>>>((const __nv_bfloat16*)xhat, colsum, M,
                                                      (int)d);
    MM25_LAUNCH_CHECK();
  }
  return MM25_OK;
}

extern "C" int mm25_rownorm_bwd(const float* g, const void* xhat, int xhat_dtype,
                                const float* inv_norm, void* dx, int out_dtype, int64_t B,
                                int64_t M, int64_t d, int normalize, const float* gscale,
                                int accumulate, void* stream) {
  if (B <= 0 || M <= 0 || d <= 0) return MM25_ERR_SHAPE;
  if (!g || !dx || (normalize && (!xhat || !inv_norm))) return MM25_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = B * M;
  dim3 grid((unsigned)ceil_div(rows, 8));
#define RB(TH, TO)                                                                            \
  rownorm_bwd_kernel<TH, TO><<<grid, 256, 0, st>>>(g, (const TH*)xhat, inv_norm, (TO*)dx, rows, \
                                                   (int)d, normalize, gscale, accumulate)
  if (xhat_dtype == MM25_F32 && out_dtype == MM25_F32) RB(float, float);
  else if (xhat_dtype == MM25_F32 && out_dtype == MM25_BF16) RB(float, __nv_bfloat16);
  else if (xhat_dtype == MM25_BF16 && out_dtype == MM25_BF16) RB(__nv_bfloat16, __nv_bfloat16);
  else if (xhat_dtype == MM25_BF16 && out_dtype == MM25_F32) RB(__nv_bfloat16, float);
  else return MM25_ERR_DTYPE;
#undef RB
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_sum_parts(const float* parts, int64_t n, float alpha, float* out,
                              void* stream) {
  if (!parts || !out || n < 0) return MM25_ERR_ARG;
  sum_parts_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(parts, n, alpha, out);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_col_merge(const float* col_max, const float* col_sum, int64_t P, int64_t N,
                              float* col_lse, void* stream) {
  if (!col_max || !col_sum || !col_lse || P <= 0 || N <= 0) return MM25_ERR_ARG;
  col_merge_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, (cudaStream_t)stream>>>(
      col_max, col_sum, P, N, col_lse);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_clip_loss_finalize(const float* row_lse, const float* diag,
                                       const float* col_lse, const float* sx, const float* sy,
                                       int64_t B, int64_t M, int64_t N, int64_t d,
                                       int64_t n_norm, const float* scale, int scale_kind,
                                       float clamp_max, float eps_ls, float* loss,
                                       void* stream) {
  if (!row_lse || !diag || !loss || B <= 0 || M <= 0) return MM25_ERR_ARG;
  if (eps_ls != 0.f && (!sx || !sy || !scale)) return MM25_ERR_ARG;
  clip_finalize_kernel<<<(unsigned)B, 256, 0, (cudaStream_t)stream>>>(
      row_lse, diag, col_lse, sx, sy, M, N, (int)d, n_norm, scale, scale_kind, clamp_max,
      eps_ls, loss);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
