// beta_vmf.cu — scalar (per sample) pieces of the spherical samplers:
//   * Wood's rejection sampler for the vMF marginal w0  (VonMisesFisher.py:184-222)
//   * Beta(alpha, beta) draws for the PowerSpherical marginal t = 2 z - 1
//   * the implicit-reparameterisation gradient dz/dalpha of a Beta draw, restating PyTorch's
//     dirichlet_grad_one (ATen/native/Distributions.h) that torch.distributions.Beta.rsample's
//     backward evaluates — the reference's kappa-gradient goes through it.
// One thread per (sample, row); the work is O(S B), negligible next to the [S, B, d] transform.
#include "common.cuh"
#include "philox.cuh"

namespace mm25 {

constexpr uint64_t STREAM_VMF = 0x764d46;
constexpr uint64_t STREAM_BETA = 0x426574;

// sequential draws from one Philox stream
struct Draws {
  uint64_t seed, idx, stream;
  uint32_t buf[4];
  int pos;
  __device__ Draws(uint64_t s, uint64_t i, uint64_t st) : seed(s), idx(i << 20), stream(st), pos(4) {}
  __device__ uint32_t next() {
    if (pos == 4) { rng::philox4(seed, idx++, stream, buf); pos = 0; }
    return buf[pos++];
  }
  __device__ float uniform() { return rng::u01(next()); }
  __device__ float normal() {
    const float u1 = uniform(), u2 = uniform();
    return sqrtf(-2.f * logf(u1)) * cosf(6.283185307179586f * u2);
  }
};

// Marsaglia-Tsang gamma(a, 1), a >= 1 (the marginals here have a >= (d-1)/2 >= 0.5; a < 1 uses
// the boost gamma(a+1) u^(1/a))
__device__ float sample_gamma(Draws& r, float a) {
  float boost = 1.f;
  if (a < 1.f) { boost = powf(r.uniform(), 1.f / a); a += 1.f; }
  const float dd = a - 1.f / 3.f, c = 1.f / sqrtf(9.f * dd);
  for (int it = 0; it < 256; ++it) {
    float x, v;
    do { x = r.normal(); v = 1.f + c * x; } while (v <= 0.f);
    v = v * v * v;
    const float u = r.uniform();
    if (u < 1.f - 0.0331f * x * x * x * x) return boost * dd * v;
    if (logf(u) < 0.5f * x * x + dd * (1.f - v + logf(v))) return boost * dd * v;
  }
  return boost * dd;
}

__device__ float sample_beta(Draws& r, float a, float b) {
  const float ga = sample_gamma(r, a), gb = sample_gamma(r, b);
  return ga / (ga + gb);
}

// ------------------------------------------------------------------------------------------
// vMF marginal.  Constants in fp64 then rounded to fp32, as numpy scalars meet fp32 tensors in
// the reference (:201-204, 217-219).  accept iff (d-1) log t0 - t0 + dd - log u >= 0.
// Injected mode: per row b, proposals prop_eps[b, 0:P], prop_u[b, 0:P] are scanned in order and
// the first S accepted become w[0:S, b] (what the reference does with its first batch of
// 10 S proposals); count[b] = number accepted (< S -> the remaining w are NaN).
// ------------------------------------------------------------------------------------------
struct WoodConst { float c1, c2, c3, dd, dm1; };

__device__ WoodConst wood_const(double kappa, int d) {
  const double dm1 = (double)(d - 1);
  const double kmr = sqrt(4.0 * kappa * kappa + dm1 * dm1);
  const double bb = (kmr - 2.0 * kappa) / dm1;
  const double aa = (kmr + 2.0 * kappa + dm1) / 4.0;
  const double dd = (4.0 * aa * bb) / (1.0 + bb) - dm1 * log(dm1);
  WoodConst w;
  w.c1 = (float)(1.0 + bb); w.c2 = (float)(1.0 - bb); w.c3 = (float)(2.0 * aa * bb);
  w.dd = (float)dd; w.dm1 = (float)dm1;
  return w;
}

__device__ __forceinline__ bool wood_accept(const WoodConst& w, float eps, float u, float* w0) {
  const float den = 1.f - w.c2 * eps;
  *w0 = (1.f - w.c1 * eps) / den;
  const float t0 = w.c3 / den;
  const float det = w.dm1 * logf(t0) - t0 + w.dd - logf(u);
  return det >= 0.f;
}

__global__ void __launch_bounds__(128)
vmf_w_injected_kernel(const float* __restrict__ kappa, int d, int64_t S, int64_t B,
                      const float* __restrict__ prop_eps, const float* __restrict__ prop_u,
                      int64_t P, float* __restrict__ w_out, int* __restrict__ count) {
  const int64_t b = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (b >= B) return;
  const WoodConst w = wood_const((double)kappa[b], d);
  int64_t n = 0;
  for (int64_t p = 0; p < P && n < S; ++p) {
    float w0;
    if (wood_accept(w, prop_eps[b * P + p], prop_u[b * P + p], &w0)) w_out[n++ * B + b] = w0;
  }
  if (count) count[b] = (int)n;
  for (; n < S; ++n) w_out[n * B + b] = NAN;
}

__global__ void __launch_bounds__(128)
vmf_w_philox_kernel(const float* __restrict__ kappa, int d, int64_t S, int64_t B, uint64_t seed,
                    float* __restrict__ w_out) {
  const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= S * B) return;
  const int64_t b = i % B;
  const WoodConst w = wood_const((double)kappa[b], d);
  Draws r(seed, (uint64_t)i, STREAM_VMF);
  const float a = 0.5f * (float)(d - 1);
  float w0 = 0.f;
  for (int it = 0; it < 4096; ++it) {
    const float eps = sample_beta(r, a, a);
    const float u = r.uniform();
    if (wood_accept(w, eps, u, &w0)) break;
  }
  w_out[i] = w0;
}

__global__ void __launch_bounds__(128)
beta_sample_kernel(const float* __restrict__ alpha, const float* __restrict__ beta, int64_t S,
                   int64_t B, uint64_t seed, float* __restrict__ z) {
  const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= S * B) return;
  const int64_t b = i % B;
  Draws r(seed, (uint64_t)i, STREAM_BETA);
  z[i] = sample_beta(r, alpha[b], beta[b]);
}

// ------------------------------------------------------------------------------------------
// dz/dalpha of z ~ Beta(alpha, beta) by implicit reparameterisation:
//   grad = (1 - z) * dirichlet_grad_one(z, alpha, alpha + beta)
// (torch _Dirichlet_backward with the 2-component concentration (alpha, beta) and an upstream
// gradient on the first component only).  Restatement of ATen/native/Distributions.h with
// scalar_t = float, accscalar_t = double (the CPU instantiation the reference runs).
// ------------------------------------------------------------------------------------------
__device__ double polevl6(double z, const double* A) {
  double r = A[0];
  for (int i = 1; i <= 6; ++i) r = r * z + A[i];
  return r;
}

template <typename T> __device__ T digamma_one(T x) {
  const double PSI_10 = 2.25175258906672110764;
  if (x == 0) return INFINITY;
  double additional = 0;
  if (x < 0) {
    if (x == floor(x)) return INFINITY;
    additional = -M_PI / tan(M_PI * (double)x);
    x = 1 - x;
  }
  double result = 0;
  while (x < 10) { result -= 1 / (double)x; x += 1; }
  if (x == 10) return (T)(result + PSI_10 + additional);
  const double A[] = {8.33333333333333333333E-2, -2.10927960927960927961E-2,
                      7.57575757575757575758E-3, -4.16666666666666666667E-3,
                      3.96825396825396825397E-3, -8.33333333333333333333E-3,
                      8.33333333333333333333E-2};
  double y = 0;
  if (x < 1.0e17f) {
    const double z = 1.0 / ((double)x * (double)x);
    y = z * polevl6(z, A);
  }
  return (T)(result + log((double)x) - (0.5 / (double)x) - y + additional);
}

__device__ float beta_grad_alpha_small(float x, float alpha, float beta) {
  const float factor = digamma_one<float>(alpha) - digamma_one<float>(alpha + beta) - logf(x);
  float numer = 1;
  float series = numer / alpha * (factor + 1 / alpha);
  for (int i = 1; i <= 10; ++i) {
    const float ci = (float)i;
    numer *= (ci - beta) * x / ci;
    const float denom = alpha + ci;
    series += numer / denom * (factor + 1 / denom);
  }
  const float result = x * powf(1 - x, -beta) * series;
  return isnan(result) ? 0.f : result;
}

__device__ float beta_grad_beta_small(float x, float alpha, float beta) {
  const float factor = digamma_one<float>(alpha + beta) - digamma_one<float>(beta);
  float numer = 1, betas = 1, dbetas = 0, series = factor / alpha;
  for (int i = 1; i <= 8; ++i) {
    const float ci = (float)i;
    numer *= -x / ci;
    dbetas = dbetas * (beta - ci) + betas;
    betas = betas * (beta - ci);
    series += numer / (alpha + ci) * (dbetas + factor * betas);
  }
  const float result = -powf(1 - x, 1 - beta) * series;
  return isnan(result) ? 0.f : result;
}

__device__ float beta_grad_alpha_mid(double x, double alpha, double beta) {
  const double total = alpha + beta;
  const double mean = alpha / total;
  const double std = sqrt(alpha * beta / (total + 1)) / total;
  if (mean - 0.1 * std <= x && x <= mean + 0.1 * std) {
    const double poly = 47 * x * (beta * beta) * (beta * beta) + alpha * (
                        (43 + 20 * (16 + 27 * beta) * x) * (beta * beta) * beta + alpha * (
                        3 * (59 + 180 * beta - 90 * x) * (beta * beta) + alpha * (
                        (453 + 1620 * beta * (1 - x) - 455 * x) * beta + alpha * (
                        8 * (1 - x) * (135 * beta - 11)))));
    const double prefactor_num = (1 + 12 * alpha) * (1 + 12 * beta) / (total * total);
    const double prefactor_den = 12960 * alpha * alpha * alpha * beta * beta * (1 + 12 * total);
    return (float)(prefactor_num / (1 - x) * poly / prefactor_den);
  }
  const double prefactor = -x / sqrt(2 * alpha * beta / total);
  const double stirling = (1 + 1 / (12 * alpha) + 1 / (288 * alpha * alpha)) *
                          (1 + 1 / (12 * beta) + 1 / (288 * beta * beta)) /
                          (1 + 1 / (12 * total) + 1 / (288 * total * total));
  const double term1_num = 2 * (alpha * alpha) * (x - 1) + alpha * beta * (x - 1) - x * (beta * beta);
  const double axbx = alpha * (x - 1) + beta * x;
  const double term1_den = sqrt(2 * alpha / beta) * pow(total, 1.5) * axbx * axbx;
  const double term1 = term1_num / term1_den;
  const double term2 = 0.5 * log(alpha / (total * x));
  const double term3_num = sqrt(8 * alpha * beta / total);
  const double term3_den = beta * x + alpha * (x - 1);
  const double term3 = term3_num / term3_den;
  const double term4_base = beta * log(beta / (total * (1 - x))) + alpha * log(alpha / (total * x));
  const double term4 = pow(term4_base, -1.5);
  const double term1234 = term1 + term2 * (term3 + (x < mean ? term4 : -term4));
  return (float)(stirling * prefactor * term1234);
}

__constant__ double kDirC[2][3][3][4] = {
    {{{1.003668233, -0.01061107488, -0.0657888334, 0.01201642863},
      {0.6336835991, -0.3557432599, 0.05486251648, -0.001465281033},
      {-0.03276231906, 0.004474107445, 0.002429354597, -0.0001557569013}},
     {{0.221950385, -0.3187676331, 0.01799915743, 0.01074823814},
      {-0.2951249643, 0.06219954479, 0.01535556598, 0.001550077057},
      {0.02155310298, 0.004170831599, 0.001292462449, 6.976601077e-05}},
     {{-0.05980841433, 0.008441916499, 0.01085618172, 0.002319392565},
      {0.02911413504, 0.01400243777, -0.002721828457, 0.000751041181},
      {0.005900514878, -0.001936558688, -9.495446725e-06, 5.385558597e-05}}},
    {{{1, -0.02924021934, -0.04438342661, 0.007285809825},
      {0.6357567472, -0.3473456711, 0.05454656494, -0.002407477521},
      {-0.03301322327, 0.004845219414, 0.00231480583, -0.0002307248149}},
     {{0.5925320577, -0.1757678135, 0.01505928619, 0.000564515273},
      {0.1014815858, -0.06589186703, 0.01272886114, -0.0007316646956},
      {-0.007258481865, 0.001096195486, 0.0003934994223, -4.12701925e-05}},
     {{0.06469649321, -0.0236701437, 0.002902096474, -5.896963079e-05},
      {0.001925008108, -0.002869809258, 0.0008000589141, -6.063713228e-05},
      {-0.0003477407336, 6.959756487e-05, 1.097287507e-05, -1.650964693e-06}}},
};

__device__ float dirichlet_grad_one(float x, float alpha, float total) {
  const double x_ = x, alpha_ = alpha, total_ = total;
  const float beta = total - alpha;
  const double beta_ = total_ - alpha_;
  const float boundary = total * x * (1 - x);
  if (x <= 0.5f && boundary < 2.5f) return beta_grad_alpha_small(x, alpha, beta);
  if (x >= 0.5f && boundary < 0.75f) return -beta_grad_beta_small(1 - x, beta, alpha);
  if (alpha > 6 && beta > 6) return beta_grad_alpha_mid(x_, alpha_, beta_);
  const double u = log(x_);
  const double a = log(alpha_) - u;
  const double b = log(total_) - a;
  const double pow_u[3] = {1, u, u * u};
  const double pow_a[3] = {1, a, a * a};
  double p = 0.0, q = 0.0;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const double ua = pow_u[i] * pow_a[j];
      p += ua * (kDirC[0][i][j][0] + b * (kDirC[0][i][j][1] + b * (kDirC[0][i][j][2] + b * kDirC[0][i][j][3])));
      q += ua * (kDirC[1][i][j][0] + b * (kDirC[1][i][j][1] + b * (kDirC[1][i][j][2] + b * kDirC[1][i][j][3])));
    }
  const double approx = x_ * ((double)digamma_one<float>(total) - (double)digamma_one<float>(alpha)) / beta_;
  return (float)(p / q * approx);
}

__global__ void __launch_bounds__(128)
beta_grad_kernel(const float* __restrict__ z, const float* __restrict__ alpha,
                 const float* __restrict__ beta, int64_t S, int64_t B, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= S * B) return;
  const int64_t b = i % B;
  const float x = z[i], a = alpha[b], bt = beta[b];
  out[i] = (1.f - x) * dirichlet_grad_one(x, a, a + bt);
}

}  // namespace mm25

using namespace mm25;

extern "C" int mm25_vmf_sample_w(const float* kappa, int64_t d, int64_t S, int64_t B,
                                 uint64_t seed, const float* prop_eps, const float* prop_u,
                                 int64_t P, float* w_out, int* count, void* stream) {
  if (S <= 0 || B <= 0 || d < 2) return MM25_ERR_SHAPE;
  if (!kappa || !w_out) return MM25_ERR_ARG;
  if ((prop_eps == nullptr) != (prop_u == nullptr)) return MM25_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (prop_eps) {
    if (P <= 0) return MM25_ERR_SHAPE;
    vmf_w_injected_kernel<<<(unsigned)ceil_div(B, 128), 128, 0, st>>>(
        kappa, (int)d, S, B, prop_eps, prop_u, P, w_out, count);
  } else {
    vmf_w_philox_kernel<<<(unsigned)ceil_div(S * B, 128), 128, 0, st>>>(kappa, (int)d, S, B, seed,
                                                                       w_out);
  }
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_beta_sample(const float* alpha, const float* beta, int64_t S, int64_t B,
                                uint64_t seed, float* z, void* stream) {
  if (S <= 0 || B <= 0) return MM25_ERR_SHAPE;
  if (!alpha || !beta || !z) return MM25_ERR_ARG;
  beta_sample_kernel<<<(unsigned)ceil_div(S * B, 128), 128, 0, (cudaStream_t)stream>>>(
      alpha, beta, S, B, seed, z);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_beta_grad(const float* z, const float* alpha, const float* beta, int64_t S,
                              int64_t B, float* dz_dalpha, void* stream) {
  if (S <= 0 || B <= 0) return MM25_ERR_SHAPE;
  if (!z || !alpha || !beta || !dz_dalpha) return MM25_ERR_ARG;
  beta_grad_kernel<<<(unsigned)ceil_div(S * B, 128), 128, 0, (cudaStream_t)stream>>>(
      z, alpha, beta, S, B, dz_dalpha);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
