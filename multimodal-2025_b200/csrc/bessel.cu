// bessel.cu — numerically stable log I_nu(kappa) on the device (fp64 arithmetic), replacing the
// reference's GPU->CPU round trip through np.vectorize(mpmath.besseli)
// (losses/distributions/VonMisesFisher.py:19-62; same routine in vmf-lib/models.py:9-66).
//
//   nu >= 16 : uniform (Debye) large-order expansion, valid for all kappa > 0
//              I_nu(nu z) ~ e^{nu eta} / (sqrt(2 pi nu) (1+z^2)^{1/4}) * sum_k U_k(t) / nu^k,
//              t = 1/sqrt(1+z^2), eta = sqrt(1+z^2) + log(z / (1 + sqrt(1+z^2))), k <= 5
//   nu <  16 : ascending series (kappa <= 200) or Hankel large-argument expansion (kappa > 200)
// Checked against mpmath 1.3.0 (the reference's pinned version) in tests: abs error < 1e-8
// relative to max(1, |log I|), far below the fp32 rounding the reference applies afterwards.
#include "common.cuh"

namespace mm25 {

__device__ double log_bessel_debye(double nu, double x) {
  const double z = x / nu;
  const double r = sqrt(1.0 + z * z);
  const double t = 1.0 / r;
  const double eta = r + log(z / (1.0 + r));
  const double t2 = t * t;
  const double u1 = t * (3.0 - 5.0 * t2) / 24.0;
  const double u2 = t2 * (81.0 + t2 * (-462.0 + 385.0 * t2)) / 1152.0;
  const double u3 = t * t2 * (30375.0 + t2 * (-369603.0 + t2 * (765765.0 - 425425.0 * t2))) /
                    414720.0;
  const double u4 = t2 * t2 *
                    (4465125.0 +
                     t2 * (-94121676.0 +
                           t2 * (349922430.0 + t2 * (-446185740.0 + 185910725.0 * t2)))) /
                    39813120.0;
  const double u5 =
      t * t2 * t2 *
      (1519035525.0 +
       t2 * (-49286948607.0 +
             t2 * (284499769554.0 +
                   t2 * (-614135872350.0 + t2 * (566098157625.0 - 188699385875.0 * t2))))) /
      6688604160.0;
  const double inu = 1.0 / nu;
  const double s = 1.0 + inu * (u1 + inu * (u2 + inu * (u3 + inu * (u4 + inu * u5))));
  return nu * eta - 0.5 * log(2.0 * M_PI * nu) - 0.5 * log(r) + log(s);
}

__device__ double log_bessel_series(double nu, double x) {
  const double q = 0.25 * x * x;
  double term = 1.0, s = 1.0;
  for (int k = 1; k < 100000; ++k) {
    term *= q / ((double)k * (nu + (double)k));
    s += term;
    if (term < 1e-17 * s) break;
  }
  return nu * log(0.5 * x) - lgamma(nu + 1.0) + log(s);
}

__device__ double log_bessel_hankel(double nu, double x) {
  const double mu = 4.0 * nu * nu;
  double term = 1.0, s = 1.0;
  for (int k = 1; k < 80; ++k) {
    const double o = (double)(2 * k - 1);
    const double nt = -term * (mu - o * o) / ((double)k * 8.0 * x);
    if (fabs(nt) > fabs(term)) break;
    term = nt;
    s += term;
    if (fabs(term) < 1e-17 * fabs(s)) break;
  }
  return x - 0.5 * log(2.0 * M_PI * x) + log(s);
}

__device__ double log_bessel_i(double nu, double x) {
  if (!(x > 0.0)) return (x == 0.0) ? (nu == 0.0 ? 0.0 : -INFINITY) : NAN;
  if (nu >= 16.0) return log_bessel_debye(nu, x);
  if (x <= 200.0) return log_bessel_series(nu, x);
  return log_bessel_hankel(nu, x);
}

__global__ void __launch_bounds__(128) log_bessel_kernel(const double* __restrict__ kappa,
                                                         int64_t n, double nu,
                                                         double* __restrict__ out0,
                                                         double* __restrict__ out1) {
  const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= n) return;
  const double x = kappa[i];
  if (out0) out0[i] = log_bessel_i(nu, x);
  if (out1) out1[i] = log_bessel_i(nu + 1.0, x);
}

}  // namespace mm25

extern "C" int mm25_log_bessel(const double* kappa, int64_t n, double nu, double* log_i_nu,
                               double* log_i_nu1, void* stream) {
  if (n < 0 || !kappa || (!log_i_nu && !log_i_nu1)) return MM25_ERR_ARG;
  if (!(nu >= 0.0)) return MM25_ERR_ARG;
  if (n == 0) return MM25_OK;
  mm25::log_bessel_kernel<<<(unsigned)mm25::ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(
      kappa, n, nu, log_i_nu, log_i_nu1);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
