// samplers.cu — HBM-bound warp-per-row kernels of the vCLIP sampling path.
//   ProjectedNormal.rsample          losses/distributions/ProjectedNormal.py:41-49
//   [t, sqrt(1-t^2) v] + Householder VonMisesFisher.py:180-183,224-236; power_spherical
//                                    _TTransform / _HouseholderRotationTransform
// One warp owns one output row [d] (forward) or one distribution row b over all S samples
// (backward, so the parameter gradients reduce in registers / shared memory without atomics).
// Rows are processed in groups of 4 consecutive elements per lane: 16-byte loads / stores when
// d % 4 == 0, scalar otherwise.  Noise is either injected by the caller (parity mode: the
// deterministic transform is compared with the reference given identical draws) or generated
// in-kernel with Philox4x32-7 keyed by (seed, row, element) and regenerated in the backward.
#include "common.cuh"
#include "philox.cuh"

namespace mm25 {

struct F4 { float v[4]; };

__device__ __forceinline__ F4 load4(const float* __restrict__ p, int k0, int d, bool vec) {
  F4 r;
  if (vec) {
    const float4 t = *reinterpret_cast<const float4*>(p + k0);
    r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) r.v[e] = (k0 + e < d) ? p[k0 + e] : 0.f;
  }
  return r;
}
__device__ __forceinline__ void store4(float* __restrict__ p, int k0, int d, bool vec,
                                       const F4& r) {
  if (vec) {
    *reinterpret_cast<float4*>(p + k0) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (k0 + e < d) p[k0 + e] = r.v[e];
  }
}

constexpr uint64_t STREAM_PN = 0x504e;      // ProjectedNormal noise
constexpr uint64_t STREAM_SPHERE = 0x5348;  // tangent noise of the spherical samplers

// ------------------------------------------------------------------------------------------
// ProjectedNormal: y[s,b,:] = (mu_b + sigma_b * eps) / ||.||       (no eps in the norm, :49)
// ------------------------------------------------------------------------------------------
// MAXG > 0: the row (<= 128 MAXG elements) stays in registers between the norm pass and the
// write pass, so parameters are read and noise generated once; MAXG == 0: generic two-pass.
template <int MAXG>
__global__ void __launch_bounds__(256)
projnormal_fwd_kernel(const float* __restrict__ mu, const float* __restrict__ sigma,
                      const float* __restrict__ eps_in, uint64_t seed, int64_t S, int64_t B,
                      int d, float* __restrict__ out, float* __restrict__ inv_norm) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= S * B) return;
  const int64_t b = row % B;
  const bool vec = (d & 3) == 0;
  const float* mr = mu + b * d;
  const float* sr = sigma + b * d;
  const int ngroups = (d + 3) >> 2;
  float ss = 0.f;
  F4 cache[MAXG > 0 ? MAXG : 1];
  auto value = [&](int g) {
    const int k0 = g * 4;
    const F4 m = load4(mr, k0, d, vec), sg = load4(sr, k0, d, vec);
    F4 e;
    if (eps_in) e = load4(eps_in + row * d, k0, d, vec);
    else rng::normal4(seed, (uint64_t)row * 0x100000ull + (uint64_t)g, STREAM_PN, e.v);
    F4 v;
#pragma unroll
    for (int i = 0; i < 4; ++i) v.v[i] = (k0 + i < d) ? fmaf(sg.v[i], e.v[i], m.v[i]) : 0.f;
    return v;
  };
  if (MAXG > 0) {
#pragma unroll
    for (int c = 0; c < MAXG; ++c) {
      const int g = lane + 32 * c;
      if (g < ngroups) {
        cache[c] = value(g);
#pragma unroll
        for (int i = 0; i < 4; ++i) ss = fmaf(cache[c].v[i], cache[c].v[i], ss);
      }
    }
  } else {
    for (int g = lane; g < ngroups; g += 32) {
      const F4 v = value(g);
#pragma unroll
      for (int i = 0; i < 4; ++i) ss = fmaf(v.v[i], v.v[i], ss);
    }
  }
  ss = warp_sum(ss);
  const float inv = 1.f / sqrtf(ss);
  if (MAXG > 0) {
#pragma unroll
    for (int c = 0; c < MAXG; ++c) {
      const int g = lane + 32 * c;
      if (g < ngroups) {
        F4 y;
#pragma unroll
        for (int i = 0; i < 4; ++i) y.v[i] = cache[c].v[i] * inv;
        store4(out + row * d, g * 4, d, vec, y);
      }
    }
  } else {
    for (int g = lane; g < ngroups; g += 32) {
      F4 y = value(g);
#pragma unroll
      for (int i = 0; i < 4; ++i) y.v[i] *= inv;
      store4(out + row * d, g * 4, d, vec, y);
    }
  }
  if (lane == 0 && inv_norm) inv_norm[row] = inv;
}

// backward: dv = (g - y (y.g)) * inv;  dmu_b = sum_s dv;  dsigma_b = sum_s dv * eps.
// One warp per b; the [d] accumulators live in shared memory (2 d floats per warp).
__global__ void __launch_bounds__(256)
projnormal_bwd_kernel(const float* __restrict__ gout, const float* __restrict__ y,
                      const float* __restrict__ inv_norm, const float* __restrict__ eps_in,
                      uint64_t seed, int64_t S, int64_t B, int d, float* __restrict__ dmu,
                      float* __restrict__ dsigma) {
  extern __shared__ float acc_sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t b = (int64_t)blockIdx.x * 8 + warp;
  if (b >= B) return;
  const bool vec = (d & 3) == 0;
  const int ngroups = (d + 3) >> 2;
  const int dpad = ngroups * 4;
  float* am = acc_sm + (size_t)warp * 2 * dpad;
  float* as = am + dpad;
  for (int k = lane; k < dpad; k += 32) { am[k] = 0.f; as[k] = 0.f; }
  __syncwarp();
  // two samples per iteration: twice the bytes in flight per warp (the loop is latency-bound:
  // each sample needs a warp reduction between its two passes)
  for (int64_t s = 0; s < S; s += 2) {
    const bool two = s + 1 < S;
    const int64_t row0 = s * B + b, row1 = two ? row0 + B : row0;
    const float* g0 = gout + row0 * d;
    const float* y0 = y + row0 * d;
    const float* g1 = gout + row1 * d;
    const float* y1 = y + row1 * d;
    float dot0 = 0.f, dot1 = 0.f;
    for (int g = lane; g < ngroups; g += 32) {
      const F4 ga = load4(g0, g * 4, d, vec), ya = load4(y0, g * 4, d, vec);
      const F4 gb = load4(g1, g * 4, d, vec), yb = load4(y1, g * 4, d, vec);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        dot0 = fmaf(ga.v[i], ya.v[i], dot0);
        dot1 = fmaf(gb.v[i], yb.v[i], dot1);
      }
    }
    dot0 = warp_sum(dot0);
    dot1 = warp_sum(dot1);
    const float inv0 = inv_norm[row0], inv1 = two ? inv_norm[row1] : 0.f;
    for (int g = lane; g < ngroups; g += 32) {
      const int k0 = g * 4;
      const F4 ga = load4(g0, k0, d, vec), ya = load4(y0, k0, d, vec);
      const F4 gb = load4(g1, k0, d, vec), yb = load4(y1, k0, d, vec);
      F4 e0, e1;
      if (eps_in) {
        e0 = load4(eps_in + row0 * d, k0, d, vec);
        e1 = load4(eps_in + row1 * d, k0, d, vec);
      } else {
        rng::normal4(seed, (uint64_t)row0 * 0x100000ull + (uint64_t)g, STREAM_PN, e0.v);
        rng::normal4(seed, (uint64_t)row1 * 0x100000ull + (uint64_t)g, STREAM_PN, e1.v);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float dv0 = (ga.v[i] - ya.v[i] * dot0) * inv0;
        const float dv1 = (gb.v[i] - yb.v[i] * dot1) * inv1;   // inv1 = 0 on the odd tail
        am[k0 + i] += dv0 + dv1;
        as[k0 + i] = fmaf(dv1, e1.v[i], fmaf(dv0, e0.v[i], as[k0 + i]));
      }
    }
  }
  __syncwarp();
  for (int g = lane; g < ngroups; g += 32) {
    F4 a, c;
#pragma unroll
    for (int i = 0; i < 4; ++i) { a.v[i] = am[g * 4 + i]; c.v[i] = as[g * 4 + i]; }
    if (dmu) store4(dmu + b * d, g * 4, d, vec, a);
    if (dsigma) store4(dsigma + b * d, g * 4, d, vec, c);
  }
}

// ------------------------------------------------------------------------------------------
// Spherical samplers (vMF and PowerSpherical share the geometry):
//   x = [t, c * v / nv],  c = sqrt(1 - t^2) (optionally clamped), nv = max(||v||, v_clamp) + v_add
//   u = (e1 - loc) / den, den = max(||e1 - loc||, h_clamp) + h_add;   y = x - 2 (x.u) u
// vMF  (VonMisesFisher.py:181-183,224-236): v_clamp = h_clamp = 1e-10, adds 0, no clamp on c.
// PowerSpherical (App. B): v_add = h_add = omt2_min = 1e-7, clamps 0.
// The tangent noise of output element k >= 1 is v_in[row, k-1] (injected) or Philox keyed by k.
// ------------------------------------------------------------------------------------------
struct SphereCfg {
  float v_clamp, v_add, h_clamp, h_add, omt2_min;  // omt2_min < 0: no clamp on 1 - t^2
};

__device__ __forceinline__ F4 sphere_noise4(const float* __restrict__ v_in, int64_t row, int k0,
                                            int d, uint64_t seed, int g) {
  F4 e;
  if (v_in) {
    const float* vr = v_in + row * (int64_t)(d - 1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = k0 + i;
      e.v[i] = (k >= 1 && k < d) ? vr[k - 1] : 0.f;
    }
  } else {
    rng::normal4(seed, (uint64_t)row * 0x100000ull + (uint64_t)g, STREAM_SPHERE, e.v);
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (k0 + i < 1 || k0 + i >= d) e.v[i] = 0.f;
  }
  return e;
}

template <int MAXG>
__global__ void __launch_bounds__(256)
sphere_fwd_kernel(const float* __restrict__ t_in, const float* __restrict__ loc,
                  const float* __restrict__ v_in, uint64_t seed, int64_t S, int64_t B, int d,
                  SphereCfg cfg, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= S * B) return;
  const int64_t b = row % B;
  const bool vec = (d & 3) == 0;
  const float* mr = loc + b * d;
  const int ngroups = (d + 3) >> 2;
  float vv = 0.f, vmu = 0.f, mm = 0.f;
  F4 ce[MAXG > 0 ? MAXG : 1], cm[MAXG > 0 ? MAXG : 1];
  auto accumulate = [&](const F4& e, const F4& m, int k0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      vv = fmaf(e.v[i], e.v[i], vv);
      vmu = fmaf(e.v[i], m.v[i], vmu);
      if (k0 + i >= 1) mm = fmaf(m.v[i], m.v[i], mm);
    }
  };
  if (MAXG > 0) {
#pragma unroll
    for (int c = 0; c < MAXG; ++c) {
      const int g = lane + 32 * c;
      if (g < ngroups) {
        cm[c] = load4(mr, g * 4, d, vec);
        ce[c] = sphere_noise4(v_in, row, g * 4, d, seed, g);
        accumulate(ce[c], cm[c], g * 4);
      }
    }
  } else {
    for (int g = lane; g < ngroups; g += 32)
      accumulate(sphere_noise4(v_in, row, g * 4, d, seed, g), load4(mr, g * 4, d, vec), g * 4);
  }
  vv = warp_sum(vv); vmu = warp_sum(vmu); mm = warp_sum(mm);
  const float mu0 = mr[0];
  const float t = t_in[row];
  const float omt2 = 1.f - t * t;
  const float c = sqrtf(cfg.omt2_min >= 0.f ? fmaxf(omt2, cfg.omt2_min) : omt2);
  const float cn = c / (fmaxf(sqrtf(vv), cfg.v_clamp) + cfg.v_add);
  const float w0 = 1.f - mu0;
  const float den = fmaxf(sqrtf(fmaf(w0, w0, mm)), cfg.h_clamp) + cfg.h_add;
  const float xw = t * w0 - cn * vmu;      // x . (e1 - loc)
  const float f = 2.f * xw / (den * den);  // y = x - f * (e1 - loc)
  auto emit = [&](const F4& e, const F4& m, int k0) {
    F4 y;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = k0 + i;
      const float x = (k == 0) ? t : cn * e.v[i];
      const float w = (k == 0) ? w0 : -m.v[i];
      y.v[i] = fmaf(-f, w, x);
    }
    store4(out + row * d, k0, d, vec, y);
  };
  if (MAXG > 0) {
#pragma unroll
    for (int c2 = 0; c2 < MAXG; ++c2) {
      const int g = lane + 32 * c2;
      if (g < ngroups) emit(ce[c2], cm[c2], g * 4);
    }
  } else {
    for (int g = lane; g < ngroups; g += 32)
      emit(sphere_noise4(v_in, row, g * 4, d, seed, g), load4(mr, g * 4, d, vec), g * 4);
  }
}

// backward: one warp per b over all S samples.  dloc [B,d] (sum over s), dt [S,B].
__global__ void __launch_bounds__(256)
sphere_bwd_kernel(const float* __restrict__ gout, const float* __restrict__ y,
                  const float* __restrict__ t_in, const float* __restrict__ loc, int64_t S,
                  int64_t B, int d, SphereCfg cfg, float* __restrict__ dloc,
                  float* __restrict__ dt) {
  extern __shared__ float acc_sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t b = (int64_t)blockIdx.x * 8 + warp;
  if (b >= B) return;
  const bool vec = (d & 3) == 0;
  const int ngroups = (d + 3) >> 2;
  const int dpad = ngroups * 4;
  float* A = acc_sm + (size_t)warp * dpad;  // sum_s dL/du
  for (int k = lane; k < dpad; k += 32) A[k] = 0.f;
  const float* mr = loc + b * d;
  float mm = 0.f;
  for (int g = lane; g < ngroups; g += 32) {
    const F4 m = load4(mr, g * 4, d, vec);
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (g * 4 + i >= 1) mm = fmaf(m.v[i], m.v[i], mm);
  }
  mm = warp_sum(mm);
  const float w0 = 1.f - mr[0];
  const float nw = sqrtf(fmaf(w0, w0, mm));
  const float den = fmaxf(nw, cfg.h_clamp) + cfg.h_add;
  const float iden = 1.f / den;
  const float uu = (nw * iden) * (nw * iden);
  __syncwarp();
  // two samples per iteration (latency-bound loop: two warp reductions per sample)
  for (int64_t s = 0; s < S; s += 2) {
    const bool two = s + 1 < S;
    const int64_t row0 = s * B + b, row1 = two ? row0 + B : row0;
    const float* g0 = gout + row0 * d;
    const float* y0 = y + row0 * d;
    const float* g1 = gout + row1 * d;
    const float* y1 = y + row1 * d;
    float yu0 = 0.f, gu0 = 0.f, yu1 = 0.f, gu1 = 0.f;
    for (int g = lane; g < ngroups; g += 32) {
      const int k0 = g * 4;
      const F4 ga = load4(g0, k0, d, vec), ya = load4(y0, k0, d, vec);
      const F4 gb = load4(g1, k0, d, vec), yb = load4(y1, k0, d, vec);
      const F4 m = load4(mr, k0, d, vec);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float u = ((k0 + i == 0) ? w0 : -m.v[i]) * iden;
        yu0 = fmaf(ya.v[i], u, yu0); gu0 = fmaf(ga.v[i], u, gu0);
        yu1 = fmaf(yb.v[i], u, yu1); gu1 = fmaf(gb.v[i], u, gu1);
      }
    }
    yu0 = warp_sum(yu0); gu0 = warp_sum(gu0);
    yu1 = warp_sum(yu1); gu1 = warp_sum(gu1);
    const float keep1 = two ? 1.f : 0.f;  // the odd tail contributes nothing to A
    const float xu0 = yu0 * (1.f - 2.f * uu), xu1 = yu1 * (1.f - 2.f * uu);
    float sgx0 = 0.f, gx00 = 0.f, sgx1 = 0.f, gx01 = 0.f;
    for (int g = lane; g < ngroups; g += 32) {
      const int k0 = g * 4;
      const F4 ga = load4(g0, k0, d, vec), ya = load4(y0, k0, d, vec);
      const F4 gb = load4(g1, k0, d, vec), yb = load4(y1, k0, d, vec);
      const F4 m = load4(mr, k0, d, vec);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = k0 + i;
        if (k < d) {
          const float u = ((k == 0) ? w0 : -m.v[i]) * iden;
          const float xa = fmaf(-2.f * yu0, u, ya.v[i]);
          const float gxa = fmaf(-2.f * gu0, u, ga.v[i]);
          const float xb = fmaf(-2.f * yu1, u, yb.v[i]);
          const float gxb = fmaf(-2.f * gu1, u, gb.v[i]);
          A[k] += -2.f * (xu0 * ga.v[i] + gu0 * xa) - 2.f * keep1 * (xu1 * gb.v[i] + gu1 * xb);
          if (k == 0) { gx00 = gxa; gx01 = gxb; }
          else { sgx0 = fmaf(gxa, xa, sgx0); sgx1 = fmaf(gxb, xb, sgx1); }
        }
      }
    }
    sgx0 = warp_sum(sgx0); sgx1 = warp_sum(sgx1);
    gx00 = __shfl_sync(0xffffffffu, gx00, 0);
    gx01 = __shfl_sync(0xffffffffu, gx01, 0);
    if (lane == 0 && dt) {
      const float t0 = t_in[row0];
      const float o0 = 1.f - t0 * t0;
      const bool c0 = cfg.omt2_min >= 0.f && !(o0 > cfg.omt2_min);
      dt[row0] = c0 ? gx00 : gx00 - t0 / o0 * sgx0;
      if (two) {
        const float t1 = t_in[row1];
        const float o1 = 1.f - t1 * t1;
        const bool c1 = cfg.omt2_min >= 0.f && !(o1 > cfg.omt2_min);
        dt[row1] = c1 ? gx01 : gx01 - t1 / o1 * sgx1;
      }
    }
  }
  __syncwarp();
  if (dloc) {
    // u = w / den(w):  dL/dw = A / den - w (w.A) / (den^2 nw) when the norm is not clamped
    float wa = 0.f;
    for (int g = lane; g < ngroups; g += 32) {
      const F4 m = load4(mr, g * 4, d, vec);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = g * 4 + i;
        if (k < d) wa = fmaf((k == 0) ? w0 : -m.v[i], A[k], wa);
      }
    }
    wa = warp_sum(wa);
    const float corr = (nw > cfg.h_clamp && nw > 0.f) ? wa * iden * iden / nw : 0.f;
    for (int g = lane; g < ngroups; g += 32) {
      const F4 m = load4(mr, g * 4, d, vec);
      F4 o;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = g * 4 + i;
        const float w = (k == 0) ? w0 : -m.v[i];
        const float dw = (k < d) ? A[k] * iden - w * corr : 0.f;
        o.v[i] = -dw;  // w = e1 - loc
      }
      store4(dloc + b * d, g * 4, d, vec, o);
    }
  }
}

}  // namespace mm25

using namespace mm25;

extern "C" int mm25_sample_projnormal_fwd(const float* mu, const float* sigma,
                                          const float* eps_in, uint64_t seed, int64_t S,
                                          int64_t B, int64_t d, float* out, float* inv_norm,
                                          void* stream) {
  if (S <= 0 || B <= 0 || d <= 0 || d > (1 << 20)) return MM25_ERR_SHAPE;
  if (!mu || !sigma || !out) return MM25_ERR_ARG;
  const unsigned grid = (unsigned)ceil_div(S * B, 8);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t gpl = ceil_div(ceil_div(d, 4), 32);  // 4-element groups per lane
#define PN(G) projnormal_fwd_kernel<G><<<grid, 256, 0, st>>>(mu, sigma, eps_in, seed, S, B, \
                                                             (int)d, out, inv_norm)
  void* tt_ = timing_begin(MM25_T_SAMPLER, st);
  if (gpl <= 1) PN(1); else if (gpl <= 2) PN(2); else if (gpl <= 4) PN(4);
  else if (gpl <= 8) PN(8); else PN(0);
  timing_end(tt_, st);
#undef PN
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_sample_projnormal_bwd(const float* gout, const float* y,
                                          const float* inv_norm, const float* eps_in,
                                          uint64_t seed, int64_t S, int64_t B, int64_t d,
                                          float* dmu, float* dsigma, void* stream) {
  if (S <= 0 || B <= 0 || d <= 0 || d > 4096) return MM25_ERR_SHAPE;
  if (!gout || !y || !inv_norm) return MM25_ERR_ARG;
  const size_t sh = sizeof(float) * 8 * 2 * (size_t)((d + 3) / 4 * 4);
  MM25_CUDA_CHECK(cudaFuncSetAttribute(projnormal_bwd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
  void* tt_ = timing_begin(MM25_T_SAMPLER, (cudaStream_t)stream);
  projnormal_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, sh, (cudaStream_t)stream>>>(
      gout, y, inv_norm, eps_in, seed, S, B, (int)d, dmu, dsigma);
  timing_end(tt_, (cudaStream_t)stream);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_sphere_sample_fwd(const float* t, const float* loc, const float* v_in,
                                      uint64_t seed, int64_t S, int64_t B, int64_t d,
                                      float v_clamp, float v_add, float h_clamp, float h_add,
                                      float omt2_min, float* out, void* stream) {
  if (S <= 0 || B <= 0 || d < 2 || d > (1 << 20)) return MM25_ERR_SHAPE;
  if (!t || !loc || !out) return MM25_ERR_ARG;
  SphereCfg cfg{v_clamp, v_add, h_clamp, h_add, omt2_min};
  const unsigned grid = (unsigned)ceil_div(S * B, 8);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t gpl = ceil_div(ceil_div(d, 4), 32);
#define SP(G) sphere_fwd_kernel<G><<<grid, 256, 0, st>>>(t, loc, v_in, seed, S, B, (int)d, cfg, out)
  void* tt_ = timing_begin(MM25_T_SAMPLER, st);
  if (gpl <= 1) SP(1); else if (gpl <= 2) SP(2); else if (gpl <= 4) SP(4);
  else if (gpl <= 8) SP(8); else SP(0);
  timing_end(tt_, st);
#undef SP
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}

extern "C" int mm25_sphere_sample_bwd(const float* gout, const float* y, const float* t,
                                      const float* loc, int64_t S, int64_t B, int64_t d,
                                      float v_clamp, float v_add, float h_clamp, float h_add,
                                      float omt2_min, float* dloc, float* dt, void* stream) {
  if (S <= 0 || B <= 0 || d < 2 || d > 8192) return MM25_ERR_SHAPE;
  if (!gout || !y || !t || !loc) return MM25_ERR_ARG;
  SphereCfg cfg{v_clamp, v_add, h_clamp, h_add, omt2_min};
  const size_t sh = sizeof(float) * 8 * (size_t)((d + 3) / 4 * 4);
  MM25_CUDA_CHECK(cudaFuncSetAttribute(sphere_bwd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
  void* tt_ = timing_begin(MM25_T_SAMPLER, (cudaStream_t)stream);
  sphere_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, sh, (cudaStream_t)stream>>>(
      gout, y, t, loc, S, B, (int)d, cfg, dloc, dt);
  timing_end(tt_, (cudaStream_t)stream);
  MM25_LAUNCH_CHECK();
  return MM25_OK;
}
