// Counter-based RNG, the model's distributions and the integer weight arithmetic, device side.
//
// Distributions follow the reference's model: prior U(0,1) (Model/bouncing_object.jl:91), Gaussian transition
// and observation noise (bouncing_object.jl:111,125), truncated-normal proposal (Inference/particle_filter.jl:23,
// Utilities/truncatednorm.jl:11-23).  The reference draws from Julia's unseeded global RNG; here every draw is
// Philox4x32-10 keyed by the run seed with counter (particle, step, move, stream), so a run is reproducible and
// independent of how particles are spread over threads or GPUs.
// Integer weights / thresholds define systematic resampling so that indices are bit-exact for any summation
// order and any GPU count: only +, *, fma and exact power-of-two scaling are used.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace smcphys {

enum { ST_PRIOR = 0, ST_TRANSITION = 1, ST_PROPOSAL = 2, ST_RESAMPLE = 3, ST_PREDICT = 4, ST_RESAMPLE_MULTI = 5, ST_MCMC = 6 };

__host__ __device__ __forceinline__ void philox4x32_10(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2,
                                                        uint32_t c3, uint32_t out[4]) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int i = 0; i < 10; i++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// uniform in (0,1) on a 24-bit grid (exact in float32)
__host__ __device__ __forceinline__ double u01(uint32_t bits) {
  return ((double)(bits >> 8) + 0.5) * (1.0 / 16777216.0);
}

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, double& z0, double& z1) {
  double r = sqrt(-2.0 * log(u01(a)));
  double th = 6.283185307179586476925286766559 * u01(b);
  double sn, cs;
  sincos(th, &sn, &cs);
  z0 = r * cs;
  z1 = r * sn;
}

__device__ __forceinline__ double norm_cdf(double x) { return 0.5 * erfc(-x * 0.70710678118654752440); }

// inverse normal CDF: Acklam's rational approximation + one Halley step
__device__ __forceinline__ double norm_icdf(double p) {
  const double a0 = -3.969683028665376e+01, a1 = 2.209460984245205e+02, a2 = -2.759285104469687e+02,
               a3 = 1.383577518672690e+02, a4 = -3.066479806614716e+01, a5 = 2.506628277459239e+00;
  const double b0 = -5.447609879822406e+01, b1 = 1.615858368580409e+02, b2 = -1.556989798598866e+02,
               b3 = 6.680131188771972e+01, b4 = -1.328068155288572e+01;
  const double c0 = -7.784894002430293e-03, c1 = -3.223964580411365e-01, c2 = -2.400758277161838e+00,
               c3 = -2.549732539343734e+00, c4 = 4.374664141464968e+00, c5 = 2.938163982698783e+00;
  const double d0 = 7.784695709041462e-03, d1 = 3.224671290700398e-01, d2 = 2.445134137142996e+00,
               d3 = 3.754408661907416e+00;
  double x;
  if (p < 0.02425) {
    double q = sqrt(-2 * log(p));
    x = (((((c0 * q + c1) * q + c2) * q + c3) * q + c4) * q + c5) / ((((d0 * q + d1) * q + d2) * q + d3) * q + 1);
  } else if (p > 1 - 0.02425) {
    double q = sqrt(-2 * log(1 - p));
    x = -(((((c0 * q + c1) * q + c2) * q + c3) * q + c4) * q + c5) / ((((d0 * q + d1) * q + d2) * q + d3) * q + 1);
  } else {
    double q = p - 0.5, r = q * q;
    x = (((((a0 * r + a1) * r + a2) * r + a3) * r + a4) * r + a5) * q /
        (((((b0 * r + b1) * r + b2) * r + b3) * r + b4) * r + 1);
  }
  double e = norm_cdf(x) - p;
  double u = e * 2.50662827463100050242 * exp(x * x / 2);
  return x - u / (1 + x * u / 2);
}

// TruncNorm(mu, sigma, lo, hi): Distributions.Truncated(Normal(mu, sigma), lo, hi), truncatednorm.jl:11-23
__device__ __forceinline__ double truncnorm_logpdf(double x, double mu, double sigma, double lo, double hi) {
  if (x < lo || x > hi) return -INFINITY;
  double z = (x - mu) / sigma;
  double lz = log(norm_cdf((hi - mu) / sigma) - norm_cdf((lo - mu) / sigma));
  return -0.5 * z * z - log(sigma) - 0.91893853320467274178 - lz;
}
__device__ __forceinline__ double truncnorm_sample(double u, double mu, double sigma, double lo, double hi) {
  double a = norm_cdf((lo - mu) / sigma), b = norm_cdf((hi - mu) / sigma);
  double x = mu + sigma * norm_icdf(a + u * (b - a));
  return fmin(fmax(x, lo), hi);
}

// Proposal draw and Metropolis-Hastings correction in one go: x' ~ TruncNorm(mu, sigma, lo, hi) by inverse CDF and
// log q(mu | x') - log q(x' | mu) = log Z(mu) - log Z(x'), Z(m) = Phi((hi - m) / sigma) - Phi((lo - m) / sigma): the
// Gaussian kernels of the forward and the backward density are the same number ((x' - mu)^2 = (mu - x')^2) and cancel,
// only the truncation normalisers differ.  Equal to truncnorm_logpdf(mu; x') - truncnorm_logpdf(x'; mu) up to rounding
// (1e-16), at two erfc and some arithmetic less per move; Z(mu) is shared with the sampler.
__device__ __forceinline__ double truncnorm_propose(double u, double mu, double sigma, double lo, double hi,
                                                    double* log_ratio) {
  const double a = norm_cdf((lo - mu) / sigma), b = norm_cdf((hi - mu) / sigma);
  double x = mu + sigma * norm_icdf(a + u * (b - a));
  x = fmin(fmax(x, lo), hi);
  const double a1 = norm_cdf((lo - x) / sigma), b1 = norm_cdf((hi - x) / sigma);
  *log_ratio = log(b - a) - log(b1 - a1);
  return x;
}

// The same with the normaliser of the current value carried by the caller: Z(mu) only changes when a move is accepted,
// and then it is the Z(x') just computed for the ratio - the same expressions on the same numbers, so the same bits as
// recomputing.  Per move: 3 erfc + 2 log instead of 5 + 3.
struct TruncZ { double a, b, logz; };  // Phi((lo - m) / sigma), Phi((hi - m) / sigma), log(b - a)
__device__ __forceinline__ TruncZ truncnorm_z(double m, double sigma, double lo, double hi) {
  TruncZ z;
  z.a = norm_cdf((lo - m) / sigma);
  z.b = norm_cdf((hi - m) / sigma);
  z.logz = log(z.b - z.a);
  return z;
}
__device__ __forceinline__ double truncnorm_propose_z(double u, double mu, const TruncZ& zmu, double sigma, double lo,
                                                      double hi, TruncZ* znew, double* log_ratio) {
  double x = mu + sigma * norm_icdf(zmu.a + u * (zmu.b - zmu.a));
  x = fmin(fmax(x, lo), hi);
  *znew = truncnorm_z(x, sigma, lo, hi);
  *log_ratio = zmu.logz - znew->logz;
  return x;
}

#define SMC_HALF_LOG_2PI 0.91893853320467274178

// ---------------------------------------------------------------- integer weights
// exp(x), x <= 0, from IEEE-754 +, *, fma and exact scaling only: identical bits on every device and host.
__host__ __device__ __forceinline__ double det_exp(double x) {
  if (!(x > -700.0)) return 0.0;
  if (x > 0.0) x = 0.0;
  const double LOG2E = 1.4426950408889634074, LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
#ifdef __CUDA_ARCH__
  double kf = rint(__dmul_rn(x, LOG2E));
#else
  double kf = nearbyint(x * LOG2E);
#endif
  double r = fma(-kf, LN2_HI, x);
  r = fma(-kf, LN2_LO, r);
  double p = 1.0 / 6227020800.0;
  p = fma(p, r, 1.0 / 479001600.0);
  p = fma(p, r, 1.0 / 39916800.0);
  p = fma(p, r, 1.0 / 3628800.0);
  p = fma(p, r, 1.0 / 362880.0);
  p = fma(p, r, 1.0 / 40320.0);
  p = fma(p, r, 1.0 / 5040.0);
  p = fma(p, r, 1.0 / 720.0);
  p = fma(p, r, 1.0 / 120.0);
  p = fma(p, r, 1.0 / 24.0);
  p = fma(p, r, 1.0 / 6.0);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  int k = (int)kf;
  int k1 = k / 2, k2 = k - k1;
  union { uint64_t u; double d; } s1, s2;
  s1.u = (uint64_t)(1023 + k1) << 52;
  s2.u = (uint64_t)(1023 + k2) << 52;
#ifdef __CUDA_ARCH__
  return __dmul_rn(__dmul_rn(p, s1.d), s2.d);
#else
  return p * s1.d * s2.d;
#endif
}

// q = floor(exp(lw - max) * 2^39); N <= 2^24 keeps the sum below 2^63
__host__ __device__ __forceinline__ uint64_t weight_q(double lw, double mx) {
#ifdef __CUDA_ARCH__
  return (uint64_t)__dmul_rn(det_exp(__dsub_rn(lw, mx)), 549755813888.0);
#else
  return (uint64_t)(det_exp(lw - mx) * 549755813888.0);
#endif
}

// tau_k = floor((U0 + k 2^32) W / (N 2^32)), exact in 128 bits
__host__ __device__ __forceinline__ uint64_t sys_threshold(uint64_t U0, uint64_t k, uint64_t W, uint64_t N) {
  unsigned __int128 P = (unsigned __int128)(U0 + (k << 32)) * W;
  P >>= 32;
  return (uint64_t)(P / N);
}

// multinomial threshold: floor(u W / 2^64) for 64 uniform bits u
__host__ __device__ __forceinline__ uint64_t multi_threshold(uint64_t u, uint64_t W) {
  return (uint64_t)(((unsigned __int128)u * W) >> 64);
}

}  // namespace smcphys
