// models.cuh — device functors for the built-in fit models (K4).
//
//   EMG             tests/emg.py:19-49 + tests/unit_test_parallel.py:81-118   p = [alpha, mu, sigma, eta]
//   skewed Gaussian tests/unit_test_skewed_gaussian.py:43-51                  p = [norm, mean, sigm, skew (, offset)]
//
// A functor is loaded once per problem and evaluated once per abscissa sample; it returns the
// model value m and k[i] = -dm/dp_i, so that for a cost f = (y-m)^2 the Jacobian row is
// J = 2 (y-m) k  (tests/unit_test_parallel.py:105-118).  All costs of a problem share (m, k):
// only the data y differs between them, which is what lets the fused kernels evaluate the
// two exp + one erf once per (b,n) instead of once per (b,c,n).
#pragma once
#include "common.cuh"

namespace nlls {

template <typename T> __device__ __forceinline__ T t_exp(T x);
template <> __device__ __forceinline__ double t_exp<double>(double x) { return exp(x); }
template <> __device__ __forceinline__ float t_exp<float>(float x) { return expf(x); }
template <typename T> __device__ __forceinline__ T t_erf(T x);
template <> __device__ __forceinline__ double t_erf<double>(double x) { return erf(x); }
template <> __device__ __forceinline__ float t_erf<float>(float x) { return erff(x); }

template <typename T>
struct EmgModel {
  static constexpr int P = 4;
  T alpha, eta, mu, inv_s, inv_s2h, xs, ce;   // 1/sigma, 1/(2 sigma^2), eta/(sigma sqrt2), sqrt(2/pi)*eta
  __device__ __forceinline__ void load(const T* __restrict__ p) {
    alpha = p[0]; mu = p[1]; eta = p[3];
    const T sigma = p[2];
    inv_s = T(1) / sigma;
    inv_s2h = T(0.5) / (sigma * sigma);
    xs = eta / (sigma * T(1.4142135623730951));
    ce = T(0.7978845608028654) * eta;
  }
  __device__ __forceinline__ T value(T t) const {
    const T d = t - mu;
    const T G = t_exp<T>(-(d * d) * inv_s2h);                 // gauss_term, emg.py:20
    const T A = T(1) + t_erf<T>(xs * d);                      // asymm_term, emg.py:21
    return alpha * G * A;
  }
  __device__ __forceinline__ void eval(T t, T& m, T (&k)[P]) const {
    const T d = t - mu;
    const T G = t_exp<T>(-(d * d) * inv_s2h);
    const T x = xs * d;
    const T A = T(1) + t_erf<T>(x);
    const T E = t_exp<T>(-(x * x));                           // erf_term, emg.py:26
    const T q = d * inv_s;
    const T GA = G * A;
    m = alpha * GA;
    const T aG = alpha * G;
    const T inner = q * A - ce * E;                           // product rule, emg.py:31-43
    k[0] = -GA;                                               // emg.py:48-49
    k[1] = -aG * inv_s * inner;
    k[2] = k[1] * q;
    k[3] = -aG * T(0.7978845608028654) * q * E;               // emg.py:45-46
  }
};

template <typename T, int NP>
struct SkewGaussModel {
  static constexpr int P = NP;                                // 4, or 5 with a constant offset
  T norm, mean, skew, inv_s, inv_s2h, xs, amp, off;
  __device__ __forceinline__ void load(const T* __restrict__ p) {
    norm = p[0]; mean = p[1]; skew = p[3];
    const T s = p[2];
    inv_s = T(1) / s;
    inv_s2h = T(0.5) / (s * s);
    xs = skew / (s * T(1.4142135623730951));
    amp = norm / (s * T(2.5066282746310002));                 // skewed_gaussian :47
    off = (NP == 5) ? p[NP - 1] : T(0);
  }
  __device__ __forceinline__ T value(T t) const {
    const T d = t - mean;
    const T S = t_exp<T>(-(d * d) * inv_s2h);                 // :48
    const T K = T(1) + t_erf<T>(xs * d);                      // :49
    return amp * S * K + off;
  }
  __device__ __forceinline__ void eval(T t, T& m, T (&k)[P]) const {
    const T d = t - mean;
    const T S = t_exp<T>(-(d * d) * inv_s2h);
    const T x = xs * d;
    const T K = T(1) + t_erf<T>(x);
    const T E = t_exp<T>(-(x * x));
    const T q = d * inv_s;
    const T SK = S * K;
    m = amp * SK + off;
    const T aS = amp * S;
    const T cs = T(0.7978845608028654) * skew;
    k[0] = -SK * inv_s * T(0.3989422804014327);
    k[1] = -aS * inv_s * (q * K - cs * E);
    k[2] = -aS * inv_s * (K * (q * q - T(1)) - cs * q * E);
    k[3] = -aS * T(0.7978845608028654) * E * q;
    if constexpr (NP == 5) k[4] = T(-1);
  }
};

}  // namespace nlls
