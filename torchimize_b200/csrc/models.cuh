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
// fp32: one FMUL + MUFU.EX2.  The rounding of x*log2(e) makes the RELATIVE error grow like |x| * 1e-7, but every
// use here is exp(-a) with a >= 0 multiplied into the model value, so the ABSOLUTE error is bounded by
// max_a(a e^-a) * 1e-7 = 4e-8 of the peak amplitude — below one fp32 ulp of the model at its peak.
template <> __device__ __forceinline__ float t_exp<float>(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// exp(-a c) for a >= 0 with a per-problem constant c prepared once: fp32 folds the sign and log2(e) into the constant
// (ONE FMUL + MUFU.EX2 per evaluation, no negation instruction); fp64 keeps CUDA's exp on the exactly negated product.
template <typename T> struct ExpNeg;
// fp64 exp for NON-POSITIVE arguments (every call site below evaluates exp(-a), a >= 0).  CUDA's exp() costs ~30
// instructions plus ~20 moves that materialise its polynomial constants as immediates (IMAD.MOV was 18 % of the warp
// instructions of the fp64 fused kernel, profiles/r01h_fused_ncu.txt); here the coefficients sit in constant memory, so
// every DFMA reads its coefficient as a c[bank][offset] operand.  Cody-Waite reduction x = n ln2 + r, |r| <= ln2/2,
// degree-13 Taylor polynomial (remainder < 5e-18 relative), scaling by 2^n through the exponent field; < 1 ulp measured
// against torch.exp (tests/test_gpu_parity.py::test_fp64_exp_nonpos_accuracy).  Results below 2^-1020 (x < -707) are
// flushed to 0: the reference's denormal tail (< 1e-307) is absolutely negligible in every sum it enters.  NaN propagates.
__constant__ double kExpTaylor[12] = {1.6059043836821613e-10, 2.08767569878681e-09, 2.505210838544172e-08, 2.755731922398589e-07,
                                      2.7557319223985893e-06, 2.48015873015873e-05, 0.0001984126984126984, 0.001388888888888889,
                                      0.008333333333333333, 0.041666666666666664, 0.16666666666666666, 0.5};
// Branch-free (the flush and the NaN case are selects at the end): several calls in one basic block — the two exps of a
// sample, the samples of a thread — interleave their 14-deep Horner chains instead of running one after the other.
__device__ __forceinline__ double exp_nonpos(double x) {
  const double magic = 6755399441055744.0;                   // 1.5 * 2^52: fma rounds x*log2(e) to the nearest integer
  double t = fma(x, 1.4426950408889634, magic);
  const int n = __double2loint(t);
  t -= magic;
  double r = fma(t, -0.6931471803691238, x);                 // ln2 split: high part has 21 trailing zero bits (t*hi exact)
  r = fma(t, -1.9082149292705877e-10, r);
  double q = kExpTaylor[0];
#pragma unroll
  for (int i = 1; i < 12; ++i) q = fma(q, r, kExpTaylor[i]);
  q = fma(q, r, 1.0);
  q = fma(q, r, 1.0);
  const double e = __hiloint2double(__double2hiint(q) + (n << 20), __double2loint(q));
  return x >= -707.0 ? e : (x < -707.0 ? 0.0 : x);          // flushed tail; NaN propagates
}
template <> struct ExpNeg<double> {
  static __device__ __forceinline__ double prep(double c) { return c; }
  static __device__ __forceinline__ double eval(double a, double c) { return exp_nonpos(-a * c); }
  static __device__ __forceinline__ double sq(double x) { return exp_nonpos(-(x * x)); }
};
template <> struct ExpNeg<float> {
  static __device__ __forceinline__ float prep(float c) { return c * -1.4426950408889634f; }
  static __device__ __forceinline__ float eval(float a, float cp) { return ex2_approx(a * cp); }
  static __device__ __forceinline__ float sq(float x) { return ex2_approx((x * x) * -1.4426950408889634f); }
};
template <typename T> __device__ __forceinline__ T t_erf(T x);
template <> __device__ __forceinline__ double t_erf<double>(double x) { return erf(x); }
template <> __device__ __forceinline__ float t_erf<float>(float x) { return erff(x); }

// K(x) = 1 + erf(x) given E = exp(-x^2).
//   fp64: CUDA's erf (the 1e-9 parity target needs full accuracy).
//   fp32: branch-free  erfc(|x|) = E * t * Q(t),  t = 1 / (1 + 0.3 |x|),  Q of degree 8 fitted to the scaled
//         complementary error function erfcx on [0, 10.5] (max relative error 2.0e-8 in exact arithmetic,
//         4.3e-7 evaluated in fp32); K = erfc(|x|) for x < 0 (no cancellation: relative accuracy in the
//         left tail, where the reference's own fp32 `1 + erf(x)` rounds to 0) and 2 - erfc(x) otherwise.
//         ~16 instructions and no divergent branch instead of erff's two-sided ~40, and E is needed for
//         the derivative anyway.  Fit: least squares with Lawson reweighting on the relative error.
template <typename T> __device__ __forceinline__ T one_plus_erf(T x, T E);
template <> __device__ __forceinline__ double one_plus_erf<double>(double x, double) {
  // |x| >= 6: 1 - erf(|x|) < 2.2e-17 < 2^-54, so erf rounds to +-1 exactly and 1 + erf(x) is exactly 2 or 0 — the value
  // the reference's torch.erf produces.  Samples are adjacent along the abscissa, so whole warps take this exit in the
  // tails of a peak (86 % of the samples of the EMG workloads) and skip CUDA's erf (~90 instructions, one division).
  if (fabs(x) >= 6.0) return x > 0.0 ? 2.0 : 0.0;
  return 1.0 + erf(x);
}
template <> __device__ __forceinline__ float one_plus_erf<float>(float x, float E) {
  const float ax = fabsf(x);
  float t;                                         // MUFU.RCP alone (1 ulp): __frcp_rn adds a Newton step and a slow-path branch
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3f, ax, 1.0f)));
  float q = -0.12955187585482458f;
  q = fmaf(q, t, 0.4857163224001443f);
  q = fmaf(q, t, -0.5126247969354352f);
  q = fmaf(q, t, 0.5481587242455374f);
  q = fmaf(q, t, -0.09498945570912465f);
  q = fmaf(q, t, 0.215089233597969f);
  q = fmaf(q, t, 0.14829135184669887f);
  q = fmaf(q, t, 0.17072300956908976f);
  q = fmaf(q, t, 0.1691874673196955f);
  const float c = E * (t * q);                    // erfc(|x|)
  return x < 0.0f ? c : 2.0f - c;
}

template <typename T>
struct EmgModel {
  static constexpr int P = 4;
  T alpha, eta, mu, inv_s, inv_s2h, xs, ce;   // 1/sigma, ExpNeg-prepared 1/(2 sigma^2), eta/(sigma sqrt2), sqrt(2/pi)*eta
  __device__ __forceinline__ void load(const T* __restrict__ p) {
    alpha = p[0]; mu = p[1]; eta = p[3];
    const T sigma = p[2];
    inv_s = T(1) / sigma;
    inv_s2h = ExpNeg<T>::prep(T(0.5) / (sigma * sigma));
    xs = eta / (sigma * T(1.4142135623730951));
    ce = T(0.7978845608028654) * eta;
  }
  __device__ __forceinline__ T value(T t) const {
    const T d = t - mu;
    const T G = ExpNeg<T>::eval(d * d, inv_s2h);                 // gauss_term, emg.py:20
    const T x = xs * d;
    T E = T(0);
    if constexpr (sizeof(T) == 4) E = ExpNeg<T>::sq(x);
    const T A = one_plus_erf<T>(x, E);                        // asymm_term, emg.py:21
    return alpha * (G * A);                                   // same association as eval(): bit-identical m
  }
  // eval() in three phases, so that callers with several evaluations in flight (the samples of a thread, the peaks of a
  // multi-peak model) can run every phase-1 first: (1) both exps, branch-free — their dependent chains interleave;
  // (2) the erf term (fp64: early-out branch); (3) the product rule.
  __device__ __forceinline__ void phase1(T t, T& d, T& G, T& x, T& E) const {
    d = t - mu;
    G = ExpNeg<T>::eval(d * d, inv_s2h);                      // gauss_term, emg.py:20
    x = xs * d;
    E = ExpNeg<T>::sq(x);                                     // erf_term, emg.py:26
  }
  __device__ __forceinline__ void phase3(T d, T G, T E, T A, T& m, T* k) const {
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
  __device__ __forceinline__ void eval(T t, T& m, T (&k)[P]) const {
    T d, G, x, E;
    phase1(t, d, G, x, E);
    const T A = one_plus_erf<T>(x, E);                        // asymm_term, emg.py:21
    phase3(d, G, E, A, m, k);
  }
  template <int S>
  __device__ __forceinline__ void eval_n(const T (&t)[S], T (&m)[S], T (&k)[S][P]) const {
    T d[S], G[S], x[S], E[S], A[S];
#pragma unroll
    for (int i = 0; i < S; ++i) phase1(t[i], d[i], G[i], x[i], E[i]);
#pragma unroll
    for (int i = 0; i < S; ++i) A[i] = one_plus_erf<T>(x[i], E[i]);
#pragma unroll
    for (int i = 0; i < S; ++i) phase3(d[i], G[i], E[i], A[i], m[i], k[i]);
  }
};

// model_eval_n: a model's own eval_n when it has one, otherwise eval() sample by sample (user functors need only eval()).
template <typename Model, typename T, int S, typename = void>
struct EvalN {
  static __device__ __forceinline__ void run(const Model& M, const T (&t)[S], T (&m)[S], T (&k)[S][Model::P]) {
#pragma unroll
    for (int i = 0; i < S; ++i) M.eval(t[i], m[i], k[i]);
  }
};
template <typename Model, typename T, int S>
struct EvalN<Model, T, S, decltype((void)&Model::template eval_n<S>)> {
  static __device__ __forceinline__ void run(const Model& M, const T (&t)[S], T (&m)[S], T (&k)[S][Model::P]) { M.template eval_n<S>(t, m, k); }
};
template <typename Model, typename T, int S>
__device__ __forceinline__ void model_eval_n(const Model& M, const T (&t)[S], T (&m)[S], T (&k)[S][Model::P]) {
  EvalN<Model, T, S>::run(M, t, m, k);
}

// Sum of K EMG peaks, p = [alpha_1, mu_1, sigma_1, eta_1, ..., alpha_K, mu_K, sigma_K, eta_K] (the
// large-P workload of cfg-4, 4 peaks = 16 parameters).  Values are summed in peak order; the derivative block of peak
// j is EmgModel's.  Used by the materialising kernel only (nlls_model_eval -> K1 on the fp64 tensor cores for P > 8).
template <typename T, int K>
struct MultiEmgModel {
  static constexpr int P = 4 * K;
  EmgModel<T> peak[K];
  __device__ __forceinline__ void load(const T* __restrict__ p) {
#pragma unroll
    for (int j = 0; j < K; ++j) peak[j].load(p + 4 * j);
  }
  __device__ __forceinline__ T value(T t) const {
    T m = peak[0].value(t);
#pragma unroll
    for (int j = 1; j < K; ++j) m += peak[j].value(t);
    return m;
  }
  __device__ __forceinline__ void eval(T t, T& m, T (&k)[P]) const {         // the K peaks in EmgModel's three phases
    T d[K], G[K], x[K], E[K], A[K];
#pragma unroll
    for (int j = 0; j < K; ++j) peak[j].phase1(t, d[j], G[j], x[j], E[j]);
#pragma unroll
    for (int j = 0; j < K; ++j) A[j] = one_plus_erf<T>(x[j], E[j]);
#pragma unroll
    for (int j = 0; j < K; ++j) {
      T mj;
      peak[j].phase3(d[j], G[j], E[j], A[j], mj, k + 4 * j);
      m = j == 0 ? mj : m + mj;
    }
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
    inv_s2h = ExpNeg<T>::prep(T(0.5) / (s * s));
    xs = skew / (s * T(1.4142135623730951));
    amp = norm / (s * T(2.5066282746310002));                 // skewed_gaussian :47
    off = (NP == 5) ? p[NP - 1] : T(0);
  }
  __device__ __forceinline__ T value(T t) const {
    const T d = t - mean;
    const T S = ExpNeg<T>::eval(d * d, inv_s2h);                 // :48
    const T x = xs * d;
    T E = T(0);
    if constexpr (sizeof(T) == 4) E = ExpNeg<T>::sq(x);
    const T K = one_plus_erf<T>(x, E);                        // :49
    return amp * (S * K) + off;                               // same association as eval()
  }
  __device__ __forceinline__ void eval(T t, T& m, T (&k)[P]) const {
    const T d = t - mean;
    const T S = ExpNeg<T>::eval(d * d, inv_s2h);
    const T x = xs * d;
    const T E = ExpNeg<T>::sq(x);
    const T K = one_plus_erf<T>(x, E);
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
