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
// (ONE FMUL + MUFU.EX2 per evaluation, no negation instruction); fp64 evaluates exp_nonpos (below) on the exactly negated product.
template <typename T> struct ExpNeg;
// fp64 exp for NON-POSITIVE arguments (every call site below evaluates exp(-a), a >= 0).  CUDA's exp() costs ~30
// instructions plus ~20 moves that materialise its polynomial constants as immediates (IMAD.MOV was 18 % of the warp
// instructions of the fp64 fused kernel, profiles/r01h_fused_ncu.txt).  Table-driven: x = (32 k + j) ln2/32 + r with
// |r| <= ln2/64 (Cody-Waite, the high part of ln2/32 has 24 trailing zero bits so t*hi is exact), exp(x) = 2^k * T[j] * (1 + p(r))
// with p of degree 6 (remainder < 4e-18 relative) and T[j] = 2^(j/32) as a hi + lo pair (one 16-byte load from a 512-byte
// table that stays in L1): 12 FP64 instructions with a 7-deep dependent chain instead of the 18 / 16-deep of a degree-13
// polynomial on |r| <= ln2/2.  0.55 ulp worst case over [-700, 0] (C model of the same fma sequence against expl; on the GPU:
// tests/test_gpu_parity.py::test_fp64_exp_nonpos_accuracy).  Results below 2^-1020 (x < -707) are flushed to 0: the
// reference's denormal tail (< 1e-307) is absolutely negligible in every sum it enters.  NaN propagates.
__device__ __align__(16) const double2 kExp2Tab[32] = {
    {1.0, 0.0}, {1.0218971486541166, 5.109225028973444e-17},
    {1.0442737824274138, 8.551889705537965e-17}, {1.0671404006768237, -7.899853966841582e-17},
    {1.0905077326652577, -3.046782079812471e-17}, {1.1143867425958924, 1.0410278456845571e-16},
    {1.1387886347566916, 8.912812676025408e-17}, {1.1637248587775775, 3.8292048369240935e-17},
    {1.189207115002721, 3.982015231465646e-17}, {1.215247359980469, -7.712630692681488e-17},
    {1.241857812073484, 4.658027591836937e-17}, {1.2690509571917332, 2.667932131342186e-18},
    {1.2968395546510096, 2.5382502794888315e-17}, {1.3252366431597413, -2.8587312100388614e-17},
    {1.3542555469368927, 7.70094837980299e-17}, {1.383909881963832, -6.770511658794786e-17},
    {1.4142135623730951, -9.667293313452913e-17}, {1.4451808069770467, -3.0237581349939873e-17},
    {1.4768261459394993, -3.483994556892796e-17}, {1.5091644275934228, -1.016455327754295e-16},
    {1.5422108254079407, 7.949834809697621e-17}, {1.5759808451078865, -1.0136916471278304e-17},
    {1.6104903319492543, 2.4707192569797888e-17}, {1.645755478153965, -1.0125679913674773e-16},
    {1.681792830507429, 8.199010020581497e-17}, {1.718619298122478, -1.851380418263111e-17},
    {1.7562521603732995, 2.960140695448873e-17}, {1.7947090750031072, 1.8227458427912087e-17},
    {1.8340080864093424, 3.283107224245627e-17}, {1.8741676341103, -6.122763413004143e-17},
    {1.9152065613971474, -1.0619946056195963e-16}, {1.9571441241754002, 8.960767791036668e-17}};
__constant__ double kExpPoly[5] = {0.001388888888888889, 0.008333333333333333, 0.041666666666666664, 0.16666666666666666, 0.5};
// Branch-free (the flush and the NaN case are selects at the end): several calls in one basic block — the two exps of a
// sample, the samples of a thread — interleave their dependent chains instead of running one after the other.
__device__ __forceinline__ double exp_nonpos(double x) {
  const double magic = 6755399441055744.0;                   // 1.5 * 2^52: fma rounds x*32/ln2 to the nearest integer
  double t = fma(x, 46.16624130844683, magic);
  const int n = __double2loint(t);                           // 32 k + j (two's complement: j = n & 31, k = n >> 5)
  t -= magic;
  // plain indexing: forming the address as base | (j * 16) would save the 64-bit add, but module globals are only 256-byte
  // aligned at load time whatever __align__ asks for (a 512-byte request held in one process and not in the next: entries
  // j >= 16 then read entry j - 16, a factor 2^-1/2 — found by smoke()).  The index is masked: in range whatever x was.
  const double2 T = __ldg(&kExp2Tab[n & 31]);
  double r = fma(t, -0.021660849335603416, x);
  r = fma(t, -5.689487495325457e-11, r);
  double q = fma(r, kExpPoly[0], kExpPoly[1]);
#pragma unroll
  for (int i = 2; i < 5; ++i) q = fma(q, r, kExpPoly[i]);
  const double p = fma(q, r * r, r);                         // exp(r) - 1
  const double m = fma(T.x, p, T.y) + T.x;
  const double e = __hiloint2double(__double2hiint(m) + ((n >> 5) << 20), __double2loint(m));
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
//   fp64: |x| >= 6 exits early (below); else erfc(|x|) = E * g(|x|) with the scaled complementary error function g = erfcx as
//         ONE polynomial of degree 20 in s = 1 - 3 v, v = |x| / (|x| + 3)  (x in [0, 6] <-> s in [-1, 1]; Chebyshev
//         interpolant converted to the monomial basis, approximation error 3.4e-19 relative; coefficients decay like 0.23^j,
//         so Horner in s is well conditioned), K = erfc(|x|) for x < 0 and 2 - erfc(x) otherwise.  v = |x| * rcp(|x| + 3) keeps
//         RELATIVE accuracy near x = 0 (forming s from the reciprocal directly would amplify its rounding 9-fold).  The
//         reciprocal is MUFU.RCP64H + two Newton steps (|x| + 3 is in [3, 9]: no special cases, no slow-path call).
//         C model of the same fma sequence against erfl / erfcl over [-6, 6]: 1.6 ulp worst for x >= 0, absolute error
//         < 2.6e-16 for x < 0 (CUDA's erf: 2 ulp; the reference's own 1 + erf(x) for x < 0: 1e-16 absolute).  48 instructions,
//         coefficients read from constant memory, instead of CUDA erf's ~125 (of which ~60 moves that materialise its
//         polynomial constants, an exp of its own and a division with a slow path): profiles/r02r.
//   fp32: branch-free  erfc(|x|) = E * t * Q(t),  t = 1 / (1 + 0.3 |x|),  Q of degree 8 fitted to the scaled
//         complementary error function erfcx on [0, 10.5] (max relative error 2.0e-8 in exact arithmetic,
//         4.3e-7 evaluated in fp32); K = erfc(|x|) for x < 0 (no cancellation: relative accuracy in the
//         left tail, where the reference's own fp32 `1 + erf(x)` rounds to 0) and 2 - erfc(x) otherwise.
//         ~16 instructions and no divergent branch instead of erff's two-sided ~40, and E is needed for
//         the derivative anyway.  Fit: least squares with Lawson reweighting on the relative error.
__constant__ double kErfcxPoly[21] = {
    0.3215854164543175, 0.3681515648982602, 0.20143885681922427, 0.08162684513404136,
    0.023308261232669675, 0.003956151636369144, 7.012531316823998e-05, -0.00012716419852449776,
    -1.5153799232302407e-05, 4.5326059043145345e-06, 8.118056744135089e-07, -2.2754311847571347e-07,
    -3.480551916961012e-08, 1.4460927996927543e-08, 8.288178726793465e-10, -9.443620513901106e-10,
    6.175705066293916e-11, 5.1980047888428754e-11, -1.1416522779737793e-11, -1.7471767934261753e-12,
    8.081438420827511e-13};
// erfc-based core for |x| < 6 (NaN falls through it and comes out as NaN)
__device__ __forceinline__ double one_plus_erf_core(double x, double E) {
  const double ax = fabs(x);
  const double w = ax + 3.0;
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(w));
  double e = fma(-w, r, 1.0);
  r = fma(r, e, r);
  e = fma(-w, r, 1.0);
  r = fma(r, e, r);
  const double s = fma(-3.0, ax * r, 1.0);
  double g = kErfcxPoly[20];
#pragma unroll
  for (int i = 19; i >= 0; --i) g = fma(g, s, kErfcxPoly[i]);
  const double c = E * g;                                     // erfc(|x|)
  return x < 0.0 ? c : 2.0 - c;
}
template <typename T> __device__ __forceinline__ T one_plus_erf(T x, T E);
template <> __device__ __forceinline__ double one_plus_erf<double>(double x, double E) {
  // |x| >= 6: 1 - erf(|x|) < 2.2e-17 < 2^-54, so erf rounds to +-1 exactly and 1 + erf(x) is exactly 2 or 0 — the value
  // the reference's torch.erf produces.  Samples are adjacent along the abscissa, so whole warps take this exit in the
  // tails of a peak (82 % of the warp-samples of the EMG workloads).
  if (fabs(x) >= 6.0) return x > 0.0 ? 2.0 : 0.0;
  return one_plus_erf_core(x, E);
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
// value()-only form (no derivative needs E): fp64 evaluates exp(-x^2) inside the |x| < 6 branch only — the SAME expression
// eval()'s phase 1 forms, so the model value is bit-identical between value() and eval().
template <typename T> __device__ __forceinline__ T one_plus_erf_lazy(T x) { return one_plus_erf<T>(x, ExpNeg<T>::sq(x)); }
template <> __device__ __forceinline__ double one_plus_erf_lazy<double>(double x) {
  if (fabs(x) >= 6.0) return x > 0.0 ? 2.0 : 0.0;
  return one_plus_erf_core(x, ExpNeg<double>::sq(x));
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
    const T A = one_plus_erf_lazy<T>(x);                      // asymm_term, emg.py:21
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
    const T K = one_plus_erf_lazy<T>(x);                      // :49
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
