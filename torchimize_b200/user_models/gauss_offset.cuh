// Example user model for torchimize_b200.models.compile_model: Gaussian peak on a constant baseline,
//   m(x; a, mu, sigma, c) = a * exp(-(x - mu)^2 / (2 sigma^2)) + c,          p = [a, mu, sigma, c]
// The functor contract is the one of the built-in models (torchimize_b200/csrc/models.cuh): load() once per problem,
// eval() once per abscissa sample returning the model value m and k[i] = -dm/dp_i; value() returns the same bits for
// m without the derivatives (trial-cost kernels).  The cost is the reference tests' squared error used as residual,
// f = (y - m)^2 with Jacobian row 2 (y - m) k  (tests/unit_test_parallel.py:40,105-118).
namespace nlls {

template <typename T>
struct UserModel {
  static constexpr int P = 4;
  T a, mu, c, inv_s, ce;                       // 1/sigma, ExpNeg-prepared 1/(2 sigma^2)
  __device__ __forceinline__ void load(const T* __restrict__ p) {
    a = p[0]; mu = p[1]; c = p[3];
    const T s = p[2];
    inv_s = T(1) / s;
    ce = ExpNeg<T>::prep(T(0.5) / (s * s));
  }
  __device__ __forceinline__ T value(T x) const {
    const T d = x - mu;
    const T G = ExpNeg<T>::eval(d * d, ce);
    return fma(a, G, c);
  }
  __device__ __forceinline__ void eval(T x, T& m, T (&k)[P]) const {
    const T d = x - mu;
    const T G = ExpNeg<T>::eval(d * d, ce);
    m = fma(a, G, c);
    const T q = d * inv_s;                     // (x - mu) / sigma
    const T aGq = a * G * q * inv_s;           // dm/dmu
    k[0] = -G;
    k[1] = -aGq;
    k[2] = -aGq * q;                           // dm/dsigma = a G (x - mu)^2 / sigma^3
    k[3] = T(-1);
  }
};

}  // namespace nlls
