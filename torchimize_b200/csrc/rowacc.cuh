// rowacc.cuh — per-thread accumulation of one problem's weighted normal equations.
//   g[i]   += w * J[i] * f
//   H[i,j] += w * J[i] * J[j]      (upper triangle, packed row-major: (0,0),(0,1)..(0,P-1),(1,1)..)
// reference arithmetic: newton_parallel.py:40-43 (sum over n, then weighted sum over c).
#pragma once
#include "common.cuh"

namespace nlls {

template <int P> struct Tri { static constexpr int N = P * (P + 1) / 2; };

template <typename T, int P>
struct RowAcc {
  static constexpr int NH = Tri<P>::N;
  T g[P];
  T h[NH];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int i = 0; i < P; ++i) g[i] = T(0);
#pragma unroll
    for (int i = 0; i < NH; ++i) h[i] = T(0);
  }
  // one residual row: Jrow[P], residual f, cost weight w
  __device__ __forceinline__ void add(const T (&Jrow)[P], T f, T w) {
    int k = 0;
#pragma unroll
    for (int i = 0; i < P; ++i) {
      const T jw = w * Jrow[i];
      g[i] = fma(jw, f, g[i]);
#pragma unroll
      for (int j = i; j < P; ++j) { h[k] = fma(jw, Jrow[j], h[k]); ++k; }
    }
  }
  // rank-1 update with separate scale factors (fused-model path): g += sg*k, H += sh*k k^T
  __device__ __forceinline__ void add_scaled(const T (&k_)[P], T sg, T sh) {
    int k = 0;
#pragma unroll
    for (int i = 0; i < P; ++i) {
      g[i] = fma(sg, k_[i], g[i]);
      const T a = sh * k_[i];
#pragma unroll
      for (int j = i; j < P; ++j) { h[k] = fma(a, k_[j], h[k]); ++k; }
    }
  }
  __device__ __forceinline__ void warp_reduce() {
#pragma unroll
    for (int i = 0; i < P; ++i) g[i] = warp_sum(g[i]);
#pragma unroll
    for (int i = 0; i < NH; ++i) h[i] = warp_sum(h[i]);
  }
};

// Read one row of P scalars from global memory with the widest loads the row size and the
// base alignment allow.  VEC = log2 of the vector width in bytes (2: 4B scalars … 5: 32B).
template <typename T, int P>
struct RowLoad {
  static constexpr int RB = P * (int)sizeof(T);
  static constexpr int vec_bytes() { return (RB % 32 == 0) ? 32 : (RB % 16 == 0) ? 16 : (RB % 8 == 0) ? 8 : (int)sizeof(T); }
};

template <int P>
__device__ __forceinline__ void load_row(const double* __restrict__ src, double (&r)[P], bool aligned) {
  constexpr int RB = P * 8;
  if constexpr (RB % 32 == 0) {
    if (aligned) {
#pragma unroll
      for (int i = 0; i < P; i += 4)
        asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
                     : "=d"(r[i]), "=d"(r[i + 1]), "=d"(r[i + 2]), "=d"(r[i + 3]) : "l"(src + i));
      return;
    }
  } else if constexpr (RB % 16 == 0) {
    if (aligned) {
#pragma unroll
      for (int i = 0; i < P; i += 2) { const double2 v = ldg_stream2(src + i); r[i] = v.x; r[i + 1] = v.y; }
      return;
    }
  }
#pragma unroll
  for (int i = 0; i < P; ++i) r[i] = __ldg(src + i);
}

template <int P>
__device__ __forceinline__ void load_row(const float* __restrict__ src, float (&r)[P], bool aligned) {
  constexpr int RB = P * 4;
  if constexpr (RB % 32 == 0) {
    if (aligned) {
#pragma unroll
      for (int i = 0; i < P; i += 8)
        asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(r[i]), "=f"(r[i + 1]), "=f"(r[i + 2]), "=f"(r[i + 3]), "=f"(r[i + 4]), "=f"(r[i + 5]),
                       "=f"(r[i + 6]), "=f"(r[i + 7]) : "l"(src + i));
      return;
    }
  } else if constexpr (RB % 16 == 0) {
    if (aligned) {
#pragma unroll
      for (int i = 0; i < P; i += 4) { const float4 v = ldg_stream4(src + i); r[i] = v.x; r[i + 1] = v.y; r[i + 2] = v.z; r[i + 3] = v.w; }
      return;
    }
  } else if constexpr (RB % 8 == 0) {
    if (aligned) {
#pragma unroll
      for (int i = 0; i < P; i += 2) { const float2 v = ldg_stream2(src + i); r[i] = v.x; r[i + 1] = v.y; }
      return;
    }
  }
#pragma unroll
  for (int i = 0; i < P; ++i) r[i] = __ldg(src + i);
}

// Row from shared memory (the bulk-copy staged path).
template <typename T, int P>
__device__ __forceinline__ void lds_row(const T* src, T (&r)[P]) {
  constexpr int RB = P * (int)sizeof(T);
  if constexpr (RB % 16 == 0) {
    const int4* s4 = reinterpret_cast<const int4*>(src);
    int4 tmp[RB / 16];
#pragma unroll
    for (int i = 0; i < RB / 16; ++i) tmp[i] = s4[i];
    const T* t = reinterpret_cast<const T*>(tmp);
#pragma unroll
    for (int i = 0; i < P; ++i) r[i] = t[i];
  } else if constexpr (RB % 8 == 0) {
    const int2* s2 = reinterpret_cast<const int2*>(src);
    int2 tmp[RB / 8];
#pragma unroll
    for (int i = 0; i < RB / 8; ++i) tmp[i] = s2[i];
    const T* t = reinterpret_cast<const T*>(tmp);
#pragma unroll
    for (int i = 0; i < P; ++i) r[i] = t[i];
  } else {
#pragma unroll
    for (int i = 0; i < P; ++i) r[i] = src[i];
  }
}

// Write the reduced packed upper triangle as a full symmetric P x P block.
template <typename T, int P>
__device__ __forceinline__ void store_sym(T* __restrict__ Hout, const T (&h)[Tri<P>::N]) {
  int k = 0;
#pragma unroll
  for (int i = 0; i < P; ++i)
#pragma unroll
    for (int j = i; j < P; ++j) { Hout[i * P + j] = h[k]; Hout[j * P + i] = h[k]; ++k; }
}

}  // namespace nlls
