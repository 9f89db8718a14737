// common.cuh — shared device/host helpers for the sm_100a NLLS kernels.
#pragma once
#include <cuda.h>            // CUtensorMap and its enums only: the encoder is fetched at run time, nothing links against libcuda
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "../../include/nlls_b200.h"

#ifndef NLLS_MAX_P
#define NLLS_MAX_P 32
#endif

namespace nlls {

// ----------------------------------------------------------------------------- host side
void set_error(const char* fmt, ...);

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

inline bool aligned_to(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

extern int g_force_direct;    // test/bench hook (nlls_set_force_direct): 1 = never take a bulk-copy staged kernel

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// --------------------------------------------------------------------------- device side
__device__ __forceinline__ bool is_done(const nlls_ctrl* ctrl) {
  // written by nlls_iter_end in an earlier launch on the same stream: a plain load suffices
  return ctrl != nullptr && *reinterpret_cast<const volatile int*>(&ctrl->done) != 0;
}
__device__ __forceinline__ int iter_index(const nlls_ctrl* ctrl) {
  return ctrl ? *reinterpret_cast<const volatile int*>(&ctrl->iters) : 0;
}

// ---- per-problem tracking / retirement (nlls_track, include/nlls_b200.h)
__device__ __forceinline__ bool track_frozen(const nlls_track& tr, long long b) { return tr.retire && tr.pstat[b * 4] >= 0; }
__device__ __forceinline__ void track_record(const nlls_track& tr, long long b, bool conv, int reason, int step_inc, bool nonfinite,
                                             int it) {
  int* s = tr.pstat + b * 4;
  if (conv && s[0] < 0) { s[0] = it; s[1] = reason; }
  if (step_inc) s[2] += step_inc;
  if (nonfinite) s[3] |= 1;
}
// every lane of the warp has to call it: the problems that stay active go to the NEXT iteration's list (warp-aggregated)
__device__ __forceinline__ void track_append(const nlls_track& tr, int it, long long b, bool keep) {
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  if (!m) return;
  const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(tr.count + ((it + 1) & 1), __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (keep) tr.active[(size_t)((it + 1) & 1) * tr.B + base + __popc(m & ((1u << lane) - 1))] = (int)b;
}
// the problems a fused model kernel walks: all of them, or this iteration's list of active ones
struct WorkList {
  const int* list;
  long long n;
  __device__ __forceinline__ WorkList(const nlls_track& tr, int it, long long B) {
    if (tr.retire) { list = tr.active + (size_t)(it & 1) * tr.B; n = tr.count[it & 1]; }
    else { list = nullptr; n = B; }
  }
  __device__ __forceinline__ long long at(long long i) const { return list ? (long long)list[i] : i; }
};
inline nlls_track track_or_none(const nlls_track* t) { nlls_track z{}; return t ? *t : z; }

template <typename T> __device__ __forceinline__ T shfl_xor(T v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
template <typename T> __device__ __forceinline__ T shfl_down(T v, int d) { return __shfl_down_sync(0xffffffffu, v, d); }

template <typename T> __device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v += shfl_xor(v, m);
  return v;
}

// Sum V values held by every lane of a warp; returns in lane 0..(all lanes) the totals.
// Recursive halving ("transpose reduce"): ~V + 5 shuffles instead of 5 V.  After the call
// lane L holds the total of value index (L % Vp) where Vp = V rounded up to a power of two
// <= 32; callers read totals through the returned owner mapping:  total of value k sits in
// out[0] of lane k (k < V).  For V > 32 falls back to plain butterflies.
template <int V, typename T>
__device__ __forceinline__ void warp_sum_many(T (&v)[V]) {
#pragma unroll
  for (int k = 0; k < V; ++k) v[k] = warp_sum(v[k]);
}

// Transposed ("recursive halving") warp reduction of VP values held by every lane, VP a power of
// two <= 32: VP - 1 + log2(32/VP) shuffles instead of 5*VP butterflies.  On return v[0] of lane L
// holds the warp total of value index L >> log2(32/VP); lanes with (L & (32/VP - 1)) == 0 own it.
template <int VP, typename R>
__device__ __forceinline__ void warp_reduce_transpose(R (&v)[VP]) {
  static_assert(VP >= 1 && VP <= 32 && (VP & (VP - 1)) == 0, "VP must be a power of two <= 32");
  const int lane = threadIdx.x & 31;
  int m = 16;
#pragma unroll
  for (int n = VP; n > 1; n >>= 1, m >>= 1) {
    const bool upper = (lane & m) != 0;
#pragma unroll
    for (int i = 0; i < n / 2; ++i) {
      const R keep = upper ? v[i + n / 2] : v[i];
      const R send = upper ? v[i] : v[i + n / 2];
      v[i] = keep + shfl_xor(send, m);
    }
  }
#pragma unroll
  for (; m > 0; m >>= 1) v[0] += shfl_xor(v[0], m);
}
template <int V> struct Pow2Ceil { static constexpr int value = V <= 1 ? 1 : V <= 2 ? 2 : V <= 4 ? 4 : V <= 8 ? 8 : V <= 16 ? 16 : 32; };

// product rounded on its own (never contracted into an FMA): the reference materialises f = r*r before summing it,
// and every kernel that evaluates the same residual has to produce the same bits whatever the compiler fuses
// Packed fp32 pairs (sm_100: add / mul / fma .f32x2 -> SASS FADD2 / FMUL2 / FFMA2): ONE issue slot for two lanes' worth of fp32
// arithmetic, each lane rounded exactly like the scalar instruction (the multiply is never contracted).  Same flop rate as
// the scalar pipe; what it saves is instruction issue, which is what bounds the fp32 fused kernels.
struct f32x2 { unsigned long long v; };
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ f32x2 neg2(f32x2 a) { f32x2 r; r.v = a.v ^ 0x8000000080000000ull; return r; }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }

// streaming (read-once) global loads: bypass L1 allocation
__device__ __forceinline__ double ldg_stream(const double* p) {
  double r;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ float ldg_stream(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ double2 ldg_stream2(const double* p) {
  double2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ldg_stream4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float2 ldg_stream2(const float* p) {
  float2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// streaming stores (written once, read by a later kernel)
__device__ __forceinline__ void stg_stream(double* p, double v) {
  asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void stg_stream(float* p, float v) {
  asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

// ---------------------------------------------------------------- mbarrier / bulk-copy (TMA)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  // try_wait suspends in hardware for a bounded time; the spin bound turns a lost transaction
  // (a bug) into a trap instead of a hung GPU.
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)
    if (spins > (1u << 24)) __trap();
}
// 1-D bulk async copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP).
// dst, src 16-byte aligned; bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// 3-D tiled tensor copy global -> shared through a tensor map (SASS: UTMALDG.3D): ONE instruction moves a box of
// [c2: 1][c1: rows][c0: samples] elements; completion (always the full box's bytes) on an mbarrier.  dst 128-byte aligned.
__device__ __forceinline__ void tma_load_3d(void* dst_smem, const CUtensorMap* tm, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
          smem_u32(dst_smem)),
      "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

// The same three operations on 32-bit shared-window addresses (smem_u32) the caller computed once: a producer that issues a
// copy per tile keeps `ring` and `bars` as two integers instead of converting generic pointers on every tile.
__device__ __forceinline__ void mbar_expect_tx_s(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s_s(uint32_t dst, const void* src_gmem, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src_gmem),
               "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tma_load_3d_s(uint32_t dst, const CUtensorMap* tm, int c0, int c1, int c2, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
               "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
               : "memory");
}

// ------------------------------------------------------------------ small dense fp64 solves
// Solve A x = b for a symmetric P x P system held fully in registers (A is destroyed).
// Cholesky first; if a pivot is not strictly positive (or not finite) the original matrix is
// re-solved by LU with partial pivoting.  Returns false when LU meets an exactly zero /
// non-finite pivot (result then carries inf/NaN, like an unpivoted QR would).
template <int P>
__device__ __forceinline__ bool spd_solve(const double (&A0)[P][P], const double (&b0)[P], double (&x)[P]) {
  double L[P][P];
  bool ok = true;
#pragma unroll
  for (int j = 0; j < P; ++j) {
    double d = A0[j][j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
    ok = ok && (d > 0.0) && (d < __longlong_as_double(0x7ff0000000000000LL));
    const double dj = sqrt(d);
    L[j][j] = dj;
    const double inv = 1.0 / dj;
#pragma unroll
    for (int i = j + 1; i < P; ++i) {
      double s = A0[i][j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
      L[i][j] = s * inv;
    }
  }
  if (ok) {
    double yv[P];
#pragma unroll
    for (int i = 0; i < P; ++i) {
      double s = b0[i];
#pragma unroll
      for (int k = 0; k < i; ++k) s -= L[i][k] * yv[k];
      yv[i] = s / L[i][i];
    }
#pragma unroll
    for (int i = P - 1; i >= 0; --i) {
      double s = yv[i];
#pragma unroll
      for (int k = i + 1; k < P; ++k) s -= L[k][i] * x[k];
      x[i] = s / L[i][i];
    }
    return true;
  }
  // ---- LU with partial pivoting (rare path; kept out of the unrolled fast path's way)
  double A[P][P], b[P];
#pragma unroll
  for (int i = 0; i < P; ++i) {
    b[i] = b0[i];
#pragma unroll
    for (int j = 0; j < P; ++j) A[i][j] = A0[i][j];
  }
  bool nonsing = true;
#pragma unroll
  for (int k = 0; k < P; ++k) {
    int piv = k;
    double best = fabs(A[k][k]);
#pragma unroll
    for (int i = k + 1; i < P; ++i) {
      const double a = fabs(A[i][k]);
      if (a > best) { best = a; piv = i; }
    }
#pragma unroll
    for (int i = k + 1; i < P; ++i) {
      if (i == piv) {
#pragma unroll
        for (int j = 0; j < P; ++j) { const double t = A[k][j]; A[k][j] = A[i][j]; A[i][j] = t; }
        const double t = b[k]; b[k] = b[i]; b[i] = t;
      }
    }
    const double d = A[k][k];
    nonsing = nonsing && (d != 0.0) && (fabs(d) < __longlong_as_double(0x7ff0000000000000LL));
    const double inv = 1.0 / d;
#pragma unroll
    for (int i = k + 1; i < P; ++i) {
      const double m = A[i][k] * inv;
#pragma unroll
      for (int j = k + 1; j < P; ++j) A[i][j] -= m * A[k][j];
      b[i] -= m * b[k];
    }
  }
#pragma unroll
  for (int i = P - 1; i >= 0; --i) {
    double s = b[i];
#pragma unroll
    for (int k = i + 1; k < P; ++k) s -= A[i][k] * x[k];
    x[i] = s / A[i][i];
  }
  return nonsing;
}

// Runtime-P variant working on a caller-provided scratch matrix (shared/local memory),
// row-major with leading dimension ld and element stride `es` (for interleaved layouts).
__device__ inline bool spd_solve_rt(int P, double* A, int ld, int es, double* b, double* x, double* Awork) {
  // copy for LU fallback
  for (int i = 0; i < P; ++i)
    for (int j = 0; j < P; ++j) Awork[(i * ld + j) * es] = A[(i * ld + j) * es];
  bool ok = true;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  for (int j = 0; j < P; ++j) {        // in-place Cholesky, lower triangle of A
    double d = A[(j * ld + j) * es];
    for (int k = 0; k < j; ++k) { const double l = A[(j * ld + k) * es]; d -= l * l; }
    ok = ok && (d > 0.0) && (d < INF);
    const double dj = sqrt(d);
    A[(j * ld + j) * es] = dj;
    const double inv = 1.0 / dj;
    for (int i = j + 1; i < P; ++i) {
      double s = A[(i * ld + j) * es];
      for (int k = 0; k < j; ++k) s -= A[(i * ld + k) * es] * A[(j * ld + k) * es];
      A[(i * ld + j) * es] = s * inv;
    }
  }
  if (ok) {
    for (int i = 0; i < P; ++i) {
      double s = b[i * es];
      for (int k = 0; k < i; ++k) s -= A[(i * ld + k) * es] * x[k * es];
      x[i * es] = s / A[(i * ld + i) * es];
    }
    for (int i = P - 1; i >= 0; --i) {
      double s = x[i * es];
      for (int k = i + 1; k < P; ++k) s -= A[(k * ld + i) * es] * x[k * es];
      x[i * es] = s / A[(i * ld + i) * es];
    }
    return true;
  }
  double* M = Awork;
  bool nonsing = true;
  for (int k = 0; k < P; ++k) {
    int piv = k;
    double best = fabs(M[(k * ld + k) * es]);
    for (int i = k + 1; i < P; ++i) {
      const double a = fabs(M[(i * ld + k) * es]);
      if (a > best) { best = a; piv = i; }
    }
    if (piv != k) {
      for (int j = 0; j < P; ++j) {
        const double t = M[(k * ld + j) * es]; M[(k * ld + j) * es] = M[(piv * ld + j) * es]; M[(piv * ld + j) * es] = t;
      }
      const double t = b[k * es]; b[k * es] = b[piv * es]; b[piv * es] = t;
    }
    const double d = M[(k * ld + k) * es];
    nonsing = nonsing && (d != 0.0) && (fabs(d) < INF);
    const double inv = 1.0 / d;
    for (int i = k + 1; i < P; ++i) {
      const double m = M[(i * ld + k) * es] * inv;
      for (int j = k + 1; j < P; ++j) M[(i * ld + j) * es] -= m * M[(k * ld + j) * es];
      b[i * es] -= m * b[k * es];
    }
  }
  for (int i = P - 1; i >= 0; --i) {
    double s = b[i * es];
    for (int k = i + 1; k < P; ++k) s -= M[(i * ld + k) * es] * x[k * es];
    x[i * es] = s / M[(i * ld + i) * es];
  }
  return nonsing;
}

}  // namespace nlls
