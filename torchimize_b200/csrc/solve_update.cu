// solve_update.cu — K2 (damped fp64 solve) and K3 (update, gain ratio, damping adaptation,
// stop tests, snapshots, batch-global termination) of the batched NLLS iteration.
//
// One thread per problem: the P x P systems live in registers (P <= 8; Cholesky with an LU
// fallback) or in an interleaved shared-memory scratch (8 < P <= 32).  The kernels are a few
// microseconds at B = 65536; what they buy is the removal of the reference's ~100 small
// launches and ~20 host synchronisations per iteration (boolean-mask indexing at
// lma:103,105,115,131-134; `converged.all()` at lma:118) — every decision stays on the device.
#include <type_traits>
#include "common.cuh"

namespace nlls {

constexpr int kUpdTPB = 128;

template <typename T> __device__ __forceinline__ bool not_finite(T v) { return !(fabs(v) <= (T)1.7976931348623157e308); }
template <> __device__ __forceinline__ bool not_finite<float>(float v) { return !(fabsf(v) <= 3.4028234663852886e38f); }
template <typename T> __device__ __forceinline__ T nan_max(T a, T b) { return (a != a) ? a : ((b != b) ? b : (a > b ? a : b)); }

// ------------------------------------------------------------------- solve dispatch
// Solves (H + diag_add*I + off_add*(1-I)) x = g in fp64.  H is TH (T or double), row-major.
template <typename TH, typename T, int P>
__device__ __forceinline__ bool solve_reg(const TH* __restrict__ Hb, const TH* __restrict__ gb, T diag_add, T off_add,
                                          bool damped, double* x_out) {
  double A[P][P], rhs[P], x[P];
#pragma unroll
  for (int i = 0; i < P; ++i) {
    rhs[i] = (double)gb[i];
#pragma unroll
    for (int j = 0; j < P; ++j) {
      TH v = Hb[i * P + j];
      if (damped) v = (TH)((T)v + (i == j ? diag_add : off_add));   // Hu formed in T, then .double() (lma:97-98)
      A[i][j] = (double)v;
    }
  }
  const bool ok = spd_solve<P>(A, rhs, x);
#pragma unroll
  for (int i = 0; i < P; ++i) x_out[i] = x[i];
  return ok;
}

template <typename TH, typename T>
__device__ __forceinline__ bool solve_any(int P, const TH* Hb, const TH* gb, T diag_add, T off_add, bool damped,
                                          double* x_out, double* scratch, int es) {
  switch (P) {
    case 1: return solve_reg<TH, T, 1>(Hb, gb, diag_add, off_add, damped, x_out);
    case 2: return solve_reg<TH, T, 2>(Hb, gb, diag_add, off_add, damped, x_out);
    case 3: return solve_reg<TH, T, 3>(Hb, gb, diag_add, off_add, damped, x_out);
    case 4: return solve_reg<TH, T, 4>(Hb, gb, diag_add, off_add, damped, x_out);
    case 5: return solve_reg<TH, T, 5>(Hb, gb, diag_add, off_add, damped, x_out);
    case 6: return solve_reg<TH, T, 6>(Hb, gb, diag_add, off_add, damped, x_out);
    case 7: return solve_reg<TH, T, 7>(Hb, gb, diag_add, off_add, damped, x_out);
    case 8: return solve_reg<TH, T, 8>(Hb, gb, diag_add, off_add, damped, x_out);
    default: break;
  }
  // runtime P: interleaved scratch, element k of this thread at scratch[k*es]
  double* A = scratch;
  double* W = A + (size_t)P * P * es;
  double* b = W + (size_t)P * P * es;
  double* x = b + (size_t)P * es;
  for (int i = 0; i < P; ++i) {
    b[i * es] = (double)gb[i];
    for (int j = 0; j < P; ++j) {
      TH v = Hb[i * P + j];
      if (damped) v = (TH)((T)v + (i == j ? diag_add : off_add));
      A[(i * P + j) * es] = (double)v;
    }
  }
  const bool ok = spd_solve_rt(P, A, P, es, b, x, W);
  for (int i = 0; i < P; ++i) x_out[i] = x[i * es];
  return ok;
}

static size_t scratch_bytes_per_thread(int P) { return P > 8 ? ((size_t)2 * P * P + 2 * P) * 8 : 0; }
static int threads_for(int P) {
  if (P <= 8) return kUpdTPB;
  int t = 32;
  while (t > 1 && scratch_bytes_per_thread(P) * t > 200 * 1024) t >>= 1;
  return t;
}

// block-wide accumulation of up to 3 per-thread 0/1 flags into ctrl->red[idx]
__device__ __forceinline__ void block_count_add(nlls_ctrl* ctrl, int f0, int f1, int f2) {
  __shared__ int cnt[3];
  if (threadIdx.x < 3) cnt[threadIdx.x] = 0;
  __syncthreads();
  const unsigned m0 = __ballot_sync(0xffffffffu, f0), m1 = __ballot_sync(0xffffffffu, f1), m2 = __ballot_sync(0xffffffffu, f2);
  if ((threadIdx.x & 31) == 0) {
    if (m0) atomicAdd(&cnt[0], __popc(m0));
    if (m1) atomicAdd(&cnt[1], __popc(m1));
    if (m2) atomicAdd(&cnt[2], __popc(m2));
  }
  __syncthreads();
  if (threadIdx.x < 3 && cnt[threadIdx.x]) atomicAdd(&ctrl->red[threadIdx.x], (double)cnt[threadIdx.x]);   // integer-valued: order-independent
}

// shared stop tests (lma:108-109, gna:78-79, gda:74-75): strict '<', NaN-propagating max (Q12)
template <typename T>
__device__ __forceinline__ void gp_tests(int P, const T* gb, int gstride, const T* hb, const T* pb, T gtol, T ptol,
                                         bool& gcon, bool& pcon) {
  T gm = fabs(gb[0]);
  T hh = T(0), pp = T(0);
  for (int i = 0; i < P; ++i) {
    if (i) gm = nan_max(gm, (T)fabs(gb[i * gstride]));
    hh = fma(hb[i], hb[i], hh);
    pp = fma(pb[i], pb[i], pp);
  }
  gcon = gm < gtol;
  pcon = sqrt(hh) < ptol * (ptol + sqrt(pp));
}

// ----------------------------------------------------------------------------- LM solve
template <typename T>
__global__ void k_lm_solve(const T* __restrict__ g, const T* __restrict__ H, const T* __restrict__ p,
                           const T* __restrict__ u, T* __restrict__ dscale, int marquardt, long long B, int P,
                           T* __restrict__ h, T* __restrict__ p_trial, nlls_ctrl* ctrl, nlls_track tr) {
  if (is_done(ctrl)) return;
  extern __shared__ double scratch[];
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  if (track_frozen(tr, b)) return;                   // retired problem: no solve, no damping update
  const T* Hb = H + b * P * P;
  T dd = dscale[b * 2], doff = dscale[b * 2 + 1];
  if (marquardt) {                                   // lma:88 — D <- D * max_p max(H_pp, D_pp)   (Q6)
    T m = nan_max(Hb[0], dd);
    for (int i = 1; i < P; ++i) m = nan_max(m, nan_max(Hb[i * P + i], dd));
    dd = dd * m;
    doff = doff * m;                                 // 0 * inf -> NaN exactly like the reference's dense D
    dscale[b * 2] = dd;
    dscale[b * 2 + 1] = doff;
  }
  const T ub = u[b];
  double x[NLLS_MAX_P];
  const bool ok = solve_any<T, T>(P, Hb, g + b * P, ub * dd, ub * doff, true, x, scratch + threadIdx.x, blockDim.x);
  if (!ok && ctrl) ctrl->singular = 1;
  for (int i = 0; i < P; ++i) {
    const T hi = -(T)x[i];                           // lma:98
    h[b * P + i] = hi;
    p_trial[b * P + i] = p[b * P + i] + hi;          // argument of the trial cost (lma:99)
  }
}

// ---------------------------------------------------------------------------- LM update
template <typename T>
__global__ void k_lm_update(T* __restrict__ p, const T* __restrict__ h, const T* __restrict__ g, T* __restrict__ u,
                            T* __restrict__ v, const double* __restrict__ stats, const double* __restrict__ stat_h,
                            int meth, T rho1, T rho2, T beta, T gama, double ftol, T ptol, T gtol, long long B, int P,
                            T* __restrict__ p_hist, T* __restrict__ rho_out, nlls_ctrl* ctrl, nlls_track tr) {
  if (is_done(ctrl)) return;
  const int it = iter_index(ctrl);
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int nc_open = 0, accepted = 0, nc_closed = 0;
  bool keep = false;
  if (b < B && track_frozen(tr, b)) {                // retired: the snapshot repeats the frozen iterate
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = p[b * P + i];
    }
  } else if (b < B) {
    const T* hb = h + b * P;
    const T* gb = g + b * P;
    T* pb = p + b * P;
    T ub = u[b], vb = v[b];
    // gain ratio (lma:100-103); the numerator is sum_c (sum_n f)^2 differences (Q3)
    const T num = (T)stats[b * 4 + 0] - (T)stat_h[b];
    T den = T(0);
    for (int i = 0; i < P; ++i) den = fma(hb[i], ub * hb[i] - gb[i], den);
    T rho = num / den;
    const T INF = (T)__longlong_as_double(0x7ff0000000000000LL);
    if (den == T(0)) rho = (num > T(0)) ? INF : -INF;
    if (rho_out) rho_out[b] = rho;
    if (meth == NLLS_LM_LEVENBERG) {                 // lma:124-136 (Q9): NaN / zero rho changes nothing
      if (rho > T(0)) {
        const T t = T(2) * rho - T(1);
        const T fac = T(1) - t * t * t;
        const T third = T(1) / T(3);
        ub *= (fac > third || fac != fac) ? fac : third;
        vb = T(2);
      } else if (rho < T(0)) {
        ub *= vb;
        vb *= T(2);
      }
    } else {                                         // lma:139-152
      if (rho < rho1) ub *= beta;
      if (rho > rho2) ub /= gama;
    }
    u[b] = ub;
    v[b] = vb;
    accepted = rho > T(0);
    if (accepted)
      for (int i = 0; i < P; ++i) pb[i] += hb[i];    // lma:105, in place on the caller's p (Q1)
    bool gcon, pcon;
    gp_tests<T>(P, gb, 1, hb, pb, gtol, ptol, gcon, pcon);   // p already updated (Q2)
    const bool fcon = (it > 0) && (stats[b * 4 + 1] < ftol * ftol * stats[b * 4 + 2]);   // lma:110
    nc_open = !(gcon || pcon || fcon);
    nc_closed = !(gcon || pcon);
    if (tr.pstat) {                                  // the problem's own test: fcon only with an accepted step of its own
      const bool own = gcon || pcon || (fcon && accepted);
      bool bad = false;
      for (int i = 0; i < P; ++i) bad = bad || not_finite(hb[i]);
      track_record(tr, b, own, (gcon ? 1 : 0) | (pcon ? 2 : 0) | ((fcon && accepted) ? 4 : 0), accepted, bad, it);
      if (tr.retire) { nc_open = nc_closed = !own; keep = !own; }
    }
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = pb[i];   // lma:116
    }
  }
  if (tr.retire) track_append(tr, it, b, keep);
  if (ctrl) block_count_add(ctrl, nc_open, accepted, nc_closed);
}

// ------------------------------------------------------------- LM update, speculative schedule
// The fused path evaluates normal equations AND cost at the trial point in one pass (nlls_model_lm_trial); this
// kernel then only chooses, per problem, between the current and the trial set.  Decisions are taken on the
// CURRENT g / stats (pre-step point, lma:108-110) before anything is replaced.
template <typename T>
__global__ void k_lm_update_spec(T* __restrict__ p, const T* __restrict__ h, T* __restrict__ g, T* __restrict__ H,
                                 const T* __restrict__ g_t, const T* __restrict__ H_t, T* __restrict__ u,
                                 T* __restrict__ v, double* __restrict__ stats, const double* __restrict__ stats_t,
                                 unsigned char* __restrict__ sel, int meth, T rho1, T rho2, T beta, T gama, double ftol,
                                 T ptol, T gtol, long long B, int P, T* __restrict__ p_hist, T* __restrict__ rho_out,
                                 nlls_ctrl* ctrl, nlls_track tr) {
  if (is_done(ctrl)) return;
  const int it = iter_index(ctrl);
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int nc_open = 0, accepted = 0, nc_closed = 0;
  bool keep = false;
  if (b < B && track_frozen(tr, b)) {                // retired: the snapshot repeats the frozen iterate
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = p[b * P + i];
    }
  } else if (b < B) {
    const T* hb = h + b * P;
    T* gb = g + b * P;
    T* pb = p + b * P;
    T ub = u[b], vb = v[b];
    const T num = (T)stats[b * 4 + 0] - (T)stats_t[b * 4 + 0];   // lma:100 (Q3)
    T den = T(0);
    for (int i = 0; i < P; ++i) den = fma(hb[i], ub * hb[i] - gb[i], den);
    T rho = num / den;
    const T INF = (T)__longlong_as_double(0x7ff0000000000000LL);
    if (den == T(0)) rho = (num > T(0)) ? INF : -INF;
    if (rho_out) rho_out[b] = rho;
    if (meth == NLLS_LM_LEVENBERG) {
      if (rho > T(0)) {
        const T t = T(2) * rho - T(1);
        const T fac = T(1) - t * t * t;
        const T third = T(1) / T(3);
        ub *= (fac > third || fac != fac) ? fac : third;
        vb = T(2);
      } else if (rho < T(0)) {
        ub *= vb;
        vb *= T(2);
      }
    } else {
      if (rho < rho1) ub *= beta;
      if (rho > rho2) ub /= gama;
    }
    u[b] = ub;
    v[b] = vb;
    accepted = rho > T(0);
    if (accepted)
      for (int i = 0; i < P; ++i) pb[i] += hb[i];
    bool gcon, pcon;
    gp_tests<T>(P, gb, 1, hb, pb, gtol, ptol, gcon, pcon);
    const bool fcon = (it > 0) && (stats[b * 4 + 1] < ftol * ftol * stats[b * 4 + 2]);
    nc_open = !(gcon || pcon || fcon);
    nc_closed = !(gcon || pcon);
    if (tr.pstat) {                                  // the problem's own test: fcon only with an accepted step of its own
      const bool own = gcon || pcon || (fcon && accepted);
      bool bad = false;
      for (int i = 0; i < P; ++i) bad = bad || not_finite(hb[i]);
      track_record(tr, b, own, (gcon ? 1 : 0) | (pcon ? 2 : 0) | ((fcon && accepted) ? 4 : 0), accepted, bad, it);
      if (tr.retire) { nc_open = nc_closed = !own; keep = !own; }
    }
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = pb[i];
    }
    if (accepted) {                                  // the trial point becomes the current point
      for (int i = 0; i < P; ++i) gb[i] = g_t[b * P + i];
      T* Hb = H + b * P * P;
      const T* Ht = H_t + b * P * P;
      for (int i = 0; i < P * P; ++i) Hb[i] = Ht[i];
      stats[b * 4 + 0] = stats_t[b * 4 + 0];
      stats[b * 4 + 1] = stats_t[b * 4 + 1];
      stats[b * 4 + 2] = stats_t[b * 4 + 2];
      sel[b] ^= 1;
    } else {
      stats[b * 4 + 1] = 0.0;                        // f unchanged: sum (f_prev - f)^2 == 0 at the next fcon
    }
  }
  if (tr.retire) track_append(tr, it, b, keep);
  if (ctrl) block_count_add(ctrl, nc_open, accepted, nc_closed);
}

// ---------------------------------------------------------------------------- GN update
template <typename T>
__global__ void k_gn_update(T* __restrict__ p, const T* __restrict__ g, const T* __restrict__ H,
                            const double* __restrict__ stats, T l, double ftol, T ptol, T gtol, int with_masks,
                            long long B, int P, T* __restrict__ h_out, T* __restrict__ p_hist, nlls_ctrl* ctrl, nlls_track tr) {
  if (is_done(ctrl)) return;
  extern __shared__ double scratch[];
  const int it = iter_index(ctrl);
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int nc = 0;
  bool keep = false;
  if (b < B && track_frozen(tr, b)) {                // retired: the snapshot repeats the frozen iterate
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = p[b * P + i];
    }
  } else if (b < B) {
    double x[NLLS_MAX_P];
    const bool ok = solve_any<T, T>(P, H + b * P * P, g + b * P, T(0), T(0), false, x, scratch + threadIdx.x, blockDim.x);
    if (!ok && ctrl) ctrl->singular = 1;             // gna:74-75 warns on an infinite condition number
    T hb[NLLS_MAX_P];
    T* pb = p + b * P;
    for (int i = 0; i < P; ++i) {
      hb[i] = (T)x[i];                               // gna:140
      pb[i] -= l * hb[i];                            // gna:141 (in place, Q1)
      if (h_out) h_out[b * P + i] = hb[i];
    }
    if (with_masks) {
      bool gcon, pcon;
      gp_tests<T>(P, g + b * P, 1, hb, pb, gtol, ptol, gcon, pcon);
      const bool fcon = (it > 0) && (stats[b * 4 + 1] < ftol * ftol * stats[b * 4 + 2]);   // gna:80
      nc = !(gcon || pcon || fcon);
      if (tr.pstat) {
        bool bad = false;
        for (int i = 0; i < P; ++i) bad = bad || not_finite(hb[i]);
        track_record(tr, b, !nc, (gcon ? 1 : 0) | (pcon ? 2 : 0) | (fcon ? 4 : 0), 1, bad, it);
        keep = tr.retire && nc;
      }
    }
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = pb[i];   // gna:86
    }
  }
  if (tr.retire) track_append(tr, it, b, keep);
  if (ctrl && with_masks) block_count_add(ctrl, nc, 0, 0);
}

// ---------------------------------------------------------------------------- GD update
template <typename T>
__global__ void k_gd_update(T* __restrict__ p, const double* __restrict__ g64, const double* __restrict__ H64,
                            const double* __restrict__ stats, T l, double ftol, T ptol, T gtol, int with_masks, long long B,
                            int P, T* __restrict__ h_out, T* __restrict__ p_hist, nlls_ctrl* ctrl, nlls_track tr) {
  if (is_done(ctrl)) return;
  extern __shared__ double scratch[];
  const int it = iter_index(ctrl);
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int nc = 0;
  bool keep = false;
  if (b < B && track_frozen(tr, b)) {                // retired: the snapshot repeats the frozen iterate
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = p[b * P + i];
    }
  } else if (b < B) {
    double x[NLLS_MAX_P];
    const bool ok = solve_any<double, double>(P, H64 + b * P * P, g64 + b * P, 0.0, 0.0, false, x, scratch + threadIdx.x, blockDim.x);
    if (!ok && ctrl) ctrl->singular = 1;
    T hb[NLLS_MAX_P], gb[NLLS_MAX_P];
    T* pb = p + b * P;
    for (int i = 0; i < P; ++i) {
      hb[i] = (T)x[i];                               // gda:139
      pb[i] -= l * hb[i];                            // gda:146
      gb[i] = hb[i] / l;                             // gda:71
      if (h_out) h_out[b * P + i] = hb[i];
    }
    if (with_masks) {
      bool gcon, pcon;
      gp_tests<T>(P, gb, 1, hb, pb, gtol, ptol, gcon, pcon);
      nc = !(gcon || pcon);
      if (tr.pstat) {                                // own test: fcon on the problem's own sums (gda:76 sums over the batch)
        const bool fcon = (it > 0) && (stats[b * 4 + 1] < ftol * ftol * stats[b * 4 + 2]);
        const bool own = gcon || pcon || fcon;
        bool bad = false;
        for (int i = 0; i < P; ++i) bad = bad || not_finite(hb[i]);
        track_record(tr, b, own, (gcon ? 1 : 0) | (pcon ? 2 : 0) | (fcon ? 4 : 0), 1, bad, it);
        if (tr.retire) { nc = !own; keep = !own; }
      }
    }
    if (p_hist) {
      T* snap = p_hist + ((long long)it * B + b) * P;
      for (int i = 0; i < P; ++i) snap[i] = pb[i];
    }
  }
  if (tr.retire) track_append(tr, it, b, keep);
  if (ctrl && with_masks) block_count_add(ctrl, nc, 0, 0);
}

// Deterministic batch-global sums of stats[:,1] and stats[:,2] into red[2], red[3] (gda:76, Q5):
// one CTA, fixed per-thread strides, fixed tree.
__global__ void __launch_bounds__(1024) k_reduce_stats(const double* __restrict__ stats, long long B, nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  __shared__ double s1[1024], s2[1024];
  double a = 0.0, c = 0.0;
  for (long long b = threadIdx.x; b < B; b += 1024) { a += stats[b * 4 + 1]; c += stats[b * 4 + 2]; }
  s1[threadIdx.x] = a;
  s2[threadIdx.x] = c;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if (threadIdx.x < s) { s1[threadIdx.x] += s1[threadIdx.x + s]; s2[threadIdx.x] += s2[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { ctrl->red[2] += s1[0]; ctrl->red[3] += s2[0]; }
}

__global__ void k_iter_end(nlls_ctrl* ctrl, int solver, double ftol, nlls_track tr) {
  if (ctrl->done) return;
  const int it = ctrl->iters;
  if (tr.retire) tr.count[it & 1] = 0;               // this iteration's list is consumed; it is the one after next's append target
  bool done;
  if (solver == NLLS_SOLVER_LM) {
    const bool gate = ctrl->red[1] > 0.0;            // (rho > 0).sum() > 0  (lma:110, Q10)
    done = (gate ? ctrl->red[0] : ctrl->red[2]) == 0.0;
  } else if (solver == NLLS_SOLVER_GD) {
    const bool fcon = (it > 0) && (ctrl->red[2] < ftol * ftol * ctrl->red[3]);   // batch-global (Q5)
    done = fcon || ctrl->red[0] == 0.0;
  } else {
    done = ctrl->red[0] == 0.0;
  }
  ctrl->iters = it + 1;
  ctrl->done = done ? 1 : 0;
  ctrl->red[0] = ctrl->red[1] = ctrl->red[2] = ctrl->red[3] = 0.0;
}

// Multi-GPU "exact" mode without a collective library call: the batch-global stop test needs the SUM over the ranks of
// the four fp64 reductions of ONE iteration (32 bytes).  Every rank owns a mailbox in peer-mapped (symmetric) memory,
// [4 slots][world][8 doubles]; thread t of this one-block kernel stores this rank's sums and then a stamp straight into
// rank t's mailbox over NVLink (P2P stores, data -> __threadfence_system -> stamp), then spins on the stamp of rank t in
// its OWN mailbox, reads that rank's sums, and thread 0 adds them in rank order (identical bits on every rank) before the
// usual decision.  slot = 2 * (epoch & 1) + (iteration & 1): a rank can run at most one iteration ahead of a peer (it
// needs that peer's stamp of the previous iteration to get there), and at most one SOLVE ahead (its first iteration of
// solve e+1 needs the peer's stamp of that iteration, which the peer's stream posts only after it finished solve e), so
// neither consecutive iterations nor consecutive solves ever share a slot.  The stamp carries the per-solve epoch so
// that a slot left over from an earlier solve can never match.  A peer that does not answer within timeout_ns sets the
// sticky ctrl->error and ctrl->done (every later kernel of the solve returns at once; the host raises after its flag
// read) instead of trapping: a trap would destroy the context and every result of the rank that waited.
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void k_iter_end_peer(nlls_ctrl* ctrl, int solver, double ftol, double* const* __restrict__ mbox, int rank, int world,
                                long long epoch, unsigned long long timeout_ns) {
  if (ctrl->done) return;
  const int it = ctrl->iters, slot = (int)((epoch & 1) << 1) | (it & 1), t = threadIdx.x;
  const double stamp = (double)(epoch * 1048576LL + it + 1);
  __shared__ double part[4][32];
  __shared__ int timed_out;
  if (t == 0) timed_out = 0;
  __syncthreads();
  if (t < world) {
    volatile double* dst = mbox[t] + ((size_t)slot * world + rank) * 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = ctrl->red[i];
    __threadfence_system();
    dst[4] = stamp;
    volatile double* src = mbox[rank] + ((size_t)slot * world + t) * 8;
    const unsigned long long t0 = global_ns();
    bool ok = true;
    while (src[4] != stamp) {
      if (global_ns() - t0 > timeout_ns) { ok = false; break; }
      __nanosleep(64);
    }
    if (!ok) timed_out = 1;
    __threadfence_system();
#pragma unroll
    for (int i = 0; i < 4; ++i) part[i][t] = ok ? src[i] : 0.0;
  }
  __syncthreads();
  if (t != 0) return;
  if (timed_out) {                                   // sticky error; the solve ends here on this rank
    ctrl->error = NLLS_CTRL_E_PEER_TIMEOUT;
    ctrl->done = 1;
    return;
  }
  double red[4] = {0.0, 0.0, 0.0, 0.0};
  for (int r = 0; r < world; ++r)
#pragma unroll
    for (int i = 0; i < 4; ++i) red[i] += part[i][r];
  bool done;
  if (solver == NLLS_SOLVER_LM) {
    const bool gate = red[1] > 0.0;
    done = (gate ? red[0] : red[2]) == 0.0;
  } else if (solver == NLLS_SOLVER_GD) {
    const bool fcon = (it > 0) && (red[2] < ftol * ftol * red[3]);
    done = fcon || red[0] == 0.0;
  } else {
    done = red[0] == 0.0;
  }
  ctrl->iters = it + 1;
  ctrl->done = done ? 1 : 0;
  ctrl->red[0] = ctrl->red[1] = ctrl->red[2] = ctrl->red[3] = 0.0;
}

// =========================================================== warp-cooperative solve, 8 < P <= 32
// One warp per problem, lane i owns row i of the (damped) matrix in REGISTERS.  Left-looking
// Cholesky with fully unrolled static register indices; L[j][k] reaches the other lanes by a
// register shuffle, so the factorisation touches no memory: P(P-1)/2 x (2 SHFL + DFMA) plus one
// sqrt and one reciprocal per column.  Forward substitution is column-oriented on registers; the
// transposed access of the back substitution goes through a per-warp shared-memory copy of L
// (leading dimension PP + 1: conflict-free for both the row writes and the column reads).
// The r01d profile of the thread-per-problem kernel it replaces showed one resident warp per SM
// (139 KB of interleaved scratch per 32 problems): profiles/r01d_cfg4.txt.
// A non-positive / non-finite pivot sends the ORIGINAL matrix (kept in shared memory) through LU
// with partial pivoting on lane 0 (rare path: Marquardt overflow, singular Hessians).
__device__ inline bool lu_solve_rt(int P, double* M, int ld, double* b, double* x) {
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  bool nonsing = true;
  for (int k = 0; k < P; ++k) {
    int piv = k;
    double best = fabs(M[k * ld + k]);
    for (int i = k + 1; i < P; ++i) {
      const double a = fabs(M[i * ld + k]);
      if (a > best) { best = a; piv = i; }
    }
    if (piv != k) {
      for (int j = 0; j < P; ++j) { const double t = M[k * ld + j]; M[k * ld + j] = M[piv * ld + j]; M[piv * ld + j] = t; }
      const double t = b[k]; b[k] = b[piv]; b[piv] = t;
    }
    const double d = M[k * ld + k];
    nonsing = nonsing && (d != 0.0) && (fabs(d) < INF);
    const double inv = 1.0 / d;
    for (int i = k + 1; i < P; ++i) {
      const double m = M[i * ld + k] * inv;
      for (int j = k + 1; j < P; ++j) M[i * ld + j] -= m * M[k * ld + j];
      b[i] -= m * b[k];
    }
  }
  for (int i = P - 1; i >= 0; --i) {
    double sacc = b[i];
    for (int k = i + 1; k < P; ++k) sacc -= M[i * ld + k] * x[k];
    x[i] = sacc / M[i * ld + i];
  }
  return nonsing;
}

template <int PP>
__device__ __forceinline__ bool warp_chol_solve(double (&row)[PP], double rhs, double* Ls, int lane, double& x_out) {
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int LD = PP + 1;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);
  bool ok = true;
  double myinv = 1.0;
#pragma unroll
  for (int j = 0; j < PP; ++j) {
    double sacc = row[j];
#pragma unroll
    for (int k = 0; k < j; ++k) sacc = fma(-row[k], __shfl_sync(FULL, row[k], j), sacc);
    const double d = __shfl_sync(FULL, sacc, j);
    ok = ok && (d > 0.0) && (d < INF);
    const double dj = sqrt(d);
    const double inv = 1.0 / dj;
    if (lane == j) myinv = inv;
    row[j] = (lane == j) ? dj : sacc * inv;       // lanes < j: upper-triangle slot, never read
  }
  if (!ok) return false;                          // warp-uniform
  if (lane < PP) {
#pragma unroll
    for (int j = 0; j < PP; ++j) Ls[lane * LD + j] = row[j];
  }
  double bb = rhs;                                // L y = b
#pragma unroll
  for (int k = 0; k < PP; ++k) {
    if (lane == k) bb *= myinv;
    const double yk = __shfl_sync(FULL, bb, k);
    if (lane > k) bb = fma(-row[k], yk, bb);
  }
  __syncwarp();
#pragma unroll
  for (int k = PP - 1; k >= 0; --k) {             // L^T x = y
    if (lane == k) bb *= myinv;
    const double xk = __shfl_sync(FULL, bb, k);
    if (lane < k) bb = fma(-Ls[k * LD + lane], xk, bb);
  }
  __syncwarp();
  x_out = bb;
  return true;
}

enum { W_LM = 0, W_GN = 1, W_GD = 2 };
struct WideArgs {
  const void *g, *H, *u;
  void *p, *dscale, *h, *p_trial, *h_out, *p_hist;
  const double* stats;
  double l, ftol, ptol, gtol;
  long long B;
  int P, marquardt, with_masks;
  nlls_ctrl* ctrl;
  nlls_track tr;
};
constexpr int kWideWarps = 4;
template <int PP> constexpr size_t wide_smem_bytes() { return (size_t)kWideWarps * (2 * PP * (PP + 1) + 2 * PP) * 8; }

template <typename T> __device__ __forceinline__ T warp_nan_max(T v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v = nan_max(v, shfl_xor(v, m));
  return v;
}

template <typename T, int PP, int MODE>
__global__ void __launch_bounds__(kWideWarps * 32) k_solve_wide(WideArgs a) {
  if (is_done(a.ctrl)) return;
  using TH = typename std::conditional<MODE == W_GD, double, T>::type;
  constexpr int LD = PP + 1;
  extern __shared__ double sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* M = sm + (size_t)warp * (2 * PP * LD + 2 * PP);
  double* Ls = M + PP * LD;
  double* bs = Ls + PP * LD;
  double* xs = bs + PP;
  const int it = iter_index(a.ctrl);
  const int P = a.P;
  const bool act = lane < P;
  T* p = reinterpret_cast<T*>(a.p);
  int nc = 0, sing = 0;
  for (long long b = (long long)blockIdx.x * kWideWarps + warp; b < a.B; b += (long long)gridDim.x * kWideWarps) {
    if (track_frozen(a.tr, b)) {                     // retired problem (warp-uniform): only its snapshot repeats
      if (MODE != W_LM && act && a.p_hist) reinterpret_cast<T*>(a.p_hist)[((long long)it * a.B + b) * P + lane] = p[b * P + lane];
      continue;
    }
    const TH* Hb = reinterpret_cast<const TH*>(a.H) + b * P * P;
    T diag_add = T(0), off_add = T(0);
    if constexpr (MODE == W_LM) {
      T* ds = reinterpret_cast<T*>(a.dscale);
      T dd = ds[b * 2], doff = ds[b * 2 + 1];
      if (a.marquardt) {                             // lma:88 — D <- D * max_p max(H_pp, D_pp)   (Q6)
        T m = act ? nan_max((T)Hb[lane * P + lane], dd) : dd;
        m = warp_nan_max(m);
        dd = dd * m;
        doff = doff * m;
        __syncwarp();
        if (lane == 0) { ds[b * 2] = dd; ds[b * 2 + 1] = doff; }
      }
      const T ub = reinterpret_cast<const T*>(a.u)[b];
      diag_add = ub * dd;
      off_add = ub * doff;
    }
    double row[PP];
#pragma unroll
    for (int j = 0; j < PP; ++j) {
      double v = (lane == j) ? 1.0 : 0.0;            // identity padding keeps the factorisation well defined
      if (act && j < P) {
        TH hv = Hb[j * P + lane];                    // H is symmetric: column read = coalesced row read
        if constexpr (MODE == W_LM) hv = (TH)((T)hv + (lane == j ? diag_add : off_add));   // Hu formed in T (lma:97-98)
        v = (double)hv;
      }
      row[j] = v;
      if (lane < PP) M[lane * LD + j] = v;
    }
    const double rhs = act ? (double)reinterpret_cast<const TH*>(a.g)[b * P + lane] : 0.0;
    if (lane < PP) bs[lane] = rhs;
    __syncwarp();
    double x = 0.0;
    bool ok = warp_chol_solve<PP>(row, rhs, Ls, lane, x);
    if (!ok) {
      int ns = 0;
      if (lane == 0) ns = lu_solve_rt(P, M, LD, bs, xs) ? 1 : 0;
      ns = __shfl_sync(0xffffffffu, ns, 0);
      __syncwarp();
      x = act ? xs[lane] : 0.0;
      ok = ns != 0;
      __syncwarp();
    }
    if (!ok) sing = 1;
    if constexpr (MODE == W_LM) {
      if (act) {
        const T hi = -(T)x;                          // lma:98
        reinterpret_cast<T*>(a.h)[b * P + lane] = hi;
        reinterpret_cast<T*>(a.p_trial)[b * P + lane] = p[b * P + lane] + hi;
      }
    } else {
      const T l = (T)a.l;
      const T hb = act ? (T)x : T(0);
      T pb = T(0), gb = T(0);
      if (act) {
        pb = p[b * P + lane] - l * hb;               // gna:141 / gda:146 (in place, Q1)
        p[b * P + lane] = pb;
        if (a.h_out) reinterpret_cast<T*>(a.h_out)[b * P + lane] = hb;
        gb = (MODE == W_GD) ? hb / l : (T)reinterpret_cast<const TH*>(a.g)[b * P + lane];
        if (a.p_hist) reinterpret_cast<T*>(a.p_hist)[((long long)it * a.B + b) * P + lane] = pb;
      }
      if (a.with_masks) {
        T ga = fabs(gb);
        const T ga0 = __shfl_sync(0xffffffffu, ga, 0);   // padding lanes repeat lane 0 (identity of the NaN-propagating max)
        ga = act ? ga : ga0;
        const T gm = warp_nan_max(ga);
        const T hh = warp_sum(hb * hb), pp = warp_sum(pb * pb);
        const bool gcon = gm < (T)a.gtol;
        const bool pcon = sqrt(hh) < (T)a.ptol * ((T)a.ptol + sqrt(pp));
        bool fcon = false;
        if constexpr (MODE == W_GN) fcon = (it > 0) && (a.stats[b * 4 + 1] < a.ftol * a.ftol * a.stats[b * 4 + 2]);
        bool conv = gcon || pcon || fcon;
        if (a.tr.pstat) {                            // the problem's own test (GD: fcon on its own sums, gda:76 sums the batch)
          const bool fown = (MODE == W_GD) ? ((it > 0) && (a.stats[b * 4 + 1] < a.ftol * a.ftol * a.stats[b * 4 + 2])) : fcon;
          const bool own = gcon || pcon || fown;
          const bool bad = __any_sync(0xffffffffu, not_finite(hb));
          if (lane == 0) {
            track_record(a.tr, b, own, (gcon ? 1 : 0) | (pcon ? 2 : 0) | (fown ? 4 : 0), 1, bad, it);
            if (a.tr.retire && !own) a.tr.active[(size_t)((it + 1) & 1) * a.tr.B + atomicAdd(a.tr.count + ((it + 1) & 1), 1)] = (int)b;
          }
          if (a.tr.retire) conv = own;
        }
        nc += !conv;
      }
    }
  }
  if (a.ctrl && lane == 0) {
    if (sing) a.ctrl->singular = 1;
    if (nc) atomicAdd(&a.ctrl->red[0], (double)nc);  // integer-valued: order-independent
  }
}

template <typename T, int MODE>
static int launch_wide(const WideArgs& a, cudaStream_t st, const char* what) {
  auto go = [&](auto kern, size_t smem) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int per_sm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kWideWarps * 32, smem);
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)kNumSMs * per_sm;
    const long long need = (a.B + kWideWarps - 1) / kWideWarps;
    if (grid > need) grid = need;
    kern<<<(int)grid, kWideWarps * 32, smem, st>>>(a);
  };
  if (a.P <= 16) go(k_solve_wide<T, 16, MODE>, wide_smem_bytes<16>());
  else go(k_solve_wide<T, 32, MODE>, wide_smem_bytes<32>());
  return check_launch(what);
}

template <typename K>
static void allow_smem(K kern, size_t smem) {
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

}  // namespace nlls

using namespace nlls;

static int check_bp(int dtype, long long B, int P) {
  if (dtype != NLLS_F32 && dtype != NLLS_F64) { set_error("dtype %d is not NLLS_F32/NLLS_F64", dtype); return NLLS_E_DTYPE; }
  if (B < 0 || P < 1 || P > NLLS_MAX_P) { set_error("unsupported shape B=%lld P=%d", B, P); return NLLS_E_SHAPE; }
  return 0;
}

extern "C" int nlls_lm_solve(int dtype, const void* g, const void* H, const void* p, const void* u, void* dscale,
                             int marquardt, long long B, int P, void* h, void* p_trial, nlls_ctrl* ctrl, const nlls_track* track,
                             void* stream) {
  const nlls_track tr = track_or_none(track);
  if (int e = check_bp(dtype, B, P)) return e;
  if (B == 0) return 0;
  if (!g || !H || !p || !u || !dscale || !h || !p_trial) { set_error("nlls_lm_solve: NULL tensor"); return NLLS_E_NULL; }
  if (P > 8 && !g_force_direct) {
    WideArgs a{};
    a.g = g; a.H = H; a.u = u; a.p = const_cast<void*>(p); a.dscale = dscale; a.h = h; a.p_trial = p_trial;
    a.B = B; a.P = P; a.marquardt = marquardt; a.ctrl = ctrl; a.tr = tr;
    return dtype == NLLS_F64 ? launch_wide<double, W_LM>(a, (cudaStream_t)stream, "nlls_lm_solve(wide)")
                             : launch_wide<float, W_LM>(a, (cudaStream_t)stream, "nlls_lm_solve(wide)");
  }
  const int tpb = threads_for(P);
  const size_t smem = scratch_bytes_per_thread(P) * tpb;
  const int grid = (int)((B + tpb - 1) / tpb);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == NLLS_F64) {
    allow_smem(k_lm_solve<double>, smem);
    k_lm_solve<double><<<grid, tpb, smem, st>>>((const double*)g, (const double*)H, (const double*)p, (const double*)u,
                                                (double*)dscale, marquardt, B, P, (double*)h, (double*)p_trial, ctrl, tr);
  } else {
    allow_smem(k_lm_solve<float>, smem);
    k_lm_solve<float><<<grid, tpb, smem, st>>>((const float*)g, (const float*)H, (const float*)p, (const float*)u,
                                               (float*)dscale, marquardt, B, P, (float*)h, (float*)p_trial, ctrl, tr);
  }
  return check_launch("nlls_lm_solve");
}

extern "C" int nlls_lm_update(int dtype, void* p, const void* h, const void* g, void* u, void* v, const double* stats,
                              const double* stat_h, int meth, double rho1, double rho2, double beta, double gama,
                              double ftol, double ptol, double gtol, long long B, int P, void* p_hist, void* rho_out,
                              nlls_ctrl* ctrl, const nlls_track* track, void* stream) {
  const nlls_track tr = track_or_none(track);
  if (int e = check_bp(dtype, B, P)) return e;
  if (B == 0) return 0;
  if (!p || !h || !g || !u || !v || !stats || !stat_h) { set_error("nlls_lm_update: NULL tensor"); return NLLS_E_NULL; }
  const int grid = (int)((B + kUpdTPB - 1) / kUpdTPB);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == NLLS_F64)
    k_lm_update<double><<<grid, kUpdTPB, 0, st>>>((double*)p, (const double*)h, (const double*)g, (double*)u, (double*)v,
                                                  stats, stat_h, meth, rho1, rho2, beta, gama, ftol, ptol, gtol, B, P,
                                                  (double*)p_hist, (double*)rho_out, ctrl, tr);
  else
    k_lm_update<float><<<grid, kUpdTPB, 0, st>>>((float*)p, (const float*)h, (const float*)g, (float*)u, (float*)v, stats,
                                                 stat_h, meth, (float)rho1, (float)rho2, (float)beta, (float)gama, ftol,
                                                 (float)ptol, (float)gtol, B, P, (float*)p_hist, (float*)rho_out, ctrl, tr);
  return check_launch("nlls_lm_update");
}

extern "C" int nlls_gn_update(int dtype, void* p, const void* g, const void* H, const double* stats, double l, double ftol,
                              double ptol, double gtol, int with_masks, long long B, int P, void* h_out, void* p_hist,
                              nlls_ctrl* ctrl, const nlls_track* track, void* stream) {
  const nlls_track tr = track_or_none(track);
  if (int e = check_bp(dtype, B, P)) return e;
  if (B == 0) return 0;
  if (!p || !g || !H || (with_masks && !stats)) { set_error("nlls_gn_update: NULL tensor"); return NLLS_E_NULL; }
  if (P > 8 && !g_force_direct) {
    WideArgs a{};
    a.g = g; a.H = H; a.p = p; a.h_out = h_out; a.p_hist = p_hist; a.stats = stats; a.l = l; a.ftol = ftol; a.ptol = ptol;
    a.gtol = gtol; a.B = B; a.P = P; a.with_masks = with_masks && ctrl; a.ctrl = ctrl; a.tr = tr;
    return dtype == NLLS_F64 ? launch_wide<double, W_GN>(a, (cudaStream_t)stream, "nlls_gn_update(wide)")
                             : launch_wide<float, W_GN>(a, (cudaStream_t)stream, "nlls_gn_update(wide)");
  }
  const int tpb = threads_for(P);
  const size_t smem = scratch_bytes_per_thread(P) * tpb;
  const int grid = (int)((B + tpb - 1) / tpb);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == NLLS_F64) {
    allow_smem(k_gn_update<double>, smem);
    k_gn_update<double><<<grid, tpb, smem, st>>>((double*)p, (const double*)g, (const double*)H, stats, l, ftol, ptol, gtol,
                                                 with_masks, B, P, (double*)h_out, (double*)p_hist, ctrl, tr);
  } else {
    allow_smem(k_gn_update<float>, smem);
    k_gn_update<float><<<grid, tpb, smem, st>>>((float*)p, (const float*)g, (const float*)H, stats, (float)l, ftol,
                                                (float)ptol, (float)gtol, with_masks, B, P, (float*)h_out, (float*)p_hist,
                                                ctrl, tr);
  }
  return check_launch("nlls_gn_update");
}

extern "C" int nlls_gd_update(int dtype, void* p, const double* g64, const double* H64, const double* stats, double l,
                              double ftol, double ptol, double gtol, int with_masks, long long B, int P, void* h_out,
                              void* p_hist, nlls_ctrl* ctrl, const nlls_track* track, void* stream) {
  const nlls_track tr = track_or_none(track);
  if (int e = check_bp(dtype, B, P)) return e;
  if (B == 0) return 0;
  if (!p || !g64 || !H64 || (with_masks && (!stats || !ctrl))) { set_error("nlls_gd_update: NULL tensor"); return NLLS_E_NULL; }
  const int tpb = threads_for(P);
  const size_t smem = scratch_bytes_per_thread(P) * tpb;
  const int grid = (int)((B + tpb - 1) / tpb);
  cudaStream_t st = (cudaStream_t)stream;
  if (P > 8 && !g_force_direct) {
    WideArgs a{};
    a.g = g64; a.H = H64; a.p = p; a.h_out = h_out; a.p_hist = p_hist; a.stats = stats; a.l = l; a.ftol = ftol; a.ptol = ptol;
    a.gtol = gtol; a.B = B; a.P = P; a.with_masks = with_masks; a.ctrl = ctrl; a.tr = tr;
    if (int e = dtype == NLLS_F64 ? launch_wide<double, W_GD>(a, st, "nlls_gd_update(wide)")
                                  : launch_wide<float, W_GD>(a, st, "nlls_gd_update(wide)")) return e;
  } else if (dtype == NLLS_F64) {
    allow_smem(k_gd_update<double>, smem);
    k_gd_update<double><<<grid, tpb, smem, st>>>((double*)p, g64, H64, stats, l, ftol, ptol, gtol, with_masks, B, P, (double*)h_out,
                                                 (double*)p_hist, ctrl, tr);
  } else {
    allow_smem(k_gd_update<float>, smem);
    k_gd_update<float><<<grid, tpb, smem, st>>>((float*)p, g64, H64, stats, (float)l, ftol, (float)ptol, (float)gtol, with_masks, B, P,
                                                (float*)h_out, (float*)p_hist, ctrl, tr);
  }
  if (int e = check_launch("nlls_gd_update")) return e;
  if (with_masks && !tr.retire) {                   // retire mode: fcon is per problem (k_gd_update), no batch sums
    k_reduce_stats<<<1, 1024, 0, st>>>(stats, B, ctrl);
    return check_launch("nlls_gd_update(reduce)");
  }
  return 0;
}

// ---------------------------------------------------------------------------- residual statistics of ONE problem
// Single-problem solvers (functions/single/*): sums over a residual vector of any length, spread over the whole GPU.
// Block k writes out[2k] = partial sum of a^2 and out[2k+1] = partial sum of (b - a)^2 (0 when b is NULL) in fp64;
// fixed grid and fixed per-block order: the host adds the kSumsqBlocks partials, so the result is deterministic.
constexpr int kSumsqBlocks = 1184, kSumsqTPB = 256;     // 148 SMs x 8

template <typename T>
__global__ void __launch_bounds__(kSumsqTPB)
k_sumsq(const T* __restrict__ a, const T* __restrict__ b, long long n, double* __restrict__ out) {
  double s0 = 0.0, s1 = 0.0;
  for (long long i = (long long)blockIdx.x * kSumsqTPB + threadIdx.x; i < n; i += (long long)kSumsqBlocks * kSumsqTPB) {
    const double x = (double)a[i];
    s0 = fma(x, x, s0);
    if (b) { const double d = (double)(T)(b[i] - a[i]); s1 = fma(d, d, s1); }   // difference rounded in T like the reference
  }
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  __shared__ double red[2][kSumsqTPB / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = s0; red[1][warp] = s1; }
  __syncthreads();
  if (threadIdx.x < 2) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kSumsqTPB / 32; ++k) t += red[threadIdx.x][k];
    out[2 * blockIdx.x + threadIdx.x] = t;
  }
}

extern "C" int nlls_sumsq_partials(void) { return kSumsqBlocks; }

extern "C" int nlls_sumsq(int dtype, const void* a, const void* b, long long n, double* out, void* stream) {
  if (dtype != NLLS_F32 && dtype != NLLS_F64) { set_error("dtype %d is not NLLS_F32/NLLS_F64", dtype); return NLLS_E_DTYPE; }
  if (n < 0) { set_error("nlls_sumsq: n=%lld", n); return NLLS_E_SHAPE; }
  if (!a || !out) { set_error("nlls_sumsq: NULL tensor"); return NLLS_E_NULL; }
  if (dtype == NLLS_F64) k_sumsq<double><<<kSumsqBlocks, kSumsqTPB, 0, (cudaStream_t)stream>>>((const double*)a, (const double*)b, n, out);
  else k_sumsq<float><<<kSumsqBlocks, kSumsqTPB, 0, (cudaStream_t)stream>>>((const float*)a, (const float*)b, n, out);
  return check_launch("nlls_sumsq");
}

extern "C" int nlls_iter_end(nlls_ctrl* ctrl, int solver, double ftol, const nlls_track* track, void* stream) {
  if (!ctrl) { set_error("nlls_iter_end: NULL ctrl"); return NLLS_E_NULL; }
  k_iter_end<<<1, 1, 0, (cudaStream_t)stream>>>(ctrl, solver, ftol, track_or_none(track));
  return check_launch("nlls_iter_end");
}

extern "C" int nlls_iter_end_peer(nlls_ctrl* ctrl, int solver, double ftol, void* const* mailboxes, int rank, int world,
                                  long long epoch, double timeout_s, void* stream) {
  if (!ctrl || !mailboxes) { set_error("nlls_iter_end_peer: NULL pointer"); return NLLS_E_NULL; }
  if (world < 1 || world > 32 || rank < 0 || rank >= world) { set_error("nlls_iter_end_peer: rank %d of %d", rank, world); return NLLS_E_SHAPE; }
  if (!(timeout_s > 0.0)) timeout_s = 60.0;
  k_iter_end_peer<<<1, 32, 0, (cudaStream_t)stream>>>(ctrl, solver, ftol, (double* const*)mailboxes, rank, world, epoch,
                                                      (unsigned long long)(timeout_s * 1e9));
  return check_launch("nlls_iter_end_peer");
}

extern "C" int nlls_lm_update_spec(int dtype, void* p, const void* h, void* g, void* H, const void* g_t, const void* H_t,
                                   void* u, void* v, double* stats, const double* stats_t, unsigned char* sel, int meth,
                                   double rho1, double rho2, double beta, double gama, double ftol, double ptol, double gtol,
                                   long long B, int P, void* p_hist, void* rho_out, nlls_ctrl* ctrl, const nlls_track* track,
                                   void* stream) {
  const nlls_track tr = track_or_none(track);
  if (int e = check_bp(dtype, B, P)) return e;
  if (B == 0) return 0;
  if (!p || !h || !g || !H || !g_t || !H_t || !u || !v || !stats || !stats_t || !sel) {
    set_error("nlls_lm_update_spec: NULL tensor"); return NLLS_E_NULL;
  }
  const int grid = (int)((B + kUpdTPB - 1) / kUpdTPB);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == NLLS_F64)
    k_lm_update_spec<double><<<grid, kUpdTPB, 0, st>>>((double*)p, (const double*)h, (double*)g, (double*)H, (const double*)g_t,
                                                       (const double*)H_t, (double*)u, (double*)v, stats, stats_t, sel, meth,
                                                       rho1, rho2, beta, gama, ftol, ptol, gtol, B, P, (double*)p_hist,
                                                       (double*)rho_out, ctrl, tr);
  else
    k_lm_update_spec<float><<<grid, kUpdTPB, 0, st>>>((float*)p, (const float*)h, (float*)g, (float*)H, (const float*)g_t,
                                                      (const float*)H_t, (float*)u, (float*)v, stats, stats_t, sel, meth,
                                                      (float)rho1, (float)rho2, (float)beta, (float)gama, ftol, (float)ptol,
                                                      (float)gtol, B, P, (float*)p_hist, (float*)rho_out, ctrl, tr);
  return check_launch("nlls_lm_update_spec");
}
