// fused_models.cu — K4: residual + Jacobian of the built-in models fused into the normal-equation
// reduction, so that J [B,C,N,P] is never materialised and the iteration is not bound by Python
// callables (north_star item 4).
//
// Per problem b the model (m, k = -dm/dp) is evaluated ONCE per abscissa sample n and shared by
// all C costs; with r_c = y_c - m:
//     f_c = r_c^2,  J_c = 2 r_c k
//     g   = sum_n k   * (2 sum_c w_c r_c^3)
//     H   = sum_n k k^T (4 sum_c w_c r_c^2)
// (newton_parallel.py:40-43 applied to tests/unit_test_parallel.py:53-65,105-118).
//
// The stop test fcon (lma:110, gna:80) needs sum (f_prev - f)^2 sample by sample.  f_prev[b,c,n] was produced one
// iteration earlier as mul_rn(y_c - m_prev, y_c - m_prev): it depends on the data y_c (in the stage anyway) and ONE
// scalar m_prev[b,n] shared by all C costs.  The kernels therefore keep the MODEL VALUE m [B,N] between iterations
// (ping-pong), not the residuals f [B,C,N], and rebuild f_prev with one subtraction and one multiplication per
// (c,n) — the same two roundings that produced it, so the sums are bit-identical to those of stored residuals.
// HBM traffic per problem-iteration: y once (s*Cy*N), m_prev read + m written (2*s*N): s*N*(C+2) instead of 3*s*C*N.
#include <cstdlib>
#include <type_traits>
#include "models.cuh"
#include "rowacc.cuh"

// Plugin build (torchimize_b200.models.compile_model): the same kernels and the same entry points, instantiated for ONE
// user-supplied device functor nlls::UserModel<T> instead of the built-in models.  The functor follows models.cuh:
//   static constexpr int P;  void load(const T* p);  T value(T x) const;  void eval(T x, T& m, T (&k)[P]) const  (k = -dm/dp)
// and may use every helper of models.cuh (ExpNeg, one_plus_erf, ...).  value() and eval() have to return the same bits for m.
#ifdef NLLS_USER_MODEL_HEADER
#include NLLS_USER_MODEL_HEADER
namespace nlls { int g_force_direct = 0; }
extern "C" int nlls_user_model_params(void) { return nlls::UserModel<double>::P; }
extern "C" int nlls_set_force_direct(int on) { const int old = nlls::g_force_direct; nlls::g_force_direct = on; return old; }
#endif

namespace nlls {

constexpr int kFusedTPB = 128;
constexpr int kFusedS = 4;       // abscissa samples per thread per tile

template <typename T>
__device__ __forceinline__ T ld_cached(const T* p) { return __ldg(p); }
__device__ __forceinline__ bool aligned_to_dev(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// ------------------------------------------------------------------ fused normal equations
template <typename T, typename Model, int TPB, int S>
__global__ void __launch_bounds__(TPB)
k_model_normal_eq(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                  long long ycs, const T* __restrict__ w, int wlen, long long B, int C, int N, T* mA, T* mB,
                  T* __restrict__ g, T* __restrict__ H, double* __restrict__ stats, const nlls_ctrl* ctrl,
                  const unsigned char* __restrict__ sel, nlls_track tr) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2;
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  // sel == NULL: buffers alternate with the iteration parity.  sel != NULL (speculative LM trial evaluation):
  // problem b's current model values live in (sel[b] ? mB : mA) and the trial values go to the other buffer.
  T* m_out = (it & 1) ? mB : mA;
  const T* m_prev = (it > 0 && mA) ? ((it & 1) ? mA : mB) : nullptr;
  extern __shared__ double smem[];
  double* red = smem;                  // [NW][V]
  double* sfc = smem + NW * V;         // [C][TPB] per-thread per-cost sums of f
  double* sfw = sfc + (size_t)C * TPB; // [C][NW]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (long long wi = blockIdx.x; wi < WL.n; wi += gridDim.x) {
    const long long b = WL.at(wi);
    if (sel) { const bool sb = sel[b] != 0; m_prev = sb ? mB : mA; m_out = sb ? mA : mB; }
    Model M;
    M.load(p + b * P);
    RowAcc<T, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    for (int c = 0; c < C; ++c) sfc[c * TPB + tid] = 0.0;
    const T* yb = y + b * ybs;
    for (int n0 = 0; n0 < N; n0 += TPB * S) {
      T m[S], mp[S], k[S][P], s2[S], s3[S];
#pragma unroll
      for (int i = 0; i < S; ++i) {
        const int n = n0 + i * TPB + tid;
        M.eval(ld_cached(x + min(n, N - 1)), m[i], k[i]);
        s2[i] = T(0);
        s3[i] = T(0);
        mp[i] = T(0);
        if (n < N) {
          if (m_prev) mp[i] = ldg_stream(m_prev + b * N + n);
          if (m_out) stg_stream(m_out + b * N + n, m[i]);
        }
      }
#pragma unroll 2
      for (int c = 0; c < C; ++c) {
        const T wc = w[wlen == 1 ? 0 : c];
        const T* yc = yb + c * ycs;
        double sfl = 0.0;
#pragma unroll
        for (int i = 0; i < S; ++i) {
          const int n = n0 + i * TPB + tid;
          if (n < N) {
            const T yv = ld_cached(yc + n);
            const T r = yv - m[i];
            const T fv = mul_rn(r, r);
            if (m_prev) {                                    // f_prev rebuilt from the stored model value (bit-identical)
              const T rp = yv - mp[i];
              const T d = mul_rn(rp, rp) - fv;
              sd = fma((double)d, (double)d, sd);
            }
            sq = fma((double)fv, (double)fv, sq);
            sfl += (double)fv;
            const T wf = wc * fv;
            s2[i] += wf;
            s3[i] = fma(wf, r, s3[i]);
          }
        }
        sfc[c * TPB + tid] += sfl;
      }
#pragma unroll
      for (int i = 0; i < S; ++i) {
        const int n = n0 + i * TPB + tid;
        if (n < N) acc.add_scaled(k[i], T(2) * s3[i], T(4) * s2[i]);
      }
    }
    acc.warp_reduce();
    sd = warp_sum(sd);
    sq = warp_sum(sq);
    for (int c = 0; c < C; ++c) {
      const double s = warp_sum(sfc[c * TPB + tid]);
      if (lane == 0) sfw[c * NW + warp] = s;
    }
    if (lane == 0) {
      double* r = red + warp * V;
#pragma unroll
      for (int i = 0; i < P; ++i) r[i] = (double)acc.g[i];
#pragma unroll
      for (int i = 0; i < NH; ++i) r[P + i] = (double)acc.h[i];
      r[P + NH] = sd;
      r[P + NH + 1] = sq;
    }
    __syncthreads();
    if (tid < V) {
      double s = 0.0;
#pragma unroll
      for (int kk = 0; kk < NW; ++kk) s += red[kk * V + tid];
      if (tid < P) {
        g[b * P + tid] = (T)s;
      } else if (tid < P + NH) {
        int kk = tid - P, i = 0;
        while (kk >= P - i) { kk -= P - i; ++i; }
        const int j = i + kk;
        H[b * P * P + i * P + j] = (T)s;
        H[b * P * P + j * P + i] = (T)s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    } else if (tid == V) {
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
        for (int kk = 0; kk < NW; ++kk) s += sfw[c * NW + kk];
        tot = fma(s, s, tot);
      }
      stats[b * 4 + 0] = tot;
      stats[b * 4 + 3] = 0.0;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ shared tile ring (bulk-copy producer)
// One elected thread streams, NSTAGE tiles ahead and across problem boundaries, up to two row
// sets per tile into a shared-memory stage: set A = nA rows a[b*a_bs + r*a_rs + n0 .. +rows],
// set B = nB rows likewise (nB = 0 when absent).  Stage layout: [nA][TN] then [nB][TN] elements.
template <typename T, int TN, int NSTAGE>
struct TileRing {
  T* ring;
  uint64_t* bars;
  const T *a, *bsrc;
  long long a_bs, a_rs, b_bs, b_rs;
  int nA, nB, N, stage_elems;
  const CUtensorMap* tm; // set A through ONE tensor copy per tile (N % TN == 0), or NULL = one bulk copy per row
  const int* wlist;      // work list (retire mode) or NULL = every problem
  long long wn, pw;      // work items, producer: next work index
  long long pb;          // producer: next problem
  int pn0, pst;          // producer: next sample offset, next stage
  int st;                // consumer: current stage
  unsigned ph;           // consumer: current parity
  __device__ __forceinline__ void init(unsigned char* smem, const T* a_, long long a_bs_, long long a_rs_, int nA_,
                                       const T* b_, long long b_bs_, long long b_rs_, int nB_, const WorkList& WL, int N_,
                                       const CUtensorMap* tm_ = nullptr) {
    tm = tm_;
    a = a_; a_bs = a_bs_; a_rs = a_rs_; nA = nA_;
    bsrc = b_; b_bs = b_bs_; b_rs = b_rs_; nB = b_ ? nB_ : 0;
    wlist = WL.list; wn = WL.n; N = N_;
    stage_elems = (nA_ + nB_) * TN;                          // fixed layout even when set B is absent
    ring = reinterpret_cast<T*>(smem);
    size_t off = ((size_t)NSTAGE * stage_elems * sizeof(T) + 15) & ~(size_t)15;
    bars = reinterpret_cast<uint64_t*>(smem + off);
    pw = blockIdx.x; pb = pw < wn ? (wlist ? (long long)wlist[pw] : pw) : 0; pn0 = 0; pst = 0; st = 0; ph = 0;
    if (threadIdx.x == 0) {
#pragma unroll
      for (int s = 0; s < NSTAGE; ++s) mbar_init(&bars[s], 1);
      mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
      for (int s = 0; s < NSTAGE; ++s) issue();
    }
  }
  static __host__ __device__ size_t bytes(int nA_, int nB_) {   // ring + barriers, 16-byte granular
    return (((size_t)NSTAGE * (nA_ + nB_) * TN * sizeof(T) + 15) & ~(size_t)15) + 64;
  }
  __device__ __forceinline__ void issue() {                  // thread 0 only, after the stage was handed back
    if (pw >= wn) return;
    const int rows = min(TN, N - pn0);
    const unsigned nbytes = (unsigned)rows * (unsigned)sizeof(T);
    T* dst = ring + (size_t)pst * stage_elems;
    uint64_t* bar = &bars[pst];
    mbar_expect_tx(bar, nbytes * (unsigned)(nA + nB));
    if (tm) {
      tma_load_3d(dst, tm, pn0, 0, (int)pb, bar);
      dst += nA * TN;
    } else {
      const T* src = a + pb * a_bs + pn0;
      for (int r = 0; r < nA; ++r, src += a_rs, dst += TN) bulk_g2s(dst, src, nbytes, bar);
    }
    if (nB) {
      const T* s2 = bsrc + pb * b_bs + pn0;
      for (int r = 0; r < nB; ++r, s2 += b_rs, dst += TN) bulk_g2s(dst, s2, nbytes, bar);
    }
    pn0 += TN;
    if (pn0 >= N) { pn0 = 0; pw += gridDim.x; pb = pw < wn ? (wlist ? (long long)wlist[pw] : pw) : 0; }
    pst = (pst + 1 == NSTAGE) ? 0 : pst + 1;
  }
  __device__ __forceinline__ const T* wait() {               // all threads; returns the stage base
    mbar_wait(&bars[st], ph);
    return ring + (size_t)st * stage_elems;
  }
  __device__ __forceinline__ void advance() { if (++st == NSTAGE) { st = 0; ph ^= 1u; } }
};

// ------------------------------------------------------------------ fused normal equations, bulk-copy staged
// Same arithmetic as k_model_normal_eq, but the data rows y[b,c,n0:n0+TN] and the row m_prev[b,n0:n0+TN]
// of a tile are brought into shared memory by 1-D bulk async copies (cp.async.bulk -> UBLKCP,
// completion on one mbarrier per stage) issued NSTAGE tiles ahead by one elected thread, across
// problem boundaries.  The r01a profile of the direct kernel showed 73 % of its stalls on the
// scoreboard of the y / previous-residual loads (profiles/r01a_fused_baseline.txt): the loads sat in the
// dependent chain of each sample.  Here the bytes in flight per SM no longer depend on registers
// or occupancy.  TN = TPB * S samples per tile.  One __syncthreads per tile (stage hand-back) and
// one per problem (cross-warp sums, double-buffered); the per-problem warp sums use the
// transposed reduction (16 shuffles for P = 4 instead of 80).
// CT > 0: the number of costs is a compile-time constant (dispatch: C <= 8): the cost loop is fully unrolled, the cost
// weights and the per-cost sums of f live in registers (no shared-memory read-modify-write per cost and tile) and the
// C * S residual chains of a tile are independent instruction streams for the scheduler.  CT == 0: any C at run time.
template <typename T, typename Model, int TPB, int S, int NSTAGE, int CT = 0, int MINB = 0>
__global__ void __launch_bounds__(TPB, MINB > 0 ? MINB : 512 / TPB)
k_model_normal_eq_staged(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                         long long ycs, const T* __restrict__ w, int wlen, long long B, int C_, int N, T* mA, T* mB,
                         T* __restrict__ g, T* __restrict__ H, double* __restrict__ stats, const nlls_ctrl* ctrl,
                         const unsigned char* __restrict__ sel, nlls_track tr, const __grid_constant__ CUtensorMap tmy,
                         int use_tm) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2, TN = TPB * S;
  constexpr int VP = Pow2Ceil<V>::value, OWN = 32 / VP;      // transposed reduction: lane L owns value L / OWN
  static_assert(V <= 32, "model with too many parameters for the transposed reduction");
  const int C = CT > 0 ? CT : C_;
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  T* m_out = (it & 1) ? mB : mA;
  const T* m_prev = (it > 0 && mA) ? ((it & 1) ? mA : mB) : nullptr;
  const bool has_prev = m_prev != nullptr || sel != nullptr;  // sel: per-problem buffer roles (see k_model_normal_eq)
  const int Cy = (ycs == 0) ? 1 : C;                         // one shared data row, or one per cost
  const int stage_elems = (Cy + 1) * TN;                     // [Cy][TN] data rows then [TN] previous model values
  extern __shared__ __align__(128) unsigned char smem_raw[];
  T* ring = reinterpret_cast<T*>(smem_raw);
  size_t off = (size_t)NSTAGE * stage_elems * sizeof(T);
  off = (off + 15) & ~(size_t)15;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + off);
  double* red = reinterpret_cast<double*>(smem_raw + off + 8 * 8);   // [2][NW][VP]
  double* sfc = red + 2 * NW * VP;                           // [2][C][TPB] per-thread per-cost sums of f
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncthreads();

  // ---- producer state (thread 0): next tile to issue = (problem pb, sample offset pn0) into stage pst.  Everything that
  // changes only with the problem (its previous-model row) or never (shared-window addresses, the byte count of a full tile)
  // is kept in registers: the per-tile path is two adds, the barrier arm and the copies.
  long long pw = blockIdx.x, pb = pw < WL.n ? WL.at(pw) : 0;   // producer: work index, problem
  int pn0 = 0;
  unsigned char psel = (sel && tid == 0 && pw < WL.n) ? sel[pb] : (unsigned char)0;
  const uint32_t ring_s = smem_u32(ring), bars_s = smem_u32(bars);
  constexpr uint32_t kTileBytes = (uint32_t)TN * (uint32_t)sizeof(T);
  const uint32_t stage_bytes = (uint32_t)stage_elems * (uint32_t)sizeof(T);
  const uint32_t full_tx = kTileBytes * (uint32_t)(Cy + (has_prev ? 1 : 0));
  uint32_t pdst = ring_s, pbar = bars_s;                       // stage pst: data address, barrier address
  int pst = 0;
  auto prev_row = [&]() -> const T* {                          // previous model values of problem pb (sel: its own buffer role)
    return has_prev ? (sel ? (psel ? mB : mA) : m_prev) + pb * (long long)N : nullptr;
  };
  const T* msrc = prev_row();
  auto issue = [&]() {
    if (pw >= WL.n) return;
    if (use_tm) {                                            // N % TN == 0: every tile is full; one tensor copy of the [Cy][TN] box
      mbar_expect_tx_s(pbar, full_tx);
      tma_load_3d_s(pdst, &tmy, pn0, 0, (int)pb, pbar);
      if (has_prev) bulk_g2s_s(pdst + (uint32_t)Cy * kTileBytes, msrc + pn0, kTileBytes, pbar);
    } else {
      const int rows = min(TN, N - pn0);
      const unsigned bytes = (unsigned)rows * (unsigned)sizeof(T);
      mbar_expect_tx_s(pbar, bytes * (unsigned)(Cy + (has_prev ? 1 : 0)));
      const T* src = y + pb * ybs + pn0;
      uint32_t dst = pdst;
      for (int c = 0; c < Cy; ++c, src += ycs, dst += kTileBytes) bulk_g2s_s(dst, src, bytes, pbar);
      if (has_prev) bulk_g2s_s(dst, msrc + pn0, bytes, pbar);
    }
    pn0 += TN;
    if (pn0 >= N) {
      pn0 = 0;
      pw += gridDim.x;
      pb = pw < WL.n ? WL.at(pw) : 0;
      if (sel && pw < WL.n) psel = sel[pb];                     // buffer role of the NEXT problem: in flight until its first tile
      msrc = prev_row();
    }
    if (++pst == NSTAGE) { pst = 0; pdst = ring_s; pbar = bars_s; }
    else { pdst += stage_bytes; pbar += 8; }
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) issue();
  }

  int st = 0, pbuf = 0;
  unsigned ph = 0;
  for (long long wi = blockIdx.x; wi < WL.n; wi += gridDim.x, pbuf ^= 1) {
    const long long b = WL.at(wi);
    Model M;
    M.load(p + b * P);
    if (tid == 0 && wi + gridDim.x < WL.n) prefetch_l1(p + WL.at(wi + gridDim.x) * P);   // next problem's parameters
    RowAcc<T, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    double* my_sfc = sfc + (size_t)pbuf * C * TPB + tid;
    [[maybe_unused]] double sfr[CT > 0 ? CT : 1];             // CT: per-cost sums of f in registers
    if constexpr (CT > 0) {
#pragma unroll
      for (int c = 0; c < CT; ++c) sfr[c] = 0.0;
    } else {
      for (int c = 0; c < C; ++c) my_sfc[c * TPB] = 0.0;
    }
    if (sel) m_out = sel[b] ? mA : mB;
    T* mo_b = m_out ? m_out + b * (long long)N : nullptr;
    // FULL tiles (n0 + TN <= N) carry no per-sample bounds predicates: the tile body is one straight-line block in which
    // the S model evaluations and the C * S residual chains interleave.
    auto tile = [&](int n0, auto full_tag) {
      constexpr bool FULL = decltype(full_tag)::value;
      T xv[S], m[S], k[S][P], s2[S], s3[S];
#pragma unroll
      for (int i = 0; i < S; ++i) {
        const int n = n0 + i * TPB + tid;
        xv[i] = ld_cached(x + (FULL ? n : min(n, N - 1)));
        s2[i] = T(0);
        s3[i] = T(0);
      }
      model_eval_n<Model, T, S>(M, xv, m, k);               // overlaps the wait below
      if (mo_b) {
#pragma unroll
        for (int i = 0; i < S; ++i)
          if (FULL || n0 + i * TPB + tid < N) stg_stream(mo_b + n0 + i * TPB + tid, m[i]);
      }
      mbar_wait(&bars[st], ph);
      const T* yc = ring + (size_t)st * stage_elems + tid;
      const int ystep = (Cy == 1) ? 0 : TN;
      T mp[S];                                               // without a previous iterate: m itself (f_prev - f == 0 exactly)
#pragma unroll
      for (int i = 0; i < S; ++i) mp[i] = has_prev ? yc[Cy * TN + i * TPB] : m[i];   // rows past N: stale, masked below
      auto cost = [&](const T* yrow, T wc) -> double {        // one cost of this tile; returns the thread's sum of f
        double sfl = 0.0;
#pragma unroll
        for (int i = 0; i < S; ++i) {
          if (FULL || n0 + i * TPB + tid < N) {
            const T yv = yrow[i * TPB];
            const T r = yv - m[i];
            const T fv = mul_rn(r, r);
            const T rp = yv - mp[i];                         // f_prev rebuilt from the stored model value (bit-identical)
            const T d = mul_rn(rp, rp) - fv;
            sd = fma((double)d, (double)d, sd);
            sq = fma((double)fv, (double)fv, sq);
            sfl += (double)fv;
            const T wf = wc * fv;
            s2[i] += wf;
            s3[i] = fma(wf, r, s3[i]);
          }
        }
        return sfl;
      };
      if constexpr (CT > 0) {
#pragma unroll
        for (int c = 0; c < CT; ++c) sfr[c] += cost(yc + c * ystep, w[wlen == 1 ? 0 : c]);
      } else {
#pragma unroll 2
        for (int c = 0; c < C; ++c, yc += ystep) my_sfc[c * TPB] += cost(yc, w[wlen == 1 ? 0 : c]);
      }
#pragma unroll
      for (int i = 0; i < S; ++i)
        if (FULL || n0 + i * TPB + tid < N) acc.add_scaled(k[i], T(2) * s3[i], T(4) * s2[i]);
    };
    for (int n0 = 0; n0 < N; n0 += TN) {
      if (n0 + TN <= N) tile(n0, std::true_type{});
      else tile(n0, std::false_type{});
      if (n0 + TN < N) {
        if constexpr (NW == 1) __syncwarp();                 // every thread is done with this stage (one warp: no CTA barrier)
        else __syncthreads();
        if (tid == 0) issue();
      }
      if (++st == NSTAGE) { st = 0; ph ^= 1u; }
    }
    if constexpr (CT > 0) {
#pragma unroll
      for (int c = 0; c < CT; ++c) my_sfc[c * TPB] = sfr[c];
    }
    // ---- per-problem sums: transposed warp reduction, then fp64 across warps through smem
    double v[VP];
#pragma unroll
    for (int i = 0; i < P; ++i) v[i] = (double)acc.g[i];
#pragma unroll
    for (int i = 0; i < NH; ++i) v[P + i] = (double)acc.h[i];
    v[P + NH] = sd;
    v[P + NH + 1] = sq;
#pragma unroll
    for (int i = V; i < VP; ++i) v[i] = 0.0;
    warp_reduce_transpose<VP>(v);
    double* myred = red + (size_t)pbuf * NW * VP;
    if ((lane & (OWN - 1)) == 0) myred[warp * VP + lane / OWN] = v[0];
    __syncthreads();                                         // last tile's stage is free; red and sfc are complete
    if (tid == 0) issue();
    auto write_sums = [&]() {                                // threads 0 .. V-1: g, H, stats[1..2]
      double s = 0.0;
#pragma unroll
      for (int kk = 0; kk < NW; ++kk) s += myred[kk * VP + tid];
      if (tid < P) {
        g[b * P + tid] = (T)s;
      } else if (tid < P + NH) {
        int kk = tid - P, i = 0;
        while (kk >= P - i) { kk -= P - i; ++i; }
        const int j = i + kk;
        H[b * P * P + i * P + j] = (T)s;
        H[b * P * P + j * P + i] = (T)s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    };
    auto write_cost = [&]() {                                // one whole warp: sum_c (sum_n f)^2 from the per-thread sums
      const double* col = sfc + (size_t)pbuf * C * TPB;
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
#pragma unroll
        for (int kk = 0; kk < NW; ++kk) s += col[c * TPB + kk * 32 + lane];
        s = warp_sum(s);
        tot = fma(s, s, tot);
      }
      if (lane == 0) {
        stats[b * 4 + 0] = tot;
        stats[b * 4 + 3] = 0.0;
      }
    };
    if constexpr (NW == 1) {                                 // single-warp CTA: the same warp does both, one after the other
      write_cost();
      if (tid < V) write_sums();
    } else {
      if (tid < V) write_sums();
      else if (warp == NW - 1) write_cost();
    }
  }
}

// ------------------------------------------------------------------ fused normal equations, staged, 128-bit lanes
// Same tiles, same bulk-copy producer and same per-problem epilogue as k_model_normal_eq_staged, but a thread owns
// VEC = 16 / sizeof(T) ADJACENT samples of a tile (4 fp32, 2 fp64) instead of S samples TPB apart: one 128-bit LDS per
// data row, one for the previous model values, one 128-bit STG of m and one 128-bit load of the abscissa per thread and tile
// (the r01d profile of the fp32 kernel counted ~70 thread instructions per (cost, sample), a third of them address
// arithmetic, per-sample bounds predicates and fp32->fp64 conversions: profiles/r01d_fused_weak.txt, cfg2gnf32).
// Bulk-copy eligibility already makes N a multiple of VEC, so ONE predicate per thread and tile covers the tail.
// fp32 keeps the per-tile partial sums of the statistics in fp32 (C*VEC terms; the reference sums them in fp32
// throughout) and folds them into the fp64 per-problem sums once per tile.
template <typename T> struct Lane16;
template <> struct Lane16<float> {
  static constexpr int VEC = 4;
  static __device__ __forceinline__ void lds(const float* p, float (&v)[4]) {
    const float4 q = *reinterpret_cast<const float4*>(p); v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  }
  static __device__ __forceinline__ void ldg(const float* p, float (&v)[4]) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(p)); v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  }
  static __device__ __forceinline__ void stg(float* p, const float (&v)[4]) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
  }
};
template <> struct Lane16<double> {
  static constexpr int VEC = 2;
  static __device__ __forceinline__ void lds(const double* p, double (&v)[2]) {
    const double2 q = *reinterpret_cast<const double2*>(p); v[0] = q.x; v[1] = q.y;
  }
  static __device__ __forceinline__ void ldg(const double* p, double (&v)[2]) {
    const double2 q = __ldg(reinterpret_cast<const double2*>(p)); v[0] = q.x; v[1] = q.y;
  }
  static __device__ __forceinline__ void stg(double* p, const double (&v)[2]) {
    asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v[0]), "d"(v[1]) : "memory");
  }
};

// CT > 0: compile-time number of costs (as in k_model_normal_eq_staged): unrolled cost loop, per-cost sums in registers.
template <typename T, typename Model, int TPB, int NSTAGE, int CT = 0>
__global__ void __launch_bounds__(TPB, 512 / TPB)
k_model_normal_eq_vec(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                      long long ycs, const T* __restrict__ w, int wlen, long long B, int C_, int N, T* mA, T* mB,
                      T* __restrict__ g, T* __restrict__ H, double* __restrict__ stats, const nlls_ctrl* ctrl,
                      const unsigned char* __restrict__ sel, nlls_track tr, const __grid_constant__ CUtensorMap tmy, int use_tm) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2, VEC = Lane16<T>::VEC, TN = TPB * VEC;
  constexpr int VP = Pow2Ceil<V>::value, OWN = 32 / VP;
  static_assert(V <= 32, "model with too many parameters for the transposed reduction");
  const int C = CT > 0 ? CT : C_;
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  T* m_out = (it & 1) ? mB : mA;
  const T* m_prev = (it > 0 && mA) ? ((it & 1) ? mA : mB) : nullptr;
  const bool has_prev = m_prev != nullptr || sel != nullptr;
  const int Cy = (ycs == 0) ? 1 : C;
  const int stage_elems = (Cy + 1) * TN;                     // [Cy][TN] data rows then [TN] previous model values
  extern __shared__ __align__(128) unsigned char smem_raw[];
  T* ring = reinterpret_cast<T*>(smem_raw);
  size_t off = (size_t)NSTAGE * stage_elems * sizeof(T);
  off = (off + 15) & ~(size_t)15;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + off);
  double* red = reinterpret_cast<double*>(smem_raw + off + 8 * 8);   // [2][NW][VP]
  double* sfc = red + 2 * NW * VP;                           // [2][C][TPB] per-thread per-cost sums of f
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncthreads();

  long long pw = blockIdx.x, pb = pw < WL.n ? WL.at(pw) : 0;   // producer: work index, problem
  int pn0 = 0, pst = 0;
  unsigned char psel = (sel && tid == 0 && pw < WL.n) ? sel[pb] : (unsigned char)0;
  auto issue = [&]() {
    if (pw >= WL.n) return;
    const int rows = min(TN, N - pn0);
    const unsigned bytes = (unsigned)rows * (unsigned)sizeof(T);
    T* dst = ring + (size_t)pst * stage_elems;
    uint64_t* bar = &bars[pst];
    mbar_expect_tx(bar, bytes * (unsigned)(Cy + (has_prev ? 1 : 0)));
    if (use_tm) {                                            // N % TN == 0: one tensor copy of the [Cy][TN] box (UTMALDG)
      tma_load_3d(dst, &tmy, pn0, 0, (int)pb, bar);
      dst += Cy * TN;
    } else {
      const T* src = y + pb * ybs + pn0;
      for (int c = 0; c < Cy; ++c, src += ycs, dst += TN) bulk_g2s(dst, src, bytes, bar);
    }
    if (has_prev) bulk_g2s(dst, (sel ? (psel ? mB : mA) : m_prev) + pb * (long long)N + pn0, bytes, bar);
    pn0 += TN;
    if (pn0 >= N) {
      pn0 = 0;
      pw += gridDim.x;
      pb = pw < WL.n ? WL.at(pw) : 0;
      if (sel && pw < WL.n) psel = sel[pb];                     // buffer role of the NEXT problem: in flight until its first tile
    }
    pst = (pst + 1 == NSTAGE) ? 0 : pst + 1;
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) issue();
  }

  int st = 0, pbuf = 0;
  unsigned ph = 0;
  const int ystep = (Cy == 1) ? 0 : TN;
  for (long long wi = blockIdx.x; wi < WL.n; wi += gridDim.x, pbuf ^= 1) {
    const long long b = WL.at(wi);
    Model M;
    M.load(p + b * P);
    if (tid == 0 && wi + gridDim.x < WL.n) prefetch_l1(p + WL.at(wi + gridDim.x) * P);
    RowAcc<T, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    double* my_sfc = sfc + (size_t)pbuf * C * TPB + tid;
    [[maybe_unused]] double sfr[CT > 0 ? CT : 1];
    if constexpr (CT > 0) {
#pragma unroll
      for (int c = 0; c < CT; ++c) sfr[c] = 0.0;
    } else {
      for (int c = 0; c < C; ++c) my_sfc[c * TPB] = 0.0;
    }
    if (sel) m_out = sel[b] ? mA : mB;
    T* mo_b = m_out ? m_out + b * (long long)N : nullptr;
    for (int n0 = 0; n0 < N; n0 += TN) {
      const int n = n0 + tid * VEC;
      const bool live = n < N;                               // N % VEC == 0: the whole lane group is in or out
      T m[VEC], k[VEC][P], s2[VEC], s3[VEC];
      if (live) {
        T xv[VEC];
        Lane16<T>::ldg(x + n, xv);
#pragma unroll
        for (int j = 0; j < VEC; ++j) { M.eval(xv[j], m[j], k[j]); s2[j] = T(0); s3[j] = T(0); }   // overlaps the wait below
        if (mo_b) Lane16<T>::stg(mo_b + n, m);
      }
      mbar_wait(&bars[st], ph);
      if (live) {
        const T* yc = ring + (size_t)st * stage_elems + tid * VEC;
        T sdt = T(0), sqt = T(0);
        T mp[VEC];
        Lane16<T>::lds(yc + Cy * TN, mp);                     // without a previous iterate: stale bytes, replaced by m below
#pragma unroll
        for (int j = 0; j < VEC; ++j) mp[j] = has_prev ? mp[j] : m[j];   // f_prev - f == 0 exactly
        auto cost = [&](const T* yrow, T wc) -> T {            // one cost of this tile; returns the thread's sum of f
          T yv[VEC], r[VEC], fv[VEC];
          Lane16<T>::lds(yrow, yv);
#pragma unroll
          for (int j = 0; j < VEC; ++j) { r[j] = yv[j] - m[j]; fv[j] = mul_rn(r[j], r[j]); }
#pragma unroll
          for (int j = 0; j < VEC; ++j) { const T rp = yv[j] - mp[j]; const T d = mul_rn(rp, rp) - fv[j]; sdt = fma(d, d, sdt); }
          T sfl = T(0);
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            sqt = fma(fv[j], fv[j], sqt);
            sfl += fv[j];
            const T wf = wc * fv[j];
            s2[j] += wf;
            s3[j] = fma(wf, r[j], s3[j]);
          }
          return sfl;
        };
        if constexpr (CT > 0 && sizeof(T) == 4 && VEC == 4) {
          // fp32, compile-time C: the cost loop on PACKED pairs of samples (FADD2 / FMUL2 / FFMA2) — 24 instructions per cost
          // instead of 44.  Every per-sample value (r, f, the per-sample sums s2, s3) is rounded exactly as in the scalar
          // loop and the thread's sum of f is added in the same order (the trial-cost kernel must reproduce it bit for bit);
          // only the tile's partial sums of (f_prev - f)^2 and f^2 are kept per pair lane and added at the end.
          f32x2 nm[2], nmp[2], s2p[2], s3p[2], sdp = pack2(0.f, 0.f), sqp = pack2(0.f, 0.f);
          const f32x2 minus1 = pack2(-1.f, -1.f);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            nm[h] = neg2(pack2(m[2 * h], m[2 * h + 1]));
            nmp[h] = neg2(pack2(mp[2 * h], mp[2 * h + 1]));
            s2p[h] = pack2(0.f, 0.f);
            s3p[h] = pack2(0.f, 0.f);
          }
#pragma unroll
          for (int c = 0; c < CT; ++c) {
            const ulonglong2 yq = *reinterpret_cast<const ulonglong2*>(yc + c * ystep);
            const f32x2 yv[2] = {{yq.x}, {yq.y}};
            const float wc1 = w[wlen == 1 ? 0 : c];
            const f32x2 wc = pack2(wc1, wc1);
            float sfl = 0.f;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const f32x2 r = add2(yv[h], nm[h]);
              const f32x2 fv = mul2(r, r);
              const f32x2 rp = add2(yv[h], nmp[h]);
              const f32x2 d = fma2(fv, minus1, mul2(rp, rp));     // rn(rn(rp * rp) - fv): the multiplication by -1 is exact
              sdp = fma2(d, d, sdp);
              sqp = fma2(fv, fv, sqp);
              float f0, f1;
              unpack2(fv, f0, f1);
              sfl += f0;
              sfl += f1;
              const f32x2 wf = mul2(wc, fv);
              s2p[h] = add2(s2p[h], wf);
              s3p[h] = fma2(wf, r, s3p[h]);
            }
            sfr[c] += (double)sfl;
          }
          float a0, a1;
#pragma unroll
          for (int h = 0; h < 2; ++h) { unpack2(s2p[h], s2[2 * h], s2[2 * h + 1]); unpack2(s3p[h], s3[2 * h], s3[2 * h + 1]); }
          unpack2(sdp, a0, a1);
          sdt = a0 + a1;
          unpack2(sqp, a0, a1);
          sqt = a0 + a1;
        } else if constexpr (CT > 0) {
#pragma unroll
          for (int c = 0; c < CT; ++c) sfr[c] += (double)cost(yc + c * ystep, w[wlen == 1 ? 0 : c]);
        } else {
#pragma unroll 2
          for (int c = 0; c < C; ++c) my_sfc[c * TPB] += (double)cost(yc + c * ystep, w[wlen == 1 ? 0 : c]);
        }
        sd += (double)sdt;
        sq += (double)sqt;
#pragma unroll
        for (int j = 0; j < VEC; ++j) acc.add_scaled(k[j], T(2) * s3[j], T(4) * s2[j]);
      }
      if (n0 + TN < N) {
        __syncthreads();
        if (tid == 0) issue();
      }
      if (++st == NSTAGE) { st = 0; ph ^= 1u; }
    }
    if constexpr (CT > 0) {
#pragma unroll
      for (int c = 0; c < CT; ++c) my_sfc[c * TPB] = sfr[c];
    }
    double v[VP];
#pragma unroll
    for (int i = 0; i < P; ++i) v[i] = (double)acc.g[i];
#pragma unroll
    for (int i = 0; i < NH; ++i) v[P + i] = (double)acc.h[i];
    v[P + NH] = sd;
    v[P + NH + 1] = sq;
#pragma unroll
    for (int i = V; i < VP; ++i) v[i] = 0.0;
    warp_reduce_transpose<VP>(v);
    double* myred = red + (size_t)pbuf * NW * VP;
    if ((lane & (OWN - 1)) == 0) myred[warp * VP + lane / OWN] = v[0];
    __syncthreads();
    if (tid == 0) issue();
    auto write_sums = [&]() {                                // threads 0 .. V-1: g, H, stats[1..2]
      double s = 0.0;
#pragma unroll
      for (int kk = 0; kk < NW; ++kk) s += myred[kk * VP + tid];
      if (tid < P) {
        g[b * P + tid] = (T)s;
      } else if (tid < P + NH) {
        int kk = tid - P, i = 0;
        while (kk >= P - i) { kk -= P - i; ++i; }
        const int j = i + kk;
        H[b * P * P + i * P + j] = (T)s;
        H[b * P * P + j * P + i] = (T)s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    };
    auto write_cost = [&]() {                                // one whole warp: sum_c (sum_n f)^2 from the per-thread sums
      const double* col = sfc + (size_t)pbuf * C * TPB;
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
#pragma unroll
        for (int kk = 0; kk < NW; ++kk) s += col[c * TPB + kk * 32 + lane];
        s = warp_sum(s);
        tot = fma(s, s, tot);
      }
      if (lane == 0) {
        stats[b * 4 + 0] = tot;
        stats[b * 4 + 3] = 0.0;
      }
    };
    if constexpr (NW == 1) {                                 // single-warp CTA: the same warp does both, one after the other
      write_cost();
      if (tid < V) write_sums();
    } else {
      if (tid < V) write_sums();
      else if (warp == NW - 1) write_cost();
    }
  }
}

// ------------------------------------------------------------------ fused normal equations, warp per problem
// For problems with little work each (C*N of a few thousand samples: cfg-3's N = 1000, C = 1) the CTA-per-problem
// kernel above spends a fifth of its issue slots in CTA barriers and a quarter in the per-problem epilogue (r01d
// profile: profiles/r01d_fused_weak.txt).  Here every WARP owns whole problems: a private ring of NSTAGE chunks of
// CH samples (bulk async copies issued by lane 0, one mbarrier per stage, __syncwarp hand-back), 32 samples per
// step and lane, one transposed warp reduction per problem and direct global writes — no CTA barrier, no
// cross-warp shared-memory reduction.  fp32 keeps its per-block partial sums and the warp reduction in fp32 (the
// reference sums everything in fp32); the cross-block sums of the statistics stay fp64.
// MODE: WPP_NEQ normal equations (newton_parallel.py:40-43 fused with the model), WPP_GDA cost-summed tall system
// (gda:134-139), WPP_COST trial cost (lma:99-100).  prevA/prevB are the [B,N] ping-pong buffers of the model value m (NEQ:
// f_prev is rebuilt from it, see the head of this file) or of the cost-summed f_s (GDA); outg/outH are T (NEQ) or fp64 (GDA).
enum { WPP_NEQ = 0, WPP_GDA = 1, WPP_COST = 2 };

template <typename T, typename Model, int MODE, int SB, int NSTAGE, int NW, bool C1, bool VL = false, int MINB = 1>
__global__ void __launch_bounds__(NW * 32, MINB)
k_model_wpp(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs, long long ycs,
            const T* __restrict__ w, int wlen, long long B, int C_, int N, T* prevA, T* prevB, void* __restrict__ outg,
            void* __restrict__ outH, double* __restrict__ stats, const nlls_ctrl* ctrl, int CH,
            const unsigned char* __restrict__ sel, nlls_track tr) {
  if (is_done(ctrl)) return;
  using TA = typename std::conditional<MODE == WPP_GDA, double, T>::type;      // accumulator type of g / H
  constexpr int P = Model::P, NH = Tri<P>::N, V = P + NH;
  constexpr int VP = Pow2Ceil<V>::value, OWN = 32 / VP;
  static_assert(V <= 32, "model with too many parameters for the transposed reduction");
  // VL ("vector lanes", single-cost problems only): a lane owns SB = 16/sizeof(T) ADJACENT samples of a block of 32*SB, so
  // the abscissa, the data row and the previous residuals arrive with one 128-bit load each and f leaves with one
  // 128-bit store per lane and block; otherwise lane L owns samples L, L+32, ... (scalar accesses).
  static_assert(!VL || (C1 && SB * sizeof(T) == 16), "vector lanes: single cost, SB = 16 / sizeof(T)");
  constexpr int ISTR = VL ? 1 : 32;          // distance between a lane's consecutive samples
  const int C = C1 ? 1 : C_;                 // C1: single cost — no cost loop, per-cost sums in a register
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  T* f_out = (MODE == WPP_COST) ? nullptr : ((it & 1) ? prevB : prevA);
  const T* f_prev = (MODE != WPP_COST && it > 0 && prevA) ? ((it & 1) ? prevA : prevB) : nullptr;
  const bool has_prev = f_prev != nullptr || sel != nullptr;  // sel (NEQ only): per-problem buffer roles
  const int Cy = (C1 || ycs == 0) ? 1 : C;
  const int CP = (MODE == WPP_COST) ? 0 : 1;                                   // rows of "previous" data per chunk (m or f_s)
  const int stage_elems = (Cy + CP) * CH;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int l0 = VL ? lane * SB : lane;      // a lane's first sample within a block
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const size_t ring_bytes = (size_t)NSTAGE * stage_elems * sizeof(T);          // CH % 64 == 0: 16-byte granular
  const bool use_sfc = !C1 && MODE != WPP_GDA;
  const size_t per_warp = ring_bytes + 64 + (use_sfc ? (size_t)C * 32 * 8 : 0);
  unsigned char* mine = smem_raw + (size_t)warp * per_warp;
  T* ring = reinterpret_cast<T*>(mine);
  uint64_t* bars = reinterpret_cast<uint64_t*>(mine + ring_bytes);
  double* sfc = reinterpret_cast<double*>(mine + ring_bytes + 64);             // [C][32] per-cost sums of f
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  const long long wid = (long long)blockIdx.x * NW + warp, nwarps = (long long)gridDim.x * NW;
  const int QN = (N + CH - 1) / CH;
  const long long prev_bs = (long long)N;                                      // problem stride of prev / f_out

  long long fw = wid, fb = fw < WL.n ? WL.at(fw) : 0;   // producer cursor: work index / problem of the next chunk to request
  int fq = 0;
  unsigned fst = 0;
  unsigned char fsel = (sel && lane == 0 && fw < WL.n) ? sel[fb] : (unsigned char)0;
  auto issue = [&]() {
    if (fw >= WL.n) return;
    if (lane == 0) {
      const int n0 = fq * CH;
      const unsigned bytes = (unsigned)min(CH, N - n0) * (unsigned)sizeof(T);
      T* dst = ring + (size_t)fst * stage_elems;
      mbar_expect_tx(&bars[fst], bytes * (unsigned)(Cy + (has_prev ? CP : 0)));
      const T* src = y + fb * ybs + n0;
      for (int c = 0; c < Cy; ++c, src += ycs, dst += CH) bulk_g2s(dst, src, bytes, &bars[fst]);
      if (has_prev) {
        const T* ps = (sel ? (fsel ? prevB : prevA) : f_prev) + fb * prev_bs + n0;
        for (int c = 0; c < CP; ++c, ps += N, dst += CH) bulk_g2s(dst, ps, bytes, &bars[fst]);
      }
    }
    fst = (fst + 1 == NSTAGE) ? 0 : fst + 1;
    if (++fq == QN) {
      fq = 0;
      fw += nwarps;
      fb = fw < WL.n ? WL.at(fw) : 0;
      if (sel && lane == 0 && fw < WL.n) fsel = sel[fb];        // next problem's buffer role, loaded a whole problem ahead of its use
    }
  };
#pragma unroll
  for (int s = 0; s < NSTAGE; ++s) issue();

  unsigned st = 0, ph = 0;
  const T w0 = (MODE == WPP_COST) ? T(1) : w[0];
  for (long long wi = wid; wi < WL.n; wi += nwarps) {
    const long long b = WL.at(wi);
    Model M;
    M.load(p + b * P);
    RowAcc<TA, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0, sf1 = 0.0;
    if (use_sfc)
      for (int c = 0; c < C; ++c) sfc[c * 32 + lane] = 0.0;
    if (sel) f_out = sel[b] ? prevA : prevB;
    T* fo_b = f_out ? f_out + b * prev_bs : nullptr;
    for (int q = 0; q < QN; ++q) {
      const int n0 = q * CH, rows = min(CH, N - n0);
      const T* sy = ring + (size_t)st * stage_elems + l0;
      const T* sp = sy + Cy * CH;
      const T* xq = x + n0 + l0;
      T* fo_q = fo_b ? fo_b + n0 + l0 : nullptr;
      const int ystep = (Cy == 1) ? 0 : CH;
      bool waited = false;
      T sd_t = T(0), sq_t = T(0), sf_t = T(0);                                 // fp32: per-chunk partial sums in T
      auto block = [&](int j0, auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
        T m[SB], k[SB][MODE == WPP_COST ? 1 : P];
        [[maybe_unused]] T xl[SB], yl[SB], pl[SB], fl[SB];   // VL: the lane's 16 bytes of x, y, f_prev, f
        if constexpr (VL) {
          if (FULL || j0 + l0 < rows) Lane16<T>::ldg(xq + j0, xl);           // rows % SB == 0: a lane is all in or all out
          else {
#pragma unroll
            for (int i = 0; i < SB; ++i) xl[i] = T(0);
          }
        }
#pragma unroll
        for (int i = 0; i < SB; ++i) {
          T xv;
          if constexpr (VL) xv = xl[i];
          else xv = ld_cached(FULL ? xq + j0 + i * 32 : x + min(n0 + j0 + i * 32 + lane, N - 1));
          if constexpr (MODE == WPP_COST) m[i] = M.value(xv);
          else M.eval(xv, m[i], k[i]);
        }
        if constexpr (MODE == WPP_NEQ) {                                      // the model value is what the next iteration reads back
          if (fo_q) {
            if constexpr (VL) { if (FULL || j0 + l0 < rows) Lane16<T>::stg(fo_q + j0, m); }
            else {
#pragma unroll
              for (int i = 0; i < SB; ++i)
                if (FULL || j0 + i * 32 + lane < rows) stg_stream(fo_q + j0 + i * 32, m[i]);
            }
          }
        }
        if (!waited) { mbar_wait(&bars[st], ph); waited = true; }             // first block's model evaluation overlaps the wait
        const T* yc = sy + j0;
        const T* pc = sp + j0;
        T* fo = fo_q ? fo_q + j0 : nullptr;
        if constexpr (VL) {
          Lane16<T>::lds(yc, yl);                                             // stale past `rows`, masked below
          if (MODE != WPP_COST && has_prev) Lane16<T>::lds(pc, pl);
        }
        if constexpr (MODE == WPP_COST) {
          if constexpr (C1) {
#pragma unroll
            for (int i = 0; i < SB; ++i)
              if (FULL || j0 + i * ISTR + l0 < rows) { const T r = (VL ? yl[i] : yc[i * 32]) - m[i]; const T fv = mul_rn(r, r); sf_t += fv; }
          } else {
            for (int c = 0; c < C; ++c, yc += ystep) {
              T sfl = T(0);
#pragma unroll
              for (int i = 0; i < SB; ++i)
                if (FULL || j0 + i * 32 + lane < rows) { const T r = yc[i * 32] - m[i]; const T fv = mul_rn(r, r); sfl += fv; }
              sfc[c * 32 + lane] += (double)sfl;
            }
          }
        } else if constexpr (MODE == WPP_GDA) {
          T s1[SB], fs[SB];
#pragma unroll
          for (int i = 0; i < SB; ++i) { s1[i] = T(0); fs[i] = T(0); }
          for (int c = 0; c < C; ++c, yc += ystep) {
            const T wc = C1 ? w0 : w[wlen == 1 ? 0 : c];
#pragma unroll
            for (int i = 0; i < SB; ++i) {
              const T r = (VL ? yl[i] : yc[i * 32]) - m[i];  // rows past the chunk hold stale data; masked below
              s1[i] = fma(wc, r, s1[i]);
              fs[i] = fma(wc, r * r, fs[i]);                 // gda:136
            }
          }
#pragma unroll
          for (int i = 0; i < SB; ++i) {
            if (FULL || j0 + i * ISTR + l0 < rows) {
              double jd[P];
#pragma unroll
              for (int qq = 0; qq < P; ++qq) jd[qq] = (double)(T(2) * s1[i] * k[i][qq]);   // gda:137 (cost-summed row, in T)
              acc.add(jd, (double)fs[i], 1.0);
              if constexpr (!VL) { if (fo) fo[i * 32] = fs[i]; }
              sq_t = fma(fs[i], fs[i], sq_t);
              if (has_prev) { const T d = (VL ? pl[i] : pc[i * 32]) - fs[i]; sd_t = fma(d, d, sd_t); }
            }
          }
          if constexpr (VL) { if (fo && (FULL || j0 + l0 < rows)) Lane16<T>::stg(fo, fs); }
        } else if constexpr (C1) {
#pragma unroll
          for (int i = 0; i < SB; ++i) {
            if (FULL || j0 + i * ISTR + l0 < rows) {
              const T yv = VL ? yl[i] : yc[i * 32];
              const T r = yv - m[i];
              const T fv = mul_rn(r, r);
              if (has_prev) {                                 // f_prev rebuilt from the stored model value (bit-identical)
                const T rp = yv - (VL ? pl[i] : pc[i * 32]);
                const T d = mul_rn(rp, rp) - fv;
                sd_t = fma(d, d, sd_t);
              }
              sq_t = fma(fv, fv, sq_t);
              sf_t += fv;
              const T wf = w0 * fv;
              acc.add_scaled(k[i], T(2) * (wf * r), T(4) * wf);
            }
          }
        } else {
          T s2[SB], s3[SB];
#pragma unroll
          for (int i = 0; i < SB; ++i) { s2[i] = T(0); s3[i] = T(0); }
          T mp[SB];
#pragma unroll
          for (int i = 0; i < SB; ++i) mp[i] = has_prev ? pc[i * 32] : T(0);   // rows past the chunk: stale, masked below
          for (int c = 0; c < C; ++c, yc += ystep) {
            const T wc = w[wlen == 1 ? 0 : c];
            T sfl = T(0);
#pragma unroll
            for (int i = 0; i < SB; ++i) {
              if (FULL || j0 + i * 32 + lane < rows) {
                const T yv = yc[i * 32];
                const T r = yv - m[i];
                const T fv = mul_rn(r, r);
                if (has_prev) {
                  const T rp = yv - mp[i];
                  const T d = mul_rn(rp, rp) - fv;
                  sd_t = fma(d, d, sd_t);
                }
                sq_t = fma(fv, fv, sq_t);
                sfl += fv;
                const T wf = wc * fv;
                s2[i] += wf;
                s3[i] = fma(wf, r, s3[i]);
              }
            }
            sfc[c * 32 + lane] += (double)sfl;
          }
#pragma unroll
          for (int i = 0; i < SB; ++i)
            if (FULL || j0 + i * 32 + lane < rows) acc.add_scaled(k[i], T(2) * s3[i], T(4) * s2[i]);
        }
      };
      int j0 = 0;
      for (; j0 + 32 * SB <= rows; j0 += 32 * SB) block(j0, std::true_type{});
      if (j0 < rows) block(j0, std::false_type{});
      sd += (double)sd_t;
      sq += (double)sq_t;
      if constexpr (C1) sf1 += (double)sf_t;
      if (++st == NSTAGE) { st = 0; ph ^= 1u; }
      __syncwarp();                          // every lane is done with this stage before it is refilled
      issue();
    }
    // ---- per-problem sums: one transposed warp reduction, totals written straight to global memory
    double tot = 0.0;
    if constexpr (MODE != WPP_GDA) {
      if constexpr (C1) {
        const double ssum = warp_sum(sf1);
        tot = ssum * ssum;
      } else {
        for (int c = 0; c < C; ++c) {
          const double ssum = warp_sum(sfc[c * 32 + lane]);
          tot = fma(ssum, ssum, tot);
        }
      }
    }
    if constexpr (MODE == WPP_COST) {
      if (lane == 0) stats[b] = tot;         // stat_out [B]
    } else {
      TA v[VP];
#pragma unroll
      for (int i = 0; i < P; ++i) v[i] = acc.g[i];
#pragma unroll
      for (int i = 0; i < NH; ++i) v[P + i] = acc.h[i];
#pragma unroll
      for (int i = V; i < VP; ++i) v[i] = TA(0);
      warp_reduce_transpose<VP>(v);
      sd = warp_sum(sd);
      sq = warp_sum(sq);
      TA* g = reinterpret_cast<TA*>(outg);
      TA* H = reinterpret_cast<TA*>(outH);
      const int vi = lane / OWN;
      if ((lane & (OWN - 1)) == 0 && vi < V) {
        if (vi < P) {
          g[b * P + vi] = v[0];
        } else {
          int kk = vi - P, i = 0;
          while (kk >= P - i) { kk -= P - i; ++i; }
          const int j = i + kk;
          H[b * P * P + i * P + j] = v[0];
          H[b * P * P + j * P + i] = v[0];
        }
      }
      if (lane == 0) {
        stats[b * 4 + 0] = tot;
        stats[b * 4 + 1] = sd;
        stats[b * 4 + 2] = sq;
        stats[b * 4 + 3] = 0.0;
      }
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------ fused GD flavour
template <typename T, typename Model, int TPB>
__global__ void __launch_bounds__(TPB)
k_model_normal_eq_gda(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                      long long ycs, const T* __restrict__ w, int wlen, long long B, int C, int N, T* fsA, T* fsB,
                      double* __restrict__ g64, double* __restrict__ H64, double* __restrict__ stats,
                      const nlls_ctrl* ctrl, nlls_track tr) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2;
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  T* __restrict__ fs_out = (it & 1) ? fsB : fsA;
  const T* __restrict__ fs_prev = (it > 0 && fsA) ? ((it & 1) ? fsA : fsB) : nullptr;
  __shared__ double red[NW * V];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (long long wi = blockIdx.x; wi < WL.n; wi += gridDim.x) {
    const long long b = WL.at(wi);
    Model M;
    M.load(p + b * P);
    RowAcc<double, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    const T* yb = y + b * ybs;
    for (int n = tid; n < N; n += TPB) {
      T m, k[P];
      M.eval(ld_cached(x + n), m, k);
      T s1 = T(0), fs = T(0);
      for (int c = 0; c < C; ++c) {
        const T wc = w[wlen == 1 ? 0 : c];
        const T r = ld_cached(yb + c * ycs + n) - m;
        s1 = fma(wc, r, s1);
        fs = fma(wc, r * r, fs);                               // gda:136
      }
      double jd[P];
#pragma unroll
      for (int i = 0; i < P; ++i) jd[i] = (double)(T(2) * s1 * k[i]);   // gda:137 (cost-summed row, in T)
      acc.add(jd, (double)fs, 1.0);
      if (fs_out) fs_out[b * N + n] = fs;
      sq = fma((double)fs, (double)fs, sq);
      if (fs_prev) { const T d = fs_prev[b * N + n] - fs; sd = fma((double)d, (double)d, sd); }
    }
    acc.warp_reduce();
    sd = warp_sum(sd);
    sq = warp_sum(sq);
    if (lane == 0) {
      double* r = red + warp * V;
#pragma unroll
      for (int i = 0; i < P; ++i) r[i] = acc.g[i];
#pragma unroll
      for (int i = 0; i < NH; ++i) r[P + i] = acc.h[i];
      r[P + NH] = sd;
      r[P + NH + 1] = sq;
    }
    __syncthreads();
    if (tid < V) {
      double s = 0.0;
#pragma unroll
      for (int kk = 0; kk < NW; ++kk) s += red[kk * V + tid];
      if (tid < P) {
        g64[b * P + tid] = s;
      } else if (tid < P + NH) {
        int kk = tid - P, i = 0;
        while (kk >= P - i) { kk -= P - i; ++i; }
        const int j = i + kk;
        H64[b * P * P + i * P + j] = s;
        H64[b * P * P + j * P + i] = s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    } else if (tid == V) {
      stats[b * 4 + 0] = 0.0;
      stats[b * 4 + 3] = 0.0;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ fused trial cost (lma:99-100)
template <typename T, typename Model, int TPB, int S>
__global__ void __launch_bounds__(TPB)
k_model_cost_stat(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                  long long ycs, long long B, int C, int N, double* __restrict__ out, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32;
  extern __shared__ double smem[];
  double* sfc = smem;                  // [C][TPB]
  double* sfw = sfc + (size_t)C * TPB; // [C][NW]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    Model M;
    M.load(p + b * P);
    for (int c = 0; c < C; ++c) sfc[c * TPB + tid] = 0.0;
    const T* yb = y + b * ybs;
    for (int n0 = 0; n0 < N; n0 += TPB * S) {
      T m[S];
#pragma unroll
      for (int i = 0; i < S; ++i) m[i] = M.value(ld_cached(x + min(n0 + i * TPB + tid, N - 1)));
      for (int c = 0; c < C; ++c) {
        const T* yc = yb + c * ycs;
        double sfl = 0.0;
#pragma unroll
        for (int i = 0; i < S; ++i) {
          const int n = n0 + i * TPB + tid;
          if (n < N) { const T r = ld_cached(yc + n) - m[i]; sfl += (double)mul_rn(r, r); }
        }
        sfc[c * TPB + tid] += sfl;
      }
    }
    for (int c = 0; c < C; ++c) {
      const double s = warp_sum(sfc[c * TPB + tid]);
      if (lane == 0) sfw[c * NW + warp] = s;
    }
    __syncthreads();
    if (tid == 0) {
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
        for (int kk = 0; kk < NW; ++kk) s += sfw[c * NW + kk];
        tot = fma(s, s, tot);
      }
      out[b] = tot;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ fused trial cost, bulk-copy staged
template <typename T, typename Model, int TPB, int S, int NSTAGE>
__global__ void __launch_bounds__(TPB)
k_model_cost_stat_staged(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                         long long ycs, long long B, int C, int N, double* __restrict__ out, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, TN = TPB * S;
  const int Cy = (ycs == 0) ? 1 : C;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  using Ring = TileRing<T, TN, NSTAGE>;
  double* sfc = reinterpret_cast<double*>(smem_raw + Ring::bytes(Cy, 0));    // [2][C][TPB]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  Ring R;
  R.init(smem_raw, y, ybs, ycs, Cy, nullptr, 0, 0, 0, WorkList(nlls_track{}, 0, B), N);
  int pbuf = 0;
  for (long long b = blockIdx.x; b < B; b += gridDim.x, pbuf ^= 1) {
    Model M;
    M.load(p + b * P);
    if (tid == 0 && b + gridDim.x < B) prefetch_l1(p + (b + gridDim.x) * P);
    double* my_sfc = sfc + (size_t)pbuf * C * TPB + tid;
    for (int c = 0; c < C; ++c) my_sfc[c * TPB] = 0.0;
    for (int n0 = 0; n0 < N; n0 += TN) {
      T m[S];
#pragma unroll
      for (int i = 0; i < S; ++i) m[i] = M.value(ld_cached(x + min(n0 + i * TPB + tid, N - 1)));
      const T* yc = R.wait() + tid;
      const int ystep = (Cy == 1) ? 0 : TN;
      for (int c = 0; c < C; ++c, yc += ystep) {
        double sfl = 0.0;
#pragma unroll
        for (int i = 0; i < S; ++i)
          if (n0 + i * TPB + tid < N) { const T r = yc[i * TPB] - m[i]; sfl += (double)mul_rn(r, r); }
        my_sfc[c * TPB] += sfl;
      }
      if (n0 + TN < N) {
        __syncthreads();
        if (tid == 0) R.issue();
      }
      R.advance();
    }
    __syncthreads();
    if (tid == 0) R.issue();
    if (warp == NW - 1) {
      const double* col = sfc + (size_t)pbuf * C * TPB;
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
#pragma unroll
        for (int kk = 0; kk < NW; ++kk) s += col[c * TPB + kk * 32 + lane];
        s = warp_sum(s);
        tot = fma(s, s, tot);
      }
      if (lane == 0) out[b] = tot;
    }
  }
}

// ------------------------------------------------------------------ fused trial cost, staged, 128-bit lanes
// Sibling of k_model_normal_eq_vec: same sample-to-thread map and the same order of every sum that reaches
// stats[0], so the two-pass and the speculative LM schedules stay bit-identical whichever kernel family runs.
template <typename T, typename Model, int TPB, int NSTAGE>
__global__ void __launch_bounds__(TPB)
k_model_cost_stat_vec(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                      long long ycs, long long B, int C, int N, double* __restrict__ out, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, VEC = Lane16<T>::VEC, TN = TPB * VEC;
  const int Cy = (ycs == 0) ? 1 : C;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  using Ring = TileRing<T, TN, NSTAGE>;
  double* sfc = reinterpret_cast<double*>(smem_raw + Ring::bytes(Cy, 0));    // [2][C][TPB]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  Ring R;
  R.init(smem_raw, y, ybs, ycs, Cy, nullptr, 0, 0, 0, WorkList(nlls_track{}, 0, B), N);
  int pbuf = 0;
  const int ystep = (Cy == 1) ? 0 : TN;
  for (long long b = blockIdx.x; b < B; b += gridDim.x, pbuf ^= 1) {
    Model M;
    M.load(p + b * P);
    if (tid == 0 && b + gridDim.x < B) prefetch_l1(p + (b + gridDim.x) * P);
    double* my_sfc = sfc + (size_t)pbuf * C * TPB + tid;
    for (int c = 0; c < C; ++c) my_sfc[c * TPB] = 0.0;
    for (int n0 = 0; n0 < N; n0 += TN) {
      const int n = n0 + tid * VEC;
      const bool live = n < N;
      T m[VEC];
      if (live) {
        T xv[VEC];
        Lane16<T>::ldg(x + n, xv);
#pragma unroll
        for (int j = 0; j < VEC; ++j) m[j] = M.value(xv[j]);
      }
      const T* yc = R.wait() + tid * VEC;
      if (live) {
        T yv[VEC];
        Lane16<T>::lds(yc, yv);
        for (int c = 0; c < C; ++c) {
          if (c > 0 && ystep) { yc += ystep; Lane16<T>::lds(yc, yv); }
          T sfl = T(0);
#pragma unroll
          for (int j = 0; j < VEC; ++j) { const T r = yv[j] - m[j]; sfl += mul_rn(r, r); }
          my_sfc[c * TPB] += (double)sfl;
        }
      }
      if (n0 + TN < N) {
        __syncthreads();
        if (tid == 0) R.issue();
      }
      R.advance();
    }
    __syncthreads();
    if (tid == 0) R.issue();
    if (warp == NW - 1) {
      const double* col = sfc + (size_t)pbuf * C * TPB;
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
#pragma unroll
        for (int kk = 0; kk < NW; ++kk) s += col[c * TPB + kk * 32 + lane];
        s = warp_sum(s);
        tot = fma(s, s, tot);
      }
      if (lane == 0) out[b] = tot;
    }
  }
}

// ------------------------------------------------------------------ fused GD flavour, bulk-copy staged
// CT > 0: compile-time number of costs (unrolled cost loop); MINB: CTAs per SM the register allocation aims at.
template <typename T, typename Model, int TPB, int S, int NSTAGE, int CT = 0, int MINB = 0>
__global__ void __launch_bounds__(TPB, MINB > 0 ? MINB : 1)
k_model_normal_eq_gda_staged(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs,
                             long long ycs, const T* __restrict__ w, int wlen, long long B, int C_, int N, T* fsA, T* fsB,
                             double* __restrict__ g64, double* __restrict__ H64, double* __restrict__ stats,
                             const nlls_ctrl* ctrl, nlls_track tr, const __grid_constant__ CUtensorMap tmy, int use_tm) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2, TN = TPB * S;
  constexpr int VP = Pow2Ceil<V>::value, OWN = 32 / VP;
  static_assert(V <= 32, "model with too many parameters for the transposed reduction");
  const int C = CT > 0 ? CT : C_;
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  T* __restrict__ fs_out = (it & 1) ? fsB : fsA;
  const T* __restrict__ fs_prev = (it > 0 && fsA) ? ((it & 1) ? fsA : fsB) : nullptr;
  const bool has_prev = fs_prev != nullptr;
  const int Cy = (ycs == 0) ? 1 : C;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  using Ring = TileRing<T, TN, NSTAGE>;
  double* red = reinterpret_cast<double*>(smem_raw + Ring::bytes(Cy, 1));    // [2][NW][VP]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  Ring R;
  R.init(smem_raw, y, ybs, ycs, Cy, fs_prev, N, 0, 1, WL, N, use_tm ? &tmy : nullptr);
  int pbuf = 0;
  for (long long wi = blockIdx.x; wi < WL.n; wi += gridDim.x, pbuf ^= 1) {
    const long long b = WL.at(wi);
    Model M;
    M.load(p + b * P);
    if (tid == 0 && wi + gridDim.x < WL.n) prefetch_l1(p + WL.at(wi + gridDim.x) * P);
    RowAcc<double, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    auto tile = [&](int n0, auto full_tag) {                 // FULL tiles: no per-sample bounds predicates (straight-line body)
      constexpr bool FULL = decltype(full_tag)::value;
      T xv[S], m[S], k[S][P];
#pragma unroll
      for (int i = 0; i < S; ++i) xv[i] = ld_cached(x + (FULL ? n0 + i * TPB + tid : min(n0 + i * TPB + tid, N - 1)));
      model_eval_n<Model, T, S>(M, xv, m, k);
      const T* sy = R.wait() + tid;
      const T* sp = sy + Cy * TN;
      const int ystep = (Cy == 1) ? 0 : TN;
      T s1[S], fs[S];
#pragma unroll
      for (int i = 0; i < S; ++i) { s1[i] = T(0); fs[i] = T(0); }
      auto cost = [&](const T* yc, T wc) {
#pragma unroll
        for (int i = 0; i < S; ++i) {
          const T r = yc[i * TPB] - m[i];                    // rows past N hold stale data; masked below
          s1[i] = fma(wc, r, s1[i]);
          fs[i] = fma(wc, r * r, fs[i]);                     // gda:136
        }
      };
      if constexpr (CT > 0) {
#pragma unroll
        for (int c = 0; c < CT; ++c) cost(sy + c * ystep, w[wlen == 1 ? 0 : c]);
      } else {
        for (int c = 0; c < C; ++c) cost(sy + c * ystep, w[wlen == 1 ? 0 : c]);
      }
#pragma unroll
      for (int i = 0; i < S; ++i) {
        const int n = n0 + i * TPB + tid;
        if (FULL || n < N) {
          double jd[P];
#pragma unroll
          for (int q = 0; q < P; ++q) jd[q] = (double)(T(2) * s1[i] * k[i][q]);   // gda:137 (cost-summed row, in T)
          acc.add(jd, (double)fs[i], 1.0);
          if (fs_out) fs_out[b * N + n] = fs[i];
          sq = fma((double)fs[i], (double)fs[i], sq);
          const T d = (has_prev ? sp[i * TPB] : fs[i]) - fs[i];   // no previous iterate: 0 exactly
          sd = fma((double)d, (double)d, sd);
        }
      }
    };
    for (int n0 = 0; n0 < N; n0 += TN) {
      if (n0 + TN <= N) tile(n0, std::true_type{});
      else tile(n0, std::false_type{});
      if (n0 + TN < N) {
        __syncthreads();
        if (tid == 0) R.issue();
      }
      R.advance();
    }
    double v[VP];
#pragma unroll
    for (int i = 0; i < P; ++i) v[i] = acc.g[i];
#pragma unroll
    for (int i = 0; i < NH; ++i) v[P + i] = acc.h[i];
    v[P + NH] = sd;
    v[P + NH + 1] = sq;
#pragma unroll
    for (int i = V; i < VP; ++i) v[i] = 0.0;
    warp_reduce_transpose<VP>(v);
    double* myred = red + (size_t)pbuf * NW * VP;
    if ((lane & (OWN - 1)) == 0) myred[warp * VP + lane / OWN] = v[0];
    __syncthreads();
    if (tid == 0) R.issue();
    if (tid < V) {
      double s = 0.0;
#pragma unroll
      for (int kk = 0; kk < NW; ++kk) s += myred[kk * VP + tid];
      if (tid < P) {
        g64[b * P + tid] = s;
      } else if (tid < P + NH) {
        int kk = tid - P, i = 0;
        while (kk >= P - i) { kk -= P - i; ++i; }
        const int j = i + kk;
        H64[b * P * P + i * P + j] = s;
        H64[b * P * P + j * P + i] = s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    } else if (tid == V) {
      stats[b * 4 + 0] = 0.0;
      stats[b * 4 + 3] = 0.0;
    }
  }
}

// ------------------------------------------------------------------ materialising callables
// A warp (or `wpp` warps) per problem: the per-problem constants (Model::load: divisions) are computed once per lane and
// problem, lanes stride over the abscissa; writes f[b,c,n] and J[b,c,n,:] for every cost c.
template <typename T, typename Model>
__global__ void __launch_bounds__(256)
k_model_eval(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs, long long ycs,
             long long B, int C, int N, T* __restrict__ f_out, T* __restrict__ J_out, const nlls_ctrl* ctrl, int wpp) {
  if (is_done(ctrl)) return;
  constexpr int P = Model::P;
  const int lane = threadIdx.x & 31;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5, items = B * (long long)wpp;
  const bool j16 = J_out != nullptr && aligned_to_dev(J_out, 16);
  for (long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < items; w += nwarps) {
    const long long b = w / wpp;
    const int part = (int)(w - b * wpp);
    Model M;
    M.load(p + b * P);
    for (int n = part * 32 + lane; n < N; n += 32 * wpp) {
      T m, k[P];
      if (J_out) M.eval(ld_cached(x + n), m, k);
      else m = M.value(ld_cached(x + n));
      for (int c = 0; c < C; ++c) {
        const T r = ld_cached(y + b * ybs + c * ycs + n) - m;
        const long long row = (b * C + c) * (long long)N + n;
        f_out[row] = r * r;
        if (J_out) {
          const T c2 = T(2) * r;
          T* Jr = J_out + row * P;
          if constexpr ((P * sizeof(T)) % 16 == 0) {         // rows are 16-byte granular: 128-bit stores
            if (j16) {
              constexpr int VEC = Lane16<T>::VEC;
#pragma unroll
              for (int i = 0; i < P; i += VEC) {
                T v[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) v[j] = c2 * k[i + j];
                Lane16<T>::stg(Jr + i, v);
              }
              continue;
            }
          }
#pragma unroll
          for (int i = 0; i < P; ++i) Jr[i] = c2 * k[i];
        }
      }
    }
  }
}

// ------------------------------------------------------------------ fused normal equations, 8 < P <= 16, fp64 tensor cores
// Multi-peak models (cfg-4: four EMG peaks, P = 16) have too many accumulators for the register-resident kernels above
// (g and the packed triangle of H would be 152 fp64 values per thread).  Here a warp owns batches of 32 consecutive
// samples: lane L evaluates sample L once (model value m and k = -dm/dp, all peaks), loops the costs for the residuals
// (m written, m_prev read and f_prev rebuilt from it, statistics, s2 = 4 sum_c w_c r_c^2, s3 = 2 sum_c w_c r_c^3 as in k_model_normal_eq) and drops
// its k row into a per-warp shared tile [32][16] (row stride 160 B: conflict-free 128-bit reads).  The 32 x 16 tile is
// then reduced on the fp64 tensor cores exactly like K1's k_normal_eq_dmma_vec (mma.sync.m8n8k4.f64 -> DMMA): lane
// (gi, k) reads the two columns {2 gi, 2 gi + 1} of row 4u + k; H += (s2 k) k^T takes 3 DMMAs per 4 rows, g += s3 k
// two DFMAs.  J [B,C,N,P] (17 GB at cfg-4) is never written or read: per problem-iteration the kernel moves s N (C + 2)
// bytes instead of the tensor path's s C N (2 P + 4).
__device__ __forceinline__ void dmma_884_acc(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// GDA = true: gradient_descent_parallel's cost-summed tall system (gda:134-139): the tile rows are j = 2 s1 k rounded in T
// (s1 = sum_c w_c r_c), unscaled in the MMA; g64 = sum_n j f_s with f_s = sum_c w_c r_c^2; outputs fp64; fA / fB are the
// [B,N] ping-pong buffers of f_s (GDA) or of the model value m (otherwise); stats = {0, sum (f_s,prev - f_s)^2, sum f_s^2, 0}.
template <typename T, typename Model, int NW, int MINB, bool GDA = false>
__global__ void __launch_bounds__(NW * 32, MINB)
k_model_normal_eq_mma(const T* __restrict__ p, const T* __restrict__ x, const T* __restrict__ y, long long ybs, long long ycs,
                      const T* __restrict__ w, int wlen, long long B, int C, int N, T* fA, T* fB, void* __restrict__ gout_,
                      void* __restrict__ Hout_, double* __restrict__ stats, const nlls_ctrl* ctrl,
                      const unsigned char* __restrict__ sel, nlls_track tr) {
  using TO = typename std::conditional<GDA, double, T>::type;
  TO* __restrict__ gout = reinterpret_cast<TO*>(gout_);
  TO* __restrict__ Hout = reinterpret_cast<TO*>(Hout_);
  if (is_done(ctrl)) return;
  constexpr int P = Model::P, NB = 2, PP = 16, RW = PP * PP + PP + 2, TPB = NW * 32, ROWD = 20;
  static_assert(P <= PP, "tensor-core fused kernel: P <= 16 (tile columns past P are zero)");
  extern __shared__ double smem_d[];
  double* red = smem_d;                          // [NW][RW]: H (canonical entries), g, {sd, sq}
  double* tiles = red + NW * RW;                 // [NW][32][ROWD]
  double* csum = tiles + NW * 32 * ROWD;         // [C][TPB] per-thread per-cost sums of f
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int k = lane & 3, gi = lane >> 2;
  double* tile = tiles + warp * 32 * ROWD;
  const int it = iter_index(ctrl);
  const WorkList WL(tr, it, B);
  const int NBATCH = (N + 31) >> 5;

  for (long long wi = blockIdx.x; wi < WL.n; wi += gridDim.x) {
    const long long b = WL.at(wi);
    T* f_out = (it & 1) ? fB : fA;
    const T* f_prev = (it > 0 && fA) ? ((it & 1) ? fA : fB) : nullptr;
    if (sel) { const bool s1 = sel[b] != 0; f_out = s1 ? fA : fB; f_prev = s1 ? fB : fA; }
    const long long fstride = (long long)N;                            // m [B,N] / f_s [B,N]
    if (f_out) f_out += b * fstride;
    if (f_prev) f_prev += b * fstride;
    Model M;
    M.load(p + b * P);
    double Ht[3][2], Gt[NB];
#pragma unroll
    for (int i = 0; i < 3; ++i) Ht[i][0] = Ht[i][1] = 0.0;
    Gt[0] = Gt[1] = 0.0;
    double sd = 0.0, sq = 0.0;
    if (!GDA)
      for (int c = 0; c < C; ++c) csum[c * TPB + tid] = 0.0;
    for (int bt = warp; bt < NBATCH; bt += NW) {
      const int n = (bt << 5) + lane;
      const bool live = n < N;
      T m = T(0), kk[P];
#pragma unroll
      for (int i = 0; i < P; ++i) kk[i] = T(0);
      T s2 = T(0), s3 = T(0);
      if (live) {
        M.eval(ld_cached(x + n), m, kk);
        const T* yb = y + b * ybs + n;
        if constexpr (GDA) {
          T s1 = T(0), fs = T(0);
          for (int c = 0; c < C; ++c) {
            const T wc = w[wlen == 1 ? 0 : c];
            const T r = ld_cached(yb + c * ycs) - m;
            s1 = fma(wc, r, s1);
            fs = fma(wc, r * r, fs);                          // gda:136
          }
          if (f_out) f_out[n] = fs;
          if (f_prev) { const T d = ldg_stream(f_prev + n) - fs; sd = fma((double)d, (double)d, sd); }
          sq = fma((double)fs, (double)fs, sq);
          const T c2 = T(2) * s1;
#pragma unroll
          for (int i = 0; i < P; ++i) kk[i] = c2 * kk[i];     // gda:137: the cost-summed Jacobian row, rounded in T
          s3 = fs;
        } else {
          if (f_out) stg_stream(f_out + n, m);
          const T mp = f_prev ? ldg_stream(f_prev + n) : T(0);
          for (int c = 0; c < C; ++c) {
            const T wc = w[wlen == 1 ? 0 : c];
            const T yv = ld_cached(yb + c * ycs);
            const T r = yv - m;
            const T fv = mul_rn(r, r);
            if (f_prev) { const T rp = yv - mp; const T d = mul_rn(rp, rp) - fv; sd = fma((double)d, (double)d, sd); }
            sq = fma((double)fv, (double)fv, sq);
            csum[c * TPB + tid] += (double)fv;
            const T wf = wc * fv;
            s2 += wf;
            s3 = fma(wf, r, s3);
          }
        }
      }
      const double s2d = GDA ? 1.0 : (double)(T(4) * s2), s3d = GDA ? (double)s3 : (double)(T(2) * s3);
#pragma unroll
      for (int i = 0; i < PP; i += 2) {
        double2 v;
        v.x = i < P ? (double)kk[i < P ? i : 0] : 0.0;
        v.y = i + 1 < P ? (double)kk[i + 1 < P ? i + 1 : 0] : 0.0;
        *reinterpret_cast<double2*>(tile + lane * ROWD + i) = v;
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int row = 4 * u + k;
        const double2 v = *reinterpret_cast<const double2*>(tile + row * ROWD + NB * gi);
        const double s2k = __shfl_sync(0xffffffffu, s2d, row), s3k = __shfl_sync(0xffffffffu, s3d, row);
        const double a0 = GDA ? v.x : v.x * s2k, a1 = GDA ? v.y : v.y * s2k;
        dmma_884_acc(Ht[0][0], Ht[0][1], a0, v.x);      // block (0,0)
        dmma_884_acc(Ht[1][0], Ht[1][1], a0, v.y);      // block (0,1)
        dmma_884_acc(Ht[2][0], Ht[2][1], a1, v.y);      // block (1,1)
        Gt[0] = fma(v.x, s3k, Gt[0]);
        Gt[1] = fma(v.y, s3k, Gt[1]);
      }
      __syncwarp();                                // the tile is rewritten by the next batch
    }
    // ---- warp partials -> shared memory; entry (i, j) of block (I, Jb) is H[NB*gi + I][NB*gj + Jb]  (as in K1)
#pragma unroll
    for (int I = 0; I < NB; ++I) { Gt[I] += shfl_xor(Gt[I], 1); Gt[I] += shfl_xor(Gt[I], 2); }
    sd = warp_sum(sd);
    sq = warp_sum(sq);
    double* r = red + warp * RW;
    {
      int idx = 0;
#pragma unroll
      for (int I = 0; I < NB; ++I)
#pragma unroll
        for (int Jb = I; Jb < NB; ++Jb, ++idx) {
          r[(NB * gi + I) * PP + NB * (2 * k) + Jb] = Ht[idx][0];
          r[(NB * gi + I) * PP + NB * (2 * k + 1) + Jb] = Ht[idx][1];
        }
    }
    if (k == 0) {
#pragma unroll
      for (int I = 0; I < NB; ++I) r[PP * PP + NB * gi + I] = Gt[I];
    }
    if (lane == 0) { r[PP * PP + PP] = sd; r[PP * PP + PP + 1] = sq; }
    __syncthreads();
    for (int e = tid; e < P * P; e += TPB) {
      const int i = e / P, j = e - i * P;
      const int lo = min(i, j), hi = max(i, j);
      const int off = ((lo % NB) <= (hi % NB)) ? lo * PP + hi : hi * PP + lo;
      double sacc = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) sacc += red[ww * RW + off];
      Hout[b * P * P + e] = (TO)sacc;
    }
    if (tid < P) {
      double sacc = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) sacc += red[ww * RW + PP * PP + tid];
      gout[b * P + tid] = (TO)sacc;
    }
    if (warp == NW - 1) {
      double tot = 0.0;
      for (int c = 0; c < (GDA ? 0 : C); ++c) {
        double sacc = 0.0;
#pragma unroll
        for (int ww = 0; ww < NW; ++ww) sacc += csum[c * TPB + ww * 32 + lane];
        sacc = warp_sum(sacc);
        tot = fma(sacc, sacc, tot);
      }
      if (lane == 0) {
        double s1 = 0.0, s2t = 0.0;
#pragma unroll
        for (int ww = 0; ww < NW; ++ww) { s1 += red[ww * RW + PP * PP + PP]; s2t += red[ww * RW + PP * PP + PP + 1]; }
        stats[b * 4 + 0] = tot; stats[b * 4 + 1] = s1; stats[b * 4 + 2] = s2t; stats[b * 4 + 3] = 0.0;
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------ dispatch
// Tensor map of the data rows y viewed as [B][Cy][N] (strides y_bstride, y_cstride, 1) with a box of [1][Cy][TN]: the whole
// data part of a tile stage arrives with ONE tensor copy instead of Cy bulk copies.  The encoder is a host-side function of the
// driver, fetched once through the runtime (no link against libcuda); false = not available / shape not encodable.
static bool make_rows_tensor_map(CUtensorMap* out, size_t es, const void* y, long long B, int Cy, int N, long long ybs,
                                 long long ycs, int TN) {
  typedef CUresult (*encode_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static encode_t encode = nullptr;
  static int tried = 0;
  if (!tried) {
    tried = 1;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      encode = (encode_t)fn;
    else
      cudaGetLastError();
  }
  if (!encode || TN > 256 || Cy > 256 || ((size_t)TN * es) % 16) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)N, (cuuint64_t)Cy, (cuuint64_t)B};
  const cuuint64_t strides[2] = {(cuuint64_t)((Cy > 1 ? ycs : (long long)N) * (long long)es), (cuuint64_t)(ybs * (long long)es)};
  const cuuint32_t box[3] = {(cuuint32_t)TN, (cuuint32_t)Cy, 1u}, estr[3] = {1u, 1u, 1u};
  if (strides[0] % 16 || strides[1] % 16 || strides[0] == 0 || strides[1] == 0) return false;
  return encode(out, es == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(y), dims,
                strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int use_tensor_maps() {                                   // NLLS_TENSOR_MAP=0: keep the per-row bulk copies (A/B hook)
  static int v = -1;
  if (v < 0) { const char* e = getenv("NLLS_TENSOR_MAP"); v = e ? atoi(e) : 1; }
  return v;
}

template <typename K>
static int grid_for(K kernel, int tpb, size_t smem, long long B) {
  int per_sm = 0;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, tpb, smem);
  if (per_sm < 1) per_sm = 1;
  int dev = 0, sms = kNumSMs;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long grid = (long long)sms * per_sm;
  if (grid > B) grid = B;
  return (int)grid;
}

struct ModelArgs {
  const void *p, *x, *y, *w;
  long long ybs, ycs, B;
  int wlen, C, N;
  void *fA, *fB, *g, *H;             // fA / fB: [B,N] ping-pong buffers (model value m, or GD's cost-summed f_s)
  double *g64, *H64, *stats, *stat_out;
  void *f_out, *J_out;
  const unsigned char* sel;          // speculative LM: per-problem current-buffer flags (NULL otherwise)
  nlls_track tr;                     // per-problem tracking / retirement (zeroed: none)
  const nlls_ctrl* ctrl;
  cudaStream_t st;
};

enum Op { OP_NEQ, OP_NEQ_GDA, OP_COST, OP_EVAL };

// tuning hook: NLLS_FUSED_VARIANT selects the (threads, samples per thread) shape of the staged fused kernel
static int fused_variant() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("NLLS_FUSED_VARIANT"); v = e ? atoi(e) : 0; }
  return v;
}

// default kernel family of the CTA-per-problem fused kernels: 128-bit lanes for fp32, strided map for fp64
template <typename T> static bool vec_family() { return sizeof(T) == 4 && fused_variant() == 0; }

// returns 0 when launched, 1 when the shape is not eligible for bulk copies, or a launch error
template <typename T, typename Model, int TPB2, int S2, int CT = 0, int MINB = 0>
static int launch_staged(const ModelArgs& a) {
  constexpr int P = Model::P, V = P + Tri<P>::N + 2, NW2 = TPB2 / 32, TN = TPB2 * S2, VP = Pow2Ceil<V>::value;
  const size_t es = sizeof(T);
  const int Cy = a.ycs == 0 ? 1 : a.C;
  const size_t stage = (size_t)(Cy + 1) * TN * es;
  const size_t tail = 16 + 64 + ((size_t)2 * NW2 * VP + (size_t)2 * a.C * TPB2) * 8;
  const bool ok16 = aligned_to(a.y, 16) && (a.ybs * es) % 16 == 0 && (a.ycs * es) % 16 == 0 && ((size_t)a.N * es) % 16 == 0 &&
                    (!a.fA || (aligned_to(a.fA, 16) && aligned_to(a.fB, 16)));
  if (!ok16) return 1;
  if (CT > 0 && a.C != CT) return 1;
  // as many CTAs per SM as registers allow (64K regs / (128 regs x TPB2)); 3 stages if they still fit, else 2
  const int want_ctas = MINB > 0 ? MINB : 512 / TPB2;
  const bool three = (size_t)want_ctas * (3 * stage + tail + 1024) <= 226 * 1024;
  const size_t smem2 = (three ? 3 : 2) * stage + tail;
  if (smem2 > 220 * 1024) return 1;
  CUtensorMap tmy{};
  const int use_tm = (a.N % TN == 0) && use_tensor_maps() && ((size_t)TN * es) % 128 == 0 &&
                     make_rows_tensor_map(&tmy, es, a.y, a.B, Cy, a.N, a.ybs, a.ycs, TN);
  auto launch = [&](auto kern2) {
    const int grid2 = grid_for(kern2, TPB2, smem2, a.B);
    kern2<<<grid2, TPB2, smem2, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen,
                                        a.B, a.C, a.N, (T*)a.fA, (T*)a.fB, (T*)a.g, (T*)a.H, a.stats, a.ctrl, a.sel, a.tr, tmy,
                                        use_tm);
  };
  if (three) launch(k_model_normal_eq_staged<T, Model, TPB2, S2, 3, CT, MINB>);
  else launch(k_model_normal_eq_staged<T, Model, TPB2, S2, 2, CT, MINB>);
  return check_launch("nlls_model_normal_eq(staged)");
}


// fp64 default: up to 8 costs as a compile-time constant, SINGLE-WARP CTAs with 4 samples per thread, 12 CTAs per SM (168
// registers): 20 independent residual chains per thread and tile, and no CTA barrier to wait in (10 % of the stalls of the
// 64-thread shape, profiles/r02n).  cfg-2 GN fp64 (profiles/r02e, r02k): run-time C at 128 x 2: 1.065 ms; compile-time C at
// 128 x 2 (128 registers): 0.998 ms; at 64 x 4: 0.929 ms (0.923 on the r02k box); at 32 x 4: 0.894 ms on that box.
template <typename T, typename Model>
static int launch_staged_ct(const ModelArgs& a) {
  if constexpr (sizeof(T) == 8) {
    switch (a.C) {
      case 1: return launch_staged<T, Model, 32, 4, 1, 12>(a);
      case 2: return launch_staged<T, Model, 32, 4, 2, 12>(a);
      case 3: return launch_staged<T, Model, 32, 4, 3, 12>(a);
      case 4: return launch_staged<T, Model, 32, 4, 4, 12>(a);
      case 5: return launch_staged<T, Model, 32, 4, 5, 12>(a);
      case 6: return launch_staged<T, Model, 32, 4, 6, 12>(a);
      case 7: return launch_staged<T, Model, 32, 4, 7, 12>(a);
      case 8: return launch_staged<T, Model, 32, 4, 8, 12>(a);
      default: break;
    }
  }
  return launch_staged<T, Model, 128, 2>(a);
}

// 128-bit-lane staged kernel: same eligibility as launch_staged plus a 16-byte aligned abscissa
template <typename T, typename Model, int TPB2, int CT = 0>
static int launch_vec(const ModelArgs& a) {
  if (CT > 0 && a.C != CT) return 1;
  constexpr int P = Model::P, V = P + Tri<P>::N + 2, NW2 = TPB2 / 32, TN = TPB2 * Lane16<T>::VEC, VP = Pow2Ceil<V>::value;
  const size_t es = sizeof(T);
  const int Cy = a.ycs == 0 ? 1 : a.C;
  const size_t stage = (size_t)(Cy + 1) * TN * es;
  const size_t tail = 16 + 64 + ((size_t)2 * NW2 * VP + (size_t)2 * a.C * TPB2) * 8;
  const bool ok16 = aligned_to(a.y, 16) && aligned_to(a.x, 16) && (a.ybs * es) % 16 == 0 && (a.ycs * es) % 16 == 0 &&
                    ((size_t)a.N * es) % 16 == 0 && (!a.fA || (aligned_to(a.fA, 16) && aligned_to(a.fB, 16)));
  if (!ok16 || (NW2 > 1 && NW2 * 32 < V + 32)) return 1;
  const int want_ctas = 512 / TPB2;
  const bool three = (size_t)want_ctas * (3 * stage + tail + 1024) <= 226 * 1024;
  const size_t smem2 = (three ? 3 : 2) * stage + tail;
  if (smem2 > 220 * 1024) return 1;
  CUtensorMap tmy{};
  const int use_tm = (a.N % TN == 0) && use_tensor_maps() && ((size_t)TN * es) % 128 == 0 &&
                     make_rows_tensor_map(&tmy, es, a.y, a.B, Cy, a.N, a.ybs, a.ycs, TN);
  auto launch = [&](auto kern2) {
    const int grid2 = grid_for(kern2, TPB2, smem2, a.B);
    kern2<<<grid2, TPB2, smem2, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen,
                                        a.B, a.C, a.N, (T*)a.fA, (T*)a.fB, (T*)a.g, (T*)a.H, a.stats, a.ctrl, a.sel, a.tr, tmy,
                                        use_tm);
  };
  if (three) launch(k_model_normal_eq_vec<T, Model, TPB2, 3, CT>);
  else launch(k_model_normal_eq_vec<T, Model, TPB2, 2, CT>);
  return check_launch("nlls_model_normal_eq(vec)");
}

// tuning hook: NLLS_WPP_VL=0 keeps the scalar lane map of the warp-per-problem kernels
static bool wpp_vector_lanes() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("NLLS_WPP_VL"); v = e ? atoi(e) : 1; }
  return v != 0;
}

// warp-per-problem variant: chunk size CH (multiple of 64 samples) chosen so that a stage is ~2-4 KB
template <typename T, typename Model, int MODE>
static int launch_wpp(const ModelArgs& a) {
  constexpr int NW = 4, SB = 2, NSTAGE = 3;
  const size_t es = sizeof(T);
  const int Cy = a.ycs == 0 ? 1 : a.C;
  const int CP = MODE == WPP_COST ? 0 : 1;
  const bool ok16 = aligned_to(a.y, 16) && (a.ybs * es) % 16 == 0 && (a.ycs * es) % 16 == 0 && ((size_t)a.N * es) % 16 == 0 &&
                    (!a.fA || (aligned_to(a.fA, 16) && aligned_to(a.fB, 16)));
  if (!ok16) return 1;
  const bool c1 = a.C == 1;
  // single-cost problems with a 16-byte aligned abscissa: vector lanes (blocks of 32 * 16/sizeof(T) samples)
  const bool vl = c1 && aligned_to(a.x, 16) && wpp_vector_lanes();
  static int chmax = -1, minb = -1;                           // tuning hooks: NLLS_WPP_CHMAX (samples per chunk), NLLS_WPP_MINB (CTAs per SM)
  if (chmax < 0) { const char* e = getenv("NLLS_WPP_CHMAX"); chmax = e ? atoi(e) : 512; }
  if (minb < 0) { const char* e = getenv("NLLS_WPP_MINB"); minb = e ? atoi(e) : 0; }
  int CH = vl ? 128 : 64;
  while (CH < chmax && CH < a.N && (size_t)(Cy + CP) * (CH * 2) * es <= 4096) CH *= 2;
  const bool use_sfc = !c1 && MODE != WPP_GDA;
  const size_t per_warp = (size_t)NSTAGE * (Cy + CP) * CH * es + 64 + (use_sfc ? (size_t)a.C * 32 * 8 : 0);
  const size_t smem = per_warp * NW;
  if (smem > 200 * 1024) return 1;
  void* og = MODE == WPP_GDA ? (void*)a.g64 : a.g;
  void* oH = MODE == WPP_GDA ? (void*)a.H64 : a.H;
  double* os = MODE == WPP_COST ? a.stat_out : a.stats;
  auto go = [&](auto kern) {
    const int grid = grid_for(kern, NW * 32, smem, (a.B + NW - 1) / NW);
    kern<<<grid, NW * 32, smem, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen, a.B,
                                        a.C, a.N, (T*)a.fA, (T*)a.fB, og, oH, os, a.ctrl, CH, MODE == WPP_NEQ ? a.sel : nullptr,
                                        MODE == WPP_COST ? nlls_track{} : a.tr);
  };
#ifndef NLLS_USER_MODEL_HEADER
  if (vl && minb == 5) go(k_model_wpp<T, Model, MODE, 16 / (int)sizeof(T), NSTAGE, NW, true, true, 5>);
  else if (vl && minb == 6) go(k_model_wpp<T, Model, MODE, 16 / (int)sizeof(T), NSTAGE, NW, true, true, 6>);
  else if (vl && minb == 8) go(k_model_wpp<T, Model, MODE, 16 / (int)sizeof(T), NSTAGE, NW, true, true, 8>);
  else
#endif
  if (vl) go(k_model_wpp<T, Model, MODE, 16 / (int)sizeof(T), NSTAGE, NW, true, true>);
  else if (c1) go(k_model_wpp<T, Model, MODE, SB, NSTAGE, NW, true>);
  else go(k_model_wpp<T, Model, MODE, SB, NSTAGE, NW, false>);
  return check_launch("nlls_model(wpp)");
}

// fp32 default: 128-bit lanes, single-warp CTAs (no CTA barrier: 21 % of the stalls of the 64-thread shape, profiles/r02n;
// cfg-2 fp32 0.527 -> 0.486 ms, profiles/r02o), up to 8 costs as a compile-time constant
template <typename T, typename Model>
static int launch_vec_ct(const ModelArgs& a) {
  if constexpr (sizeof(T) == 4) {
    switch (a.C) {
      case 1: return launch_vec<T, Model, 32, 1>(a);
      case 2: return launch_vec<T, Model, 32, 2>(a);
      case 3: return launch_vec<T, Model, 32, 3>(a);
      case 4: return launch_vec<T, Model, 32, 4>(a);
      case 5: return launch_vec<T, Model, 32, 5>(a);
      case 6: return launch_vec<T, Model, 32, 6>(a);
      case 7: return launch_vec<T, Model, 32, 7>(a);
      case 8: return launch_vec<T, Model, 32, 8>(a);
      default: break;
    }
  }
  return launch_vec<T, Model, 64>(a);
}

template <typename T>
static bool rows16(const ModelArgs& a) {
  const size_t es = sizeof(T);
  return aligned_to(a.y, 16) && (a.ybs * es) % 16 == 0 && (a.ycs * es) % 16 == 0 && ((size_t)a.N * es) % 16 == 0;
}

template <typename T, typename Model, int TPB2, int S2>
static int launch_cost_staged(const ModelArgs& a) {
  constexpr int TN = TPB2 * S2;
  if (!rows16<T>(a)) return 1;
  const int Cy = a.ycs == 0 ? 1 : a.C;
  const size_t tail = (size_t)2 * a.C * TPB2 * 8;
  const int want_ctas = 1024 / TPB2;                         // cost-only evaluation: ~64 registers per thread
  const bool three = (size_t)want_ctas * (TileRing<T, TN, 3>::bytes(Cy, 0) + tail + 1024) <= 226 * 1024;
  const size_t smem2 = (three ? TileRing<T, TN, 3>::bytes(Cy, 0) : TileRing<T, TN, 2>::bytes(Cy, 0)) + tail;
  if (smem2 > 220 * 1024) return 1;
  auto launch = [&](auto kern2) {
    const int grid2 = grid_for(kern2, TPB2, smem2, a.B);
    kern2<<<grid2, TPB2, smem2, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, a.B, a.C, a.N, a.stat_out,
                                        a.ctrl);
  };
  if (three) launch(k_model_cost_stat_staged<T, Model, TPB2, S2, 3>);
  else launch(k_model_cost_stat_staged<T, Model, TPB2, S2, 2>);
  return check_launch("nlls_model_cost_stat(staged)");
}

template <typename T, typename Model, int TPB2>
static int launch_cost_vec(const ModelArgs& a) {
  constexpr int TN = TPB2 * Lane16<T>::VEC;
  if (!rows16<T>(a) || !aligned_to(a.x, 16)) return 1;
  const int Cy = a.ycs == 0 ? 1 : a.C;
  const size_t tail = (size_t)2 * a.C * TPB2 * 8;
  const int want_ctas = 1024 / TPB2;
  const bool three = (size_t)want_ctas * (TileRing<T, TN, 3>::bytes(Cy, 0) + tail + 1024) <= 226 * 1024;
  const size_t smem2 = (three ? TileRing<T, TN, 3>::bytes(Cy, 0) : TileRing<T, TN, 2>::bytes(Cy, 0)) + tail;
  if (smem2 > 220 * 1024) return 1;
  auto launch = [&](auto kern2) {
    const int grid2 = grid_for(kern2, TPB2, smem2, a.B);
    kern2<<<grid2, TPB2, smem2, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, a.B, a.C, a.N, a.stat_out,
                                        a.ctrl);
  };
  if (three) launch(k_model_cost_stat_vec<T, Model, TPB2, 3>);
  else launch(k_model_cost_stat_vec<T, Model, TPB2, 2>);
  return check_launch("nlls_model_cost_stat(vec)");
}

template <typename T, typename Model, int TPB2, int S2, int CT = 0, int MINB = 0>
static int launch_gda_staged(const ModelArgs& a) {
  constexpr int P = Model::P, V = P + Tri<P>::N + 2, NW2 = TPB2 / 32, TN = TPB2 * S2, VP = Pow2Ceil<V>::value;
  if (CT > 0 && a.C != CT) return 1;
  if (!rows16<T>(a) || (a.fA && !(aligned_to(a.fA, 16) && aligned_to(a.fB, 16)))) return 1;
  const int Cy = a.ycs == 0 ? 1 : a.C;
  const size_t tail = (size_t)2 * NW2 * VP * 8;
  const int want_ctas = MINB > 0 ? MINB : 512 / TPB2;
  const bool three = (size_t)want_ctas * (TileRing<T, TN, 3>::bytes(Cy, 1) + tail + 1024) <= 226 * 1024;
  const size_t smem2 = (three ? TileRing<T, TN, 3>::bytes(Cy, 1) : TileRing<T, TN, 2>::bytes(Cy, 1)) + tail;
  if (smem2 > 220 * 1024) return 1;
  CUtensorMap tmy{};
  const int use_tm = (a.N % TN == 0) && use_tensor_maps() && ((size_t)TN * sizeof(T)) % 128 == 0 &&
                     make_rows_tensor_map(&tmy, sizeof(T), a.y, a.B, Cy, a.N, a.ybs, a.ycs, TN);
  auto launch = [&](auto kern2) {
    const int grid2 = grid_for(kern2, TPB2, smem2, a.B);
    kern2<<<grid2, TPB2, smem2, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen, a.B,
                                        a.C, a.N, (T*)a.fA, (T*)a.fB, a.g64, a.H64, a.stats, a.ctrl, a.tr, tmy, use_tm);
  };
  if (three) launch(k_model_normal_eq_gda_staged<T, Model, TPB2, S2, 3, CT, MINB>);
  else launch(k_model_normal_eq_gda_staged<T, Model, TPB2, S2, 2, CT, MINB>);
  return check_launch("nlls_model_normal_eq_gda(staged)");
}

template <typename T, typename Model>
static int launch_eval(const ModelArgs& a) {
  auto kern = k_model_eval<T, Model>;
  const long long target_warps = 148LL * 8 * 8;              // 148 SMs x 8 CTAs x 8 warps
  long long wpp = a.B >= target_warps ? 1 : (target_warps + a.B - 1) / a.B;   // warps per problem when the batch is small
  const long long max_wpp = (a.N + 63) / 64;
  if (wpp > max_wpp) wpp = max_wpp;
  if (wpp < 1) wpp = 1;
  long long blocks = (a.B * wpp + 7) / 8;
  if (blocks > 148LL * 8) blocks = 148LL * 8;
  kern<<<(int)blocks, 256, 0, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, a.B, a.C, a.N,
                                      (T*)a.f_out, (T*)a.J_out, a.ctrl, (int)wpp);
  return check_launch("nlls_model_eval");
}

template <typename T, typename Model>
static int run_op(Op op, const ModelArgs& a) {
  constexpr int TPB = kFusedTPB, S = kFusedS, NW = TPB / 32, P = Model::P, V = P + Tri<P>::N + 2;
  switch (op) {
    case OP_NEQ: {
      const size_t smem = ((size_t)NW * V + (size_t)a.C * TPB + (size_t)a.C * NW) * 8;
      if (smem > 200 * 1024) { set_error("fused model path: too many costs (C=%d)", a.C); return NLLS_E_SHAPE; }
      if (!g_force_direct) {   // bulk-copy staged variant whenever every data row is 16-byte addressable
        int rc = 1;
        // single-cost problems of a few thousand samples: warp per problem (cfg-3: 2.5 -> 1.8 ms, profiles/r01e)
        if (fused_variant() == 4 || (fused_variant() == 0 && a.C == 1 && a.N <= 4096)) {
          rc = launch_wpp<T, Model, WPP_NEQ>(a);
          if (rc <= 0) return rc;
        }
        switch (fused_variant()) {
#ifndef NLLS_USER_MODEL_HEADER                               // tuning variants: built-in models only (plugin compile time)
          case 1: rc = launch_staged<T, Model, 256, 2>(a); break;
          case 2: rc = launch_staged<T, Model, 64, 2>(a); break;
          case 5: rc = launch_vec<T, Model, 128>(a); break;
          case 6: rc = launch_vec<T, Model, 64>(a); break;
          case 7: rc = launch_vec<T, Model, 256>(a); break;
          case 8: rc = launch_staged<T, Model, 128, 2, 5, 4>(a); break;       // compile-time C = 5 at other CTA shapes
          case 10: rc = launch_staged<T, Model, 128, 2, 5, 3>(a); break;
          case 12: rc = launch_staged<T, Model, 64, 4, 5, 6>(a); break;       // the round-2 64 x 4 shape
          case 13: rc = launch_staged<T, Model, 32, 8, 5, 8>(a); break;
          case 15: rc = launch_vec<T, Model, 32, 5>(a); break;                // fp32 lanes, single-warp CTAs
          case 16: rc = launch_vec<T, Model, 128, 5>(a); break;
          case 11: rc = launch_staged<T, Model, 128, 2, 0, 0>(a); break;      // run-time C (the C > 8 kernel)
#endif
          default:
            // fp32: 128-bit lanes, 64-thread CTAs (cfg-2 fp32 1.01 -> 0.76 ms, profiles/r01f); fp64 is HBM-bound either
            // way and measures a little better with the strided map at 4 CTAs/SM (r01 sweeps)
            if (vec_family<T>()) rc = launch_vec_ct<T, Model>(a);
            if (rc == 1) rc = launch_staged_ct<T, Model>(a);
            break;
        }
        if (rc <= 0) return rc;                              // 0 launched, < 0 error, 1 = shape not eligible
      }
      auto kern = k_model_normal_eq<T, Model, TPB, S>;
      const int grid = grid_for(kern, TPB, smem, a.B);
      kern<<<grid, TPB, smem, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen, a.B,
                                      a.C, a.N, (T*)a.fA, (T*)a.fB, (T*)a.g, (T*)a.H, a.stats, a.ctrl, a.sel, a.tr);
      return check_launch("nlls_model_normal_eq");
    }
    case OP_NEQ_GDA: {
      if (!g_force_direct) {
        if (fused_variant() == 4 || (fused_variant() == 0 && a.C == 1 && a.N <= 4096)) {
          const int rcw = launch_wpp<T, Model, WPP_GDA>(a);
          if (rcw <= 0) return rcw;
        }
        int rc = 1;
        if constexpr (sizeof(T) == 8) {                        // fp64: the NEQ kernel's shape (single-warp CTAs, 4 samples per thread, compile-time C)
          switch (fused_variant() == 0 ? a.C : 0) {
            case 1: rc = launch_gda_staged<T, Model, 32, 4, 1, 12>(a); break;
            case 2: rc = launch_gda_staged<T, Model, 32, 4, 2, 12>(a); break;
            case 3: rc = launch_gda_staged<T, Model, 32, 4, 3, 12>(a); break;
            case 4: rc = launch_gda_staged<T, Model, 32, 4, 4, 12>(a); break;
            case 5: rc = launch_gda_staged<T, Model, 32, 4, 5, 12>(a); break;
            case 6: rc = launch_gda_staged<T, Model, 32, 4, 6, 12>(a); break;
            case 7: rc = launch_gda_staged<T, Model, 32, 4, 7, 12>(a); break;
            case 8: rc = launch_gda_staged<T, Model, 32, 4, 8, 12>(a); break;
            default: break;
          }
        }
        if (rc == 1) rc = launch_gda_staged<T, Model, 128, 2>(a);
        if (rc <= 0) return rc;
      }
      auto kern = k_model_normal_eq_gda<T, Model, TPB>;
      const int grid = grid_for(kern, TPB, 0, a.B);
      kern<<<grid, TPB, 0, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen, a.B, a.C,
                                   a.N, (T*)a.fA, (T*)a.fB, a.g64, a.H64, a.stats, a.ctrl, a.tr);
      return check_launch("nlls_model_normal_eq_gda");
    }
    case OP_COST: {
      const size_t smem = ((size_t)a.C * TPB + (size_t)a.C * NW) * 8;
      if (smem > 200 * 1024) { set_error("fused model path: too many costs (C=%d)", a.C); return NLLS_E_SHAPE; }
      if (!g_force_direct) {
        if (fused_variant() == 4 || (fused_variant() == 0 && a.C == 1 && a.N <= 4096)) {
          const int rcw = launch_wpp<T, Model, WPP_COST>(a);
          if (rcw <= 0) return rcw;
        }
        int rc = 1;
#ifndef NLLS_USER_MODEL_HEADER
        if (fused_variant() == 15) rc = launch_cost_vec<T, Model, 32>(a);
        else if (fused_variant() == 5) rc = launch_cost_vec<T, Model, 128>(a);
        else if (fused_variant() == 7) rc = launch_cost_vec<T, Model, 256>(a);
        else if (fused_variant() == 6) rc = launch_cost_vec<T, Model, 64>(a);
        else
#endif
        if (vec_family<T>()) rc = a.C <= 8 ? launch_cost_vec<T, Model, 32>(a) : launch_cost_vec<T, Model, 64>(a);
        // same sample-to-thread map as the normal-equation kernel of the same shape (launch_staged_ct): the two-pass and
        // the speculative LM schedules see the same bits in stats[0]
        if (rc == 1 && sizeof(T) == 8 && a.C <= 8) rc = launch_cost_staged<T, Model, 32, 4>(a);
        if (rc == 1) rc = launch_cost_staged<T, Model, 128, 2>(a);
        if (rc <= 0) return rc;
      }
      auto kern = k_model_cost_stat<T, Model, TPB, S>;
      const int grid = grid_for(kern, TPB, smem, a.B);
      kern<<<grid, TPB, smem, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, a.B, a.C, a.N, a.stat_out,
                                      a.ctrl);
      return check_launch("nlls_model_cost_stat");
    }
    case OP_EVAL: return launch_eval<T, Model>(a);
  }
  return NLLS_E_MODEL;
}

// models with 6 < P <= 16: materialising kernel + the tensor-core fused normal equations (GN, speculative LM, GD);
// no two-pass trial-cost kernel
template <typename T, typename Model>
static int run_wide(Op op, const ModelArgs& a, const char* name) {
  if (op == OP_EVAL) return launch_eval<T, Model>(a);
  if (op == OP_NEQ || op == OP_NEQ_GDA) {
    constexpr int RW = 16 * 16 + 16 + 2;
    const bool gda = op == OP_NEQ_GDA;
    auto go = [&](auto kern, int NWv) {
      const int TPB = NWv * 32;
      const size_t smem = ((size_t)NWv * RW + (size_t)NWv * 32 * 20 + (size_t)a.C * TPB) * 8;
      if (smem > 200 * 1024) { set_error("fused model path: too many costs (C=%d)", a.C); return (int)NLLS_E_SHAPE; }
      const int grid = grid_for(kern, TPB, smem, a.B);
      kern<<<grid, TPB, smem, a.st>>>((const T*)a.p, (const T*)a.x, (const T*)a.y, a.ybs, a.ycs, (const T*)a.w, a.wlen, a.B,
                                      a.C, a.N, (T*)a.fA, (T*)a.fB, gda ? (void*)a.g64 : a.g, gda ? (void*)a.H64 : a.H, a.stats,
                                      a.ctrl, a.sel, a.tr);
      return check_launch(gda ? "nlls_model_normal_eq_gda(mma)" : "nlls_model_normal_eq(mma)");
    };
    if (gda) return go(k_model_normal_eq_mma<T, Model, 4, 6, true>, 4);
    switch (fused_variant()) {                               // tuning hook (occupancy against registers)
#ifndef NLLS_USER_MODEL_HEADER
      case 1: return go(k_model_normal_eq_mma<T, Model, 4, 5>, 4);   // 96 registers, 20 warps/SM: 6.08 ms at cfg-4
      case 2: return go(k_model_normal_eq_mma<T, Model, 4, 4>, 4);   // 128 registers, 16 warps/SM: 6.14 ms
      case 3: return go(k_model_normal_eq_mma<T, Model, 4, 3>, 4);   // 163 registers, 12 warps/SM: 6.48 ms (198 regs, 4 warps: 8.4 ms)
#endif
      // the kernel is bound by the latency of the fp64 model math: occupancy beats registers (profiles/r01j): 80 registers
      // (a few spills), 24 warps/SM: 5.93 ms
      default: return go(k_model_normal_eq_mma<T, Model, 4, 6>, 4);
    }
  }
  set_error("%s (P=%d): no two-pass trial-cost kernel (nlls_model_cost_stat) for P > 6; use nlls_model_lm_trial", name, Model::P);
  return NLLS_E_MODEL;
}

// Build-time split (Makefile): NLLS_FUSED_PART == 1 compiles only the fp32 instantiations, NLLS_FUSED_PART == 0 everything
// else plus the entry points; undefined (plugin build) = one translation unit.
template <typename T>
int run_model(int model, int P, Op op, const ModelArgs& a) {
#ifdef NLLS_USER_MODEL_HEADER
  if (model == NLLS_MODEL_USER && P == UserModel<T>::P) {
    if constexpr (UserModel<T>::P <= 6) return run_op<T, UserModel<T>>(op, a);        // register-resident fused kernels
    else return run_wide<T, UserModel<T>>(op, a, "compiled model");                   // 6 < P <= 16: tensor-core fused kernel
  }
  set_error("plugin library: model id %d / P=%d does not match the compiled functor (id %d, P=%d)", model, P, NLLS_MODEL_USER,
            UserModel<T>::P);
  return NLLS_E_MODEL;
#else
  if (model == NLLS_MODEL_EMG && P == 4) return run_op<T, EmgModel<T>>(op, a);
  if (model == NLLS_MODEL_SKEWGAUSS && P == 4) return run_op<T, SkewGaussModel<T, 4>>(op, a);
  if (model == NLLS_MODEL_SKEWGAUSS && P == 5) return run_op<T, SkewGaussModel<T, 5>>(op, a);
  if (model == NLLS_MODEL_MULTI_EMG && P == 8) return run_wide<T, MultiEmgModel<T, 2>>(op, a, "2-peak EMG");
  if (model == NLLS_MODEL_MULTI_EMG && P == 12) return run_wide<T, MultiEmgModel<T, 3>>(op, a, "3-peak EMG");
  if (model == NLLS_MODEL_MULTI_EMG && P == 16) return run_wide<T, MultiEmgModel<T, 4>>(op, a, "4-peak EMG");
  set_error("unknown model %d with P=%d (EMG: P=4, skewed Gaussian: P=4|5, multi-peak EMG: P=8|12|16)", model, P);
  return NLLS_E_MODEL;
#endif
}

#if defined(NLLS_FUSED_PART) && NLLS_FUSED_PART == 1
template int run_model<float>(int, int, Op, const ModelArgs&);
#elif defined(NLLS_FUSED_PART)
extern template int run_model<float>(int, int, Op, const ModelArgs&);
#endif

#if !defined(NLLS_FUSED_PART) || NLLS_FUSED_PART == 0
static int dispatch(int model, int dtype, int P, Op op, const ModelArgs& a) {
  if (dtype != NLLS_F32 && dtype != NLLS_F64) { set_error("dtype %d is not NLLS_F32/NLLS_F64", dtype); return NLLS_E_DTYPE; }
  if (a.B < 0 || a.C < 1 || a.N < 1 || (op <= OP_NEQ_GDA && a.wlen != 1 && a.wlen != a.C)) {
    set_error("unsupported shape B=%lld C=%d N=%d wlen=%d", a.B, a.C, a.N, a.wlen);
    return NLLS_E_SHAPE;
  }
  if (a.B == 0) return 0;
  if (!a.p || !a.x || !a.y) { set_error("fused model path: NULL tensor"); return NLLS_E_NULL; }
  return dtype == NLLS_F64 ? run_model<double>(model, P, op, a) : run_model<float>(model, P, op, a);
}

#endif

}  // namespace nlls

#if !defined(NLLS_FUSED_PART) || NLLS_FUSED_PART == 0
using namespace nlls;

extern "C" int nlls_model_eval(int model, int dtype, const void* p, int P, const void* x, const void* y, long long y_bstride,
                               long long y_cstride, long long B, int C, int N, void* f_out, void* J_out,
                               const nlls_ctrl* ctrl, void* stream) {
  ModelArgs a{};
  a.p = p; a.x = x; a.y = y; a.ybs = y_bstride; a.ycs = y_cstride; a.B = B; a.C = C; a.N = N; a.wlen = 1;
  a.f_out = f_out; a.J_out = J_out; a.ctrl = ctrl; a.st = (cudaStream_t)stream;
  if (!f_out) { set_error("nlls_model_eval: NULL f_out"); return NLLS_E_NULL; }
  return dispatch(model, dtype, P, OP_EVAL, a);
}

extern "C" int nlls_model_normal_eq(int model, int dtype, const void* p, int P, const void* x, const void* y,
                                    long long y_bstride, long long y_cstride, const void* wvec, int wlen, long long B, int C,
                                    int N, void* m_pp0, void* m_pp1, void* g, void* H, double* stats, const nlls_ctrl* ctrl,
                                    const nlls_track* track, void* stream) {
  ModelArgs a{};
  a.tr = track_or_none(track);
  a.p = p; a.x = x; a.y = y; a.w = wvec; a.ybs = y_bstride; a.ycs = y_cstride; a.B = B; a.C = C; a.N = N; a.wlen = wlen;
  a.fA = m_pp0; a.fB = m_pp1; a.g = g; a.H = H; a.stats = stats; a.ctrl = ctrl; a.st = (cudaStream_t)stream;
  if (!wvec || !g || !H || !stats || (m_pp0 && !m_pp1)) { set_error("nlls_model_normal_eq: NULL tensor"); return NLLS_E_NULL; }
  return dispatch(model, dtype, P, OP_NEQ, a);
}

extern "C" int nlls_model_normal_eq_gda(int model, int dtype, const void* p, int P, const void* x, const void* y,
                                        long long y_bstride, long long y_cstride, const void* wvec, int wlen, long long B,
                                        int C, int N, void* fsum_pp0, void* fsum_pp1, double* g64, double* H64, double* stats,
                                        const nlls_ctrl* ctrl, const nlls_track* track, void* stream) {
  ModelArgs a{};
  a.tr = track_or_none(track);
  a.p = p; a.x = x; a.y = y; a.w = wvec; a.ybs = y_bstride; a.ycs = y_cstride; a.B = B; a.C = C; a.N = N; a.wlen = wlen;
  a.fA = fsum_pp0; a.fB = fsum_pp1; a.g64 = g64; a.H64 = H64; a.stats = stats; a.ctrl = ctrl; a.st = (cudaStream_t)stream;
  if (!wvec || !g64 || !H64 || !stats || (fsum_pp0 && !fsum_pp1)) { set_error("nlls_model_normal_eq_gda: NULL tensor"); return NLLS_E_NULL; }
  return dispatch(model, dtype, P, OP_NEQ_GDA, a);
}

extern "C" int nlls_model_cost_stat(int model, int dtype, const void* p_trial, int P, const void* x, const void* y,
                                    long long y_bstride, long long y_cstride, long long B, int C, int N, double* stat_out,
                                    const nlls_ctrl* ctrl, void* stream) {
  ModelArgs a{};
  a.p = p_trial; a.x = x; a.y = y; a.ybs = y_bstride; a.ycs = y_cstride; a.B = B; a.C = C; a.N = N; a.wlen = 1;
  a.stat_out = stat_out; a.ctrl = ctrl; a.st = (cudaStream_t)stream;
  if (!stat_out) { set_error("nlls_model_cost_stat: NULL stat_out"); return NLLS_E_NULL; }
  return dispatch(model, dtype, P, OP_COST, a);
}

extern "C" int nlls_model_lm_trial(int model, int dtype, const void* p_trial, int P, const void* x, const void* y,
                                   long long y_bstride, long long y_cstride, const void* wvec, int wlen, long long B, int C,
                                   int N, void* m_buf0, void* m_buf1, const unsigned char* sel, void* g_t, void* H_t,
                                   double* stats_t, const nlls_ctrl* ctrl, const nlls_track* track, void* stream) {
  ModelArgs a{};
  a.tr = track_or_none(track);
  a.p = p_trial; a.x = x; a.y = y; a.w = wvec; a.ybs = y_bstride; a.ycs = y_cstride; a.B = B; a.C = C; a.N = N; a.wlen = wlen;
  a.fA = m_buf0; a.fB = m_buf1; a.g = g_t; a.H = H_t; a.stats = stats_t; a.sel = sel; a.ctrl = ctrl; a.st = (cudaStream_t)stream;
  if (!wvec || !g_t || !H_t || !stats_t || !m_buf0 || !m_buf1 || !sel) { set_error("nlls_model_lm_trial: NULL tensor"); return NLLS_E_NULL; }
  return dispatch(model, dtype, P, OP_NEQ, a);
}
#endif  // NLLS_FUSED_PART
