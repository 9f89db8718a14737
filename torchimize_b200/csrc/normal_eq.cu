// normal_eq.cu — K1: fused, wvec-weighted J^T W J / J^T W f reduction over materialised
// [B,C,N,P] Jacobians (the tensor path), plus the cost-summed GD flavour and the cost statistic.
//
// Replaces newton_parallel.py:40-43 (4 einsums, J read twice, gc/Hc round trips), the rho
// numerator's f pass (lma:100) and the fcon passes over f / f_prev (lma:110, gna:80, gda:76)
// with ONE streaming pass.  HBM-bound: s*C*N*(P+2) algorithmic bytes per problem.
//
// Two kernels for P <= 8:
//   k_normal_eq_staged  persistent CTAs; every warp runs its own ring of 1-D bulk async copies
//                       (cp.async.bulk -> SASS UBLKCP, completion on an mbarrier) so that
//                       S-1 chunks of J/f/f_prev are always in flight per warp, independent of
//                       register pressure; rows are consumed from shared memory.
//   k_normal_eq_direct  fallback for shapes the bulk copy cannot take (16-byte alignment of
//                       every (b,c) segment): 256/128-bit read-only global loads.
// and k_normal_eq_wide for 8 < P <= 32, which splits the P x P triangle over 4 lanes per row.
#include "rowacc.cuh"

namespace nlls {

// ============================================================================ direct
template <typename T, int P, int TPB>
__global__ void __launch_bounds__(TPB)
k_normal_eq_direct(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fprev,
                   const T* __restrict__ w, int wlen, long long B, int C, int N,
                   T* __restrict__ g, T* __restrict__ H, double* __restrict__ stats,
                   const nlls_ctrl* ctrl, int vec_ok) {
  if (is_done(ctrl)) return;
  constexpr int NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2;
  extern __shared__ double smem[];
  double* red = smem;            // [NW][V]
  double* sfc = smem + NW * V;   // [C][NW]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    RowAcc<T, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    for (int c = 0; c < C; ++c) {
      const T wc = w[wlen == 1 ? 0 : c];
      const long long seg = (b * C + c) * (long long)N;
      const T* Jc = J + seg * P;
      const T* fc = f + seg;
      const T* pc = fprev ? fprev + seg : nullptr;
      double sf = 0.0;
#pragma unroll 4
      for (int n = tid; n < N; n += TPB) {
        T row[P];
        load_row<P>(Jc + (long long)n * P, row, vec_ok != 0);
        const T fv = ldg_stream(fc + n);
        acc.add(row, fv, wc);
        sf += (double)fv;
        sq = fma((double)fv, (double)fv, sq);
        if (pc) {
          const T d = ldg_stream(pc + n) - fv;
          sd = fma((double)d, (double)d, sd);
        }
      }
      sf = warp_sum(sf);
      if (lane == 0) sfc[c * NW + warp] = sf;
    }
    acc.warp_reduce();
    sd = warp_sum(sd);
    sq = warp_sum(sq);
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
      for (int k = 0; k < NW; ++k) s += red[k * V + tid];
      if (tid < P) {
        g[b * P + tid] = (T)s;
      } else if (tid < P + NH) {
        // packed index -> (i,j)
        int k = tid - P, i = 0;
        while (k >= P - i) { k -= P - i; ++i; }
        const int j = i + k;
        H[b * P * P + i * P + j] = (T)s;
        H[b * P * P + j * P + i] = (T)s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    } else if (tid == V) {
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
        for (int k = 0; k < NW; ++k) s += sfc[c * NW + k];
        tot = fma(s, s, tot);
      }
      stats[b * 4 + 0] = tot;
      stats[b * 4 + 3] = 0.0;
    }
    __syncthreads();
  }
}

// ============================================================================ staged
// R rows per chunk, S stages per warp, NW warps per CTA.
template <typename T, int P, int NW, int R, int S>
struct StagedCfg {
  static constexpr int NH = Tri<P>::N;
  static constexpr int V = P + NH + 2;
  static constexpr int J_BYTES = R * P * (int)sizeof(T);
  static constexpr int F_BYTES = R * (int)sizeof(T);
  static constexpr int STAGE_BYTES = J_BYTES + 2 * F_BYTES;
  static_assert(F_BYTES % 16 == 0, "chunk rows must keep 16-byte granularity");
  static constexpr size_t RING_BYTES = (size_t)NW * S * STAGE_BYTES;
  static size_t smem_bytes(int C) {
    return RING_BYTES + (size_t)NW * S * 8 + (size_t)2 * NW * V * 8 + (size_t)2 * C * NW * 8;
  }
};

template <typename T, int P, int NW, int R, int S>
__global__ void __launch_bounds__(NW * 32)
k_normal_eq_staged(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fprev,
                   const T* __restrict__ w, int wlen, long long B, int C, int N,
                   T* __restrict__ g, T* __restrict__ H, double* __restrict__ stats,
                   const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  using Cfg = StagedCfg<T, P, NW, R, S>;
  constexpr int NH = Cfg::NH, V = Cfg::V;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned char* ring = smem_raw + (size_t)warp * S * Cfg::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + Cfg::RING_BYTES) + warp * S;
  double* red = reinterpret_cast<double*>(smem_raw + Cfg::RING_BYTES + (size_t)NW * S * 8);  // [2][NW][V]
  double* sfc = red + 2 * NW * V;                                                             // [2][C][NW]

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();

  const int QN = (N + R - 1) / R;      // chunks per (b,c) segment
  const int Q = C * QN;                // chunks per problem
  const bool has_prev = fprev != nullptr;

  // ---- producer side: flat iterator over this warp's chunks, running S stages ahead
  long long fb = blockIdx.x;
  int fq = warp;
  if (fq >= Q) fb = B;                 // more warps than chunks: this warp only joins the barriers
  unsigned fk = 0;                     // chunks issued so far
  auto issue = [&]() {
    if (fb >= B) return;
    if (lane == 0) {
      const int c = fq / QN, j = fq - c * QN;
      const int rows = min(R, N - j * R);
      const long long row0 = (fb * C + c) * (long long)N + (long long)j * R;
      const unsigned st = fk % S;
      unsigned char* dst = ring + (size_t)st * Cfg::STAGE_BYTES;
      const unsigned jb = (unsigned)rows * P * (unsigned)sizeof(T), fbyt = (unsigned)rows * (unsigned)sizeof(T);
      mbar_expect_tx(&bars[st], jb + fbyt + (has_prev ? fbyt : 0u));
      bulk_g2s(dst, J + row0 * P, jb, &bars[st]);
      bulk_g2s(dst + Cfg::J_BYTES, f + row0, fbyt, &bars[st]);
      if (has_prev) bulk_g2s(dst + Cfg::J_BYTES + Cfg::F_BYTES, fprev + row0, fbyt, &bars[st]);
    }
    ++fk;
    fq += NW;
    if (fq >= Q) { fq = warp; fb += gridDim.x; }
  };
#pragma unroll
  for (int s = 0; s < S; ++s) issue();

  // ---- consumer side
  unsigned ck = 0;                     // chunks consumed so far
  int pbuf = 0;
  for (long long b = blockIdx.x; b < B; b += gridDim.x, pbuf ^= 1) {
    RowAcc<T, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0, sf = 0.0;
    double* my_sfc = sfc + (size_t)pbuf * C * NW;
    for (int c = lane; c < C; c += 32) my_sfc[c * NW + warp] = 0.0;
    __syncwarp();
    int cur_c = -1;
    for (int q = warp; q < Q; q += NW) {
      const int c = q / QN, j = q - c * QN;
      const int rows = min(R, N - j * R);
      if (c != cur_c) {
        if (cur_c >= 0) {
          sf = warp_sum(sf);
          if (lane == 0) my_sfc[cur_c * NW + warp] = sf;
          sf = 0.0;
        }
        cur_c = c;
      }
      const T wc = w[wlen == 1 ? 0 : c];
      const unsigned st = ck % S, ph = (ck / S) & 1u;
      mbar_wait(&bars[st], ph);
      const unsigned char* src = ring + (size_t)st * Cfg::STAGE_BYTES;
      const T* sJ = reinterpret_cast<const T*>(src);
      const T* sF = reinterpret_cast<const T*>(src + Cfg::J_BYTES);
      const T* sP = reinterpret_cast<const T*>(src + Cfg::J_BYTES + Cfg::F_BYTES);
#pragma unroll
      for (int r0 = 0; r0 < R; r0 += 32) {
        const int r = r0 + lane;
        if (r < rows) {
          T row[P];
          lds_row<T, P>(sJ + r * P, row);
          const T fv = sF[r];
          acc.add(row, fv, wc);
          sf += (double)fv;
          sq = fma((double)fv, (double)fv, sq);
          if (has_prev) {
            const T d = sP[r] - fv;
            sd = fma((double)d, (double)d, sd);
          }
        }
      }
      ++ck;
      __syncwarp();                    // every lane is done with this stage before it is refilled
      issue();
    }
    if (cur_c >= 0) {
      sf = warp_sum(sf);
      if (lane == 0) my_sfc[cur_c * NW + warp] = sf;
    }
    acc.warp_reduce();
    sd = warp_sum(sd);
    sq = warp_sum(sq);
    double* myred = red + (size_t)pbuf * NW * V;
    if (lane == 0) {
      double* r = myred + warp * V;
#pragma unroll
      for (int i = 0; i < P; ++i) r[i] = (double)acc.g[i];
#pragma unroll
      for (int i = 0; i < NH; ++i) r[P + i] = (double)acc.h[i];
      r[P + NH] = sd;
      r[P + NH + 1] = sq;
    }
    __syncthreads();                   // one barrier per problem (red/sfc are double-buffered)
    if (tid < V) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < NW; ++k) s += myred[k * V + tid];
      if (tid < P) {
        g[b * P + tid] = (T)s;
      } else if (tid < P + NH) {
        int k = tid - P, i = 0;
        while (k >= P - i) { k -= P - i; ++i; }
        const int jj = i + k;
        H[b * P * P + i * P + jj] = (T)s;
        H[b * P * P + jj * P + i] = (T)s;
      } else {
        stats[b * 4 + (tid - P - NH) + 1] = s;
      }
    } else if (tid == V) {
      double tot = 0.0;
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
        for (int k = 0; k < NW; ++k) s += my_sfc[c * NW + k];
        tot = fma(s, s, tot);
      }
      stats[b * 4 + 0] = tot;
      stats[b * 4 + 3] = 0.0;
    }
  }
}

// ============================================================================== wide
// 8 < P <= PP (PP = 16 or 32): LPR = 4 lanes share a row; lane q owns the H columns
// j = q + 4*jj and accumulates the block-upper part (i < 4*(jj+1)) of each.  g entries are
// split the same way.  Rows are read straight from global memory (the 3 redundant reads per
// row hit L1).  CTA per problem, persistent over problems.
template <typename T, int PP, int TPB>
__global__ void __launch_bounds__(TPB)
k_normal_eq_wide(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fprev,
                 const T* __restrict__ w, int wlen, long long B, int C, int N, int P,
                 T* __restrict__ g, T* __restrict__ H, double* __restrict__ stats, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  constexpr int LPR = 4, NJ = PP / LPR, NW = TPB / 32, RPB = TPB / LPR;   // rows in flight per CTA
  extern __shared__ double smem[];
  double* redH = smem;                         // [NW][LPR][NJ][PP]  per-warp partial columns
  double* redg = redH + NW * LPR * NJ * PP;    // [NW][PP]
  double* reds = redg + NW * PP;               // [NW][2]
  double* sfc = reds + NW * 2;                 // [C][NW]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int q = tid & (LPR - 1), rslot = tid / LPR;

  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    T accH[NJ][PP];
    T accg[NJ];
#pragma unroll
    for (int jj = 0; jj < NJ; ++jj) {
      accg[jj] = T(0);
#pragma unroll
      for (int i = 0; i < PP; ++i) accH[jj][i] = T(0);
    }
    double sd = 0.0, sq = 0.0;
    for (int c = 0; c < C; ++c) {
      const T wc = w[wlen == 1 ? 0 : c];
      const long long seg = (b * C + c) * (long long)N;
      double sf = 0.0;
      for (int n = rslot; n < N; n += RPB) {
        T row[PP];
        const T* src = J + (seg + n) * P;
#pragma unroll
        for (int i = 0; i < PP; ++i) row[i] = (i < P) ? __ldg(src + i) : T(0);
        const T fv = __ldg(f + seg + n);
        if (q == 0) {
          sf += (double)fv;
          sq = fma((double)fv, (double)fv, sq);
          if (fprev) { const T d = __ldg(fprev + seg + n) - fv; sd = fma((double)d, (double)d, sd); }
        }
#pragma unroll
        for (int jj = 0; jj < NJ; ++jj) {
          // column j = q + LPR*jj, selected without dynamic register indexing
          T cj = row[LPR * jj];
#pragma unroll
          for (int k = 1; k < LPR; ++k) cj = (q == k) ? row[LPR * jj + k] : cj;
          const T jw = wc * cj;
          accg[jj] = fma(jw, fv, accg[jj]);
#pragma unroll
          for (int i = 0; i < LPR * (jj + 1); ++i) accH[jj][i] = fma(jw, row[i], accH[jj][i]);
        }
      }
      sf = warp_sum(sf);
      if (lane == 0) sfc[c * NW + warp] = sf;
    }
    // reduce over the 8 row slots of a warp that share q (xor 4, 8, 16)
#pragma unroll
    for (int jj = 0; jj < NJ; ++jj) {
#pragma unroll
      for (int m = LPR; m < 32; m <<= 1) accg[jj] += shfl_xor(accg[jj], m);
#pragma unroll
      for (int i = 0; i < LPR * (jj + 1); ++i) {
#pragma unroll
        for (int m = LPR; m < 32; m <<= 1) accH[jj][i] += shfl_xor(accH[jj][i], m);
      }
    }
    sd = warp_sum(sd);
    sq = warp_sum(sq);
    if (lane < LPR) {
#pragma unroll
      for (int jj = 0; jj < NJ; ++jj) {
        redg[warp * PP + lane + LPR * jj] = (double)accg[jj];
#pragma unroll
        for (int i = 0; i < PP; ++i)
          redH[((warp * LPR + lane) * NJ + jj) * PP + i] = (i < LPR * (jj + 1)) ? (double)accH[jj][i] : 0.0;
      }
    }
    if (lane == 0) { reds[warp * 2] = sd; reds[warp * 2 + 1] = sq; }
    __syncthreads();
    // final: thread t -> entry (i, j) of the P x P block, i <= j taken from column owner
    for (int e = tid; e < P * P; e += TPB) {
      const int i = e / P, j = e - i * P;
      const int lo = min(i, j), hi = max(i, j);
      const int qq = hi & (LPR - 1), jj = hi / LPR;
      double s = 0.0;
      for (int k = 0; k < NW; ++k) s += redH[((k * LPR + qq) * NJ + jj) * PP + lo];
      H[b * P * P + e] = (T)s;
    }
    if (tid < P) {
      double s = 0.0;
      for (int k = 0; k < NW; ++k) s += redg[k * PP + tid];
      g[b * P + tid] = (T)s;
    }
    if (tid == TPB - 1) {
      double s1 = 0.0, s2 = 0.0, tot = 0.0;
      for (int k = 0; k < NW; ++k) { s1 += reds[k * 2]; s2 += reds[k * 2 + 1]; }
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
        for (int k = 0; k < NW; ++k) s += sfc[c * NW + k];
        tot = fma(s, s, tot);
      }
      stats[b * 4 + 0] = tot; stats[b * 4 + 1] = s1; stats[b * 4 + 2] = s2; stats[b * 4 + 3] = 0.0;
    }
    __syncthreads();
  }
}

// ============================================================================== DMMA (8 < P <= 32)
// fp64 tensor-core path for wide rows.  H = sum_n J_n^T J_n is a (P x rows) x (rows x P) product, so
// it maps onto mma.sync.m8n8k4.f64 (SASS DMMA) with K = 4 residual rows per instruction: lane
// (gi = lane >> 2, k = lane & 3) holds J[n0 + k][8 I + gi] for every 8-column block I, and that ONE
// register is both the A fragment of block-row I (A[gi][k] = J^T) and the B fragment of block-column
// I (B[k][gi] = J).  Per 4 rows a warp issues NB coalesced 8-byte loads (each touching 4 x 64
// contiguous bytes), NB (NB + 1) / 2 DMMAs for the upper block triangle and NB DFMAs for g —
// about 3 warp instructions per row at P = 16 instead of the ~80 of the lane-split FMA kernel,
// which was bound by its redundant row loads (profiles/r01d_cfg4.txt).  fp32 inputs are widened on
// load; products and sums are fp64 (at least as accurate as the reference's fp32 bmm).
// GDA = true: rows are first summed over the costs in T (gda:136-137), outputs are the fp64 normal
// equations of the tall system, f_s is written out.
__device__ __forceinline__ void dmma_884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <typename T, int NB, int NW, int U, bool GDA>
__global__ void __launch_bounds__(NW * 32, NB == 2 ? 3 : (NB == 3 ? 2 : 1))
k_normal_eq_dmma(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fprev,
                 const T* __restrict__ w, int wlen, long long B, int C, int N, int P,
                 void* __restrict__ gout, void* __restrict__ Hout, T* __restrict__ fsum_out,
                 double* __restrict__ stats, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  constexpr int NBT = NB * (NB + 1) / 2, RW = NBT * 64 + NB * 8 + 2, TPB = NW * 32;
  extern __shared__ double smem[];
  double* red = smem;                    // [NW][RW]: block triangle of H, g, {sd, sq}
  double* sfc = smem + NW * RW;          // [C][NW] per-cost sums of f (unused for GDA)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int k = lane & 3, gi = lane >> 2;
  const int G = (N + 3) >> 2;            // 4-row groups per (b,c) segment
  bool colok[NB];
#pragma unroll
  for (int I = 0; I < NB; ++I) colok[I] = 8 * I + gi < P;

  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    double Ht[NBT][2], Gt[NB];
#pragma unroll
    for (int i = 0; i < NBT; ++i) Ht[i][0] = Ht[i][1] = 0.0;
#pragma unroll
    for (int I = 0; I < NB; ++I) Gt[I] = 0.0;
    double sd = 0.0, sq = 0.0;
    if constexpr (!GDA) {
      for (int c = 0; c < C; ++c) {
        const long long seg = (b * C + c) * (long long)N;
        const T* Jc = J + seg * P + gi;
        const T* fc = f + seg;
        const T* pc = fprev ? fprev + seg : nullptr;
        double Hc[NBT][2], Gc[NB], sf = 0.0;
#pragma unroll
        for (int i = 0; i < NBT; ++i) Hc[i][0] = Hc[i][1] = 0.0;
#pragma unroll
        for (int I = 0; I < NB; ++I) Gc[I] = 0.0;
        for (int g0 = warp; g0 < G; g0 += NW * U) {
          T v[U][NB], fv[U], pv[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int n = 4 * (g0 + u * NW) + k;
            const bool ok = n < N;
#pragma unroll
            for (int I = 0; I < NB; ++I) v[u][I] = (ok && colok[I]) ? ldg_stream(Jc + (long long)n * P + 8 * I) : T(0);
            fv[u] = ok ? ldg_stream(fc + n) : T(0);
            pv[u] = (ok && pc && gi == 0) ? ldg_stream(pc + n) : fv[u];
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            double vd[NB];
#pragma unroll
            for (int I = 0; I < NB; ++I) vd[I] = (double)v[u][I];
            int idx = 0;
#pragma unroll
            for (int I = 0; I < NB; ++I)
#pragma unroll
              for (int Jb = I; Jb < NB; ++Jb, ++idx) dmma_884(Hc[idx][0], Hc[idx][1], vd[I], vd[Jb]);
            const double fd = (double)fv[u];
#pragma unroll
            for (int I = 0; I < NB; ++I) Gc[I] = fma(vd[I], fd, Gc[I]);
            if (gi == 0) {
              sf += fd;
              sq = fma(fd, fd, sq);
              const double d = (double)(T)(pv[u] - fv[u]);
              sd = fma(d, d, sd);
            }
          }
        }
        const double wc = (double)w[wlen == 1 ? 0 : c];
#pragma unroll
        for (int i = 0; i < NBT; ++i) { Ht[i][0] = fma(wc, Hc[i][0], Ht[i][0]); Ht[i][1] = fma(wc, Hc[i][1], Ht[i][1]); }
#pragma unroll
        for (int I = 0; I < NB; ++I) Gt[I] = fma(wc, Gc[I], Gt[I]);
        sf = warp_sum(sf);
        if (lane == 0) sfc[c * NW + warp] = sf;
      }
    } else {
      const T* fsp = fprev ? fprev + b * (long long)N : nullptr;   // fsum_prev [B,N]
      for (int g0 = warp; g0 < G; g0 += NW * U) {
        T js[U][NB], fs[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          fs[u] = T(0);
#pragma unroll
          for (int I = 0; I < NB; ++I) js[u][I] = T(0);
        }
        for (int c = 0; c < C; ++c) {
          const T wc = w[wlen == 1 ? 0 : c];
          const long long seg = (b * C + c) * (long long)N;
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int n = 4 * (g0 + u * NW) + k;
            if (n < N) {
#pragma unroll
              for (int I = 0; I < NB; ++I)
                if (colok[I]) js[u][I] = fma(ldg_stream(J + (seg + n) * P + 8 * I + gi), wc, js[u][I]);
              fs[u] = fma(ldg_stream(f + seg + n), wc, fs[u]);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int n = 4 * (g0 + u * NW) + k;
          double vd[NB];
#pragma unroll
          for (int I = 0; I < NB; ++I) vd[I] = (double)js[u][I];
          int idx = 0;
#pragma unroll
          for (int I = 0; I < NB; ++I)
#pragma unroll
            for (int Jb = I; Jb < NB; ++Jb, ++idx) dmma_884(Ht[idx][0], Ht[idx][1], vd[I], vd[Jb]);
          const double fd = (double)fs[u];
#pragma unroll
          for (int I = 0; I < NB; ++I) Gt[I] = fma(vd[I], fd, Gt[I]);
          if (gi == 0 && n < N) {
            if (fsum_out) fsum_out[b * N + n] = fs[u];
            sq = fma(fd, fd, sq);
            if (fsp) { const double d = (double)(T)(fsp[n] - fs[u]); sd = fma(d, d, sd); }
          }
        }
      }
    }
    // ---- warp partials -> shared memory
#pragma unroll
    for (int I = 0; I < NB; ++I) { Gt[I] += shfl_xor(Gt[I], 1); Gt[I] += shfl_xor(Gt[I], 2); }
    sd = warp_sum(sd);
    sq = warp_sum(sq);
    double* r = red + warp * RW;
#pragma unroll
    for (int i = 0; i < NBT; ++i) { r[i * 64 + gi * 8 + 2 * k] = Ht[i][0]; r[i * 64 + gi * 8 + 2 * k + 1] = Ht[i][1]; }
    if (k == 0) {
#pragma unroll
      for (int I = 0; I < NB; ++I) r[NBT * 64 + I * 8 + gi] = Gt[I];
    }
    if (lane == 0) { r[NBT * 64 + NB * 8] = sd; r[NBT * 64 + NB * 8 + 1] = sq; }
    __syncthreads();
    for (int e = tid; e < P * P; e += TPB) {
      const int i = e / P, j = e - i * P;
      const int lo = min(i, j), hi = max(i, j);
      const int I = lo >> 3, Jb = hi >> 3;
      const int off = (I * NB - (I * (I - 1)) / 2 + (Jb - I)) * 64 + (lo & 7) * 8 + (hi & 7);
      double s = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) s += red[ww * RW + off];
      if constexpr (GDA) reinterpret_cast<double*>(Hout)[b * P * P + e] = s;
      else reinterpret_cast<T*>(Hout)[b * P * P + e] = (T)s;
    }
    if (tid < P) {
      double s = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) s += red[ww * RW + NBT * 64 + tid];
      if constexpr (GDA) reinterpret_cast<double*>(gout)[b * P + tid] = s;
      else reinterpret_cast<T*>(gout)[b * P + tid] = (T)s;
    }
    if (tid == TPB - 1) {
      double s1 = 0.0, s2 = 0.0, tot = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) { s1 += red[ww * RW + NBT * 64 + NB * 8]; s2 += red[ww * RW + NBT * 64 + NB * 8 + 1]; }
      if constexpr (!GDA) {
        for (int c = 0; c < C; ++c) {
          double s = 0.0;
          for (int ww = 0; ww < NW; ++ww) s += sfc[c * NW + ww];
          tot = fma(s, s, tot);
        }
      }
      stats[b * 4 + 0] = tot; stats[b * 4 + 1] = s1; stats[b * 4 + 2] = s2; stats[b * 4 + 3] = 0.0;
    }
    __syncthreads();
  }
}

// ---- vector-load variant of the DMMA kernel (the one cfg-4 runs): P == 2*NB-even rows whose start
// is NB*sizeof(T)-byte aligned.  The 8-column "blocks" of the fragment layout are INTERLEAVED:
// block I holds the columns {NB*gi + I}, so lane (gi, k) gets its NB fragment values for row
// n0 + k with ONE NB*sizeof(T)-byte load, and the warp's load covers 4 whole rows (4 x P x s
// contiguous bytes).  A warp owns batches of 32 consecutive rows: 8 such loads in flight per lane,
// f / f_prev read once per batch with one coalesced load (lane L <-> row L; the statistics are
// then one row per lane instead of predicated on 4 lanes) and handed to the fragment layout by
// shuffles.  r01d profile of the scalar variant: 74 instructions per 4-row step, 54 % of stalls on
// the load scoreboard (profiles/r01d_cfg4.txt).
template <typename T, int NB> struct VecRow;
template <> struct VecRow<double, 2> {
  double v[2];
  __device__ __forceinline__ void load(const double* p) { const double2 t = ldg_stream2(p); v[0] = t.x; v[1] = t.y; }
};
template <> struct VecRow<double, 4> {
  double v[4];
  __device__ __forceinline__ void load(const double* p) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
  }
};
template <> struct VecRow<float, 2> {
  float v[2];
  __device__ __forceinline__ void load(const float* p) { const float2 t = ldg_stream2(p); v[0] = t.x; v[1] = t.y; }
};
template <> struct VecRow<float, 4> {
  float v[4];
  __device__ __forceinline__ void load(const float* p) { const float4 t = ldg_stream4(p); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
};

template <typename T, int NB, int NW>
__global__ void __launch_bounds__(NW * 32, NB == 2 ? 2 : 1)
k_normal_eq_dmma_vec(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fprev,
                     const T* __restrict__ w, int wlen, long long B, int C, int N, int P,
                     T* __restrict__ gout, T* __restrict__ Hout, double* __restrict__ stats, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  constexpr int NBT = NB * (NB + 1) / 2, PP = 8 * NB, RW = PP * PP + PP + 2, TPB = NW * 32, U = 8;
  extern __shared__ double smem[];
  double* red = smem;                    // [NW][RW]: H (canonical entries), g, {sd, sq}
  double* sfc = smem + NW * RW;          // [C][NW]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int k = lane & 3, gi = lane >> 2;
  const bool colok = NB * gi < P;        // P is a multiple of NB: a lane's NB columns are all inside or all outside
  const int NBATCH = (N + 31) >> 5;

  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    double Ht[NBT][2], Gt[NB];
#pragma unroll
    for (int i = 0; i < NBT; ++i) Ht[i][0] = Ht[i][1] = 0.0;
#pragma unroll
    for (int I = 0; I < NB; ++I) Gt[I] = 0.0;
    double sd = 0.0, sq = 0.0;
    for (int c = 0; c < C; ++c) {
      const long long seg = (b * C + c) * (long long)N;
      const T* Jl = J + seg * P + (long long)k * P + NB * gi;      // this lane's element of row k of a 4-row step
      const T* fc = f + seg;
      const T* pc = fprev ? fprev + seg : nullptr;
      double Hc[NBT][2], Gc[NB], sf = 0.0;
#pragma unroll
      for (int i = 0; i < NBT; ++i) Hc[i][0] = Hc[i][1] = 0.0;
#pragma unroll
      for (int I = 0; I < NB; ++I) Gc[I] = 0.0;
      for (int bt = warp; bt < NBATCH; bt += NW) {
        const int n0 = bt << 5;
        VecRow<T, NB> v[U];
        T fl, pl;
        if (n0 + 32 <= N) {                                        // full batch: no predicates
          const T* jp = Jl + (long long)n0 * P;
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (colok) v[u].load(jp + (long long)(4 * u) * P);
            else {
#pragma unroll
              for (int I = 0; I < NB; ++I) v[u].v[I] = T(0);
            }
          }
          fl = ldg_stream(fc + n0 + lane);
          pl = pc ? ldg_stream(pc + n0 + lane) : fl;
        } else {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int n = n0 + 4 * u + k;
            if (colok && n < N) v[u].load(Jl + (long long)(n0 + 4 * u) * P);
            else {
#pragma unroll
              for (int I = 0; I < NB; ++I) v[u].v[I] = T(0);
            }
          }
          const bool okr = n0 + lane < N;
          fl = okr ? ldg_stream(fc + n0 + lane) : T(0);
          pl = (okr && pc) ? ldg_stream(pc + n0 + lane) : fl;
        }
        const double fd = (double)fl;
        sf += fd;
        sq = fma(fd, fd, sq);
        const double dd = (double)(T)(pl - fl);
        sd = fma(dd, dd, sd);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          double vd[NB];
#pragma unroll
          for (int I = 0; I < NB; ++I) vd[I] = (double)v[u].v[I];
          int idx = 0;
#pragma unroll
          for (int I = 0; I < NB; ++I)
#pragma unroll
            for (int Jb = I; Jb < NB; ++Jb, ++idx) dmma_884(Hc[idx][0], Hc[idx][1], vd[I], vd[Jb]);
          const double fk = __shfl_sync(0xffffffffu, fd, 4 * u + k);   // f of row n0 + 4u + k
#pragma unroll
          for (int I = 0; I < NB; ++I) Gc[I] = fma(vd[I], fk, Gc[I]);
        }
      }
      const double wc = (double)w[wlen == 1 ? 0 : c];
#pragma unroll
      for (int i = 0; i < NBT; ++i) { Ht[i][0] = fma(wc, Hc[i][0], Ht[i][0]); Ht[i][1] = fma(wc, Hc[i][1], Ht[i][1]); }
#pragma unroll
      for (int I = 0; I < NB; ++I) Gt[I] = fma(wc, Gc[I], Gt[I]);
      sf = warp_sum(sf);
      if (lane == 0) sfc[c * NW + warp] = sf;
    }
    // ---- warp partials -> shared memory; entry (i, j) of block (I, Jb) is H[NB*gi + I][NB*gj + Jb]
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
      // block (lo % NB, hi % NB) was computed iff lo % NB <= hi % NB; otherwise its transpose was
      const int off = ((lo % NB) <= (hi % NB)) ? lo * PP + hi : hi * PP + lo;
      double sacc = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) sacc += red[ww * RW + off];
      Hout[b * P * P + e] = (T)sacc;
    }
    if (tid < P) {
      double sacc = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) sacc += red[ww * RW + PP * PP + tid];
      gout[b * P + tid] = (T)sacc;
    }
    if (tid == TPB - 1) {
      double s1 = 0.0, s2 = 0.0, tot = 0.0;
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) { s1 += red[ww * RW + PP * PP + PP]; s2 += red[ww * RW + PP * PP + PP + 1]; }
      for (int c = 0; c < C; ++c) {
        double sacc = 0.0;
        for (int ww = 0; ww < NW; ++ww) sacc += sfc[c * NW + ww];
        tot = fma(sacc, sacc, tot);
      }
      stats[b * 4 + 0] = tot; stats[b * 4 + 1] = s1; stats[b * 4 + 2] = s2; stats[b * 4 + 3] = 0.0;
    }
    __syncthreads();
  }
}

// ======================================================================= cost statistic
// stat[b] = sum_c (sum_n f[b,c,n])^2  (lma:100).  One warp per (b), lanes stride n.
template <typename T>
__global__ void __launch_bounds__(256)
k_cost_stat(const T* __restrict__ f, long long B, int C, int N, double* __restrict__ out, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long b = warp0; b < B; b += nwarps) {
    double tot = 0.0;
    for (int c = 0; c < C; ++c) {
      const T* fc = f + (b * C + c) * (long long)N;
      double s = 0.0;
#pragma unroll 4
      for (int n = lane; n < N; n += 32) s += (double)ldg_stream(fc + n);
      s = warp_sum(s);
      tot = fma(s, s, tot);
    }
    if (lane == 0) out[b] = tot;
  }
}

// ============================================================================== GD flavour
// Cost-summed rows (gda:136-137) in T, normal equations of the tall system in fp64.
template <typename T, int P, int TPB>
__global__ void __launch_bounds__(TPB)
k_normal_eq_gda(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fsum_prev,
                const T* __restrict__ w, int wlen, long long B, int C, int N,
                double* __restrict__ g64, double* __restrict__ H64, T* __restrict__ fsum_out,
                double* __restrict__ stats, const nlls_ctrl* ctrl, int vec_ok) {
  if (is_done(ctrl)) return;
  constexpr int NW = TPB / 32, NH = Tri<P>::N, V = P + NH + 2;
  __shared__ double red[NW * V];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    RowAcc<double, P> acc;
    acc.clear();
    double sd = 0.0, sq = 0.0;
    for (int n = tid; n < N; n += TPB) {
      T js[P];
#pragma unroll
      for (int i = 0; i < P; ++i) js[i] = T(0);
      T fs = T(0);
      for (int c = 0; c < C; ++c) {
        const T wc = w[wlen == 1 ? 0 : c];
        const long long row = (b * C + c) * (long long)N + n;
        T r[P];
        load_row<P>(J + row * P, r, vec_ok != 0);
#pragma unroll
        for (int i = 0; i < P; ++i) js[i] = fma(r[i], wc, js[i]);
        fs = fma(ldg_stream(f + row), wc, fs);
      }
      double jd[P];
#pragma unroll
      for (int i = 0; i < P; ++i) jd[i] = (double)js[i];
      acc.add(jd, (double)fs, 1.0);
      if (fsum_out) fsum_out[b * N + n] = fs;
      sq = fma((double)fs, (double)fs, sq);
      if (fsum_prev) { const T d = fsum_prev[b * N + n] - fs; sd = fma((double)d, (double)d, sd); }
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
      for (int k = 0; k < NW; ++k) s += red[k * V + tid];
      if (tid < P) {
        g64[b * P + tid] = s;
      } else if (tid < P + NH) {
        int k = tid - P, i = 0;
        while (k >= P - i) { k -= P - i; ++i; }
        const int j = i + k;
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

// GD flavour for 8 < P <= 32: thread-per-row outer products do not fit in registers, so rows
// are cost-summed into shared memory tiles and the P x P products are spread over the CTA.
template <typename T, int TPB, int TILE>
__global__ void __launch_bounds__(TPB)
k_normal_eq_gda_wide(const T* __restrict__ J, const T* __restrict__ f, const T* __restrict__ fsum_prev,
                     const T* __restrict__ w, int wlen, long long B, int C, int N, int P,
                     double* __restrict__ g64, double* __restrict__ H64, T* __restrict__ fsum_out,
                     double* __restrict__ stats, const nlls_ctrl* ctrl) {
  if (is_done(ctrl)) return;
  extern __shared__ double smem[];
  double* tj = smem;                 // [TILE][P+1]  (last column = f_s)
  double* red = tj + TILE * (P + 1); // [TPB] scratch for sd / sq
  const int tid = threadIdx.x;
  const int NE = P * P + P;          // H entries then g entries
  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    double accs[(NLLS_MAX_P * NLLS_MAX_P + NLLS_MAX_P + TPB - 1) / TPB];
#pragma unroll
    for (int k = 0; k < (int)(sizeof(accs) / sizeof(double)); ++k) accs[k] = 0.0;
    double sd = 0.0, sq = 0.0;
    for (int n0 = 0; n0 < N; n0 += TILE) {
      const int rows = min(TILE, N - n0);
      for (int e = tid; e < rows * (P + 1); e += TPB) {
        const int r = e / (P + 1), i = e - r * (P + 1);
        T s = T(0);
        for (int c = 0; c < C; ++c) {
          const T wc = w[wlen == 1 ? 0 : c];
          const long long row = (b * C + c) * (long long)N + n0 + r;
          s = fma(i < P ? __ldg(J + row * P + i) : __ldg(f + row), wc, s);
        }
        tj[r * (P + 1) + i] = (double)s;
        if (i == P) {
          if (fsum_out) fsum_out[b * N + n0 + r] = s;
          sq = fma((double)s, (double)s, sq);
          if (fsum_prev) { const T d = fsum_prev[b * N + n0 + r] - s; sd = fma((double)d, (double)d, sd); }
        }
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < (int)(sizeof(accs) / sizeof(double)); ++k) {
        const int e = tid + k * TPB;
        if (e < NE) {
          const int i = (e < P * P) ? e / P : e - P * P;
          const int j = (e < P * P) ? e - i * P : P;
          double a = accs[k];
          for (int r = 0; r < rows; ++r) a = fma(tj[r * (P + 1) + i], tj[r * (P + 1) + j], a);
          accs[k] = a;
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < (int)(sizeof(accs) / sizeof(double)); ++k) {
      const int e = tid + k * TPB;
      if (e < P * P) H64[b * P * P + e] = accs[k];
      else if (e < NE) g64[b * P + (e - P * P)] = accs[k];
    }
    red[tid] = sd;
    __syncthreads();
    for (int s = TPB / 2; s > 0; s >>= 1) { if (tid < s) red[tid] += red[tid + s]; __syncthreads(); }
    const double sdt = red[0];
    __syncthreads();
    red[tid] = sq;
    __syncthreads();
    for (int s = TPB / 2; s > 0; s >>= 1) { if (tid < s) red[tid] += red[tid + s]; __syncthreads(); }
    if (tid == 0) { stats[b * 4 + 0] = 0.0; stats[b * 4 + 1] = sdt; stats[b * 4 + 2] = red[0]; stats[b * 4 + 3] = 0.0; }
    __syncthreads();
  }
}

// ============================================================================ dispatch
template <typename K>
static int occupancy_grid(K kernel, int tpb, size_t smem, long long B) {
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

int g_force_direct = 0;          // test/bench hook: NLLS_FORCE_DIRECT=1 disables the staged kernels

template <typename T, int P>
static int launch_normal_eq_p(const T* J, const T* f, const T* fprev, const T* w, int wlen, long long B, int C, int N,
                              T* g, T* H, double* stats, const nlls_ctrl* ctrl, cudaStream_t st) {
  constexpr int NH = Tri<P>::N, V = P + NH + 2;
  const size_t esz = sizeof(T);
  const bool seg16 = ((size_t)N * esz) % 16 == 0;
  const bool bulk_ok = seg16 && aligned_to(J, 16) && aligned_to(f, 16) && (!fprev || aligned_to(fprev, 16)) &&
                       ((size_t)N * P * esz) % 16 == 0 && !g_force_direct;
  if (bulk_ok) {
    constexpr int NW = 8, S = 4;
    constexpr int R = ((P + 2) * sizeof(T) * 128 <= 6144) ? 128 : 64;   // ~<= 6 KB per stage
    using Cfg = StagedCfg<T, P, NW, R, S>;
    const size_t smem = Cfg::smem_bytes(C);
    if (smem <= 220 * 1024) {
      auto kern = k_normal_eq_staged<T, P, NW, R, S>;
      const int grid = occupancy_grid(kern, NW * 32, smem, B);
      kern<<<grid, NW * 32, smem, st>>>(J, f, fprev, w, wlen, B, C, N, g, H, stats, ctrl);
      return check_launch("nlls_normal_eq(staged)");
    }
  }
  constexpr int TPB = 128;
  const size_t smem = (size_t)(TPB / 32) * V * 8 + (size_t)C * (TPB / 32) * 8;
  // widest load a P-scalar row allows; every row start is aligned to it iff the base is
  constexpr size_t RB = (size_t)P * sizeof(T);
  constexpr size_t VW = (RB % 32 == 0) ? 32 : (RB % 16 == 0) ? 16 : (RB % 8 == 0) ? 8 : sizeof(T);
  const bool vec = aligned_to(J, VW);
  auto kern = k_normal_eq_direct<T, P, TPB>;
  const int grid = occupancy_grid(kern, TPB, smem, B);
  kern<<<grid, TPB, smem, st>>>(J, f, fprev, w, wlen, B, C, N, g, H, stats, ctrl, vec ? 1 : 0);
  return check_launch("nlls_normal_eq(direct)");
}

template <typename T>
static int launch_normal_eq(const void* J, const void* f, const void* fprev, const void* w, int wlen, long long B, int C,
                            int N, int P, void* g, void* H, double* stats, const nlls_ctrl* ctrl, cudaStream_t st) {
#define NLLS_CASE(PP)                                                                                              \
  case PP:                                                                                                         \
    return launch_normal_eq_p<T, PP>((const T*)J, (const T*)f, (const T*)fprev, (const T*)w, wlen, B, C, N, (T*)g, \
                                     (T*)H, stats, ctrl, st);
  switch (P) {
    NLLS_CASE(1) NLLS_CASE(2) NLLS_CASE(3) NLLS_CASE(4) NLLS_CASE(5) NLLS_CASE(6) NLLS_CASE(7) NLLS_CASE(8)
    default: break;
  }
#undef NLLS_CASE
  if (!g_force_direct) {
    constexpr int NWD = 8;
    const int NB = (P + 7) / 8;
    const size_t smemd = ((size_t)NWD * (NB * (NB + 1) / 2 * 64 + NB * 8 + 2) + (size_t)C * NWD) * 8;
    auto go = [&](auto kern) {
      const int grid = occupancy_grid(kern, NWD * 32, smemd, B);
      kern<<<grid, NWD * 32, smemd, st>>>((const T*)J, (const T*)f, (const T*)fprev, (const T*)w, wlen, B, C, N, P, g, H,
                                          (T*)nullptr, stats, ctrl);
    };
    // vector-load variant: rows of NB-element groups, every group NB*sizeof(T)-byte aligned
    const int NBv = P <= 16 ? 2 : 4;
    if (P % NBv == 0 && aligned_to(J, NBv * sizeof(T))) {
      auto gov = [&](auto kern, int PPv) {
        const size_t smv = ((size_t)NWD * (PPv * PPv + PPv + 2) + (size_t)C * NWD) * 8;
        const int grid = occupancy_grid(kern, NWD * 32, smv, B);
        kern<<<grid, NWD * 32, smv, st>>>((const T*)J, (const T*)f, (const T*)fprev, (const T*)w, wlen, B, C, N, P, (T*)g, (T*)H,
                                          stats, ctrl);
      };
      if (NBv == 2) gov(k_normal_eq_dmma_vec<T, 2, NWD>, 16);
      else gov(k_normal_eq_dmma_vec<T, 4, NWD>, 32);
      return check_launch("nlls_normal_eq(dmma_vec)");
    }
    if (NB == 2) go(k_normal_eq_dmma<T, 2, NWD, 4, false>);
    else if (NB == 3) go(k_normal_eq_dmma<T, 3, NWD, 2, false>);
    else go(k_normal_eq_dmma<T, 4, NWD, 2, false>);
    return check_launch("nlls_normal_eq(dmma)");
  }
  constexpr int TPB = 128, NW = TPB / 32;
  if (P <= 16) {
    constexpr int PP = 16;
    const size_t smem = ((size_t)NW * 4 * (PP / 4) * PP + NW * PP + NW * 2 + (size_t)C * NW) * 8;
    auto kern = k_normal_eq_wide<T, PP, TPB>;
    const int grid = occupancy_grid(kern, TPB, smem, B);
    kern<<<grid, TPB, smem, st>>>((const T*)J, (const T*)f, (const T*)fprev, (const T*)w, wlen, B, C, N, P, (T*)g, (T*)H,
                                  stats, ctrl);
  } else {
    constexpr int PP = 32;
    const size_t smem = ((size_t)NW * 4 * (PP / 4) * PP + NW * PP + NW * 2 + (size_t)C * NW) * 8;
    auto kern = k_normal_eq_wide<T, PP, TPB>;
    const int grid = occupancy_grid(kern, TPB, smem, B);
    kern<<<grid, TPB, smem, st>>>((const T*)J, (const T*)f, (const T*)fprev, (const T*)w, wlen, B, C, N, P, (T*)g, (T*)H,
                                  stats, ctrl);
  }
  return check_launch("nlls_normal_eq(wide)");
}

template <typename T>
static int launch_normal_eq_gda(const void* J, const void* f, const void* fsp, const void* w, int wlen, long long B, int C,
                                int N, int P, double* g64, double* H64, void* fso, double* stats, const nlls_ctrl* ctrl,
                                cudaStream_t st) {
  constexpr int TPB = 128;
#define NLLS_CASE(PP)                                                                                               \
  case PP: {                                                                                                        \
    constexpr size_t RB = (size_t)PP * sizeof(T);                                                                   \
    constexpr size_t VW = (RB % 32 == 0) ? 32 : (RB % 16 == 0) ? 16 : (RB % 8 == 0) ? 8 : sizeof(T);                \
    const bool vec = aligned_to(J, VW);                                                                             \
    auto kern = k_normal_eq_gda<T, PP, TPB>;                                                                        \
    const int grid = occupancy_grid(kern, TPB, 0, B);                                                               \
    kern<<<grid, TPB, 0, st>>>((const T*)J, (const T*)f, (const T*)fsp, (const T*)w, wlen, B, C, N, g64, H64,       \
                               (T*)fso, stats, ctrl, vec ? 1 : 0);                                                  \
    return check_launch("nlls_normal_eq_gda");                                                                      \
  }
  switch (P) {
    NLLS_CASE(1) NLLS_CASE(2) NLLS_CASE(3) NLLS_CASE(4) NLLS_CASE(5) NLLS_CASE(6) NLLS_CASE(7) NLLS_CASE(8)
    default: break;
  }
#undef NLLS_CASE
  if (!g_force_direct) {
    constexpr int NWD = 8;
    const int NB = (P + 7) / 8;
    const size_t smemd = ((size_t)NWD * (NB * (NB + 1) / 2 * 64 + NB * 8 + 2)) * 8;
    auto go = [&](auto kern) {
      const int grid = occupancy_grid(kern, NWD * 32, smemd, B);
      kern<<<grid, NWD * 32, smemd, st>>>((const T*)J, (const T*)f, (const T*)fsp, (const T*)w, wlen, B, C, N, P, g64, H64,
                                          (T*)fso, stats, ctrl);
    };
    if (NB == 2) go(k_normal_eq_dmma<T, 2, NWD, 4, true>);
    else if (NB == 3) go(k_normal_eq_dmma<T, 3, NWD, 2, true>);
    else go(k_normal_eq_dmma<T, 4, NWD, 2, true>);
    return check_launch("nlls_normal_eq_gda(dmma)");
  }
  constexpr int WT = 256, TILE = 64;
  const size_t smem = ((size_t)TILE * (P + 1) + WT) * 8;
  auto kern = k_normal_eq_gda_wide<T, WT, TILE>;
  const int grid = occupancy_grid(kern, WT, smem, B);
  kern<<<grid, WT, smem, st>>>((const T*)J, (const T*)f, (const T*)fsp, (const T*)w, wlen, B, C, N, P, g64, H64, (T*)fso,
                               stats, ctrl);
  return check_launch("nlls_normal_eq_gda(wide)");
}

static int check_common(int dtype, long long B, int C, int N, int P, int wlen) {
  if (dtype != NLLS_F32 && dtype != NLLS_F64) { set_error("dtype %d is not NLLS_F32/NLLS_F64", dtype); return NLLS_E_DTYPE; }
  if (B < 0 || C < 1 || N < 1 || P < 1 || P > NLLS_MAX_P || (wlen != 1 && wlen != C)) {
    set_error("unsupported shape B=%lld C=%d N=%d P=%d wlen=%d (need 1<=P<=%d, wlen in {1,C})", B, C, N, P, wlen, NLLS_MAX_P);
    return NLLS_E_SHAPE;
  }
  return 0;
}

}  // namespace nlls

using namespace nlls;

extern "C" int nlls_set_force_direct(int on) { const int old = g_force_direct; g_force_direct = on; return old; }

extern "C" int nlls_normal_eq(int dtype, const void* J, const void* f, const void* f_prev, const void* wvec, int wlen,
                              long long B, int C, int N, int P, void* g, void* H, double* stats, const nlls_ctrl* ctrl,
                              void* stream) {
  if (int e = check_common(dtype, B, C, N, P, wlen)) return e;
  if (B == 0) return 0;
  if (!J || !f || !wvec || !g || !H || !stats) { set_error("nlls_normal_eq: NULL tensor"); return NLLS_E_NULL; }
  const size_t esz = dtype == NLLS_F64 ? 8 : 4;
  if (!aligned_to(J, esz) || !aligned_to(f, esz) || !aligned_to(g, esz) || !aligned_to(H, esz) || !aligned_to(stats, 8)) {
    set_error("nlls_normal_eq: misaligned tensor"); return NLLS_E_ALIGN;
  }
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == NLLS_F64 ? launch_normal_eq<double>(J, f, f_prev, wvec, wlen, B, C, N, P, g, H, stats, ctrl, st)
                           : launch_normal_eq<float>(J, f, f_prev, wvec, wlen, B, C, N, P, g, H, stats, ctrl, st);
}

extern "C" int nlls_normal_eq_gda(int dtype, const void* J, const void* f, const void* fsum_prev, const void* wvec, int wlen,
                                  long long B, int C, int N, int P, double* g64, double* H64, void* fsum_out, double* stats,
                                  const nlls_ctrl* ctrl, void* stream) {
  if (int e = check_common(dtype, B, C, N, P, wlen)) return e;
  if (B == 0) return 0;
  if (!J || !f || !wvec || !g64 || !H64 || !stats) { set_error("nlls_normal_eq_gda: NULL tensor"); return NLLS_E_NULL; }
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == NLLS_F64
             ? launch_normal_eq_gda<double>(J, f, fsum_prev, wvec, wlen, B, C, N, P, g64, H64, fsum_out, stats, ctrl, st)
             : launch_normal_eq_gda<float>(J, f, fsum_prev, wvec, wlen, B, C, N, P, g64, H64, fsum_out, stats, ctrl, st);
}

extern "C" int nlls_cost_stat(int dtype, const void* f, long long B, int C, int N, double* stat_out, const nlls_ctrl* ctrl,
                              void* stream) {
  if (int e = check_common(dtype, B, C, N, 1, 1)) return e;
  if (B == 0) return 0;
  if (!f || !stat_out) { set_error("nlls_cost_stat: NULL tensor"); return NLLS_E_NULL; }
  cudaStream_t st = (cudaStream_t)stream;
  const int tpb = 256;
  long long blocks = (B + 7) / 8;
  if (blocks > 148LL * 8 * 4) blocks = 148LL * 8 * 4;
  if (dtype == NLLS_F64) k_cost_stat<double><<<(int)blocks, tpb, 0, st>>>((const double*)f, B, C, N, stat_out, ctrl);
  else k_cost_stat<float><<<(int)blocks, tpb, 0, st>>>((const float*)f, B, C, N, stat_out, ctrl);
  return check_launch("nlls_cost_stat");
}
