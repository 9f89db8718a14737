/* nlls_b200.h — C ABI of the B200 (sm_100a) batched nonlinear-least-squares core.
 *
 * Drop-in boundary for ONE path of hahnec/torchimize (v0.0.16): the per-iteration step of
 * lsq_lma_parallel / lsq_gna_parallel / gradient_descent_parallel.  The reference has no
 * FFI of its own (it is pure Python over torch ops); each entry point below replaces the
 * torch-op sequence at the cited reference lines and is what a binding for this path binds
 * (INTEGRATION.md shows the ctypes stub).  Paths are relative to the reference root.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer on the current CUDA device unless marked "host";
 *   - tensors are dense row-major: p [B,P], f [B,C,N], J [B,C,N,P] (P fastest),
 *     g [B,P], H [B,P,P], wvec [wlen] with wlen == C or 1 (1 broadcasts, lma:76);
 *   - dtype: NLLS_F32 / NLLS_F64 selects the element type of every `void*` tensor;
 *     `double*` tensors are fp64 regardless (solves are always fp64, lma:98 / gna:140 / gda:139);
 *   - nothing allocates, nothing synchronises: kernels are enqueued on `stream`
 *     (a cudaStream_t passed as void*; NULL = legacy default stream);
 *   - return value: 0 on success, a cudaError_t (>0) from the launch, or a negative
 *     NLLS_E_* code for rejected arguments; nlls_last_error() gives a message;
 *   - `ctrl` (nullable) is the device-resident iteration control block below.  When
 *     non-NULL every kernel returns immediately once ctrl->done is set, which lets a host
 *     enqueue max_iter iterations ahead without reading anything back.
 */
#ifndef NLLS_B200_H
#define NLLS_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define NLLS_B200_ABI_VERSION 2   /* 2: the fused kernels keep the model value m [B,N] between iterations, not f [B,C,N];
                                     per-problem tracking / retirement (nlls_track) */

enum { NLLS_F32 = 0, NLLS_F64 = 1 };

enum {                       /* negative return codes */
  NLLS_E_DTYPE = -1,         /* dtype not NLLS_F32 / NLLS_F64 */
  NLLS_E_SHAPE = -2,         /* B,C,N,P out of the supported range (1<=P<=32, wlen in {1,C}) */
  NLLS_E_NULL = -3,          /* a required pointer is NULL */
  NLLS_E_ALIGN = -4,         /* pointer not aligned to its element size */
  NLLS_E_MODEL = -5,         /* unknown model id / parameter count mismatch */
  NLLS_E_ARCH = -6           /* not running on an sm_100 device */
};

enum { NLLS_CTRL_E_PEER_TIMEOUT = 1 /* nlls_iter_end_peer: a peer's stamp did not arrive within the timeout */ };

enum { NLLS_SOLVER_GN = 0, NLLS_SOLVER_LM = 1, NLLS_SOLVER_GD = 2 };
enum { NLLS_LM_LEVENBERG = 0, NLLS_LM_MARQUARDT = 1 };
enum { NLLS_MODEL_EMG = 0, NLLS_MODEL_SKEWGAUSS = 1,
       NLLS_MODEL_MULTI_EMG = 2 /* sum of 2..4 EMG peaks (P = 8|12|16): nlls_model_eval only */,
       NLLS_MODEL_USER = 100 /* the functor a plugin library was compiled for (torchimize_b200.models.compile_model) */ };

/* Device-resident iteration control block (48 bytes, 8-byte aligned, zero-initialise).
 * red[] are batch-global reductions accumulated by the update kernels of ONE iteration and
 * consumed + cleared by nlls_iter_end; they are plain fp64 sums so that a multi-GPU "exact"
 * run can all-reduce(SUM) the 4 doubles between the update kernel and nlls_iter_end.
 *   red[0] problems not converged (LM: assuming the fcon gate is open)
 *   red[1] LM: problems with rho > 0 (gate of fcon, lma:110)        GD: unused
 *   red[2] LM: problems not converged assuming the gate is closed    GD: sum_b sum_n (f_prev-f)^2
 *   red[3]                                                           GD: sum_b sum_n f^2          */
typedef struct nlls_ctrl {
  int done;          /* 1 once every problem of the batch satisfied a stop test (lma:118, gna:88, gda:84) */
  int iters;         /* iterations completed == snapshots written to p_hist */
  int singular;      /* sticky: some solve met a non-positive / non-finite pivot (gna:74-75 warning) */
  int error;         /* sticky: NLLS_CTRL_E_* set by a kernel (0 = none); done is set with it */
  double red[4];
} nlls_ctrl;

/* Per-problem tracking and retirement (north_star kernel 3).  SURVEY 3.7 Q1: the reference's masked copy
 * `p[~converged] = pn[~converged]` (lma:115, gna:85, gda:81) was meant to freeze converged problems and is a no-op, so
 * every problem keeps stepping until ALL have converged in the same iteration.  The default (track == NULL, or
 * retire == 0) reproduces that.  Device pointers, host struct:
 *   pstat [B,4] int32, initialise every row to {-1,0,0,0}:
 *     [0] first iteration at which the problem's OWN stop test fired, -1 = not yet.  Own test = gcon | pcon | fcon with
 *         per-problem gates: LM's fcon counts only in an iteration whose step this problem accepted (the reference's
 *         batch-global gate lma:110 made per problem), GD's fcon uses the problem's own two sums instead of gda:76's
 *         batch totals
 *     [1] the reasons at that iteration: 1 = gcon, 2 = pcon, 4 = fcon
 *     [2] LM: accepted steps so far; GN / GD: steps taken
 *     [3] flags: 1 = a non-finite step was produced
 *   retire == 0: pstat is recorded, nothing else changes.
 *   retire != 0: a problem with pstat[0] >= 0 is frozen at the iterate it had when its test fired (== the snapshot of
 *     that iteration of a reference-semantics run: per-problem arithmetic does not depend on the batch).  K2 / K3 skip it
 *     (its snapshot repeats), the fused model kernels never evaluate it again: iteration k (= ctrl->iters) walks
 *     active[k & 1][0 .. count[k & 1]) and its update kernel appends the problems that stay active to
 *     active[(k+1) & 1] (order arbitrary, results independent of it); nlls_iter_end clears the consumed count and sets
 *     ctrl->done when no problem is left — no host round-trip anywhere.  Initialise active[0] = 0..B-1, count = {B, 0}. */
typedef struct nlls_track {
  int* pstat;        /* [B,4] */
  int* active;       /* [2,B] */
  int* count;        /* [2] */
  int retire;
  int reserved;
  long long B;
} nlls_track;

/* ---- housekeeping ---------------------------------------------------------------- */
int nlls_abi_version(void);
const char* nlls_last_error(void);                 /* host string, thread-local */
int nlls_device_check(void);                       /* 0 if the current device is sm_100; else NLLS_E_ARCH */
int nlls_set_force_direct(int on);                 /* test/bench hook: 1 = K1 never takes the bulk-copy staged
                                                      kernel; returns the previous setting */

/* ---- K1: fused weighted normal equations (tensor path) ------------------------------
 * Replaces newton_parallel.py:40-43 (four einsums; J read twice) and the f re-reads of
 * lma:100,110 / gna:80:  ONE pass over J, f (and f_prev) producing
 *   g[b,p]   = sum_c w_c sum_n J[b,c,n,p] f[b,c,n]
 *   H[b,p,q] = sum_c w_c sum_n J[b,c,n,p] J[b,c,n,q]          (full symmetric P x P)
 *   stats[b] = { sum_c (sum_n f)^2  (lma:100, unweighted, Q3),
 *                sum_{c,n} (f_prev - f)^2  (0 when f_prev == NULL),
 *                sum_{c,n} f^2,  0 }                                         fp64 [B,4]
 * g, H are written in `dtype`.  Algorithmic bytes: s*C*N*(P+1) (+ s*C*N with f_prev).  */
int nlls_normal_eq(int dtype, const void* J, const void* f, const void* f_prev /*nullable*/,
                   const void* wvec, int wlen, long long B, int C, int N, int P,
                   void* g, void* H, double* stats, const nlls_ctrl* ctrl, void* stream);

/* Cost-summed variant for gradient_descent_parallel (gda:134-139): never materialises
 *   f_s[b,n] = sum_c w_c f[b,c,n],  j_s[b,n,:] = sum_c w_c J[b,c,n,:]   (summed in `dtype`)
 * but forms the fp64 normal equations of the tall N x P least-squares problem in-stream:
 *   H64 = j_s^T j_s,  g64 = j_s^T f_s                                    fp64 [B,P,P], [B,P]
 * writes f_s to fsum_out [B,N] (dtype) and stats[b] = {0, sum_n (fsum_prev-f_s)^2, sum_n f_s^2, 0}. */
int nlls_normal_eq_gda(int dtype, const void* J, const void* f, const void* fsum_prev /*nullable*/,
                       const void* wvec, int wlen, long long B, int C, int N, int P,
                       double* g64, double* H64, void* fsum_out, double* stats,
                       const nlls_ctrl* ctrl, void* stream);

/* stat_out[b] = sum_c (sum_n f[b,c,n])^2 — the trial-cost half of the gain ratio (lma:100). */
int nlls_cost_stat(int dtype, const void* f, long long B, int C, int N, double* stat_out,
                   const nlls_ctrl* ctrl, void* stream);

/* ---- K2: damped solve (lma:96-98) -------------------------------------------------
 * marquardt != 0: dscale[b,:] *= max_p max(H[b,p,p], dscale[b,0])   (lma:88, compounding, Q6)
 * dscale is [B,2]: {diagonal scale, off-diagonal state}; initialise to {1, 0}.  The second entry
 * reproduces the reference's dense D, whose zeros turn NaN once the scale overflows (0*inf).
 * Hu = H + u[b]*D ;  h = -solve_fp64(Hu, g) cast to dtype ;  p_trial = p + h.
 * Cholesky in registers, LU with partial pivoting when a pivot is not positive.  */
int nlls_lm_solve(int dtype, const void* g, const void* H, const void* p, const void* u,
                  void* dscale, int marquardt, long long B, int P,
                  void* h, void* p_trial, nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* ---- K3: update / gain ratio / damping adaptation / stop tests ----------------------
 * LM (lma:100-116, 124-152): rho from stats[b][0] - stat_h[b] and h^T(u h - g); Nielsen or
 * Marquardt update of u, v in place; p += h where rho > 0 (in place, Q1); gcon/pcon/fcon;
 * snapshot p into p_hist[ctrl->iters] ([max_iter,B,P], nullable); accumulates ctrl->red. */
int nlls_lm_update(int dtype, void* p, const void* h, const void* g, void* u, void* v,
                   const double* stats, const double* stat_h, int meth,
                   double rho1, double rho2, double beta, double gama,
                   double ftol, double ptol, double gtol,
                   long long B, int P, void* p_hist, void* rho_out /*nullable*/,
                   nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* GN (gna:139-141, 78-86): h = solve_fp64(H, g); p -= l*h; stop tests; snapshot. */
int nlls_gn_update(int dtype, void* p, const void* g, const void* H, const double* stats,
                   double l, double ftol, double ptol, double gtol, int with_masks,
                   long long B, int P, void* h_out /*nullable*/, void* p_hist /*nullable*/,
                   nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* GD (gda:139-146, 71-82): h = solve_fp64(H64, g64) (== lstsq(j_s, f_s)); p -= l*h; g = h/l;
 * gcon/pcon per problem; the batch-global fcon sums (Q5) are reduced deterministically
 * into ctrl->red[2..3].  */
int nlls_gd_update(int dtype, void* p, const double* g64, const double* H64, const double* stats,
                   double l, double ftol, double ptol, double gtol, int with_masks,
                   long long B, int P, void* h_out /*nullable*/, void* p_hist /*nullable*/,
                   nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* Closes an iteration: decides ctrl->done from ctrl->red (per `solver`), ++ctrl->iters,
 * clears red.  have to run after the update kernel (and after the all-reduce of red in a
 * multi-GPU exact run).  */
int nlls_iter_end(nlls_ctrl* ctrl, int solver, double ftol, const nlls_track* track /*nullable*/, void* stream);

/* nlls_iter_end for a batch sharded over the GPUs of one box in "exact" mode (lma:110,118, gna:88, gda:76,84 see the
 * WHOLE batch): sums ctrl->red over the ranks through peer-mapped mailboxes (NVLink P2P stores + stamp, no collective
 * library call) and then decides like nlls_iter_end.  mailboxes: device array of `world` pointers, entry r = rank r's
 * mailbox of 4*world*8 doubles (zero-initialised symmetric memory, mapped into this process; slot = 2*(epoch&1) +
 * (iteration&1)); epoch: a per-solve counter, equal on every rank.  Every rank has to launch it once per iteration.
 * A peer that stays silent for timeout_s seconds (<= 0: 60 s) sets ctrl->error = NLLS_CTRL_E_PEER_TIMEOUT and
 * ctrl->done = 1 on the waiting rank (no trap: its context and earlier results survive). */
int nlls_iter_end_peer(nlls_ctrl* ctrl, int solver, double ftol, void* const* mailboxes, int rank, int world,
                       long long epoch, double timeout_s, void* stream);

/* ---- single-problem solvers (torchimize/functions/single/*.py: lma_fun_single.py:81,99; gna:81; gda:78) ----------
 * Residual statistics of ONE problem over a vector of any length, spread over the whole GPU: block k writes
 * out[2k] = partial sum_i a[i]^2 and out[2k+1] = partial sum_i (b[i]-a[i])^2 (0 when b is NULL), fp64;
 * out holds 2 * nlls_sumsq_partials() doubles, the caller adds the partials (fixed order: deterministic). */
int nlls_sumsq_partials(void);
int nlls_sumsq(int dtype, const void* a, const void* b /*nullable*/, long long n, double* out, void* stream);

/* ---- K4: fused built-in models (tests/emg.py:19-49, tests/unit_test_parallel.py:81-118;
 *          tests/unit_test_skewed_gaussian.py:43-51) ------------------------------------
 * Data layout: x [N] shared abscissa, y [B,*,N] with strides (y_bstride, y_cstride) in
 * elements; y_cstride == 0 shares one data row between all C costs (the reference tests).
 * cost f = (y - m(p; x))^2, J = -2 (y - m) dm/dp.
 *
 * nlls_model_eval: materialises f [B,C,N] and (nullable) J [B,C,N,P] — the CUDA-backed
 * `function` / `jac_function` callables of the tensor path. */
int nlls_model_eval(int model, int dtype, const void* p, int P, const void* x, const void* y,
                    long long y_bstride, long long y_cstride, long long B, int C, int N,
                    void* f_out, void* J_out /*nullable*/, const nlls_ctrl* ctrl, void* stream);

/* nlls_model_normal_eq: K4+K1 fused — evaluates the model once per (b,n), loops the costs,
 * never materialises J.  Same outputs as nlls_normal_eq.  m_pp[2] are two [B,N] ping-pong
 * buffers of the MODEL VALUE m(p; x): iteration k reads m_prev from m_pp[(k+1)&1] (when k>0)
 * and writes m to m_pp[k&1] with k = ctrl->iters (k = 0 when ctrl == NULL).  The previous
 * residuals that fcon needs (lma:110, gna:80) are rebuilt in-kernel as
 * f_prev[b,c,n] = (y[b,c,n] - m_prev[b,n])^2 — the same two roundings that produced them, so
 * stats[1] is bit-identical to the sum over stored residuals at s*N*(C+2) bytes per
 * problem-iteration instead of 3*s*C*N.  m_pp0 may be NULL to skip that traffic (then stats[1] = 0). */
int nlls_model_normal_eq(int model, int dtype, const void* p, int P, const void* x, const void* y,
                         long long y_bstride, long long y_cstride,
                         const void* wvec, int wlen, long long B, int C, int N,
                         void* m_pp0, void* m_pp1, void* g, void* H, double* stats,
                         const nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* nlls_model_normal_eq_static: nlls_model_normal_eq for callers that cannot allocate device memory (code under CUDA graph
 * capture, serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The two [B,N] model-value ping-pong buffers (m_pp0 / m_pp1 above — the kernel's scratch) live in a
 * file-scope static __device__ array of the library (csrc/static_scratch.cu): 2 x NLLS_STATIC_SCRATCH_ELEMS x 8 bytes = 16 MiB
 * on every GPU that loads the library, held until the process exits.  The caller passes inputs and outputs only; everything else
 * (arguments, outputs, iteration parity from ctrl->iters, error codes) is nlls_model_normal_eq's, and the results are bit-identical.
 *   limit   B * N <= NLLS_STATIC_SCRATCH_ELEMS for either dtype; beyond it the call returns NLLS_E_SHAPE with a message.
 *   allocs  none: no cudaMalloc / cudaFree / allocating library routine.  The first call on each device looks the array's device
 *           address up (cudaGetSymbolAddress, with that device current) — make that first call outside stream capture.
 *   sharing every call on a device uses the SAME array, and the array keeps the last call's contents: iteration k > 0 reads the
 *           model values that the previous call (iteration k-1 of the same batch) left there; iteration 0 reads nothing, so
 *           nothing has to be cleared.  Calls that overlap on the GPU (two streams, two host threads, two solves interleaved
 *           iteration by iteration) RACE on it: one solve at a time per device. */
#define NLLS_STATIC_SCRATCH_ELEMS 1048576LL
int nlls_model_normal_eq_static(int model, int dtype, const void* p, int P, const void* x, const void* y,
                                long long y_bstride, long long y_cstride,
                                const void* wvec, int wlen, long long B, int C, int N,
                                void* g, void* H, double* stats,
                                const nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* Speculative Levenberg-Marquardt trial evaluation (fused path): ONE pass at the trial point p + h that yields
 * both halves of what lma:99-100 and the NEXT iteration's newton_parallel.py:38-43 need:
 *   stats_t[b] = { sum_c (sum_n f_t)^2  (trial cost of the gain ratio, lma:100),
 *                  sum (f_cur - f_t)^2, sum f_t^2  (the NEXT iteration's fcon sums, lma:110), 0 }
 *   g_t, H_t   = normal equations at the trial point (what the next iteration uses if the step is accepted;
 *                a rejected step keeps the current g, H — p did not move — so nothing is recomputed).
 * sel [B] (bytes): the model values m_cur [B,N] at problem b's CURRENT point live in (sel[b] ? m_buf1 : m_buf0)
 * (f_cur is rebuilt from them and y); the trial values m_t are written to the other buffer.  nlls_lm_update_spec
 * flips sel[b] on acceptance.  Same arithmetic, in the same order, as nlls_model_normal_eq at the accepted point:
 * trajectories are bit-identical to the two-pass schedule (nlls_model_normal_eq + nlls_model_cost_stat). */
int nlls_model_lm_trial(int model, int dtype, const void* p_trial, int P, const void* x, const void* y,
                        long long y_bstride, long long y_cstride, const void* wvec, int wlen,
                        long long B, int C, int N, void* m_buf0, void* m_buf1, const unsigned char* sel,
                        void* g_t, void* H_t, double* stats_t, const nlls_ctrl* ctrl,
                        const nlls_track* track /*nullable*/, void* stream);

/* K3 for the speculative schedule (lma:100-116, 124-152): as nlls_lm_update, with
 *   rho numerator  = stats[b][0] - stats_t[b][0];
 *   stop tests on the CURRENT g and stats (pre-step point, lma:108-110) — evaluated before anything is replaced;
 *   on acceptance (rho > 0): p += h, g <- g_t, H <- H_t, stats[b] <- stats_t[b], sel[b] ^= 1;
 *   on rejection: stats[b][1] <- 0 (f did not change: the next fcon sees sum (f_prev - f)^2 == 0). */
int nlls_lm_update_spec(int dtype, void* p, const void* h, void* g, void* H, const void* g_t, const void* H_t,
                        void* u, void* v, double* stats, const double* stats_t, unsigned char* sel, int meth,
                        double rho1, double rho2, double beta, double gama,
                        double ftol, double ptol, double gtol,
                        long long B, int P, void* p_hist, void* rho_out /*nullable*/,
                        nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* GD flavour of the fused kernel: outputs as nlls_normal_eq_gda; fsum_pp are [B,N] ping-pong. */
int nlls_model_normal_eq_gda(int model, int dtype, const void* p, int P, const void* x, const void* y,
                             long long y_bstride, long long y_cstride,
                             const void* wvec, int wlen, long long B, int C, int N,
                             void* fsum_pp0, void* fsum_pp1, double* g64, double* H64, double* stats,
                             const nlls_ctrl* ctrl, const nlls_track* track /*nullable*/, void* stream);

/* stat_out[b] = sum_c (sum_n f(p_trial)[b,c,n])^2 without materialising f (lma:99-100). */
int nlls_model_cost_stat(int model, int dtype, const void* p_trial, int P, const void* x, const void* y,
                         long long y_bstride, long long y_cstride, long long B, int C, int N,
                         double* stat_out, const nlls_ctrl* ctrl, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NLLS_B200_H */
