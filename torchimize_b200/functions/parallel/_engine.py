"""Host driver shared by the three batched solvers.

It owns the work buffers and the device-resident control block and enqueues, per iteration, the
C-ABI kernels of include/nlls_b200.h.  Two ways to produce the normal equations:

  tensor path  opaque Python callables materialise f [B,C,N] and J [B,C,N,P]; K1
               (nlls_normal_eq / nlls_normal_eq_gda) streams them once.  One 12-byte flag read
               per iteration tells the host when the batch is done (the reference syncs ~20x).
  fused path   both callables belong to a built-in FusedModel: nlls_model_normal_eq* evaluates
               the model in-kernel.  Iterations are enqueued ahead in groups; every kernel
               returns at once when ctrl->done is set, so the host never waits inside a group.

Semantics mirror the reference exactly: in-place mutation of the caller's p,
no freezing of converged problems, batch-global termination, per-iteration snapshots.
"""
import warnings

import torch

from ... import _lib
from ...models import bound_model_of, fused_model_of, model_owner_of
from ..jacobian import batch_jacobian_approx_t

GN, LM, GD = _lib.SOLVER_GN, _lib.SOLVER_LM, _lib.SOLVER_GD
_HIST_CHUNK_BYTES = 256 << 20
_GROUP = 4                      # fused path: iterations enqueued between flag reads
SPECULATIVE_LM = True           # fused LM: one model pass per iteration (nlls_model_lm_trial); False = two-pass schedule


_default_reduce_hook = None


def set_default_reduce_hook(hook):
    """Installs the multi-GPU 'exact' reduction hook used by solve(); returns the previous one."""
    global _default_reduce_hook
    old, _default_reduce_hook = _default_reduce_hook, hook
    return old


def bind_callables(function, jac_function, args):
    """reference: lma_fun_parallel.py:62-72 (identical in gna:53-63, gda:51-61)."""
    fun = (lambda q: function(q, *args)) if len(args) > 0 else function
    if jac_function is None:
        jac = lambda q: batch_jacobian_approx_t(q, f=fun)
    else:
        jac = lambda q: jac_function(q, *args)
    return fun, jac


def _as_work(t, dtype, what, shape=None):
    if not torch.is_tensor(t):
        raise _lib.NllsError(f'{what} must return a tensor, got {type(t).__name__}')
    _lib.require_cuda(t)
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise _lib.NllsError(f'{what} returned shape {tuple(t.shape)}, expected {tuple(shape)}')
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.detach().contiguous()


class Workspace:
    """Per-solve device buffers (allocated once; nothing is allocated inside the loop)."""

    def __init__(self, p, wvec, solver):
        _lib.require_cuda(p)
        self.dtype, self.device = p.dtype, p.device
        self.code = _lib.dtype_code(p)
        self.B, self.P = p.shape
        if not 1 <= self.P <= 32:
            raise _lib.NllsError(f'P={self.P} unsupported (1 <= P <= 32)')
        B, P, kw = self.B, self.P, dict(dtype=p.dtype, device=p.device)
        k64 = dict(dtype=torch.float64, device=p.device)
        self.solver = solver
        self.w = (torch.ones(1, **kw) if wvec is None else wvec.to(**kw)).contiguous()
        self.ctrl = torch.zeros(_lib.CTRL_BYTES // 8, **k64)
        self.ctrl_i32 = self.ctrl.view(torch.int32)
        self.stats = torch.zeros(B, 4, **k64)
        if solver == GD:
            self.g = torch.empty(B, P, **k64)
            self.H = torch.empty(B, P, P, **k64)
        else:
            self.g = torch.empty(B, P, **kw)
            self.H = torch.empty(B, P, P, **kw)
        if solver == LM:
            self.h = torch.empty(B, P, **kw)
            self.p_trial = torch.empty(B, P, **kw)
            self.stat_h = torch.empty(B, **k64)
            self.u = torch.empty(B, **kw)
            self.v = torch.empty(B, **kw)
            self.dscale = torch.empty(B, 2, **kw)
        self._pending, self._last = [], (0, 0, 0)
        self.hist = []                 # (first_iteration, tensor[chunk,B,P])
        self.hist_chunk = max(1, _HIST_CHUNK_BYTES // max(1, B * P * p.element_size()))

    # snapshots ------------------------------------------------------------------------------
    def hist_base_ptr(self, k, max_iter):
        """Device address such that snapshot k lands at base + k*B*P*s; grows the history in
        chunks so max_iter=10^4 does not allocate 10^4 snapshots up front."""
        if not self.hist or k >= self.hist[-1][0] + self.hist[-1][1].shape[0]:
            n = min(self.hist_chunk, max_iter - k)
            self.hist.append((k, torch.empty(n, self.B, self.P, dtype=self.dtype, device=self.device)))
        k0, buf = self.hist[-1]
        return buf.data_ptr() - k0 * self.B * self.P * buf.element_size()

    def hist_room(self, k):
        k0, buf = self.hist[-1]
        return k0 + buf.shape[0] - k

    def snapshots(self, iters):
        out = []
        for k0, buf in self.hist:
            n = min(buf.shape[0], iters - k0)
            if n <= 0:
                break
            part = buf[:n]
            if n * 2 < buf.shape[0]:
                part = part.clone()
            out.extend(part.unbind(0))
        return out

    def flags(self):
        """(done, iters, singular) — one small device->host read (synchronises)."""
        d, it, sing, err = self.ctrl_i32[:4].tolist()
        self.raise_on(err)
        return d, it, sing

    @staticmethod
    def raise_on(err):
        if err == _lib.CTRL_E_PEER_TIMEOUT:
            raise _lib.NllsError("mode='exact-peer': a peer rank did not post its sums within the timeout (NLLS_PEER_TIMEOUT_S); "
                                 'this rank stopped at the iteration it was waiting in — its earlier results are intact')
        if err:
            raise _lib.NllsError(f'device-side error flag {err} in the control block')

    # Fused path: the flags travel to pinned host memory behind each group of iterations and are looked at ONE group
    # late, so the device always has a full group queued while the host waits (no drain bubble between groups).  The
    # price is at most one extra group of launches whose kernels return at once (ctrl->done).
    def post_flags(self):
        ring = _pinned_ring(self.device)
        slot = self._slot = (getattr(self, '_slot', -1) + 1) % ring.shape[0]
        ring[slot].copy_(self.ctrl_i32[:4], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pending.append((slot, ev))

    def poll_flags(self, lag):
        ring = _pinned_ring(self.device)
        while len(self._pending) > lag:
            slot, ev = self._pending.pop(0)
            ev.synchronize()
            d, it, sing, err = ring[slot].tolist()
            self.raise_on(err)
            self._last = (d, it, sing)
        return self._last


_PINNED = {}


def _pinned_ring(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _PINNED:
        _PINNED[key] = torch.zeros(16, 4, dtype=torch.int32).pin_memory()
    return _PINNED[key]


def _check_p(p):
    assert len(p.shape) == 2, 'parameter tensor is supposed to have 2 dims, but has %s' % str(len(p.shape))
    _lib.require_cuda(p)
    _lib.dtype_code(p)


STOP_G, STOP_P, STOP_F = 1, 2, 4        # bits of stats['stop_reason']


@_lib.on_device_of(1)
def solve(solver, p, function, jac_function, args, wvec, ftol, ptol, gtol, max_iter, l=1.,
          lm=None, fused=True, check_every=_GROUP, reduce_hook=None, mode='exact', stats=None):
    """Runs one batched solve; returns the list of per-iteration snapshots (coeffs_list).
    mode  'exact' (default): the reference's semantics — nothing freezes, the batch stops when every problem passes a stop
          test in the same iteration.  'retire' (not in the reference; north_star kernel 3): a problem whose
          OWN stop test fired is frozen at that iterate and never evaluated again (no host round-trip: the kernels walk a
          device-resident list of active problems); the solve ends when no problem is left.  Its iterate equals the
          snapshot of that iteration of an 'exact' run bit for bit; later snapshots repeat it.
    stats optional dict, filled at return: iterations, stop_iteration [B] (first iteration at which the problem's own test
          fired, -1 = never), stop_reason [B] (bits STOP_G | STOP_P | STOP_F at that iteration), steps [B] (LM: accepted
          steps; GN / GD: steps taken), nonfinite [B] (a non-finite step was produced), mode.
    reduce_hook(red) — optional, multi-GPU 'exact' mode: called between the update kernel and
    nlls_iter_end with the 4 fp64 batch-global sums of the control block (a device tensor view)
    so that a caller can all-reduce(SUM) them over the ranks (torchimize_b200/sharding.py)."""
    _check_p(p)
    if p.shape[0] == 0 or max_iter <= 0:
        return []
    if mode not in ('exact', 'retire'):
        raise ValueError("mode must be 'exact' or 'retire'")
    if reduce_hook is None:
        reduce_hook = _default_reduce_hook
    if mode == 'retire' and reduce_hook is not None:
        raise _lib.NllsError("mode='retire' stops every problem on its own test: it has no batch-global decision to exchange "
                             "between ranks (use sharding mode='local')")
    if getattr(reduce_hook, 'peer', None) is not None:
        reduce_hook.peer.epoch += 1                          # one epoch per solve, in step on every rank
    work_p = p if p.is_contiguous() else p.contiguous()
    ws = Workspace(work_p, wvec, solver)
    # 'lm' = speculative schedule (nlls_model_lm_trial), 'lm2' = two-pass schedule (needs nlls_model_cost_stat)
    skey = {GN: 'gn', LM: 'lm' if SPECULATIVE_LM else 'lm2', GD: 'gd'}[solver]
    model = fused_model_of(function, jac_function, skey) if fused else None
    if model is None and fused and len(args) == 0:          # FusedModel.bind(x, y): the data travels with the callables
        bound = bound_model_of(function, jac_function, skey)
        if bound is not None:
            model, args = bound
    st = _lib.stream_ptr
    B, P, code = ws.B, ws.P, ws.code
    ctrl = _lib.ptr(ws.ctrl)
    state = lm.get('state') if solver == LM else None
    trk = None
    if mode == 'retire' or stats is not None:               # per-problem tracking (nlls_track): device arrays, host struct
        ws.pstat = torch.zeros(B, 4, dtype=torch.int32, device=ws.device)
        ws.pstat[:, 0] = -1
        trk = _lib.Track(pstat=_lib.ptr(ws.pstat), active=None, count=None, retire=0, reserved=0, B=B)
        if mode == 'retire':
            ws.active = torch.empty(2, B, dtype=torch.int32, device=ws.device)
            ws.active[0] = torch.arange(B, dtype=torch.int32, device=ws.device)
            ws.act_count = torch.tensor([B, 0], dtype=torch.int32, device=ws.device)
            trk.active, trk.count, trk.retire = _lib.ptr(ws.active), _lib.ptr(ws.act_count), 1
    if solver == LM:
        meth = _lib.LM_LEVENBERG if lm['meth'] == 'lev' else _lib.LM_MARQUARDT
        if state:                                # warm restart: the damping state of a previous call
            for name in ('u', 'v', 'dscale'):
                src, dst = state.get(name), getattr(ws, name)
                if not isinstance(src, torch.Tensor) or src.shape != dst.shape or src.device != dst.device:
                    raise _lib.NllsError(f"state['{name}'] must be a {tuple(dst.shape)} tensor on {dst.device}")
                dst.copy_(src)
            if state.get('meth', lm['meth']) != lm['meth']:
                raise _lib.NllsError("state was produced with another damping method ('meth')")
        else:
            ws.u.fill_(lm['tau'])                # u = tau * max diag(I)   (lma:77-78, Q8)
            ws.v.fill_(2.)
            ws.dscale[:, 0] = 1.
            ws.dscale[:, 1] = 0.

    def update(k_hist_ptr):
        if solver == GN:
            _lib.call('nlls_gn_update', code, _lib.ptr(work_p), _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(ws.stats),
                      float(l), ftol, ptol, gtol, 1, B, P, None, k_hist_ptr, ctrl, st(), track=trk)
        elif solver == GD:
            _lib.call('nlls_gd_update', code, _lib.ptr(work_p), _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(ws.stats),
                      float(l), ftol, ptol, gtol, 1, B, P, None, k_hist_ptr, ctrl, st(), track=trk)
        else:
            _lib.call('nlls_lm_update', code, _lib.ptr(work_p), _lib.ptr(ws.h), _lib.ptr(ws.g), _lib.ptr(ws.u),
                      _lib.ptr(ws.v), _lib.ptr(ws.stats), _lib.ptr(ws.stat_h), meth, float(lm['rho1']), float(lm['rho2']),
                      float(lm['beta']), float(lm['gama']), ftol, ptol, gtol, B, P, k_hist_ptr, None, ctrl, st(), track=trk)
        iter_end()

    def iter_end():
        """closes the iteration; multi-GPU exact mode: NCCL all-reduce of the sums, or the peer-memory kernel"""
        peer = getattr(reduce_hook, 'peer', None)
        if peer is not None:
            _lib.call('nlls_iter_end_peer', ctrl, solver, ftol, _lib.ptr(peer.ptrs), peer.rank, peer.world, peer.epoch,
                      float(getattr(peer, 'timeout_s', 60.)), st())
            return
        if reduce_hook is not None:
            reduce_hook(ws.ctrl[2:6])
        _lib.call('nlls_iter_end', ctrl, solver, ftol, st(), track=trk)

    def lm_solve():
        _lib.call('nlls_lm_solve', code, _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(work_p), _lib.ptr(ws.u),
                  _lib.ptr(ws.dscale), int(meth == _lib.LM_MARQUARDT), B, P, _lib.ptr(ws.h), _lib.ptr(ws.p_trial), ctrl, st(), track=trk)

    if model is not None:
        # ------------------------------------------------------------------ fused path
        x, y, ybs, ycs, C, N = model.layout(work_p, *args)
        if ws.w.numel() not in (1, C):
            raise _lib.NllsError(f'wvec has {ws.w.numel()} entries for {C} costs')
        mid = model.model_id
        # [B,N] ping-pong: the model value m (GN, LM: f_prev is rebuilt from it in-kernel) or GD's cost-summed f_s
        fpp = torch.empty(2, B, N, dtype=ws.dtype, device=ws.device)
        launched, done = 0, False
        spec = solver == LM and SPECULATIVE_LM
        if spec:
            # Speculative schedule: ONE model pass per iteration at the trial point yields the trial cost (lma:99-100)
            # AND the next iteration's normal equations; the update kernel picks per problem (nlls_lm_update_spec).
            g_t, H_t, stats_t = torch.empty_like(ws.g), torch.empty_like(ws.H), torch.empty_like(ws.stats)
            sel = torch.zeros(B, dtype=torch.uint8, device=ws.device)
            _lib.call('nlls_model_normal_eq', mid, code, _lib.ptr(work_p), P, _lib.ptr(x), _lib.ptr(y), ybs, ycs,
                      _lib.ptr(ws.w), ws.w.numel(), B, C, N, _lib.ptr(fpp[0]), _lib.ptr(fpp[1]), _lib.ptr(ws.g),
                      _lib.ptr(ws.H), _lib.ptr(ws.stats), ctrl, st(), lib=model.lib, track=trk)             # iteration 0: m(p0) -> fpp[0], sel == 0
        while spec and launched < max_iter and not done:
            hp = ws.hist_base_ptr(launched, max_iter)
            group = min(check_every, max_iter - launched, ws.hist_room(launched))
            for _ in range(group):
                lm_solve()
                _lib.call('nlls_model_lm_trial', mid, code, _lib.ptr(ws.p_trial), P, _lib.ptr(x), _lib.ptr(y), ybs, ycs,
                          _lib.ptr(ws.w), ws.w.numel(), B, C, N, _lib.ptr(fpp[0]), _lib.ptr(fpp[1]), _lib.ptr(sel),
                          _lib.ptr(g_t), _lib.ptr(H_t), _lib.ptr(stats_t), ctrl, st(), lib=model.lib, track=trk)
                _lib.call('nlls_lm_update_spec', code, _lib.ptr(work_p), _lib.ptr(ws.h), _lib.ptr(ws.g), _lib.ptr(ws.H),
                          _lib.ptr(g_t), _lib.ptr(H_t), _lib.ptr(ws.u), _lib.ptr(ws.v), _lib.ptr(ws.stats), _lib.ptr(stats_t),
                          _lib.ptr(sel), meth, float(lm['rho1']), float(lm['rho2']), float(lm['beta']), float(lm['gama']),
                          ftol, ptol, gtol, B, P, hp, None, ctrl, st(), track=trk)
                iter_end()
            launched += group
            ws.post_flags()
            done, iters, sing = ws.poll_flags(lag=1)
        while not spec and launched < max_iter and not done:
            hp = ws.hist_base_ptr(launched, max_iter)
            group = min(check_every, max_iter - launched, ws.hist_room(launched))
            for _ in range(group):
                if solver == GD:
                    _lib.call('nlls_model_normal_eq_gda', mid, code, _lib.ptr(work_p), P, _lib.ptr(x), _lib.ptr(y), ybs, ycs,
                              _lib.ptr(ws.w), ws.w.numel(), B, C, N, _lib.ptr(fpp[0]), _lib.ptr(fpp[1]), _lib.ptr(ws.g),
                              _lib.ptr(ws.H), _lib.ptr(ws.stats), ctrl, st(), lib=model.lib, track=trk)
                else:
                    _lib.call('nlls_model_normal_eq', mid, code, _lib.ptr(work_p), P, _lib.ptr(x), _lib.ptr(y), ybs, ycs,
                              _lib.ptr(ws.w), ws.w.numel(), B, C, N, _lib.ptr(fpp[0]), _lib.ptr(fpp[1]), _lib.ptr(ws.g),
                              _lib.ptr(ws.H), _lib.ptr(ws.stats), ctrl, st(), lib=model.lib, track=trk)
                if solver == LM:
                    lm_solve()
                    _lib.call('nlls_model_cost_stat', mid, code, _lib.ptr(ws.p_trial), P, _lib.ptr(x), _lib.ptr(y), ybs, ycs,
                              B, C, N, _lib.ptr(ws.stat_h), ctrl, st(), lib=model.lib)
                update(hp)
            launched += group
            ws.post_flags()
            done, iters, sing = ws.poll_flags(lag=1)
        done, iters, sing = ws.poll_flags(lag=0) if ws._pending else ws.flags()
    else:
        # ----------------------------------------------------------------- tensor path
        fun, jac = bind_callables(function, jac_function, args)
        owner = model_owner_of(function, jac_function)      # built-in / compiled model on the tensor path: one eval launch
        f_prev = None
        iters, sing = 0, 0
        for k in range(max_iter):
            if owner is not None:
                f, J = owner.cost_and_jac(work_p, *args)
            else:
                f = _as_work(fun(work_p), ws.dtype, 'function')
                if f.dim() != 3 or f.shape[0] != B:
                    raise _lib.NllsError(f'function must return [B,C,N] residuals, got {tuple(f.shape)}')
                J = None
            _, C, N = f.shape
            if J is None:
                J = _as_work(jac(work_p), ws.dtype, 'jac_function', (B, C, N, P))
            if ws.w.numel() not in (1, C):
                raise _lib.NllsError(f'wvec has {ws.w.numel()} entries for {C} costs')
            hp = ws.hist_base_ptr(k, max_iter)
            if solver == GD:
                fs = torch.empty(B, N, dtype=ws.dtype, device=ws.device)
                _lib.call('nlls_normal_eq_gda', code, _lib.ptr(J), _lib.ptr(f), _lib.ptr(f_prev), _lib.ptr(ws.w),
                          ws.w.numel(), B, C, N, P, _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(fs), _lib.ptr(ws.stats), ctrl, st())
                keep = fs
            else:
                _lib.call('nlls_normal_eq', code, _lib.ptr(J), _lib.ptr(f), _lib.ptr(f_prev), _lib.ptr(ws.w), ws.w.numel(),
                          B, C, N, P, _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(ws.stats), ctrl, st())
                keep = f
            if solver == LM:
                lm_solve()
                f_h = _as_work(fun(ws.p_trial), ws.dtype, 'function', (B, C, N))       # lma:99
                _lib.call('nlls_cost_stat', code, _lib.ptr(f_h), B, C, N, _lib.ptr(ws.stat_h), ctrl, st())
            update(hp)
            # the reference clones (lma:111, gna:81).  Our own model kernels hand out a fresh tensor per call, so the pointer
            # is enough there; an opaque user callable may write into a buffer it reuses (out=...), which would make f_prev
            # alias the next f and fire fcon at once: clone
            f_prev = keep if (owner is not None or solver == GD) else keep.clone()
            done, iters, sing = ws.flags()
            if done:
                break
    if sing and solver == GN:
        # reference: gna_fun_parallel.py:74-75 (there: cond(H) is inf, via a batched SVD every iteration)
        warnings.warn('Condition number of at least one Hessian is infinite giving NaNs. Consider a step rate l<1.')
    if work_p is not p:
        p.copy_(work_p)                                     # Q1: the caller's tensor ends as the final iterate
    if state is not None:                                   # caller-owned dict, updated in place like p
        state.update(u=ws.u.clone(), v=ws.v.clone(), dscale=ws.dscale.clone(), meth=lm['meth'], iterations=int(iters))
    if stats is not None:                                   # caller-owned dict
        stats.update(iterations=int(iters), mode=mode, stop_iteration=ws.pstat[:, 0].clone(), stop_reason=ws.pstat[:, 1].clone(),
                     steps=ws.pstat[:, 2].clone(), nonfinite=(ws.pstat[:, 3] & 1).bool())
    return ws.snapshots(iters)


@_lib.on_device_of(1)
def plain(solver, p, function, jac_function, wvec, l, max_iter):
    """Fixed-iteration variants (gna:94-117, gda:90-113): no stop tests, returns the final p.  Callables of a
    FusedModel.bind(x, y) run the fused kernels (no Jacobian tensor, nothing kept between iterations)."""
    _check_p(p)
    work_p = p if p.is_contiguous() else p.contiguous()
    ws = Workspace(work_p, wvec, solver)
    B, P, code = ws.B, ws.P, ws.code
    st = _lib.stream_ptr
    bound = bound_model_of(function, jac_function, 'gd' if solver == GD else 'gn')
    if bound is not None:
        model, margs = bound
        x, y, ybs, ycs, C, N = model.layout(work_p, *margs)
        if ws.w.numel() not in (1, C):
            raise _lib.NllsError(f'wvec has {ws.w.numel()} entries for {C} costs')
    for _ in range(max_iter):
        if bound is not None:
            name = 'nlls_model_normal_eq_gda' if solver == GD else 'nlls_model_normal_eq'
            _lib.call(name, model.model_id, code, _lib.ptr(work_p), P, _lib.ptr(x), _lib.ptr(y), ybs, ycs, _lib.ptr(ws.w),
                      ws.w.numel(), B, C, N, None, None, _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(ws.stats), None, st(), lib=model.lib)
        else:
            f = _as_work(function(work_p), ws.dtype, 'function')
            _, C, N = f.shape
            J = _as_work(jac_function(work_p), ws.dtype, 'jac_function', (B, C, N, P))
            if solver == GD:
                _lib.call('nlls_normal_eq_gda', code, _lib.ptr(J), _lib.ptr(f), None, _lib.ptr(ws.w), ws.w.numel(), B, C, N, P,
                          _lib.ptr(ws.g), _lib.ptr(ws.H), None, _lib.ptr(ws.stats), None, st())
            else:
                _lib.call('nlls_normal_eq', code, _lib.ptr(J), _lib.ptr(f), None, _lib.ptr(ws.w), ws.w.numel(), B, C, N, P,
                          _lib.ptr(ws.g), _lib.ptr(ws.H), _lib.ptr(ws.stats), None, st())
        _lib.call('nlls_gd_update' if solver == GD else 'nlls_gn_update', code, _lib.ptr(work_p), _lib.ptr(ws.g), _lib.ptr(ws.H),
                  None, float(l), 0., 0., 0., 0, B, P, None, None, None, st())
    if work_p is not p:
        p.copy_(work_p)
    return p
