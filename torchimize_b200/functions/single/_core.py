"""Single-problem solvers on the B200 kernels (torchimize/functions/single/*.py).

One problem: p [P], function(p) -> f [N], jac_function(p) -> J [N,P].  Everything of size N stays on the device and
is reduced by the batched kernels: the N rows are cut into chunks that K1 (nlls_normal_eq / nlls_normal_eq_gda) treats
as a pseudo-batch, so g = J^T f and H = J^T J of ONE long problem run at the same HBM rate as a batch of short ones;
residual sums go through nlls_sumsq and the P x P solve through K2 (nlls_lm_solve, fp64 Cholesky).  What is left on the
host is the reference's own scalar control flow on P-sized CPU tensors (gain ratio, damping update, stop tests), written
with the reference's expressions so that its dtype behaviour carries over (e.g. the float32 identity D of
lma_fun_single.py:76-77 rounds the damping term to float32 even for float64 problems).
"""
import torch

from ... import _lib
from ..jacobian import jacobian_approx_t

_CHUNK_MIN, _CHUNKS_MAX = 256, 4096


def bind(function, jac_function, args):
    """lma_fun_single.py:57-67 (same in gna:47-57, gda:47-57)."""
    fun = (lambda q: function(q, *args)) if len(args) > 0 else function
    if jac_function is None:
        jac = lambda q: jacobian_approx_t(q, f=fun)
    else:
        jac = lambda q: jac_function(q, *args)
    return fun, jac


def _work(t, dtype, what, shape=None):
    if not torch.is_tensor(t):
        raise _lib.NllsError(f'{what} must return a tensor, got {type(t).__name__}')
    _lib.require_cuda(t)
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise _lib.NllsError(f'{what} returned shape {tuple(t.shape)}, expected {tuple(shape)}')
    w = t.detach().to(dtype).contiguous()
    # a private copy (the reference clones f, lma_fun_single.py:89): a callable that reuses its output buffer must not
    # make the kept residuals alias the next evaluation
    return w.clone() if w.data_ptr() == t.data_ptr() else w


class Single:
    """Device buffers and kernel calls of one single-problem solve."""

    def __init__(self, p):
        assert p.dim() == 1, 'parameter tensor is supposed to have 1 dim, but has %s' % str(p.dim())
        _lib.require_cuda(p)
        self.code, self.dtype, self.dev, self.P = _lib.dtype_code(p), p.dtype, p.device, p.shape[0]
        if not 1 <= self.P <= 32:
            raise _lib.NllsError(f'P={self.P} unsupported (1 <= P <= 32)')
        self.nsum = _lib.load().nlls_sumsq_partials()
        self.sums = torch.empty(self.nsum, 2, dtype=torch.float64, device=self.dev)
        self.w = torch.ones(1, dtype=self.dtype, device=self.dev)
        self.one_zero = torch.tensor([[1., 0.]], dtype=self.dtype, device=self.dev)

    def chunks(self, N):
        """rows per pseudo-problem: a power of two >= 256 that keeps the pseudo-batch <= 4096 (full chunks + one tail)."""
        R = _CHUNK_MIN
        while N // R > _CHUNKS_MAX:
            R *= 2
        return R

    def normal_eq(self, f, J, tall=False):
        """g = J^T f [P], H = J^T J [P,P] (tall: fp64 normal equations of the N x P least-squares problem, gda:62)."""
        N, P = J.shape
        R = self.chunks(N)
        nb, tail = N // R, N % R
        odt = torch.float64 if tall else self.dtype
        parts = nb + (1 if tail else 0)
        g = torch.empty(parts, P, dtype=odt, device=self.dev)
        H = torch.empty(parts, P, P, dtype=odt, device=self.dev)
        st = torch.empty(parts, 4, dtype=torch.float64, device=self.dev)
        es = f.element_size()
        for k0, B, n, off in ((0, nb, R, 0), (nb, 1 if tail else 0, tail, nb * R)):
            if B == 0:
                continue
            Jp, fp = J.data_ptr() + off * P * es, f.data_ptr() + off * es
            if tall:
                _lib.call('nlls_normal_eq_gda', self.code, Jp, fp, None, _lib.ptr(self.w), 1, B, 1, n, P, _lib.ptr(g[k0:]),
                          _lib.ptr(H[k0:]), None, _lib.ptr(st[k0:]), None, _lib.stream_ptr())
            else:
                _lib.call('nlls_normal_eq', self.code, Jp, fp, None, _lib.ptr(self.w), 1, B, 1, n, P, _lib.ptr(g[k0:]),
                          _lib.ptr(H[k0:]), _lib.ptr(st[k0:]), None, _lib.stream_ptr())
        return g.sum(0), H.sum(0)

    def sumsq(self, a, b=None):
        """(sum a^2, sum (b-a)^2) as Python floats (fp64 sums)."""
        _lib.call('nlls_sumsq', self.code, _lib.ptr(a), _lib.ptr(b), a.numel(), _lib.ptr(self.sums), _lib.stream_ptr())
        s = self.sums.sum(0).tolist()
        return s[0], s[1]

    def solve(self, g, H, p, damp):
        """h = -solve(H + damp I, g) in fp64 (K2), p + h; g/H may be fp64 (tall system) while p keeps its dtype."""
        code = _lib.F64 if g.dtype == torch.float64 else _lib.F32
        pw = p if p.dtype == g.dtype else p.to(g.dtype)
        u = torch.full((1,), float(damp), dtype=g.dtype, device=self.dev)
        ds = self.one_zero if g.dtype == self.dtype else self.one_zero.to(g.dtype)
        h = torch.empty(1, self.P, dtype=g.dtype, device=self.dev)
        pt = torch.empty(1, self.P, dtype=g.dtype, device=self.dev)
        _lib.call('nlls_lm_solve', code, _lib.ptr(g), _lib.ptr(H), _lib.ptr(pw), _lib.ptr(u), _lib.ptr(ds), 0, 1, self.P,
                  _lib.ptr(h), _lib.ptr(pt), None, _lib.stream_ptr())
        return h[0]


def stop(gc, hc, pc, ff, sd, ftol, ptol, gtol, f_gate=True):
    """gcon / pcon / fcon (lma_fun_single.py:97-100, gna:79-81, gda:76-78) on CPU tensors; ff = sum f^2, sd = sum (f_prev-f)^2."""
    gcon = max(abs(gc)) < gtol
    pcon = (hc ** 2).sum() ** .5 < ptol * (ptol + (pc ** 2).sum() ** .5)
    fcon = sd < (ftol ** 2) * ff if f_gate else False
    return bool(gcon) or bool(pcon) or bool(fcon)


@_lib.on_device_of(0)
def lsq_lma(p, function, jac_function, args, ftol, ptol, gtol, tau, meth, rho1, rho2, bet, gam, max_iter):
    S = Single(p)
    fun, jac = bind(function, jac_function, args)
    dt, P = S.dtype, S.P
    f = _work(fun(p), dt, 'function')
    N = f.shape[0]
    j = _work(jac(p), dt, 'jac_function', (N, P))
    g, H = S.normal_eq(f, j)
    gc, Hc = g.cpu(), H.cpu()
    ff, _ = S.sumsq(f)
    u, v = tau * torch.max(torch.diag(Hc)), 2                    # lma_fun_single.py:72-73
    out = [p]
    pd = p.detach()
    marquardt = meth != 'lev'
    while len(out) < max_iter:
        # Damping added to the diagonal: u times the scale of the reference's D (lma_fun_single.py:76-78).  D starts as a
        # float32 identity there, so scale and product are rounded to float32 whatever the problem's dtype — kept, since the
        # trajectories depend on it.  Levenberg: scale 1.  Marquardt: the largest of max diag(H) and 1.
        scale = Hc.diagonal().max().clamp_min(1.).to(torch.float32) if marquardt else torch.ones((), dtype=torch.float32)
        damp = u.to(torch.float32) * scale
        h = S.solve(g[None], H[None], pd[None], damp)
        hc = h.cpu()
        p_try = pd + h
        f_h = _work(fun(p_try), dt, 'function', (N,))
        fh2, _ = S.sumsq(f_h)
        # gain ratio (lma:80-82): cost decrease over the decrease the damped quadratic model predicts; a model that predicts
        # no decrease gives +-inf by the sign of the actual change
        predicted = hc @ (u * hc - gc)
        actual = torch.tensor(ff - fh2, dtype=dt)                # sums in fp64 here
        if predicted > 0:
            rho = actual / predicted
        else:
            rho = torch.tensor(float('inf') if actual > 0 else -float('inf'))
        accepted = bool(rho > 0)
        f_new = f
        if accepted:                                             # lma:83-87
            pd = p_try
            f_new = f_h                                          # fun(p) at the accepted point is the trial evaluation
            j = _work(jac(pd), dt, 'jac_function', (N, P))
            g, H = S.normal_eq(f_new, j)
            gc, Hc = g.cpu(), H.cpu()
        out.append(pd.clone())
        ff_new, sd = S.sumsq(f_new, f)                           # lma:89-90, 99
        f, ff = f_new, ff_new
        if not marquardt:                                        # Nielsen's rule (lma:91-92), factor in rho's dtype as there
            if accepted:
                shrink = (1 - (2 * rho - 1) ** 3).clamp_min(1 / 3)
                u, v = u * shrink, 2
            else:
                u, v = u * v, v * 2
        elif rho < rho1:                                         # Marquardt's rule (lma:93-94)
            u = u * bet
        elif rho > rho2:
            u = u / gam
        if stop(gc, hc, pd.cpu(), ff, sd, ftol, ptol, gtol, f_gate=bool(rho > 0)):
            break
    return out


@_lib.on_device_of(0)
def _newton_like(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter, tall):
    """lsq_gna (gna_fun_single.py:59-84: h = -l solve(H, g)) and gradient_descent (gda_fun_single.py:58-81: h = -l lstsq(j, f),
    here through the fp64 normal equations of the tall system)."""
    S = Single(p)
    fun, jac = bind(function, jac_function, args)
    dt, P = S.dtype, S.P
    f = _work(fun(p), dt, 'function')
    N = f.shape[0]
    j = _work(jac(p), dt, 'jac_function', (N, P))
    g, H = S.normal_eq(f, j, tall)
    ff, _ = S.sumsq(f)
    out = [p]
    pd = p.detach()
    while len(out) < max_iter:
        h = (l * S.solve(g[None], H[None], pd[None], 0.)).to(dt)
        pd = pd + h
        out.append(pd.clone())
        f_new = _work(fun(pd), dt, 'function', (N,))
        j = _work(jac(pd), dt, 'jac_function', (N, P))
        g, H = S.normal_eq(f_new, j, tall)
        ff, sd = S.sumsq(f_new, f)
        f = f_new
        gc = g.cpu().to(dt)                                      # gda:74: g = j^T f in the problem's dtype
        if stop(gc, h.cpu(), pd.cpu(), ff, sd, ftol, ptol, gtol):
            break
    return out


def lsq_gna(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter):
    return _newton_like(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter, tall=False)


def gradient_descent(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter):
    return _newton_like(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter, tall=True)
