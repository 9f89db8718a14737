"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's single-problem solvers
(torchimize/functions/single/{lma,gna,gda}_fun_single.py), built from the same torch primitives so that the dtype
quirks of the reference carry over (e.g. its float32 identity D, lma_fun_single.py:73-74, rounds the damping term
u*D to float32 even for float64 problems).  Pinned against the imported reference by tests/golden/single.npz
(tests/golden/make_golden_single.py); checked in tests/test_oracle_golden.py.

Differences to the batched solvers that matter for parity: p is [P], f [N], J [N,P]; the
returned list STARTS with the initial p; u = tau * max diag(H); the gain ratio is guarded against a non-positive
denominator; the normal equations are only re-evaluated after an accepted step."""
import torch


def _bind(function, jac_function, args):
    """lma_fun_single.py:57-67 (same in gna:47-57, gda:47-57); autograd fallback = functions/jacobian.py:23-37."""
    fun = (lambda q: function(q, *args)) if len(args) > 0 else function
    if jac_function is None:
        def jac(q):
            try:
                return torch.autograd.functional.jacobian(fun, q, vectorize=True)
            except RuntimeError:
                return torch.autograd.functional.jacobian(fun, q, strict=True, vectorize=False)
    else:
        jac = lambda q: jac_function(q, *args)
    return fun, jac


def _stops(g, h, p, f, f_prev, ftol, ptol, gtol, f_gate=True):
    """gcon / pcon / fcon of lma_fun_single.py:97-100 (gna:79-81, gda:76-78)."""
    gcon = max(abs(g)) < gtol
    pcon = (h ** 2).sum() ** .5 < ptol * (ptol + (p ** 2).sum() ** .5)
    fcon = ((f_prev - f) ** 2).sum() < ((ftol * f) ** 2).sum() if f_gate else False
    return bool(gcon) or bool(pcon) or bool(fcon)


def lsq_lma(p, function, jac_function=None, args=(), ftol=1e-8, ptol=1e-8, gtol=1e-8, tau=1e-3, meth='lev',
            rho1=.25, rho2=.75, bet=2, gam=3, max_iter=100):
    """lma_fun_single.py:23-104."""
    fun, jac = _bind(function, jac_function, args)
    f, j = fun(p), jac(p)
    g, H = j.T @ f, j.T @ j                                     # :69-71
    u, v = tau * torch.max(torch.diag(H)), 2                    # :72-73
    out = [p]
    while len(out) < max_iter:
        D = torch.eye(j.shape[1], device=j.device)              # default dtype (float32), :76
        D *= 1 if meth == 'lev' else torch.max(torch.maximum(H.diagonal(), D.diagonal()))   # :77
        h = -torch.linalg.lstsq(H + u * D, g, rcond=None, driver=None)[0]                  # :78
        f_h = fun(p + h)
        den = h @ (u * h - g)                                   # :80
        num = f @ f - f_h @ f_h                                 # :81
        rho = num / den if den > 0 else torch.inf if num > 0 else -torch.inf               # :82
        if rho > 0:                                             # :83-87
            p = p + h
            j = jac(p)
            g, H = j.T @ fun(p), j.T @ j
        out.append(p.clone())
        f_prev, f = f.clone(), fun(p)                           # :89-90
        if meth == 'lev':                                       # :91-94
            u, v = (u * torch.max(torch.tensor([1 / 3, 1 - (2 * rho - 1) ** 3])), 2) if rho > 0 else (u * v, v * 2)
        else:
            u = u * bet if rho < rho1 else u / gam if rho > rho2 else u
        if _stops(g, h, p, f, f_prev, ftol, ptol, gtol, f_gate=bool(rho > 0)):
            break
    return out


def lsq_gna(p, function, jac_function=None, args=(), ftol=1e-8, ptol=1e-8, gtol=1e-8, l=1., max_iter=100):
    """gna_fun_single.py:23-87."""
    fun, jac = _bind(function, jac_function, args)
    f, j = fun(p), jac(p)
    g, H = j.T @ f, j.T @ j
    out = [p]
    while len(out) < max_iter:
        h = -l * torch.linalg.lstsq(H, g, rcond=None, driver=None)[0]                       # :64
        p = p + h
        out.append(p.clone())
        f_prev, f = f.clone(), fun(p)
        j = jac(p)
        g, H = j.T @ f, j.T @ j
        if _stops(g, h, p, f, f_prev, ftol, ptol, gtol):
            break
    return out


def gradient_descent(p, function, jac_function=None, args=(), ftol=1e-8, ptol=1e-8, gtol=1e-8, l=1., max_iter=100):
    """gda_fun_single.py:23-84: the step is the least-squares solution of the tall system j h = f."""
    fun, jac = _bind(function, jac_function, args)
    f, j = fun(p), jac(p)
    out = [p]
    while len(out) < max_iter:
        h = -l * torch.linalg.lstsq(j, f, rcond=None, driver=None)[0]                       # :62
        p = p + h
        out.append(p.clone())
        f_prev, f = f.clone(), fun(p)
        j = jac(p)
        g = j.T @ f
        if _stops(g, h, p, f, f_prev, ftol, ptol, gtol):
            break
    return out
