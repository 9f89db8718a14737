"""TEST INFRASTRUCTURE ONLY — CPU restatement (torch CPU ops, any float dtype) of the
reference's batched nonlinear-least-squares path.  Not product code: see oracle/__init__.py.

Parity status: PINNED.  ``tests/golden/make_golden.py`` ran the unmodified reference
(hahnec/torchimize v0.0.16, imported from /root/reference in the build container,
torch 2.11.0 CPU) on the reference's own test fixtures and on seeded synthetic batches and
committed inputs + per-iteration outputs under ``tests/golden/*.npz``;
``tests/test_oracle_golden.py`` checks every function below against those vectors.
The reference's tests hold no golden vectors of their own.

The arithmetic of this path lives in PyTorch (requirements.txt:1 pins ``torch>=1.11.0``,
no lock): einsum, linalg.lstsq, erf/exp.  The restatement therefore uses the same torch
CPU primitives, so on the same machine it reproduces the reference to the last bit, while
being laid out as explicit steps that expose the intermediates (g, H, h, rho, u, masks)
the CUDA kernels are tested against.

Shapes: p [B,P]; f [B,C,N]; J [B,C,N,P]; wvec [C] or [1].
Every quirk id (Q1..Q13) refers to SURVEY.md §3.7.
"""
from dataclasses import dataclass, field
from typing import Callable, List, Optional

import torch


# --------------------------------------------------------------------------- pieces
def normal_equations(f, J, wvec):
    """g = sum_c w_c sum_n J f ; H = sum_c w_c sum_n J^T J.
    reference: torchimize/functions/parallel/newton_parallel.py:40-43"""
    gc = torch.einsum('bcnp,bcnp->bcp', J, f[..., None])
    Hc = torch.einsum('bcnp,bcni->bcpi', J, J)
    g = torch.einsum('bcp,c->bp', gc, wvec)
    H = torch.einsum('bcpi,c->bpi', Hc, wvec)
    return g, H


def cost_sum_stat(f):
    """sum_c (sum_n f)^2  — the quantity the reference's gain ratio uses (Q3).
    reference: lma_fun_parallel.py:100 (einsum 'bcp,bci->bc' then .sum(1))"""
    return torch.einsum('bcp,bci->bc', f, f).sum(1)


def solve_fp64(A, b, out_dtype):
    """P x P (or tall N x P) least-squares solve, always in fp64 (Q7).
    reference: lma_fun_parallel.py:98, gna_fun_parallel.py:140, gda_fun_parallel.py:139"""
    return torch.linalg.lstsq(A.double(), b.double(), rcond=None, driver=None)[0].to(dtype=out_dtype)


def stop_masks(g, h, p_after, f, f_prev, ftol, ptol, gtol, fcon_enabled=True):
    """gcon / pcon / fcon.  reference: lma:108-110, gna:78-80, gda:74-76 (Q2, Q12).
    For a 2-D f (GDA) the fcon sums collapse to one batch-global scalar (Q5)."""
    gcon = torch.max(abs(g), dim=-1)[0] < gtol
    pcon = (h ** 2).sum(-1) ** .5 < ptol * (ptol + (p_after ** 2).sum(-1) ** .5)
    if fcon_enabled and f_prev.shape == f.shape:
        fcon = ((f_prev - f) ** 2).sum((-1, -2)) < ((ftol * f) ** 2).sum((-1, -2))
    else:
        fcon = torch.zeros_like(gcon)
    return gcon, pcon, fcon


def levenberg_update(rho, u, v):
    """Nielsen damping update, in place (Q9).  reference: lma_fun_parallel.py:124-136"""
    ones = torch.ones_like(u)
    up, dn = rho > 0, rho < 0
    u[up] *= torch.maximum(ones / 3, 1 - (2 * rho - 1) ** 3)[up]
    u[dn] *= v[dn]
    v[up] = 2 * ones[up]
    v[dn] *= 2
    return u, v


def marquardt_update(rho, u, rho1, rho2, beta, gama):
    """Marquardt damping update, in place (Q9).  reference: lma_fun_parallel.py:139-152"""
    u[rho < rho1] *= beta
    u[rho > rho2] /= gama
    return u


# ---------------------------------------------------------------------------- trace
@dataclass
class Trace:
    """Everything one solve produced.  p_list mirrors the reference's return value."""
    p_list: List[torch.Tensor] = field(default_factory=list)
    g: List[torch.Tensor] = field(default_factory=list)
    H: List[torch.Tensor] = field(default_factory=list)
    h: List[torch.Tensor] = field(default_factory=list)
    rho: List[torch.Tensor] = field(default_factory=list)
    u: List[torch.Tensor] = field(default_factory=list)
    converged: List[torch.Tensor] = field(default_factory=list)

    @property
    def iterations(self):
        return len(self.p_list)


def _bind(function, jac_function, args):
    # reference: lma:62-72 / gna:53-63 / gda:51-61
    fun = (lambda q: function(q, *args)) if len(args) > 0 else function
    if jac_function is None:
        jac = lambda q: autograd_batch_jacobian(q, fun)
    else:
        jac = lambda q: jac_function(q, *args)
    return fun, jac


def autograd_batch_jacobian(p, fun):
    """[B,C,N,P] Jacobian via one autograd pass over the batch-summed function.
    reference: torchimize/functions/jacobian.py:36-57"""
    summed = lambda q: torch.sum(fun(q), axis=0)
    try:
        js = torch.autograd.functional.jacobian(summed, p, vectorize=True)
    except RuntimeError:
        js = torch.autograd.functional.jacobian(summed, p, strict=True, vectorize=False)
    return js.permute(2, 0, 1, 3)          # [C,N,B,P] -> [B,C,N,P]


# -------------------------------------------------------------------------- solvers
def lma(p, function, jac_function=None, args=(), wvec=None, ftol=1e-8, ptol=1e-8, gtol=1e-8,
        tau=1e-3, meth='lev', rho1=.25, rho2=.75, beta=2, gama=3, max_iter=100) -> Trace:
    """Batched Levenberg-Marquardt.  reference: lma_fun_parallel.py:24-121.
    Mutates p in place like the reference (Q1)."""
    fun, jac = _bind(function, jac_function, args)
    assert p.dim() == 2
    B, P = p.shape
    wvec = torch.ones(1, dtype=p.dtype, device=p.device) if wvec is None else wvec
    kw = dict(dtype=p.dtype, device=p.device)      # reference: every state tensor lives on p.device (lma:76-79)
    dscale = torch.ones(B, **kw)                   # D is always dscale * I  (lma:77,84,88)
    eye = torch.eye(P, **kw)[None]
    u = tau * torch.ones(B, **kw)                  # Q8
    v = 2 * torch.ones(B, **kw)
    inf2 = torch.tensor([-torch.inf, torch.inf], **kw)
    tr = Trace()
    f_prev = torch.zeros(1, dtype=p.dtype, device=p.device)
    while tr.iterations < max_iter:
        f, J = fun(p), jac(p)
        g, H = normal_equations(f, J, wvec)
        if meth != 'lev':                           # Q6: scale compounds every iteration
            diagH = H.diagonal(dim1=1, dim2=2)
            dscale = dscale * torch.max(torch.maximum(diagH, dscale[:, None].expand(-1, P)), dim=1)[0]
        D = dscale[:, None, None] * eye
        Hu = H + u[:, None, None] * D
        h = -solve_fp64(Hu, g, p.dtype)
        f_h = fun(p + h)
        num = cost_sum_stat(f) - cost_sum_stat(f_h)
        den = torch.einsum('bnp,bni->bi', h[..., None], (u[:, None] * h - g)[..., None])[..., 0]  # lma:101
        rho = num / den
        zero = den == 0
        rho[zero] = inf2[(num > 0).type(torch.int64)][zero]
        if meth == 'lev':
            levenberg_update(rho, u, v)
        else:
            marquardt_update(rho, u, rho1, rho2, beta, gama)
        acc = rho > 0
        p[acc] += h[acc]                            # in place on caller's p (Q1)
        gcon, pcon, fcon = stop_masks(g, h, p, f, f_prev, ftol, ptol, gtol,
                                      fcon_enabled=bool(acc.sum() > 0))   # Q10
        f_prev = f.clone()
        conv = gcon | pcon | fcon
        tr.p_list.append(p.clone()); tr.g.append(g); tr.H.append(H); tr.h.append(h)
        tr.rho.append(rho.clone()); tr.u.append(u.clone()); tr.converged.append(conv)
        if conv.all():
            break
    return tr


def gna(p, function, jac_function=None, args=(), wvec=None, ftol=1e-8, ptol=1e-8, gtol=1e-8,
        l=1., max_iter=100) -> Trace:
    """Batched Gauss-Newton with stop conditions.  reference: gna_fun_parallel.py:25-91.
    The reference hands ``l`` to an unused positional slot, so the step rate is always 1 (Q4)."""
    fun, jac = _bind(function, jac_function, args)
    assert p.dim() == 2
    wvec = torch.ones(1, dtype=p.dtype, device=p.device) if wvec is None else wvec
    tr = Trace()
    f_prev = torch.zeros(1, dtype=p.dtype, device=p.device)
    while tr.iterations < max_iter:
        f, J = fun(p), jac(p)
        g, H = normal_equations(f, J, wvec)
        h = solve_fp64(H, g, p.dtype)
        p -= 1. * h                                 # Q4
        gcon, pcon, fcon = stop_masks(g, h, p, f, f_prev, ftol, ptol, gtol)
        f_prev = f.clone()
        conv = gcon | pcon | fcon
        tr.p_list.append(p.clone()); tr.g.append(g); tr.H.append(H); tr.h.append(h)
        tr.converged.append(conv)
        if conv.all():
            break
    return tr


def gna_plain(p, function, jac_function, wvec, l=1., max_iter=100):
    """Fixed-iteration Gauss-Newton; honours l.  reference: gna_fun_parallel.py:94-117"""
    for _ in range(max_iter):
        f, J = function(p), jac_function(p)
        g, H = normal_equations(f, J, wvec)
        p -= l * solve_fp64(H, g, p.dtype)
    return p


def gda_step(p, fun, jac, wvec, l):
    """Cost-summed tall least-squares step.  reference: gda_fun_parallel.py:116-148
    (the _LinAlgError identity retry is not restated: torch CPU lstsq/gelsy never raises it)."""
    fc, jc = fun(p), jac(p)
    f = torch.einsum('bcp,c->bp', fc, wvec)
    j = torch.einsum('bcpi,c->bpi', jc, wvec)
    h = solve_fp64(j, f, p.dtype)
    p -= l * h
    return f, h


def gda(p, function, jac_function=None, args=(), wvec=None, ftol=1e-8, ptol=1e-8, gtol=1e-8,
        l=1., max_iter=100) -> Trace:
    """Batched "gradient descent" (really a cost-summed Gauss-Newton/least-squares step).
    reference: gda_fun_parallel.py:23-87.  fcon is batch-global (Q5)."""
    fun, jac = _bind(function, jac_function, args)
    assert p.dim() == 2
    wvec = torch.ones(1, dtype=p.dtype, device=p.device) if wvec is None else wvec
    tr = Trace()
    f_prev = torch.zeros(1, dtype=p.dtype, device=p.device)
    while tr.iterations < max_iter:
        f, h = gda_step(p, fun, jac, wvec, l)
        g = h / l
        gcon, pcon, fcon = stop_masks(g, h, p, f, f_prev, ftol, ptol, gtol)
        f_prev = f.clone()
        conv = gcon | pcon | fcon
        tr.p_list.append(p.clone()); tr.g.append(g); tr.h.append(h); tr.converged.append(conv)
        if conv.all():
            break
    return tr


def gda_plain(p, function, jac_function, wvec, l=1., max_iter=100):
    """reference: gda_fun_parallel.py:90-113"""
    for _ in range(max_iter):
        gda_step(p, function, jac_function, wvec, l)
    return p
