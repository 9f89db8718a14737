"""gradient_descent_parallel(_plain) — drop-in for torchimize/functions/parallel/gda_fun_parallel.py:23-113."""
from typing import Callable, List, Tuple, Union

import torch

from . import _engine


def gradient_descent_parallel(
        p: torch.Tensor,
        function: Callable,
        jac_function: Callable = None,
        args: Union[Tuple, List] = (),
        wvec: torch.Tensor = None,
        ftol: float = 1e-8,
        ptol: float = 1e-8,
        gtol: float = 1e-8,
        l: float = 1.,
        max_iter: int = 100,
        *, fused: bool = True, mode: str = 'exact', stats: dict = None,
) -> List[torch.Tensor]:
    """The reference's "gradient descent": h = argmin || (sum_c w_c J_c) h - sum_c w_c f_c ||,
    p -= l h (gda:116-148).  The tall N x P least-squares problem is solved through fp64 normal
    equations formed in-stream (nlls_normal_eq_gda).  The f-tolerance test is one scalar for the
    whole batch, as in the reference (gda:76); mode='retire' tests each problem's own two sums instead.
    Keyword-only, not in the reference: ``fused`` (in-kernel model evaluation for FusedModel callables), ``mode``
    ('exact' = the reference's semantics; 'retire' = a problem whose own stop test fired is frozen at that iterate and never
    evaluated again, the solve ends when none is left — see _engine.solve), ``stats`` (a caller-owned dict that receives
    per-problem stop iteration / reason bits / step counts / non-finite flags)."""
    return _engine.solve(_engine.GD, p, function, jac_function, args, wvec, ftol, ptol, gtol, max_iter, l=l, fused=fused,
                         mode=mode, stats=stats)


def gradient_descent_parallel_plain(
        p: torch.Tensor,
        function: Callable,
        jac_function: Callable,
        wvec: torch.Tensor,
        l: float = 1.,
        max_iter: int = 100,
) -> torch.Tensor:
    """Fixed number of steps, no stop tests; returns the final p (gda:90-113)."""
    return _engine.plain(_engine.GD, p, function, jac_function, wvec, l, max_iter)
