"""lsq_gna_parallel(_plain) — drop-in for torchimize/functions/parallel/gna_fun_parallel.py:25-117."""
from typing import Callable, List, Tuple, Union

import torch

from . import _engine


def lsq_gna_parallel(
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
    """Batched Gauss-Newton with stop conditions.  NOTE: like the reference, ``l`` is accepted
    but has no effect here — gna:72 passes it into an unused positional slot, so the step rate is
    always 1.  lsq_gna_parallel_plain honours it.
    Keyword-only, not in the reference: ``fused`` (in-kernel model evaluation for FusedModel callables), ``mode``
    ('exact' = the reference's semantics; 'retire' = a problem whose own stop test fired is frozen at that iterate and never
    evaluated again, the solve ends when none is left — see _engine.solve), ``stats`` (a caller-owned dict that receives
    per-problem stop iteration / reason bits / step counts / non-finite flags)."""
    return _engine.solve(_engine.GN, p, function, jac_function, args, wvec, ftol, ptol, gtol, max_iter, l=1., fused=fused,
                         mode=mode, stats=stats)


def lsq_gna_parallel_plain(
        p: torch.Tensor,
        function: Callable,
        jac_function: Callable,
        wvec: torch.Tensor,
        l: float = 1.,
        max_iter: int = 100,
) -> torch.Tensor:
    """Fixed number of Gauss-Newton steps, no stop tests; returns the final p (gna:94-117)."""
    return _engine.plain(_engine.GN, p, function, jac_function, wvec, l, max_iter)
