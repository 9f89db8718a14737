"""lsq_lma_parallel — drop-in for torchimize/functions/parallel/lma_fun_parallel.py:24-121."""
from typing import Callable, List, Tuple, Union

import torch

from . import _engine


def lsq_lma_parallel(
        p: torch.Tensor,
        function: Callable,
        jac_function: Callable = None,
        args: Union[Tuple, List] = (),
        wvec: torch.Tensor = None,
        ftol: float = 1e-8,
        ptol: float = 1e-8,
        gtol: float = 1e-8,
        tau: float = 1e-3,
        meth: str = 'lev',
        rho1: float = .25,
        rho2: float = .75,
        beta: float = 2,
        gama: float = 3,
        max_iter: int = 100,
        *, fused: bool = True, state: dict = None, mode: str = 'exact', stats: dict = None,
) -> List[torch.Tensor]:
    """Batched Levenberg-Marquardt on the B200 core.

    Same arguments, tensor contracts (function -> B x C x N, jac_function -> B x C x N x P) and
    return value (list of per-iteration B x P snapshots; p is updated in place) as the reference.
    ``meth='lev'`` is Levenberg/Nielsen damping, anything else Marquardt (lma:83-88).
    ``fused`` (keyword-only, not in the reference): allow the in-kernel model evaluation when
    ``function``/``jac_function`` are the methods of a torchimize_b200.models.FusedModel.
    ``state`` (keyword-only, not in the reference): a caller-owned dict for warm restarts. An empty dict
    starts as the reference does (u = tau, v = 2) and receives the damping state (u, v, Marquardt scale) at return;
    passing it to the next call continues from that state, so that k calls of max_iter = n walk the same iterates as one
    call of max_iter = k n (as long as no call stopped early).
    ``mode`` / ``stats`` (keyword-only, not in the reference): see _engine.solve — 'retire' freezes a problem at the iterate
    at which its own stop test fired (fcon counts only in an iteration whose step that problem accepted)."""
    return _engine.solve(_engine.LM, p, function, jac_function, args, wvec, ftol, ptol, gtol, max_iter,
                         lm=dict(tau=tau, meth=meth, rho1=rho1, rho2=rho2, beta=beta, gama=gama, state=state), fused=fused,
                         mode=mode, stats=stats)
