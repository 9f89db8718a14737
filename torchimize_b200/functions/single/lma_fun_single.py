"""lsq_lma — drop-in for torchimize/functions/single/lma_fun_single.py:23-104."""
from typing import Callable, List, Tuple, Union

import torch

from . import _core


def lsq_lma(
        p: torch.Tensor,
        function: Callable,
        jac_function: Callable = None,
        args: Union[Tuple, List] = (),
        ftol: float = 1e-8,
        ptol: float = 1e-8,
        gtol: float = 1e-8,
        tau: float = 1e-3,
        meth: str = 'lev',
        rho1: float = .25,
        rho2: float = .75,
        bet: float = 2,
        gam: float = 3,
        max_iter: int = 100,
) -> List[torch.Tensor]:
    """Levenberg-Marquardt for ONE problem: p [P], function(p, *args) -> [N], jac_function(p, *args) -> [N,P]
    (autograd when None).  Returns the list of iterates starting with the initial p, like the reference."""
    return _core.lsq_lma(p, function, jac_function, args, ftol, ptol, gtol, tau, meth, rho1, rho2, bet, gam, max_iter)
