"""lsq_gna — drop-in for torchimize/functions/single/gna_fun_single.py:23-87."""
from typing import Callable, List, Tuple, Union

import torch

from . import _core


def lsq_gna(
        p: torch.Tensor,
        function: Callable,
        jac_function: Callable = None,
        args: Union[Tuple, List] = (),
        ftol: float = 1e-8,
        ptol: float = 1e-8,
        gtol: float = 1e-8,
        l: float = 1.,
        max_iter: int = 100,
) -> List[torch.Tensor]:
    """Gauss-Newton for ONE problem with step rate l; returns the iterates starting with the initial p."""
    return _core.lsq_gna(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter)
