"""gradient_descent — drop-in for torchimize/functions/single/gda_fun_single.py:23-84."""
from typing import Callable, List, Tuple, Union

import torch

from . import _core


def gradient_descent(
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
    """The reference's "gradient descent" for ONE problem: the step is the least-squares solution of the tall system
    J h = f scaled by l (gda:62); returns the iterates starting with the initial p."""
    return _core.gradient_descent(p, function, jac_function, args, ftol, ptol, gtol, l, max_iter)
