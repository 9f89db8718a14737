"""newton_step_parallel — drop-in for torchimize/functions/parallel/newton_parallel.py:21-45."""
from typing import Callable, Tuple

import torch

from ... import _lib
from ._engine import _as_work


@_lib.on_device_of(0)
def newton_step_parallel(p: torch.Tensor, function: Callable, jac_function: Callable, wvec: torch.Tensor
                         ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """Builds g = sum_c w_c J^T f and H = sum_c w_c J^T J with ONE fused pass over J (K1,
    nlls_normal_eq) instead of the reference's four einsums.  Returns (p, f, g, H) with p the
    same tensor object, like the reference (newton_parallel.py:45)."""
    _lib.require_cuda(p)
    f = _as_work(function(p), p.dtype, 'function')
    B, C, N = f.shape
    P = p.shape[1]
    J = _as_work(jac_function(p), p.dtype, 'jac_function', (B, C, N, P))
    w = wvec.to(dtype=p.dtype, device=p.device).contiguous()
    g = torch.empty(B, P, dtype=p.dtype, device=p.device)
    H = torch.empty(B, P, P, dtype=p.dtype, device=p.device)
    stats = torch.empty(B, 4, dtype=torch.float64, device=p.device)
    _lib.call('nlls_normal_eq', _lib.dtype_code(p), _lib.ptr(J), _lib.ptr(f), None, _lib.ptr(w), w.numel(), B, C, N, P,
              _lib.ptr(g), _lib.ptr(H), _lib.ptr(stats), None, _lib.stream_ptr())
    return p, f, g, H
