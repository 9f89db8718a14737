"""Numeric-Jacobian fallback used when ``jac_function is None``.

Boundary helper, not a kernel target: like the reference
(torchimize/functions/jacobian.py:20-57) it asks torch autograd for the exact derivative of the
user's callable; problems are independent, so one pass over the batch-summed function yields
every per-problem Jacobian."""
import torch


def _autograd_jacobian(fn, p, create_graph):
    try:
        return torch.autograd.functional.jacobian(fn, p, create_graph=create_graph, vectorize=True)
    except RuntimeError:
        return torch.autograd.functional.jacobian(fn, p, create_graph=create_graph, strict=True, vectorize=False)


def jacobian_approx_t(p, f, create_graph=False):
    """Jacobian of f at p (no batch handling); reference: jacobian.py:20-34."""
    return _autograd_jacobian(f, p, create_graph)


def batch_jacobian_approx_t(p, f, create_graph=False):
    """[B,C,N,P] Jacobian of a batched callable f(p)->[B,C,N]; reference: jacobian.py:36-57."""
    over_batch = lambda q: f(q).sum(dim=0)
    jac = _autograd_jacobian(over_batch, p, create_graph)       # [C,N,B,P]
    lead = jac.dim() - 2
    return jac.permute(lead, *range(lead), lead + 1)
