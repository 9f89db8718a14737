"""Shape validator for user callables; reference: torchimize/functions/fun_dims.py:23-63."""
import torch

from .jacobian import jacobian_approx_t


def test_fun_dims_parallel(p, function, jac_function=None, args=(), wvec=None) -> bool:
    fun = (lambda q: function(q, *args)) if len(args) > 0 else function
    jac = (lambda q: jacobian_approx_t(q, f=fun)) if jac_function is None else (lambda q: jac_function(q, *args))
    assert len(p.shape) == 2, 'parameter tensor is supposed to have 2 dims, but has %s' % str(len(p.shape))
    f = fun(p)
    assert len(f.shape) == 3, 'residual tensor is supposed to have 3 dims, but has %s' % str(len(f.shape))
    j = jac(p)
    assert len(j.shape) == 4, 'jacobian tensor is supposed to have 4 dims, but has %s' % str(len(j.shape))
    n_costs = f.shape[1]
    wvec = torch.ones(n_costs, dtype=p.dtype, device=p.device) if wvec is None else wvec
    assert len(wvec) == n_costs, 'weights vector length of %s is supposed to match cost number which is %s' % (
        str(len(wvec)), str(n_costs))
    return True


test_fun_dims_parallel.__test__ = False      # not a pytest test
