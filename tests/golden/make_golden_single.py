#!/usr/bin/env python
"""Generates tests/golden/single.npz by importing the UNMODIFIED reference from /root/reference (build container only):
inputs and the full p_list of the single-problem solvers lsq_lma (lev, marq), lsq_gna, gradient_descent on a seeded
version of the reference's own single-EMG fixture (tests/unit_test_analytical_jacobian.py:37-47 uses torch.seed();
here the noise is seeded so that the vectors are reproducible), fp64 and fp32, analytic and autograd Jacobian.

    PYTHONDONTWRITEBYTECODE=1 PYTHONPATH=/root/reference python tests/golden/make_golden_single.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, '/root/reference')
from torchimize.functions.single.lma_fun_single import lsq_lma            # noqa: E402
from torchimize.functions.single.gna_fun_single import lsq_gna            # noqa: E402
from torchimize.functions.single.gda_fun_single import gradient_descent   # noqa: E402
from tests.unit_test_analytical_jacobian import JacobianFunctionTest      # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = {
    'lma_lev': (lsq_lma, dict(meth='lev', gtol=1e-6, max_iter=39)),
    'lma_marq': (lsq_lma, dict(meth='marq', gtol=1e-6, max_iter=39)),
    'gna': (lsq_gna, dict(l=.1, gtol=1e-6, max_iter=199)),
    'gna_l025': (lsq_gna, dict(l=.25, max_iter=99)),
    'gda': (gradient_descent, dict(l=.1, gtol=1e-6, max_iter=199)),
    'gda_l025': (gradient_descent, dict(l=.25, max_iter=99)),
}


def main():
    T = JacobianFunctionTest()
    g = torch.Generator().manual_seed(3008)
    gt = torch.tensor([10, -1, 80, 5], dtype=torch.float64)
    p0 = torch.tensor([7.5, -.75, 10, 3], dtype=torch.float64)
    t = torch.linspace(-1e3, 1e3, int(2e3))                       # built in float32 as in the fixture, cast per case
    y = T.emg_model(gt, t=t) + .01 * torch.randn(len(t), generator=g, dtype=torch.float64)
    out = {'t': t.numpy(), 'y': y.numpy(), 'p0': p0.numpy(), 'gt': gt.numpy()}
    cost = lambda p, t, y: (y - T.emg_model(p, t)) ** 2
    for dn, dt in (('f64', torch.float64), ('f32', torch.float32)):
        for name, (solver, kw) in CASES.items():
            for jn, jac in (('ana', T.emg_jac), ('auto', None)):
                if jn == 'auto' and (dn == 'f32' or name not in ('lma_lev', 'gna')):
                    continue
                tt = t.to(dt)                                    # same dtype as p: no silent float32 arithmetic in the model
                pl = solver(p0.to(dt).clone(), cost, jac_function=jac, args=(tt, y.to(dt)), **kw)
                out[f'{name}__{dn}__{jn}'] = torch.stack([q.detach() for q in pl]).numpy()
                print(name, dn, jn, len(pl), pl[-1].detach().numpy())
    np.savez_compressed(os.path.join(HERE, 'single.npz'), **out)


if __name__ == '__main__':
    main()
