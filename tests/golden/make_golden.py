"""Generates tests/golden/*.npz by running the UNMODIFIED reference (hahnec/torchimize,
imported read-only from /root/reference) in the build container.  The reference cannot
travel to the GPU box, so its inputs and per-iteration outputs are committed as fixtures.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Files written (all arrays fp64 unless the key ends in _f32):
  fixtures.npz   the reference's own ParallelOptimizationTest fixtures (B=6,C=2,N=2000,P=4;
                 tests/unit_test_parallel.py:28-65) pushed through every batched solver
  synthetic.npz  seeded EMG batch B=48,C=3,N=256,P=4 with per-cost data and a non-trivial wvec
  skewgauss.npz  seeded skewed-Gaussian batch B=32,C=1,N=1000, P=4 and P=5

Key scheme: ``<case>__<dtype>__plist`` = stacked per-iteration snapshots [K,B,P]; the case's
call arguments are in CASES below and mirrored by tests/test_oracle_golden.py.
"""
import os
import sys

import numpy as np
import torch

REF = '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.dont_write_bytecode = True
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)       # reference first: its `tests` package must win over ours

from torchimize.functions import (lsq_lma_parallel, lsq_gna_parallel, lsq_gna_parallel_plain,   # noqa: E402
                                  gradient_descent_parallel, gradient_descent_parallel_plain,
                                  newton_step_parallel)
from tests.unit_test_parallel import ParallelOptimizationTest                                    # noqa: E402  (reference's tests pkg)
from oracle import models                                                                        # noqa: E402

TOL = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8)

# case name -> (solver, kwargs).  Mirrored in tests/test_oracle_golden.py::CASES.
CASES = {
    'lma_lev':   (lsq_lma_parallel, dict(meth='lev', max_iter=99, **TOL)),
    'lma_marq':  (lsq_lma_parallel, dict(meth='marq', max_iter=99, **TOL)),
    'gna':       (lsq_gna_parallel, dict(l=1., max_iter=99, **TOL)),
    'gna_l05':   (lsq_gna_parallel, dict(l=.5, max_iter=99, **TOL)),          # Q4: l ignored
    'gda':       (gradient_descent_parallel, dict(l=1., max_iter=99, **TOL)),
    'gda_l05':   (gradient_descent_parallel, dict(l=.5, max_iter=99, **TOL)),
}
PLAIN = {
    'gna_plain': (lsq_gna_parallel_plain, dict(l=.1, max_iter=25)),
    'gda_plain': (gradient_descent_parallel_plain, dict(l=.1, max_iter=25)),
}


def run_all(out, tag, p0, fun, jac, args, wvec, dtypes=(torch.float64, torch.float32), cases=CASES, plain=PLAIN):
    for dt in dtypes:
        dn = 'f64' if dt == torch.float64 else 'f32'
        a = tuple(x.to(dt) if torch.is_tensor(x) else x for x in args)
        w = wvec.to(dt)
        for name, (solver, kw) in cases.items():
            p = p0.clone().to(dt)
            try:
                pl = solver(p=p, function=fun, jac_function=jac, args=a, wvec=w, **kw)
            except RuntimeError as e:      # Marquardt scale overflow (Q6) -> NaN -> LAPACK argument error
                out[f'{tag}{name}__{dn}__raises'] = np.array(1)
                print(f'{tag}{name:10s} {dn}: reference raised {type(e).__name__}')
                continue
            assert torch.equal(pl[-1], p), 'reference mutates p in place (Q1)'
            out[f'{tag}{name}__{dn}__plist'] = torch.stack(pl).numpy()
            print(f'{tag}{name:10s} {dn}: {len(pl):3d} iterations')
        for name, (solver, kw) in plain.items():
            p = p0.clone().to(dt)
            res = solver(p=p, function=lambda q: fun(q, *a), jac_function=lambda q: jac(q, *a), wvec=w, **kw)
            out[f'{tag}{name}__{dn}__p'] = res.numpy()
        # one normal-equation build at the initial point (newton_parallel.py:21-45)
        p = p0.clone().to(dt)
        _, f, g, H = newton_step_parallel(p, lambda q: fun(q, *a), lambda q: jac(q, *a), w)
        out[f'{tag}newton__{dn}__g'] = g.numpy()
        out[f'{tag}newton__{dn}__H'] = H.numpy()
        out[f'{tag}newton__{dn}__f'] = f.numpy()


def fixtures():
    t = ParallelOptimizationTest()
    t.setUp()
    out = {'t': t.t.numpy(), 'y': t.batch_data_channels.numpy(), 'p0': t.batch_initials.numpy(),
           'gt': t.gt_params.numpy(), 'wvec': np.ones(2)}
    # the vectorised restatement must be bit-identical to the reference's loop callables
    for dt in (torch.float64, torch.float32):
        p, tt, yy = t.batch_initials.to(dt), t.t.to(dt), t.batch_data_channels.to(dt)
        assert torch.equal(t.multi_cost_batch(p, tt, yy), models.emg_cost(p, tt, yy, n_costs=2))
        assert torch.equal(t.multi_jaco_batch(p, tt, yy), models.emg_jac(p, tt, yy, n_costs=2))
    run_all(out, '', t.batch_initials, t.multi_cost_batch, t.multi_jaco_batch,
            (t.t, t.batch_data_channels), torch.ones(2, dtype=torch.float64))
    # jac_function=None path (autograd Jacobian), tolerances of the reference's tests (1e-5)
    p = t.batch_initials.clone()
    pl = lsq_lma_parallel(p=p, function=t.multi_cost_batch, jac_function=None,
                          args=(t.t.double(), t.batch_data_channels), wvec=torch.ones(2, dtype=torch.float64),
                          ftol=1e-5, ptol=1e-5, gtol=1e-5, meth='lev', max_iter=99)
    out['lma_lev_autojac__f64__plist'] = torch.stack(pl).numpy()
    np.savez_compressed(os.path.join(HERE, 'fixtures.npz'), **out)


def emg_batch(B, C, N, seed):
    """Seeded synthetic EMG batch as SURVEY.md §8d describes (t built in fp32 then cast)."""
    gen = torch.Generator().manual_seed(seed)
    base = torch.tensor([10., -1., 80., 5.], dtype=torch.float64)
    gt = base * (1 + .1 * (torch.rand(B, 4, generator=gen, dtype=torch.float64) - .5))
    p0 = gt * (1 + .2 * (torch.rand(B, 4, generator=gen, dtype=torch.float64) - .5))
    t = torch.linspace(-1e3, 1e3, N).double()
    clean = models.emg_model(gt, t)
    y = torch.stack([clean + .01 * torch.randn(B, N, generator=torch.Generator().manual_seed(seed + c),
                                                 dtype=torch.float64) for c in range(C)], dim=1)
    return gt, p0, t, y


def synthetic():
    gt, p0, t, y = emg_batch(48, 3, 256, 3008)
    w = torch.tensor([1., .5, 2.], dtype=torch.float64)
    out = {'t': t.numpy(), 'y': y.numpy(), 'p0': p0.numpy(), 'gt': gt.numpy(), 'wvec': w.numpy()}
    run_all(out, '', p0, models.emg_cost, models.emg_jac, (t, y), w)
    np.savez_compressed(os.path.join(HERE, 'synthetic.npz'), **out)


def skewgauss():
    out = {}
    for P in (4, 5):
        gen = torch.Generator().manual_seed(3006 + P)
        B, N = 32, 1000
        x = torch.linspace(-5, 5, N, dtype=torch.float64)
        base = torch.tensor([10., -1., 2., 5., 0.][:P], dtype=torch.float64)
        init = torch.tensor([7.5, -.5, .5, 2.5, .1][:P], dtype=torch.float64)
        jit = lambda s: 1 + s * (torch.rand(B, P, generator=gen, dtype=torch.float64) - .5)
        gt = base * jit(.05)
        p0 = init * jit(.05)
        y = models.skewgauss_model(gt, x) + .01 * torch.randn(B, N, generator=gen, dtype=torch.float64)
        tag = f'P{P}_'
        out.update({tag + 'x': x.numpy(), tag + 'y': y.numpy(), tag + 'p0': p0.numpy(), tag + 'gt': gt.numpy()})
        cases = {k: CASES[k] for k in ('lma_lev', 'gna', 'gda')}
        run_all(out, tag, p0, models.skewgauss_cost, models.skewgauss_jac, (x, y),
                torch.ones(1, dtype=torch.float64), cases=cases, plain={})
    np.savez_compressed(os.path.join(HERE, 'skewgauss.npz'), **out)


if __name__ == '__main__':
    torch.set_num_threads(8)
    fixtures()
    synthetic()
    skewgauss()
    for f in ('fixtures', 'synthetic', 'skewgauss'):
        print(f, os.path.getsize(os.path.join(HERE, f + '.npz')) // 1024, 'KiB')
