"""Single-problem solvers (torchimize_b200.functions.{lsq_lma, lsq_gna, gradient_descent}; reference:
torchimize/functions/single/*.py) on the B200 kernels: against the golden trajectories of the imported reference
(tests/golden/single.npz) and against the oracle on a long problem that spans many K1 chunks."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

DEV = 'cuda:0'
CASES = {
    'lma_lev': ('lsq_lma', dict(meth='lev', gtol=1e-6, max_iter=39)),
    'lma_marq': ('lsq_lma', dict(meth='marq', gtol=1e-6, max_iter=39)),
    'gna': ('lsq_gna', dict(l=.1, gtol=1e-6, max_iter=199)),
    'gna_l025': ('lsq_gna', dict(l=.25, max_iter=99)),
    'gda': ('gradient_descent', dict(l=.1, gtol=1e-6, max_iter=199)),
    'gda_l025': ('gradient_descent', dict(l=.25, max_iter=99)),
}


def _callables():
    """The single-EMG cost and analytic Jacobian as CUDA callables: the built-in EMG kernel on a batch of one."""
    from torchimize_b200.models import EMGModel
    emg = EMGModel()
    cost = lambda p, t, y: emg.cost(p[None], t, y[None])[0, 0]
    jac = lambda p, t, y: emg.jac(p[None], t, y[None])[0, 0]
    return cost, jac


@pytest.mark.gpu
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', sorted(CASES))
def test_single_solvers_against_the_reference_golden_trajectories(name, dn):
    import torchimize_b200.functions as F
    gold = dict(np.load(os.path.join(GOLDEN, 'single.npz')))
    dt = torch.float64 if dn == 'f64' else torch.float32
    t, y, p0 = (torch.as_tensor(gold[k]).to(dt).to(DEV) for k in ('t', 'y', 'p0'))
    fn, kw = CASES[name]
    cost, jac = _callables()
    want = torch.as_tensor(gold[f'{name}__{dn}__ana'])
    p_in = p0.clone()
    got = getattr(F, fn)(p_in, cost, jac, args=(t, y), **kw)
    assert got[0] is p_in                                         # the list starts with the caller's tensor (lma:74)
    got = torch.stack([q.detach().cpu() for q in got])
    if dn == 'f32':
        # fp32 trajectories sit on rounding noise (the reference's own iteration count changes between runs): finals
        if got.shape != want.shape:
            print(f'BOUNDARY FLIP single {name} f32: {got.shape[0]} iterates, reference {want.shape[0]}')
        err = float((got[-1] - want[-1]).norm() / want[-1].norm())
        assert err <= 1e-4, f'{name} f32 final {err:.2e}'
        return
    assert got.shape == want.shape, f'{name}: {got.shape[0]} iterates, reference {want.shape[0]}'
    err = float(((got - want).norm(dim=1) / want.norm(dim=1)).max())
    assert err <= 1e-9, f'{name} f64: {err:.2e}'


@pytest.mark.gpu
@pytest.mark.parametrize('fn,kw', [('lsq_lma', dict(meth='lev', max_iter=40)), ('lsq_lma', dict(meth='marq', max_iter=40)),
                                   ('lsq_gna', dict(l=.5, max_iter=60)), ('gradient_descent', dict(l=.5, max_iter=60))])
def test_single_long_problem_spans_many_chunks_and_matches_the_oracle(fn, kw):
    """N = 3,000,017 residuals (odd: full chunks + a tail, unaligned rows): K1 runs on ~2900 pseudo-problems."""
    import torchimize_b200.functions as F
    from oracle import single_oracle as S, models as M
    N = 3_000_017
    g = torch.Generator().manual_seed(5)
    t = torch.linspace(-1e3, 1e3, N, dtype=torch.float64)
    gt = torch.tensor([10., -1., 80., 5.], dtype=torch.float64)
    y = M.emg_model(gt[None], t)[0] + .01 * torch.randn(N, generator=g, dtype=torch.float64)
    p0 = torch.tensor([9., -2., 70., 4.], dtype=torch.float64)
    ref = getattr(S, fn)(p0.clone(), M.emg_single_cost, M.emg_single_jac, args=(t, y), **kw)
    cost, jac = _callables()
    got = getattr(F, fn)(p0.clone().to(DEV), cost, jac, args=(t.to(DEV), y.to(DEV)), **kw)
    assert len(got) == len(ref)
    ref, got = torch.stack(ref), torch.stack([q.cpu() for q in got])
    err = float(((got - ref).norm(dim=1) / ref.norm(dim=1)).max())
    assert err <= 1e-9, f'{fn}: {err:.2e}'
    assert torch.allclose(got[-1], gt, atol=1e-2)


@pytest.mark.gpu
def test_single_autograd_jacobian_fallback():
    import torchimize_b200.functions as F
    gold = dict(np.load(os.path.join(GOLDEN, 'single.npz')))
    t, y, p0 = (torch.as_tensor(gold[k]).double().to(DEV) for k in ('t', 'y', 'p0'))

    def cost(p, t, y):                                            # differentiable torch callable (tests/emg.py:19-21)
        a, mu, s, eta = p
        d = t - mu
        return (y - a * torch.exp(-.5 * d ** 2 / s ** 2) * (1 + torch.erf(eta * d / (s * 2 ** .5)))) ** 2

    got = F.lsq_lma(p0.clone(), cost, None, args=(t, y), meth='lev', gtol=1e-6, max_iter=39)
    want = torch.as_tensor(gold['lma_lev__f64__auto'])
    assert len(got) == want.shape[0]
    err = float(((torch.stack([q.cpu() for q in got]) - want).norm(dim=1) / want.norm(dim=1)).max())
    assert err <= 1e-8, f'{err:.2e}'


@pytest.mark.gpu
def test_single_rejects_cpu_and_batched_input():
    import torchimize_b200.functions as F
    from torchimize_b200 import _lib
    with pytest.raises(_lib.NllsError):
        F.lsq_gna(torch.zeros(4, dtype=torch.float64), lambda p: p, lambda p: torch.eye(4, dtype=torch.float64))
    with pytest.raises(AssertionError):
        F.lsq_gna(torch.zeros(2, 4, dtype=torch.float64, device=DEV), lambda p: p, None)
