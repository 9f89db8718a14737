"""mode='retire' (north_star kernel 3) and the per-solve stats dict.

The reference never freezes a converged problem: its masked copy `p[~converged] = pn[~converged]` (lma:115, gna:85, gda:81)
is a no-op because pn aliases p.  mode='retire' does what that line was meant to do, on the device: a problem whose OWN stop
test fired keeps the iterate it had at that moment and is never evaluated again.  Per-problem arithmetic does not depend on
the rest of the batch, so the contract is exact:

  * stats['stop_iteration'][b] of the retire run == that of an 'exact' run (first iteration at which b's own test fired)
  * the frozen iterate == the 'exact' run's snapshot of that iteration, bit for bit; earlier snapshots coincide, later repeat
  * the retire run ends at max_b stop_iteration + 1 iterations (or max_iter)
  * for Gauss-Newton the own test IS the reference's per-problem mask (gna:78-84): stop_iteration is checked against the
    oracle's converged masks
"""
import os
import sys

import pytest
import torch

from helpers import DT, load

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle import nlls_oracle as O, models as M                   # noqa: E402
from torchimize_b200 import _lib                                   # noqa: E402
from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel   # noqa: E402
from torchimize_b200.functions.parallel import _engine             # noqa: E402
from torchimize_b200.models import EMGModel, SkewGaussModel, MultiEMGModel   # noqa: E402

pytestmark = pytest.mark.gpu
DEV = 'cuda'
TOL = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8)
SOLVERS = {'gna': (lsq_gna_parallel, dict(max_iter=99)), 'lma': (lsq_lma_parallel, dict(max_iter=99, meth='lev')),
           'lma_marq': (lsq_lma_parallel, dict(max_iter=99, meth='marq')), 'gda': (gradient_descent_parallel, dict(max_iter=40))}


def _both(solver, kw, p0, fun, jac, args, w, **extra):
    out = {}
    for mode in ('exact', 'retire'):
        st = {}
        p = p0.clone()
        snaps = torch.stack(solver(p, fun, jac, args=args, wvec=w, mode=mode, stats=st, **TOL, **kw, **extra))
        assert torch.equal(snaps[-1], p), 'caller p must end as the final iterate (Q1)'
        out[mode] = (snaps, st)
    return out


def _check_contract(res, max_iter):
    (E, se), (R, sr) = res['exact'], res['retire']
    stop = se['stop_iteration']
    B = stop.shape[0]
    assert se['mode'] == 'exact' and sr['mode'] == 'retire'
    assert E.shape[0] == se['iterations'] and R.shape[0] == sr['iterations']
    seen = stop >= 0                      # own test fired while the exact run lasted (LM: the reference's batch-global fcon gate
    if bool(seen.all()):                  # can end an exact run before a problem's OWN test fires; GD: batch-global fcon sums)
        assert R.shape[0] == int(stop.max()) + 1, f'retire ran {R.shape[0]} iterations, last own stop at {int(stop.max())}'
        assert R.shape[0] <= E.shape[0]
    else:
        assert R.shape[0] >= min(E.shape[0], max_iter) or bool((sr['stop_iteration'] >= 0).all())
    # the same own-test decisions in both runs, as long as the exact run lasted
    assert torch.equal(sr['stop_iteration'][seen], stop[seen])
    assert torch.equal(sr['stop_reason'][seen], se['stop_reason'][seen])
    assert bool((sr['stop_reason'][seen] > 0).all()) and bool((se['stop_reason'][~seen] == 0).all())
    K = min(R.shape[0], E.shape[0]) if not bool(seen.all()) else R.shape[0]
    idx = torch.arange(B, device=E.device)
    k = torch.arange(K, device=E.device)[:, None]
    upto = torch.where(seen, stop, torch.full_like(stop, E.shape[0] - 1)).long()[None, :]   # [1,B]
    src = torch.minimum(k.expand(K, B), upto.expand(K, B))                               # snapshot index of the exact run
    want = E[src, idx[None, :].expand(K, B)]
    got = R[:K]
    same = (got == want) | (torch.isnan(got) & torch.isnan(want))
    assert bool(same.all()), 'retire-mode snapshots differ from the exact run frozen at each problem\'s own stop iteration'
    # steps: a retired problem stops counting
    assert bool((sr['steps'][seen] <= se['steps'][seen]).all())
    return stop


@pytest.mark.parametrize('path', ['fused', 'cuda_callables', 'torch_callables'])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', ['gna', 'lma', 'gda'])
def test_retire_freezes_each_problem_at_its_own_stop_iteration_synthetic(name, dn, path):
    gold = load('synthetic')
    dt = DT[dn]
    t, y, p0, w = (torch.as_tensor(gold[k]).to(dt).to(DEV) for k in ('t', 'y', 'p0', 'wvec'))
    if path == 'torch_callables':
        fun, jac, extra = M.emg_cost, M.emg_jac, {}
    else:
        emg = EMGModel()
        fun, jac, extra = emg.cost, emg.jac, ({'fused': False} if path == 'cuda_callables' else {})
    solver, kw = SOLVERS[name]
    res = _both(solver, kw, p0, fun, jac, (t, y), w, **extra)
    stop = _check_contract(res, kw['max_iter'])
    if name == 'gna' and dn == 'f64':
        # GN's own test is the reference's per-problem mask: first iteration with converged[b] in the oracle's trace
        tr = O.gna(p0.cpu().clone(), M.emg_cost, M.emg_jac, args=(t.cpu(), y.cpu()), wvec=w.cpu(), **TOL, max_iter=kw['max_iter'])
        conv = torch.stack(tr.converged)                              # [K,B]
        first = torch.where(conv.any(0), conv.int().argmax(0), torch.full((conv.shape[1],), -1))
        assert torch.equal(first.to(torch.int32), stop.cpu()), 'stop_iteration differs from the oracle\'s masks (gna:78-84)'


@pytest.mark.parametrize('name', ['gna', 'lma', 'lma_marq', 'gda'])
def test_retire_large_batch_many_ctas(name):
    """4133 problems: the active list is compacted across hundreds of CTAs every iteration (warp-aggregated appends in
    arbitrary order) — the results do not depend on that order."""
    from bench import make_batch
    B, C, N = 4096 + 37, 3, 512
    p0, t, y = make_batch(B, C, N, 77, torch.device(DEV), torch.float64)
    w = torch.ones(C, dtype=torch.float64, device=DEV)
    emg = EMGModel()
    solver, kw = SOLVERS[name]
    res = _both(solver, kw, p0, emg.cost, emg.jac, (t, y), w)
    stop = _check_contract(res, kw['max_iter'])
    again = _both(solver, kw, p0, emg.cost, emg.jac, (t, y), w)      # list order differs run to run: same bits
    assert torch.equal(again['retire'][0], res['retire'][0])
    if name in ('gna', 'lma'):                                        # (Marquardt scaling stops after a few iterations everywhere, Q6)
        assert int(stop.max()) > int(stop[stop >= 0].min()), 'every problem stopped in the same iteration: nothing was retired early'


@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', ['gna', 'lma', 'gda'])
def test_retire_warp_per_problem_kernels_skew_gauss(name, dn):
    """single-cost problems (cfg-3 shape at test size): the warp-per-problem kernels walk the active list"""
    gold = load('skewgauss')
    dt = DT[dn]
    x, y, p0 = (torch.as_tensor(gold['P5_' + k]).to(dt).to(DEV) for k in ('x', 'y', 'p0'))
    sg = SkewGaussModel()
    solver, kw = SOLVERS[name]
    res = _both(solver, kw, p0.repeat(9, 1), sg.cost, sg.jac, (x, y.repeat(9, 1)), torch.ones(1, dtype=dt, device=DEV))
    _check_contract(res, kw['max_iter'])


@pytest.mark.parametrize('name', ['gna', 'lma', 'gda'])
def test_retire_tensor_core_fused_kernel_multi_peak(name):
    """P = 8 (two EMG peaks): the DMMA fused kernel and the warp-cooperative solve / update kernels (P > 8 is the same code)"""
    sys.path.insert(0, os.path.join(ROOT, 'scripts'))
    from bench_configs import multi_emg_batch
    for K in (2, 3):
        p0, t, y, _ = multi_emg_batch(300, K, 1000, torch.float64)
        model = MultiEMGModel(K)
        solver, kw = SOLVERS[name]
        res = _both(solver, dict(kw, max_iter=40), p0, model.cost, model.jac, (t, y), torch.ones(1, dtype=torch.float64, device=DEV))
        _check_contract(res, 40)


def test_stats_dict_in_exact_mode_does_not_change_the_trajectory():
    gold = load('fixtures')
    dt = torch.float64
    t, y, p0 = (torch.as_tensor(gold[k]).to(dt).to(DEV) for k in ('t', 'y', 'p0'))
    emg = EMGModel(n_costs=2)
    w = torch.ones(2, dtype=dt, device=DEV)
    for solver, kw in (SOLVERS['gna'], SOLVERS['lma'], SOLVERS['gda']):
        a = torch.stack(solver(p0.clone(), emg.cost, emg.jac, args=(t, y), wvec=w, **TOL, **kw))
        st = {}
        b = torch.stack(solver(p0.clone(), emg.cost, emg.jac, args=(t, y), wvec=w, stats=st, **TOL, **kw))
        assert torch.equal(a, b)
        assert st['iterations'] == a.shape[0] and st['stop_iteration'].shape == (6,) and not bool(st['nonfinite'].any())
        assert int(st['stop_iteration'].max()) <= a.shape[0] - 1
        if solver is lsq_gna_parallel:          # GN's own test is the reference's mask: the run ends when all fire together
            assert bool((st['stop_iteration'] >= 0).all()) and int(st['stop_iteration'].max()) == a.shape[0] - 1
        for it, bit in zip(st['stop_iteration'].tolist(), st['stop_reason'].tolist()):
            assert (bit & (_engine.STOP_G | _engine.STOP_P | _engine.STOP_F)) if it >= 0 else bit == 0


def test_stats_flag_non_finite_steps():
    """a NaN data sample makes every step of that problem non-finite: flagged, not raised (reported, not raising)"""
    gold = load('synthetic')
    dt = torch.float64
    t, y, p0, w = (torch.as_tensor(gold[k]).to(dt).to(DEV) for k in ('t', 'y', 'p0', 'wvec'))
    y = y.clone()
    y[3, 1, 17] = float('nan')
    emg = EMGModel()
    st = {}
    lsq_gna_parallel(p0.clone(), emg.cost, emg.jac, args=(t, y), wvec=w, stats=st, max_iter=5, **TOL)
    assert bool(st['nonfinite'][3]) and int(st['nonfinite'].sum()) == 1


def test_retire_rejects_exact_sharding_hooks():
    gold = load('fixtures')
    dt = torch.float64
    t, y, p0 = (torch.as_tensor(gold[k]).to(dt).to(DEV) for k in ('t', 'y', 'p0'))
    emg = EMGModel(n_costs=2)
    old = _engine.set_default_reduce_hook(lambda red: None)
    try:
        with pytest.raises(_lib.NllsError, match='retire'):
            lsq_gna_parallel(p0.clone(), emg.cost, emg.jac, args=(t, y), wvec=torch.ones(2, dtype=dt, device=DEV), mode='retire')
    finally:
        _engine.set_default_reduce_hook(old)
    with pytest.raises(ValueError):
        lsq_gna_parallel(p0.clone(), emg.cost, emg.jac, args=(t, y), wvec=torch.ones(2, dtype=dt, device=DEV), mode='freeze')
