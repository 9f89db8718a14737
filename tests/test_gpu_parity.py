"""GPU (-m gpu): parity of the CUDA path, called through the C ABI, against
  (1) the oracle on the same seeded inputs (kernel-level intermediates g, H, stats, f, J), and
  (2) the golden vectors the unmodified reference produced (tests/golden/*.npz): every
      per-iteration snapshot and the iteration count.
Tolerances (north_star): <= 1e-9 relative in fp64, <= 1e-4 in fp32; iteration counts identical
in fp64, boundary flips reported in fp32."""
import pytest
import torch

from helpers import CASES, PLAIN, DT, RTOL, TOL, CHAOTIC, load, compare_plist, compare_chaotic, rel_err
from oracle import nlls_oracle as O
from oracle import models as M

pytestmark = pytest.mark.gpu

from torchimize_b200 import _lib                                   # noqa: E402
from torchimize_b200.functions import (lsq_lma_parallel, lsq_gna_parallel, lsq_gna_parallel_plain,   # noqa: E402
                                       gradient_descent_parallel, gradient_descent_parallel_plain,
                                       newton_step_parallel, test_fun_dims_parallel)
from torchimize_b200.models import EMGModel, SkewGaussModel       # noqa: E402

DEV = 'cuda'
SOLVERS = {'lma': lsq_lma_parallel, 'gna': lsq_gna_parallel, 'gda': gradient_descent_parallel,
           'gna_plain': lsq_gna_parallel_plain, 'gda_plain': gradient_descent_parallel_plain}


def cu(a, dt):
    return torch.as_tensor(a).to(device=DEV, dtype=dt)


# ---------------------------------------------------------------------------- K1 kernels
def _rand_problem(B, C, N, P, dt, seed=0):
    g = torch.Generator().manual_seed(seed)
    J = torch.randn(B, C, N, P, generator=g, dtype=torch.float64).to(dt)
    f = torch.randn(B, C, N, generator=g, dtype=torch.float64).to(dt)
    fp = torch.randn(B, C, N, generator=g, dtype=torch.float64).to(dt)
    w = (torch.rand(C, generator=g, dtype=torch.float64) + .5).to(dt)
    return J, f, fp, w


def _run_k1(J, f, fp, w, force_direct=False):
    B, C, N, P = J.shape
    dt = J.dtype
    Jc, fc, wc = J.to(DEV).contiguous(), f.to(DEV).contiguous(), w.to(DEV)
    fpc = None if fp is None else fp.to(DEV).contiguous()
    g = torch.empty(B, P, dtype=dt, device=DEV)
    H = torch.empty(B, P, P, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    lib = _lib.load()
    old = lib.nlls_set_force_direct(1 if force_direct else 0)
    try:
        _lib.call('nlls_normal_eq', _lib.dtype_code(Jc), _lib.ptr(Jc), _lib.ptr(fc), _lib.ptr(fpc), _lib.ptr(wc), wc.numel(),
                  B, C, N, P, _lib.ptr(g), _lib.ptr(H), _lib.ptr(st), None, _lib.stream_ptr())
        torch.cuda.synchronize()
    finally:
        lib.nlls_set_force_direct(old)
    return g.cpu(), H.cpu(), st.cpu()


K1_SHAPES = [  # B, C, N, P
    (7, 2, 2000, 4), (5, 5, 1024, 4), (9, 1, 1000, 5), (3, 3, 130, 3), (4, 2, 257, 4), (6, 1, 64, 1),
    (3, 2, 96, 2), (5, 1, 512, 6), (4, 2, 300, 7), (3, 1, 1024, 8), (300, 1, 40, 4), (3, 1, 4096, 16),
    (2, 2, 200, 12), (2, 1, 100, 9), (2, 1, 128, 24), (1, 1, 7, 4), (2, 70, 16, 4),
    (2, 3, 333, 17), (2, 2, 250, 32), (3, 1, 1023, 16), (2, 2, 5, 10),
]


@pytest.mark.parametrize('force_direct', [False, True])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('shape', K1_SHAPES)
def test_normal_eq_matches_oracle(shape, dn, force_direct):
    import os
    if not force_direct and os.environ.get('NLLS_FORCE_DIRECT') == '1':
        pytest.skip('staged kernel disabled by NLLS_FORCE_DIRECT')
    dt = DT[dn]
    J, f, fp, w = _rand_problem(*shape, dt, seed=sum(shape))
    g, H, st = _run_k1(J, f, fp, w, force_direct)
    g0, H0 = O.normal_equations(f.double(), J.double(), w.double())     # fp64 truth of the same inputs
    tol = 1e-12 if dn == 'f64' else 2e-5
    scale_g = g0.abs().max()
    scale_H = H0.abs().max()
    assert (g.double() - g0).abs().max() <= tol * scale_g
    assert (H.double() - H0).abs().max() <= tol * scale_H
    assert torch.equal(H, H.transpose(1, 2)), 'H must be exactly symmetric'
    s0 = O.cost_sum_stat(f.double())
    s1 = ((fp.double() - f.double()) ** 2).sum((1, 2))
    s2 = (f.double() ** 2).sum((1, 2))
    stol = 1e-12 if dn == 'f64' else 1e-5
    assert torch.allclose(st[:, 0], s0, rtol=stol, atol=stol * float(s0.abs().max()))
    assert torch.allclose(st[:, 1], s1, rtol=stol)
    assert torch.allclose(st[:, 2], s2, rtol=stol)


def test_normal_eq_without_f_prev_and_broadcast_weight():
    J, f, _, _ = _rand_problem(5, 3, 256, 4, torch.float64, 1)
    w = torch.ones(1, dtype=torch.float64)                              # wvec=None path: ones(1) broadcast (lma:76)
    g, H, st = _run_k1(J, f, None, w)
    g0, H0 = O.normal_equations(f, J, w)
    assert rel_err(g, g0) < 1e-12 and rel_err(H.flatten(1), H0.flatten(1)) < 1e-12
    assert float(st[:, 1].abs().max()) == 0.0


@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('shape', [(6, 2, 500, 4), (4, 3, 1000, 5), (3, 1, 77, 3), (2, 2, 64, 12), (2, 3, 301, 16), (2, 1, 90, 25)])
def test_normal_eq_gda_matches_oracle(shape, dn):
    dt = DT[dn]
    B, C, N, P = shape
    J, f, _, w = _rand_problem(*shape, dt, seed=3)
    fsp = torch.randn(B, N, dtype=torch.float64).to(dt)
    Jc, fc, wc, fpc = J.to(DEV), f.to(DEV), w.to(DEV), fsp.to(DEV)
    g = torch.empty(B, P, dtype=torch.float64, device=DEV)
    H = torch.empty(B, P, P, dtype=torch.float64, device=DEV)
    fs = torch.empty(B, N, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    _lib.call('nlls_normal_eq_gda', _lib.dtype_code(Jc), _lib.ptr(Jc), _lib.ptr(fc), _lib.ptr(fpc), _lib.ptr(wc), C, B, C, N, P,
              _lib.ptr(g), _lib.ptr(H), _lib.ptr(fs), _lib.ptr(st), None, _lib.stream_ptr())
    f_s = torch.einsum('bcp,c->bp', f, w)                               # gda:136-137 in the input dtype
    j_s = torch.einsum('bcpi,c->bpi', J, w)
    H0 = torch.einsum('bnp,bnq->bpq', j_s.double(), j_s.double())
    g0 = torch.einsum('bnp,bn->bp', j_s.double(), f_s.double())
    tol = 1e-12 if dn == 'f64' else 1e-5
    assert (H.cpu() - H0).abs().max() <= tol * H0.abs().max()
    assert (g.cpu() - g0).abs().max() <= tol * g0.abs().max()
    assert torch.allclose(fs.cpu(), f_s, rtol=1e-12 if dn == 'f64' else 1e-5, atol=1e-6 if dn == 'f32' else 1e-14)
    assert torch.allclose(st[:, 2].cpu(), (f_s.double() ** 2).sum(1), rtol=1e-5 if dn == 'f32' else 1e-12)
    assert torch.allclose(st[:, 1].cpu(), ((fsp.double() - f_s.double()) ** 2).sum(1), rtol=1e-4 if dn == 'f32' else 1e-12)
    # the normal-equation solve equals the reference's tall lstsq (gda:139)
    h_ref = torch.linalg.lstsq(j_s.double(), f_s.double())[0]
    h_ne = torch.linalg.solve(H.cpu(), g.cpu())
    assert rel_err(h_ne, h_ref) < (1e-9 if dn == 'f64' else 1e-6)


def test_cost_stat():
    f = torch.randn(11, 3, 777, dtype=torch.float64)
    out = torch.empty(11, dtype=torch.float64, device=DEV)
    fc = f.to(DEV)
    _lib.call('nlls_cost_stat', _lib.F64, _lib.ptr(fc), 11, 3, 777, _lib.ptr(out), None, _lib.stream_ptr())
    assert torch.allclose(out.cpu(), O.cost_sum_stat(f), rtol=1e-12)


# ------------------------------------------------------------------------ model kernels
@pytest.mark.parametrize('dn', ['f64', 'f32'])
def test_emg_callables_match_oracle(dn):
    gold = load('synthetic')
    dt = DT[dn]
    p, t, y = cu(gold['p0'], dt), cu(gold['t'], dt), cu(gold['y'], dt)
    emg = EMGModel()
    f, J = emg.cost(p, t, y), emg.jac(p, t, y)
    f0 = M.emg_cost(p.cpu(), t.cpu(), y.cpu())
    J0 = M.emg_jac(p.cpu(), t.cpu(), y.cpu())
    tol = 1e-12 if dn == 'f64' else 2e-5
    assert (f.cpu() - f0).abs().max() <= tol * f0.abs().max()
    assert (J.cpu() - J0).abs().max() <= tol * J0.abs().max()
    # shared data row + n_costs (reference's multi_cost_batch layout)
    fx = load('fixtures')
    p, t, y = cu(fx['p0'], dt), cu(fx['t'], dt), cu(fx['y'], dt)
    emg2 = EMGModel(n_costs=2)
    assert emg2.cost(p, t, y).shape == (6, 2, 2000) and emg2.jac(p, t, y).shape == (6, 2, 2000, 4)
    assert (emg2.cost(p, t, y).cpu() - M.emg_cost(p.cpu(), t.cpu(), y.cpu(), n_costs=2)).abs().max() <= tol * 1e2


@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('P', [4, 5])
def test_skewgauss_callables_match_oracle(dn, P):
    gold = load('skewgauss')
    dt = DT[dn]
    pre = f'P{P}_'
    p, x, y = cu(gold[pre + 'p0'], dt), cu(gold[pre + 'x'], dt), cu(gold[pre + 'y'], dt)
    sg = SkewGaussModel()
    f, J = sg.cost(p, x, y), sg.jac(p, x, y)
    f0 = M.skewgauss_cost(p.cpu(), x.cpu(), y.cpu())
    J0 = M.skewgauss_jac(p.cpu(), x.cpu(), y.cpu())
    tol = 1e-12 if dn == 'f64' else 2e-5
    assert (f.cpu() - f0).abs().max() <= tol * f0.abs().max()
    assert (J.cpu() - J0).abs().max() <= tol * J0.abs().max()


@pytest.mark.parametrize('force_direct', [False, True])
@pytest.mark.parametrize('n_keep', [256, 250, 251, 37])
@pytest.mark.parametrize('shared_y', [False, True])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
def test_fused_normal_eq_matches_oracle(dn, shared_y, n_keep, force_direct):
    """K4 (bulk-copy staged and direct variants; ragged N falls back to direct by itself) on two
    consecutive iterations: the second rebuilds the first's f from the model values m [B,N] kept in the
    ping-pong buffers (f_prev = (y - m_prev)^2), so stats[1] has to equal the sum over stored residuals."""
    gold = load('synthetic')
    dt = DT[dn]
    lo = (gold['t'].shape[0] - n_keep) // 2                  # centred window: the peak sits mid-range
    p, t, w = cu(gold['p0'], dt), cu(gold['t'][lo:lo + n_keep], dt), cu(gold['wvec'], dt)
    y = cu(gold['y'][:, :, lo:lo + n_keep], dt).contiguous()
    B, C, N = y.shape
    if shared_y:
        y = y[:, 0].contiguous()
        ybs, ycs = N, 0
    else:
        ybs, ycs = C * N, N
    g = torch.empty(B, 4, dtype=dt, device=DEV)
    H = torch.empty(B, 4, 4, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    fpp = torch.zeros(2, B, N, dtype=dt, device=DEV)
    ctrl = _lib.new_ctrl(DEV)
    lib = _lib.load()
    old = lib.nlls_set_force_direct(1 if force_direct else 0)
    try:
        ps = [p, (p * 1.01).contiguous()]
        outs = []
        for it in range(2):
            ctrl.view(torch.int32)[1] = it
            _lib.call('nlls_model_normal_eq', _lib.MODEL_EMG, _lib.dtype_code(p), _lib.ptr(ps[it]), 4, _lib.ptr(t), _lib.ptr(y),
                      ybs, ycs, _lib.ptr(w), C, B, C, N, _lib.ptr(fpp[0]), _lib.ptr(fpp[1]), _lib.ptr(g), _lib.ptr(H),
                      _lib.ptr(st), _lib.ptr(ctrl), _lib.stream_ptr())
            torch.cuda.synchronize()
            outs.append((g.cpu().double(), H.cpu().double(), st.cpu().clone(), fpp[it].cpu().double()))
    finally:
        lib.nlls_set_force_direct(old)
    tc, yc, wc = t.cpu().double(), y.cpu().double(), w.cpu().double()
    tol = 1e-11 if dn == 'f64' else 5e-4
    f_last = None
    for it in range(2):
        pc = ps[it].cpu().double()
        f0, J0 = M.emg_cost(pc, tc, yc, n_costs=C if shared_y else None), M.emg_jac(pc, tc, yc, n_costs=C if shared_y else None)
        g0, H0 = O.normal_equations(f0, J0, wc)
        gg, HH, ss, mm = outs[it]
        ff = ((yc.reshape(B, -1, N) - mm[:, None, :]) ** 2).expand(B, C, N)   # residuals rebuilt from the stored model values
        assert (gg - g0).abs().max() <= tol * g0.abs().max()
        assert (HH - H0).abs().max() <= tol * H0.abs().max()
        rt = 1e-11 if dn == 'f64' else 1e-3
        assert torch.allclose(ss[:, 0], O.cost_sum_stat(f0), rtol=rt)
        assert torch.allclose(ss[:, 2], (f0 ** 2).sum((1, 2)), rtol=rt)
        assert torch.allclose(mm, M.emg_model(pc, tc), rtol=rt, atol=1e-12 if dn == 'f64' else 1e-6)
        assert torch.allclose(ff, f0, rtol=rt, atol=1e-12 if dn == 'f64' else 1e-6)
        if it == 1:
            assert torch.allclose(ss[:, 1], ((f_last - ff) ** 2).sum((1, 2)), rtol=1e-9 if dn == 'f64' else 1e-3)
        f_last = ff


# ------------------------------------------------------------------------------ solvers
def _torch_emg(n_costs):
    return (lambda p, t, y: M.emg_cost(p, t, y, n_costs=n_costs)), (lambda p, t, y: M.emg_jac(p, t, y, n_costs=n_costs))


def _solve_case(gold, name, dn, fun, jac, args, prefix='', chaotic=False, gset='', **extra):
    key = f'{prefix}{name}__{dn}'
    if key + '__raises' in gold:
        pytest.skip('reference raises on this input (Marquardt overflow, Q6)')
    dt = DT[dn]
    solver, kw = CASES[name]
    p = cu(gold[prefix + 'p0'], dt).clone()
    w = cu(gold.get('wvec', [1.]), dt)
    pl = SOLVERS[solver](p, fun, jac, args=args, wvec=w, **kw, **extra)
    assert torch.equal(pl[-1], p), 'caller p must end as the final iterate (Q1)'
    if chaotic:
        return compare_chaotic(pl, gold[key + '__plist'], dn, key)
    return compare_plist(pl, gold[key + '__plist'], dn, gset + key, allow_boundary=(dn == 'f32'), spread_key=gset + key)


PATHS = ['torch_callables', 'cuda_callables', 'fused']


def _emg_path(path, n_costs):
    if path == 'torch_callables':            # opaque user callables (torch ops on the GPU) -> tensor path
        fun, jac = _torch_emg(n_costs)
        return fun, jac, {}
    emg = EMGModel(n_costs=n_costs)
    return emg.cost, emg.jac, ({'fused': False} if path == 'cuda_callables' else {})


@pytest.mark.parametrize('path', PATHS)
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', list(CASES))
def test_fixture_cases_match_reference(name, dn, path):
    """cfg-1 of BASELINE.json: the reference's own test fixtures (B=6,C=2,N=2000,P=4)."""
    gold = load('fixtures')
    dt = DT[dn]
    fun, jac, extra = _emg_path(path, 2)
    _solve_case(gold, name, dn, fun, jac, (cu(gold['t'], dt), cu(gold['y'], dt)), **extra)


@pytest.mark.parametrize('path', PATHS)
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', list(CASES))
def test_synthetic_cases_match_reference(name, dn, path):
    gold = load('synthetic')
    dt = DT[dn]
    fun, jac, extra = _emg_path(path, None)
    _solve_case(gold, name, dn, fun, jac, (cu(gold['t'], dt), cu(gold['y'], dt)), chaotic=name in CHAOTIC['synthetic'], gset='synthetic:', **extra)


@pytest.mark.parametrize('path', PATHS)
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', list(CHAOTIC['synthetic']))
def test_synthetic_gda_every_step_teacher_forced(name, dn, path):
    """The chaotic GD cases (helpers.CHAOTIC): every step p_k = step(p_{k-1}) of the reference's
    trajectory, started from the reference's own p_{k-1}, all K-1 steps solved as ONE batch."""
    gold = load('synthetic')
    dt = DT[dn]
    _, kw = CASES[name]
    traj = torch.as_tensor(gold[f'{name}__{dn}__plist'])                  # [K,B,P]
    K, B, P = traj.shape
    starts = torch.cat([torch.as_tensor(gold['p0'])[None].to(traj.dtype), traj[:-1]]).reshape(K * B, P)
    y = torch.as_tensor(gold['y'])
    y_rep = y[None].expand(K, *y.shape).reshape(K * B, *y.shape[1:])
    fun, jac, extra = _emg_path(path, None)
    p = cu(starts, dt).clone()
    out = gradient_descent_parallel(p, fun, jac, args=(cu(gold['t'], dt), cu(y_rep, dt)), wvec=cu(gold['wvec'], dt),
                                    **{**kw, 'max_iter': 1}, **extra)
    assert len(out) == 1
    e = rel_err(out[0].cpu().reshape(K, B, P), traj)
    assert e <= RTOL[dn], f'{name}__{dn}: one-step rel err {e:.3e} > {RTOL[dn]:.0e}'


@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', ['lma_lev', 'gna', 'gda'])
def test_fixture_cases_fused_direct_kernels(name, dn):
    """Same as test_fixture_cases_match_reference[fused] but with the bulk-copy staged kernels switched off,
    so the direct-load fallbacks (unaligned shapes) stay covered at solver level."""
    gold = load('fixtures')
    dt = DT[dn]
    fun, jac, extra = _emg_path('fused', 2)
    lib = _lib.load()
    old = lib.nlls_set_force_direct(1)
    try:
        _solve_case(gold, name, dn, fun, jac, (cu(gold['t'], dt), cu(gold['y'], dt)), **extra)
    finally:
        lib.nlls_set_force_direct(old)


@pytest.mark.parametrize('path', ['torch_callables', 'fused'])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', ['lma_lev', 'gna', 'gda'])
@pytest.mark.parametrize('P', [4, 5])
def test_skewgauss_cases_match_reference(name, dn, P, path):
    """cfg-3 of BASELINE.json at test size (C=1, N=1000, P=4|5)."""
    gold = load('skewgauss')
    dt = DT[dn]
    pre = f'P{P}_'
    args = (cu(gold[pre + 'x'], dt), cu(gold[pre + 'y'], dt))
    if path == 'fused':
        sg = SkewGaussModel()
        fun, jac = sg.cost, sg.jac
    else:
        fun, jac = M.skewgauss_cost, M.skewgauss_jac
    _solve_case(gold, name, dn, fun, jac, args, prefix=pre)


@pytest.mark.parametrize('path', ['lambdas', 'bound_fused'])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('name', list(PLAIN))
def test_plain_variants(name, dn, path):
    """gna:94-117, gda:90-113.  'bound_fused': the callables of FusedModel.bind(t, y) run the fused kernels (no J)."""
    gold = load('fixtures')
    dt = DT[dn]
    t, y = cu(gold['t'], dt), cu(gold['y'], dt)
    emg = EMGModel(n_costs=2)
    solver, kw = PLAIN[name]
    p = cu(gold['p0'], dt).clone()
    if path == 'bound_fused':
        bm = emg.bind(t, y)
        fun, jac = bm.cost, bm.jac
        n0 = _lib.launches
    else:
        fun, jac = (lambda q: emg.cost(q, t, y)), (lambda q: emg.jac(q, t, y))
    res = SOLVERS[solver](p, fun, jac, torch.ones(2, dtype=dt, device=DEV), **kw)
    if path == 'bound_fused':                                       # 2 launches per iteration (GD: + none, with_masks = 0): no model_eval
        assert _lib.launches - n0 in (2 * kw['max_iter'], 3 * kw['max_iter'])
    assert res is p
    assert rel_err(res.cpu(), gold[f'{name}__{dn}__p']) <= (1e-10 if dn == 'f64' else 1e-4)


def test_bound_model_callables_take_the_fused_path_in_the_conditional_solvers():
    """FusedModel.bind(t, y): data bound to the callables, args=() — same kernels, same bits as args=(t, y)."""
    gold = load('synthetic')
    dt = torch.float64
    t, y, w = cu(gold['t'], dt), cu(gold['y'], dt), cu(gold['wvec'], dt)
    emg = EMGModel()
    bm = emg.bind(t, y)
    for solver, kw in ((lsq_gna_parallel, {}), (lsq_lma_parallel, dict(meth='lev')), (gradient_descent_parallel, {})):
        a = torch.stack(solver(cu(gold['p0'], dt).clone(), emg.cost, emg.jac, args=(t, y), wvec=w, max_iter=30, **kw))
        n0 = _lib.launches
        b = torch.stack(solver(cu(gold['p0'], dt).clone(), bm.cost, bm.jac, wvec=w, max_iter=30, **kw))
        assert torch.equal(a, b) and _lib.launches - n0 < 5 * 30 + 8


@pytest.mark.parametrize('dn', ['f64', 'f32'])
def test_newton_step_parallel(dn):
    gold = load('fixtures')
    dt = DT[dn]
    t, y = cu(gold['t'], dt), cu(gold['y'], dt)
    fun, jac = _torch_emg(2)
    p = cu(gold['p0'], dt)
    out = newton_step_parallel(p, lambda q: fun(q, t, y), lambda q: jac(q, t, y), torch.ones(2, dtype=dt, device=DEV))
    assert out[0] is p                                             # aliasing, newton_parallel.py:45
    tol = 1e-12 if dn == 'f64' else 1e-5
    assert rel_err(out[2].cpu(), gold[f'newton__{dn}__g']) <= tol
    assert rel_err(out[3].cpu().flatten(1), torch.as_tensor(gold[f'newton__{dn}__H']).flatten(1)) <= tol


def test_autograd_jacobian_path():
    """jac_function=None -> autograd Jacobian of opaque callables (functions/jacobian.py:36-57)."""
    gold = load('fixtures')
    dt = torch.float64
    fun, _ = _torch_emg(2)
    p = cu(gold['p0'], dt).clone()
    pl = lsq_lma_parallel(p, fun, None, args=(cu(gold['t'], dt), cu(gold['y'], dt)), wvec=torch.ones(2, dtype=dt, device=DEV),
                          ftol=1e-5, ptol=1e-5, gtol=1e-5, meth='lev', max_iter=99)
    compare_plist(pl, gold['lma_lev_autojac__f64__plist'], 'f64', 'lma_lev_autojac')


def test_quirks_q1_q4_wvec_none_and_noncontiguous_p():
    gold = load('fixtures')
    dt = torch.float64
    t, y = cu(gold['t'], dt), cu(gold['y'], dt)
    emg = EMGModel(n_costs=2)
    # Q4: l has no effect in lsq_gna_parallel
    pa, pb = cu(gold['p0'], dt).clone(), cu(gold['p0'], dt).clone()
    la = lsq_gna_parallel(pa, emg.cost, emg.jac, args=(t, y), l=1., max_iter=40)
    lb = lsq_gna_parallel(pb, emg.cost, emg.jac, args=(t, y), l=.3, max_iter=40)
    assert len(la) == len(lb) and torch.equal(la[-1], lb[-1])
    # wvec=None == ones (lma:76) ; reference golden used ones(2)
    compare_plist(la, gold['gna__f64__plist'], 'f64', 'gna wvec=None')
    # non-contiguous caller p is still updated in place (Q1)
    big = torch.zeros(6, 8, dtype=dt, device=DEV)
    view = big[:, ::2]
    view.copy_(cu(gold['p0'], dt))
    lv = lsq_lma_parallel(view, emg.cost, emg.jac, args=(t, y), max_iter=99)
    assert torch.equal(view, lv[-1])
    compare_plist(lv, gold['lma_lev__f64__plist'], 'f64', 'lma non-contiguous p')
    assert test_fun_dims_parallel(pa, emg.cost, emg.jac, args=(t, y))


def test_reference_own_assertions_hold():
    """tests/unit_test_parallel.py:229-260 with the import swapped to this package."""
    gold = load('fixtures')
    for dt, tol in ((torch.float64, 1e-8), (torch.float32, 1e-8)):
        t, y, gt = cu(gold['t'], dt), cu(gold['y'], dt), cu(gold['gt'], dt)
        emg = EMGModel(n_costs=2)
        p = cu(gold['p0'], dt).clone()
        coeffs = lsq_lma_parallel(p=p, function=emg.cost, jac_function=emg.jac, args=(t, y),
                                  wvec=torch.ones(2, device=DEV, dtype=dt), ftol=tol, ptol=tol, gtol=tol, meth='lev', max_iter=99)
        assert torch.allclose(coeffs[-1], gt, atol=1e-1)
        eps = torch.sum(emg.cost(coeffs[-1], t, y))
        assert eps.cpu() / 6 < 1
        assert len(coeffs) < 100


def test_large_p_cfg4_shape_against_oracle():
    """cfg-4 shape family (C=1, N=4096, P=16 fp64): linear least squares, LM and GN agree with the oracle."""
    torch.manual_seed(5)
    B, C, N, P = 8, 1, 4096, 16
    A = torch.randn(B, C, N, P, dtype=torch.float64)
    xt = torch.randn(B, P, dtype=torch.float64)
    yv = torch.einsum('bcnp,bp->bcn', A, xt) + .01 * torch.randn(B, C, N, dtype=torch.float64)
    for name in ('lma', 'gna', 'gda'):
        Ad, yd = A.to(DEV), yv.to(DEV)
        fun_d = lambda p: torch.einsum('bcnp,bp->bcn', Ad, p) - yd
        jac_d = lambda p: Ad
        fun_c = lambda p: torch.einsum('bcnp,bp->bcn', A, p) - yv
        jac_c = lambda p: A
        p_d = torch.zeros(B, P, dtype=torch.float64, device=DEV)
        p_c = torch.zeros(B, P, dtype=torch.float64)
        pl = SOLVERS[name](p_d, fun_d, jac_d, max_iter=30)
        tr = getattr(O, name)(p_c, fun_c, jac_c, max_iter=30)
        compare_plist(pl, torch.stack(tr.p_list), 'f64', f'largeP {name}')


@pytest.mark.gpu
@pytest.mark.parametrize('meth', ['lev', 'marq'])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
def test_speculative_lm_schedule_is_bit_identical_to_two_pass(dn, meth):
    """nlls_model_lm_trial + nlls_lm_update_spec (one model pass per iteration) against
    nlls_model_normal_eq + nlls_model_cost_stat + nlls_lm_update: same arithmetic in the same order,
    so every snapshot has to be bit-identical (incl. rejected steps and buffer-role flips)."""
    from torchimize_b200.functions.parallel import _engine
    gold = load('synthetic')
    dt = DT[dn]
    t, y, p0, w = (cu(gold[k], dt) for k in ('t', 'y', 'p0', 'wvec'))
    p0 = p0 * (1 + .3 * torch.sin(torch.arange(p0.numel(), device=DEV, dtype=dt).reshape(p0.shape)))   # provoke rejections
    emg = EMGModel()
    outs = []
    for spec in (True, False):
        old, _engine.SPECULATIVE_LM = _engine.SPECULATIVE_LM, spec
        try:
            p = p0.clone()
            outs.append(torch.stack(lsq_lma_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, meth=meth, max_iter=40, **TOL)))
        finally:
            _engine.SPECULATIVE_LM = old
    assert outs[0].shape == outs[1].shape
    same = (outs[0] == outs[1]) | (torch.isnan(outs[0]) & torch.isnan(outs[1]))
    assert bool(same.all())
    assert bool((outs[0][1:] == outs[0][:-1]).all(dim=2).any()), 'no rejected step in the trajectory: test is vacuous'


def test_fp64_exp_nonpos_accuracy():
    """csrc/models.cuh::exp_nonpos (the fp64 exp of every built-in and compiled model) against torch.exp over its whole
    range, through the C ABI: the compiled Gaussian + offset example with a = 1, c = 0, sigma = 1/sqrt(2), mu = 0 has
    m(x) = exp(-x^2), and with y = 0 the Jacobian column of the offset is 2 (y - m) * (-1) = 2 m (an exact doubling)."""
    from torchimize_b200.models import compile_model_file
    import os
    from conftest import ROOT
    model = compile_model_file(os.path.join(ROOT, 'torchimize_b200', 'user_models', 'gauss_offset.cuh'))
    x = torch.cat([torch.linspace(0, 26.5, 200001, dtype=torch.float64), torch.tensor([26.6, 27., 30., 1e3], dtype=torch.float64)])
    p = torch.tensor([[1., 0., 0.5 ** .5, 0.]], dtype=torch.float64)
    y = torch.zeros(1, x.numel(), dtype=torch.float64)
    got = model.jac(p.to(DEV), x.to(DEV), y.to(DEV))[0, 0, :, 3].cpu() / 2
    arg = -(x * x) * (0.5 / (p[0, 2] * p[0, 2]))                 # what load()/eval() form: -(d*d) * (0.5 / sigma^2)
    want = torch.exp(arg)
    big = want > 1e-300
    rel = ((got[big] - want[big]).abs() / want[big]).max()
    assert float(rel) <= 2.3e-16, f'exp_nonpos: {float(rel):.2e} relative'      # 1 ulp = 2.2e-16
    assert float(got[~big].max()) <= 1e-300                       # flushed tail
    assert bool(torch.isnan(model.cost(torch.tensor([[1., float('nan'), 1., 0.]], dtype=torch.float64, device=DEV),
                                       x[:8].to(DEV), y[:, :8].to(DEV))).all())
