"""Multi-peak EMG (SURVEY.md §8d large-P workload, cfg-4 family): nlls_model_eval materialises f and J for the sum of
2..4 EMG peaks (P = 8, 12, 16); the solvers run the tensor path (K1 on the fp64 tensor cores for P > 8, warp-per-problem
Cholesky K2).  Parity with the oracle on the same seeded inputs."""
import pytest
import torch

DEV = 'cuda:0'


def multi_emg_batch(B, K, N, seed, dt, C=1):
    from oracle import models as M
    g = torch.Generator().manual_seed(seed)
    t = torch.linspace(-1e3, 1e3, N).double()
    centres = torch.linspace(-600, 600, K, dtype=torch.float64) if K > 1 else torch.zeros(1, dtype=torch.float64)
    base = torch.stack([torch.tensor([10., 0., 60., 3.], dtype=torch.float64) + torch.tensor([0., float(c), 0., 0.], dtype=torch.float64)
                        for c in centres]).reshape(-1)
    gt = base + base.abs().clamp(min=20.) * .05 * (torch.rand(B, 4 * K, generator=g, dtype=torch.float64) - .5)
    p0 = gt + base.abs().clamp(min=20.) * .04 * (torch.rand(B, 4 * K, generator=g, dtype=torch.float64) - .5)
    y = M.multi_emg_model(gt, t)[:, None, :] + .01 * torch.randn(B, C, N, generator=g, dtype=torch.float64)
    return p0.to(dt), t.to(dt), y.to(dt), gt


@pytest.mark.gpu
@pytest.mark.parametrize('K', [2, 3, 4])
def test_multi_emg_eval_matches_the_oracle_callables(K):
    from oracle import models as M
    from torchimize_b200.models import MultiEMGModel
    for dt, tol in ((torch.float64, 1e-12), (torch.float32, 2e-4)):
        p0, t, y, _ = multi_emg_batch(5, K, 333, 2, dt, C=2)
        model = MultiEMGModel(K)
        f, J = model.cost_and_jac(p0.to(DEV), t.to(DEV), y.to(DEV))
        fr, Jr = M.multi_emg_cost(p0, t, y), M.multi_emg_jac(p0, t, y)
        assert f.shape == fr.shape and J.shape == Jr.shape == (5, 2, 333, 4 * K)
        assert float((f.cpu() - fr).abs().max()) <= tol * float(fr.abs().max())
        assert float((J.cpu() - Jr).abs().max()) <= tol * float(Jr.abs().max())


@pytest.mark.gpu
@pytest.mark.parametrize('solver', ['lma', 'gna', 'gda'])
@pytest.mark.parametrize('K', [2, 3, 4])
def test_multi_emg_solves_match_the_oracle_fp64(solver, K):
    from oracle import nlls_oracle as O, models as M
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel
    from torchimize_b200.models import MultiEMGModel
    from helpers import compare_plist
    p0, t, y, gt = multi_emg_batch(12, K, 1024, 5, torch.float64)
    kw = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=30)
    model = MultiEMGModel(K)
    from torchimize_b200.functions import gradient_descent_parallel
    ours = dict(lma=lsq_lma_parallel, gna=lsq_gna_parallel, gda=gradient_descent_parallel)[solver]
    ref = getattr(O, solver)(p0.clone(), M.multi_emg_cost, M.multi_emg_jac, args=(t, y), **kw)
    for fused in (True, False):                                   # tensor-core fused kernel / materialised J + K1
        p = p0.clone().to(DEV)
        got = ours(p, model.cost, model.jac, args=(t.to(DEV), y.to(DEV)), fused=fused, **kw)
        compare_plist(got, torch.stack(ref.p_list), 'f64', f'{K}-peak EMG {solver} fused={fused}')
        assert float(((p.cpu() - gt).abs() / gt.abs().clamp(min=1.)).max()) < 5e-2


@pytest.mark.gpu
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('C', [1, 3])
@pytest.mark.parametrize('K', [2, 3, 4])
def test_multi_emg_tensor_core_fused_normal_equations_match_the_oracle(K, C, dn):
    """nlls_model_normal_eq for P = 8|12|16 (k_model_normal_eq_mma: DMMA over the per-warp tile of k rows) against the
    oracle's newton step on the materialised J, two consecutive iterations (the second rebuilds f_prev from the stored
    model values m [B,N]), ragged N."""
    from oracle import nlls_oracle as O, models as M
    from torchimize_b200 import _lib
    dt = torch.float64 if dn == 'f64' else torch.float32
    tol = 1e-11 if dn == 'f64' else 3e-4
    B, N, P = 7, 1000 + 13, 4 * K
    p0, t, y, _ = multi_emg_batch(B, K, N, 9, dt, C=C)
    w = torch.linspace(1., 1.4, C, dtype=dt)
    pd, td, yd, wd = (v.to(DEV) for v in (p0, t, y, w))
    fpp = torch.zeros(2, B, N, dtype=dt, device=DEV)
    g = torch.empty(B, P, dtype=dt, device=DEV)
    H = torch.empty(B, P, P, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    ctrl = _lib.new_ctrl(DEV)
    f_prev = None
    for it, scale in enumerate((1., 1.01)):
        ctrl.view(torch.int32)[1] = it                            # ctrl->iters selects the ping-pong roles
        pi = (p0 * scale).to(DEV)
        _lib.call('nlls_model_normal_eq', _lib.MODEL_MULTI_EMG, _lib.dtype_code(pi), _lib.ptr(pi), P, _lib.ptr(td), _lib.ptr(yd),
                  C * N, N, _lib.ptr(wd), C, B, C, N, _lib.ptr(fpp[0]), _lib.ptr(fpp[1]), _lib.ptr(g), _lib.ptr(H), _lib.ptr(st),
                  _lib.ptr(ctrl), _lib.stream_ptr())
        f = M.multi_emg_cost(p0 * scale, t, y)
        J = M.multi_emg_jac(p0 * scale, t, y)
        g0, H0 = O.normal_equations(f, J, w)
        assert float((g.cpu() - g0).abs().max()) <= tol * float(g0.abs().max())
        assert float((H.cpu() - H0).abs().max()) <= tol * float(H0.abs().max())
        m0 = M.multi_emg_model(p0 * scale, t)
        assert float((fpp[it & 1].cpu() - m0).abs().max()) <= tol * float(m0.abs().max())
        assert torch.allclose(st[:, 0].cpu(), O.cost_sum_stat(f).double(), rtol=1e-9 if dn == 'f64' else 1e-3)
        assert torch.allclose(st[:, 2].cpu(), (f.double() ** 2).sum((1, 2)), rtol=1e-9 if dn == 'f64' else 1e-3)
        if f_prev is not None:
            assert torch.allclose(st[:, 1].cpu(), ((f_prev - f).double() ** 2).sum((1, 2)), rtol=1e-9 if dn == 'f64' else 1e-2)
        f_prev = f


@pytest.mark.gpu
def test_multi_emg_two_pass_lm_takes_the_tensor_path_and_the_abi_says_what_exists():
    from torchimize_b200 import _lib
    from torchimize_b200.models import MultiEMGModel, fused_model_of
    model = MultiEMGModel(4)
    for key in ('gn', 'lm', 'gd'):
        assert fused_model_of(model.cost, model.jac, key) is model
    assert fused_model_of(model.cost, model.jac, 'lm2') is None   # two-pass LM schedule: nlls_model_eval + K1
    p0, t, y, _ = multi_emg_batch(2, 4, 64, 1, torch.float64)
    g = torch.empty(2, 16, dtype=torch.float64, device=DEV)
    H = torch.empty(2, 16, 16, dtype=torch.float64, device=DEV)
    st = torch.empty(2, 4, dtype=torch.float64, device=DEV)
    w = torch.ones(1, dtype=torch.float64, device=DEV)
    pd, td, yd = p0.to(DEV), t.to(DEV), y.to(DEV)
    with pytest.raises(_lib.NllsError, match='no two-pass'):
        _lib.call('nlls_model_cost_stat', _lib.MODEL_MULTI_EMG, _lib.F64, _lib.ptr(pd), 16, _lib.ptr(td), _lib.ptr(yd), 64, 0,
                  2, 1, 64, _lib.ptr(st), None, _lib.stream_ptr())
