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
@pytest.mark.parametrize('solver', ['lma', 'gna'])
@pytest.mark.parametrize('K', [2, 3, 4])
def test_multi_emg_solves_match_the_oracle_fp64(solver, K):
    from oracle import nlls_oracle as O, models as M
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel
    from torchimize_b200.models import MultiEMGModel
    from helpers import compare_plist
    p0, t, y, gt = multi_emg_batch(12, K, 1024, 5, torch.float64)
    kw = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=30)
    model = MultiEMGModel(K)
    ours = lsq_lma_parallel if solver == 'lma' else lsq_gna_parallel
    ref = getattr(O, solver)(p0.clone(), M.multi_emg_cost, M.multi_emg_jac, args=(t, y), **kw)
    p = p0.clone().to(DEV)
    got = ours(p, model.cost, model.jac, args=(t.to(DEV), y.to(DEV)), **kw)
    compare_plist(got, torch.stack(ref.p_list), 'f64', f'{K}-peak EMG {solver}')
    assert float(((p.cpu() - gt).abs() / gt.abs().clamp(min=1.)).max()) < 5e-2


@pytest.mark.gpu
def test_multi_emg_has_no_fused_kernels_and_says_so():
    from torchimize_b200 import _lib
    from torchimize_b200.models import MultiEMGModel, fused_model_of
    model = MultiEMGModel(4)
    assert fused_model_of(model.cost, model.jac) is None          # solvers take the tensor path
    p0, t, y, _ = multi_emg_batch(2, 4, 64, 1, torch.float64)
    g = torch.empty(2, 16, dtype=torch.float64, device=DEV)
    H = torch.empty(2, 16, 16, dtype=torch.float64, device=DEV)
    st = torch.empty(2, 4, dtype=torch.float64, device=DEV)
    w = torch.ones(1, dtype=torch.float64, device=DEV)
    pd, td, yd = p0.to(DEV), t.to(DEV), y.to(DEV)
    with pytest.raises(_lib.NllsError, match='no fused'):
        _lib.call('nlls_model_normal_eq', _lib.MODEL_MULTI_EMG, _lib.F64, _lib.ptr(pd), 16, _lib.ptr(td), _lib.ptr(yd), 64, 0,
                  _lib.ptr(w), 1, 2, 1, 64, None, None, _lib.ptr(g), _lib.ptr(H), _lib.ptr(st), None, _lib.stream_ptr())
