"""Parity at BASELINE.json's FULL sizes through size-independent properties (the oracle cannot run these
sizes in seconds):
  cfg-3  B=262144, C=1, N=1000, P=5   the committed golden skewed-Gaussian batch (B=32) tiled 8192x: every replica
                                      must be bit-identical to replica 0, replica 0 must match the reference's
                                      golden trajectory, and the iteration count must be the reference's
  cfg-2  B=65536, C=5, N=1024, P=4    fits recover the ground truth (the reference's own assertions); a slice solved
                                      alone for the same number of iterations and a permuted batch give identical bits
  cfg-4  B=32768, C=1, N=4096, P=16   K1 (fp64 tensor-core path) against the oracle on sampled problems, exact
                                      linearity in the weight, slice consistency; K2 residual of every system
  cfg-5  B=1 Mi, C=5, N=1024, P=4     the committed cfg-5 shaped golden batch (B=16, from the unmodified reference) tiled
                                      65536x, fp32 and fp64, LM and GN: replicas bit-identical, replica 0 on the golden
                                      trajectory with the reference's iteration count
AND, at the real cfg-2 / cfg-4 batches, a 64-problem SAMPLE of the very batch pushed through the oracle for the same K
iterations with tolerances 0 (per-problem arithmetic is batch-independent — the slice tests prove it), every snapshot
held to 1e-9 against full[:, idx].
"""
import os
import sys

import pytest
import torch

from helpers import CASES, CASES5, CHAOTIC_PREFIX, DT, RTOL, load, compare_plist, rel_err
from oracle import nlls_oracle as O
from oracle import models as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from torchimize_b200 import _lib                                   # noqa: E402
from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel   # noqa: E402
from torchimize_b200.models import EMGModel, SkewGaussModel       # noqa: E402

pytestmark = pytest.mark.gpu
DEV = 'cuda'
SOLVERS = dict(lma=lsq_lma_parallel, gna=lsq_gna_parallel, gda=gradient_descent_parallel)


# ------------------------------------------------------------------------------------------ cfg-3
@pytest.mark.parametrize('dn', ['f32', 'f64'])
@pytest.mark.parametrize('name', ['lma_lev', 'gna', 'gda'])
def test_cfg3_full_size_tiled_golden(name, dn):
    gold = load('skewgauss')
    dt = DT[dn]
    reps = 8192                                                      # 32 * 8192 = 262144 problems (BASELINE configs[2])
    x = torch.as_tensor(gold['P5_x']).to(dt).to(DEV)
    y = torch.as_tensor(gold['P5_y']).to(dt).to(DEV).repeat(reps, 1)
    p = torch.as_tensor(gold['P5_p0']).to(dt).to(DEV).repeat(reps, 1)
    assert p.shape == (262144, 5) and y.shape == (262144, 1000)
    solver_name, kw = CASES[name]
    m = SkewGaussModel()
    got = torch.stack(SOLVERS[solver_name](p, m.cost, m.jac, args=(x, y), wvec=torch.ones(1, dtype=dt, device=DEV), **kw))
    want = torch.as_tensor(gold[f'P5_{name}__{dn}__plist'])
    K = got.shape[0]
    blocks = got.view(K, reps, 32, 5)
    same = (blocks == blocks[:, :1]) | (torch.isnan(blocks) & torch.isnan(blocks[:, :1]))
    assert bool(same.all()), 'replicas of the same problem diverged'
    compare_plist(blocks[:, 0], want, dn, f'cfg3 full size {name} {dn}', allow_boundary=(dn == 'f32'), spread_key=f'P5_{name}__{dn}')


# ------------------------------------------------------------------------------------------ cfg-2
@pytest.fixture(scope='module')
def cfg2():
    from bench import make_batch
    B, C, N = 65536, 5, 1024
    p0, t, y = make_batch(B, C, N, 3008, torch.device(DEV), torch.float64)
    return p0, t, y


def _gn(p0, t, y, max_iter=80, tol=1e-8, fused=True):
    emg = EMGModel()
    p = p0.clone()
    w = torch.ones(y.shape[1], dtype=p.dtype, device=DEV)
    out = lsq_gna_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, ftol=tol, ptol=tol, gtol=tol, max_iter=max_iter, fused=fused)
    return torch.stack(out)


def test_cfg2_full_size_fit_quality_slice_and_permutation(cfg2):
    p0, t, y = cfg2
    full = _gn(p0, t, y)
    K = full.shape[0]
    assert 5 < K < 80
    # the reference's own assertions (tests/unit_test_parallel.py:221-227): close to the ground truth, small cost
    gen = torch.Generator(device=DEV).manual_seed(3008)           # first draw of bench.make_batch = the ground truth
    gt = torch.tensor([10., -1., 80., 5.], dtype=torch.float64, device=DEV) * (
        1 + .1 * (torch.rand(p0.shape[0], 4, generator=gen, dtype=torch.float64, device=DEV) - .5))
    assert torch.allclose(full[-1], gt, atol=1e-1), f'max |p - gt| = {float((full[-1] - gt).abs().max()):.3e}'
    emg = EMGModel()
    cost = emg.cost(full[-1].contiguous(), t, y).sum((1, 2))
    assert float(cost.max()) < 1.0
    # a slice iterated alone for the same K steps (tolerances 0: it never stops early) gives the same bits
    sl = slice(1000, 1000 + 4096)
    alone = _gn(p0[sl], t, y[sl], max_iter=K, tol=0.)
    assert alone.shape[0] == K and torch.equal(alone, full[:, sl])
    # permutation equivariance (batch-global termination is order independent for GN)
    perm = torch.randperm(p0.shape[0], device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
    permuted = _gn(p0[perm], t, y[perm])
    assert permuted.shape[0] == K and torch.equal(permuted, full[:, perm])


def _sample_idx(B, n=64, seed=5):
    return torch.randperm(B, generator=torch.Generator().manual_seed(seed))[:n].sort().values


@pytest.mark.parametrize('solver', ['gna', 'lma', 'gda'])
def test_cfg2_full_size_sampled_oracle(cfg2, solver):
    """64 problems of the REAL cfg-2 batch through the oracle (tolerances 0, max_iter = the K the full batch ran):
    every snapshot of those problems in the full solve within 1e-9 (gna:71-89, lma:92-119, gda:69-85)."""
    p0, t, y = cfg2
    emg = EMGModel()
    w = torch.ones(y.shape[1], dtype=p0.dtype, device=DEV)
    kw = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8)
    p = p0.clone()
    if solver == 'lma':
        full = torch.stack(lsq_lma_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, meth='lev', max_iter=40, **kw))
    elif solver == 'gda':
        full = torch.stack(gradient_descent_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, l=1., max_iter=CHAOTIC_PREFIX, **kw))   # the reference's GD map is chaotic on this kind of batch (helpers.CHAOTIC): first 16 iterations
    else:
        full = torch.stack(lsq_gna_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, max_iter=80, **kw))
    K = full.shape[0]
    idx = _sample_idx(p0.shape[0])
    ps, ts, ys, ws = p0[idx.to(DEV)].cpu(), t.cpu(), y[idx.to(DEV)].cpu(), w.cpu()
    zero = dict(ftol=0., ptol=0., gtol=0., max_iter=K)
    if solver == 'lma':
        ref = O.lma(ps, M.emg_cost, M.emg_jac, args=(ts, ys), wvec=ws, meth='lev', **zero).p_list
    elif solver == 'gda':
        ref = O.gda(ps, M.emg_cost, M.emg_jac, args=(ts, ys), wvec=ws, l=1., **zero).p_list
    else:
        ref = O.gna(ps, M.emg_cost, M.emg_jac, args=(ts, ys), wvec=ws, **zero).p_list
    assert len(ref) == K
    e = rel_err(full[:, idx.to(DEV)].cpu(), torch.stack(ref))
    assert e <= RTOL['f64'], f'cfg-2 {solver}: sampled-oracle rel err {e:.3e} over {K} snapshots'


def test_cfg2_full_size_tensor_path_equals_fused_path(cfg2):
    """Materialised J (10.7 GB) through K1 against the fused kernels: same model code, different summation order."""
    p0, t, y = cfg2
    a = _gn(p0, t, y, fused=True)
    b = _gn(p0, t, y, fused=False)
    assert a.shape == b.shape
    assert rel_err(a[-1], b[-1]) <= 1e-9


# ------------------------------------------------------------------------------------------ cfg-4
def test_cfg4_full_size_k1_k2_properties():
    B, C, N, P = 32768, 1, 4096, 16
    dt = torch.float64
    g = torch.Generator(device=DEV).manual_seed(4)
    J = torch.randn(B, C, N, P, generator=g, dtype=dt, device=DEV)
    f = torch.randn(B, C, N, generator=g, dtype=dt, device=DEV)
    fp = torch.randn(B, C, N, generator=g, dtype=dt, device=DEV)

    def k1(Jx, fx, fpx, wval):
        Bx = Jx.shape[0]
        w = torch.full((1,), wval, dtype=dt, device=DEV)
        gg = torch.empty(Bx, P, dtype=dt, device=DEV)
        H = torch.empty(Bx, P, P, dtype=dt, device=DEV)
        st = torch.empty(Bx, 4, dtype=torch.float64, device=DEV)
        _lib.call('nlls_normal_eq', _lib.F64, _lib.ptr(Jx), _lib.ptr(fx), _lib.ptr(fpx), _lib.ptr(w), 1, Bx, C, N, P, _lib.ptr(gg),
                  _lib.ptr(H), _lib.ptr(st), None, _lib.stream_ptr())
        return gg, H, st

    g1, H1, s1 = k1(J, f, fp, 1.0)
    # sampled problems against the oracle (CPU fp64)
    idx = torch.arange(0, B, 997, device=DEV)
    g0, H0 = O.normal_equations(f[idx].cpu(), J[idx].cpu(), torch.ones(1, dtype=dt))
    assert (g1[idx].cpu() - g0).abs().max() <= 1e-12 * g0.abs().max()
    assert (H1[idx].cpu() - H0).abs().max() <= 1e-12 * H0.abs().max()
    assert torch.allclose(s1[idx, 0].cpu(), O.cost_sum_stat(f[idx].cpu()), rtol=1e-11, atol=1e-9)
    assert torch.allclose(s1[idx, 1].cpu(), ((fp[idx] - f[idx]) ** 2).sum((1, 2)).cpu(), rtol=1e-12)
    assert torch.equal(H1, H1.transpose(1, 2))
    # exact linearity in a power-of-two weight, slice consistency
    g2, H2, _ = k1(J, f, fp, 2.0)
    assert torch.equal(g2, 2 * g1) and torch.equal(H2, 2 * H1)
    sl = slice(5000, 5000 + 300)
    g3, H3, s3 = k1(J[sl], f[sl], fp[sl], 1.0)
    assert torch.equal(g3, g1[sl]) and torch.equal(H3, H1[sl]) and torch.equal(s3, s1[sl])
    # K2: every damped system is solved (lma:96-98): || (H + u I) h + g || small relative to || g ||
    p = torch.zeros(B, P, dtype=dt, device=DEV)
    u = torch.full((B,), 1e-3, dtype=dt, device=DEV)
    ds = torch.zeros(B, 2, dtype=dt, device=DEV)
    ds[:, 0] = 1
    h = torch.empty(B, P, dtype=dt, device=DEV)
    pt = torch.empty(B, P, dtype=dt, device=DEV)
    ctrl = torch.zeros(_lib.CTRL_BYTES // 8, dtype=torch.float64, device=DEV)
    _lib.call('nlls_lm_solve', _lib.F64, _lib.ptr(g1), _lib.ptr(H1), _lib.ptr(p), _lib.ptr(u), _lib.ptr(ds), 0, B, P, _lib.ptr(h),
              _lib.ptr(pt), _lib.ptr(ctrl), _lib.stream_ptr())
    Hu = H1 + 1e-3 * torch.eye(P, dtype=dt, device=DEV)
    res = torch.einsum('bij,bj->bi', Hu, h) + g1
    assert float((res.norm(dim=1) / g1.norm(dim=1)).max()) <= 1e-12
    assert torch.equal(pt, p + h) and int(ctrl.view(torch.int32)[2]) == 0


def test_cfg4_full_size_end_to_end_fused_tensor_core_path():
    """BASELINE.json configs[3] end to end: lsq_lma_parallel / lsq_gna_parallel on B=32768 four-peak EMG fits
    (C=1, N=4096, P=16, fp64) through the tensor-core fused kernel (k_model_normal_eq_mma + warp Cholesky K2):
    fit quality against the ground truth, the slice property, and agreement with the materialised-J path (K1 DMMA)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'scripts'))
    from bench_configs import multi_emg_batch
    from torchimize_b200.functions import lsq_lma_parallel
    from torchimize_b200.models import MultiEMGModel
    B, K, N = 32768, 4, 4096
    p0, t, y, gt = multi_emg_batch(B, K, N, torch.float64)
    model = MultiEMGModel(K)

    def run(solver, ps, ys, fused=True, **kw):
        p = ps.clone()
        return torch.stack(solver(p, model.cost, model.jac, args=(t, ys), fused=fused, **kw))

    gn = run(lsq_gna_parallel, p0, y, ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=40)
    lm = run(lsq_lma_parallel, p0, y, ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=40, meth='lev')
    scale = gt.abs().clamp(min=1.)
    for name, out in (('gn', gn), ('lm', lm)):
        assert 5 < out.shape[0] < 40, f'{name}: {out.shape[0]} iterations'
        assert float(((out[-1] - gt).abs() / scale).max()) < 5e-2, name
        assert float(model.cost(out[-1].contiguous(), t, y).sum((1, 2)).max()) < 4.0   # ~ N * noise^2 * kurtosis-ish, well below 1e-2 per sample
    assert rel_err(gn[-1], lm[-1]) <= 1e-3                        # both stop near the same least-squares optimum (stop tests at 1e-8 on different trajectories)
    # 64 problems of this very batch through the oracle (tolerances 0, same number of iterations): every snapshot <= 1e-9
    idx = _sample_idx(B)
    ps, ts, ys = p0[idx.to(DEV)].cpu(), t.cpu(), y[idx.to(DEV)].cpu()
    for name, out, orc, extra in (('gn', gn, O.gna, {}), ('lm', lm, O.lma, dict(meth='lev'))):
        ref = orc(ps.clone(), M.multi_emg_cost, M.multi_emg_jac, args=(ts, ys), wvec=torch.ones(1, dtype=torch.float64),
                  ftol=0., ptol=0., gtol=0., max_iter=out.shape[0], **extra).p_list
        e = rel_err(out[:, idx.to(DEV)].cpu(), torch.stack(ref))
        assert e <= RTOL['f64'], f'cfg-4 {name}: sampled-oracle rel err {e:.3e} over {out.shape[0]} snapshots'
    # a slice iterated alone for the same number of steps (tolerances 0) gives the same bits
    Kg = gn.shape[0]
    sl = slice(777, 777 + 2048)
    alone = run(lsq_gna_parallel, p0[sl], y[sl], ftol=0., ptol=0., gtol=0., max_iter=Kg)
    assert alone.shape[0] == Kg and torch.equal(alone, gn[:, sl])
    # the materialised-J path (nlls_model_eval -> K1 on the fp64 tensor cores) on that slice
    tens = run(lsq_gna_parallel, p0[sl], y[sl], fused=False, ftol=0., ptol=0., gtol=0., max_iter=Kg)
    assert rel_err(tens[-1], alone[-1]) <= 1e-9


# ------------------------------------------------------------------------------------------ cfg-5
@pytest.mark.parametrize('dn', ['f32', 'f64'])
@pytest.mark.parametrize('name', list(CASES5))
def test_cfg5_shape_tiled_golden_1mi_problems(name, dn):
    """BASELINE.json configs[4] shape at B = 1 Mi problems resident on one GPU (the 4 Mi batch is 4 such shards on 4+ GPUs;
    a shard is what a rank solves): the committed 16-problem golden batch of tests/golden/make_golden_cfg5.py (unmodified
    reference, C=5, N=1024, P=4) tiled 65536x."""
    gold = load('cfg5')
    dt = DT[dn]
    reps = 65536
    nb = gold['p0'].shape[0]
    t = torch.as_tensor(gold['t']).to(dt).to(DEV)
    y = torch.as_tensor(gold['y']).to(dt).to(DEV).repeat(reps, 1, 1)
    p = torch.as_tensor(gold['p0']).to(dt).to(DEV).repeat(reps, 1)
    assert p.shape == (1 << 20, 4) and y.shape == (1 << 20, 5, 1024)
    solver_name, kw = CASES5[name]
    emg = EMGModel()
    got = torch.stack(SOLVERS[solver_name](p, emg.cost, emg.jac, args=(t, y), wvec=torch.ones(5, dtype=dt, device=DEV), **kw))
    del y
    K = got.shape[0]
    blocks = got.view(K, reps, nb, 4)
    same = (blocks == blocks[:, :1]) | (torch.isnan(blocks) & torch.isnan(blocks[:, :1]))
    assert bool(same.all()), 'replicas of the same problem diverged'
    compare_plist(blocks[:, 0], gold[f'{name}__{dn}__plist'], dn, f'cfg5 1Mi {name} {dn}', allow_boundary=(dn == 'f32'), spread_key=f'cfg5 {name}__{dn}')
