"""K2/K3 entry points called directly through the C ABI (nlls_lm_solve, nlls_gn_update, nlls_gd_update)
against the oracle's fp64 `lstsq` (lma:98, gna:140, gda:139) for every parameter count the library
accepts: register Cholesky (P <= 8), warp-cooperative Cholesky (8 < P <= 32) and — with
nlls_set_force_direct(1) — the thread-per-problem shared-memory variant; LU fallback on indefinite
systems; singular flag; Marquardt scaling (lma:88) incl. its overflow."""
import ctypes

import pytest
import torch

from helpers import DT, rel_err
from oracle import nlls_oracle as O
from torchimize_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = 'cuda'
PS = [1, 2, 3, 4, 5, 8, 9, 12, 16, 17, 24, 32]


def _spd(B, P, seed, cond=50.):
    g = torch.Generator().manual_seed(seed)
    A = torch.randn(B, 3 * P + 4, P, generator=g, dtype=torch.float64)
    H = A.transpose(1, 2) @ A
    H = (H + H.transpose(1, 2)) / 2
    H += torch.eye(P, dtype=torch.float64) * (H.diagonal(dim1=1, dim2=2).mean() / cond)
    rhs = torch.randn(B, P, generator=g, dtype=torch.float64)
    p = torch.randn(B, P, generator=g, dtype=torch.float64)
    return H, rhs, p


def _ctrl():
    return torch.zeros(_lib.CTRL_BYTES // 8, dtype=torch.float64, device=DEV)


class _Force:
    def __init__(self, on):
        self.on = on

    def __enter__(self):
        self.old = _lib.load().nlls_set_force_direct(1 if self.on else 0)

    def __exit__(self, *a):
        _lib.load().nlls_set_force_direct(self.old)


@pytest.mark.parametrize('force_direct', [False, True])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('P', PS)
def test_lm_solve_matches_lstsq(P, dn, force_direct):
    dt = DT[dn]
    B = 67
    H, rhs, p = _spd(B, P, 10 + P)
    H, rhs, p = H.to(dt), rhs.to(dt), p.to(dt)
    u = (torch.rand(B, dtype=torch.float64) * 1e-2 + 1e-3).to(dt)
    for marq in (0, 1):
        ds = torch.zeros(B, 2, dtype=dt)
        ds[:, 0] = 1.5
        Hd, gd, pd, ud, dsd = (x.to(DEV).contiguous() for x in (H, rhs, p, u, ds))
        h = torch.empty(B, P, dtype=dt, device=DEV)
        pt = torch.empty(B, P, dtype=dt, device=DEV)
        ctrl = _ctrl()
        with _Force(force_direct):
            _lib.call('nlls_lm_solve', _lib.dtype_code(Hd), _lib.ptr(gd), _lib.ptr(Hd), _lib.ptr(pd), _lib.ptr(ud), _lib.ptr(dsd),
                      marq, B, P, _lib.ptr(h), _lib.ptr(pt), _lib.ptr(ctrl), _lib.stream_ptr())
        torch.cuda.synchronize()
        # oracle: lma:88,96-98
        scale = ds[:, 0].clone()
        if marq:
            diagH = H.diagonal(dim1=1, dim2=2)
            scale = scale * torch.max(torch.maximum(diagH, scale[:, None].expand(-1, P)), dim=1)[0]
        Hu = H + (u * scale)[:, None, None] * torch.eye(P, dtype=dt)[None]
        want = -O.solve_fp64(Hu, rhs, dt)
        tol = 1e-10 if dn == 'f64' else 2e-5
        assert rel_err(h.cpu(), want) <= tol, f'P={P} marq={marq}'
        assert torch.equal(pt.cpu(), p + h.cpu())
        assert torch.allclose(dsd.cpu()[:, 0], scale, rtol=1e-6 if dn == 'f32' else 1e-14)
        assert int(ctrl.view(torch.int32)[2]) == 0


@pytest.mark.parametrize('force_direct', [False, True])
@pytest.mark.parametrize('dn', ['f64', 'f32'])
@pytest.mark.parametrize('P', PS)
@pytest.mark.parametrize('solver', ['gn', 'gd'])
def test_gn_gd_update_match_oracle(solver, P, dn, force_direct):
    dt = DT[dn]
    B, l = 45, .75
    H, rhs, p = _spd(B, P, 20 + P)
    tH = dt if solver == 'gn' else torch.float64              # GD hands fp64 normal equations (gda:139)
    H, rhs, p = H.to(tH), rhs.to(tH), p.to(dt)
    stats = torch.zeros(B, 4, dtype=torch.float64)
    stats[:, 1] = 1.0
    stats[:, 2] = 1.0                                          # fcon false
    Hd, gd, pd, sd = (x.to(DEV).contiguous() for x in (H, rhs, p, stats))
    hout = torch.empty(B, P, dtype=dt, device=DEV)
    hist = torch.full((1, B, P), float('nan'), dtype=dt, device=DEV)
    ctrl = _ctrl()
    gtol, ptol = 0.3, 1e-3                                     # gtol chosen so that some problems pass the gradient test
    with _Force(force_direct):
        _lib.call('nlls_gn_update' if solver == 'gn' else 'nlls_gd_update', _lib.dtype_code(pd), _lib.ptr(pd), _lib.ptr(gd),
                  _lib.ptr(Hd), _lib.ptr(sd), l, 1e-8, ptol, gtol, 1, B, P, _lib.ptr(hout), _lib.ptr(hist), _lib.ptr(ctrl),
                  _lib.stream_ptr())
    torch.cuda.synchronize()
    h = O.solve_fp64(H, rhs, dt)
    p_new = p - l * h
    tol = 1e-10 if dn == 'f64' else 2e-5
    assert rel_err(hout.cpu(), h) <= tol
    assert rel_err(pd.cpu(), p_new) <= tol
    assert torch.equal(hist[0], pd)
    g_test = (rhs.to(dt) if solver == 'gn' else h / l)
    gcon = g_test.abs().max(dim=1)[0] < gtol
    pcon = h.norm(dim=1) < ptol * (ptol + p_new.norm(dim=1))
    margin = ((g_test.abs().max(dim=1)[0] - gtol).abs() > 1e-4)   # ignore problems sitting on the threshold
    want_nc = int((~(gcon | pcon))[margin].sum())
    got_nc = float(ctrl[2])                                     # red[0]
    assert abs(got_nc - want_nc) <= int((~margin).sum()), (got_nc, want_nc)
    assert int(ctrl.view(torch.int32)[2]) == 0


@pytest.mark.parametrize('force_direct', [False, True])
@pytest.mark.parametrize('P', [4, 12, 16, 32])
def test_indefinite_system_takes_lu_and_singular_sets_flag(P, force_direct):
    dt = torch.float64
    B = 9
    H, rhs, p = _spd(B, P, 30 + P)
    H[::2, 0, 0] = -H[::2, 0, 0]                               # indefinite but non-singular: Cholesky must hand over to LU
    u = torch.zeros(B, dtype=dt)
    ds = torch.zeros(B, 2, dtype=dt)
    ds[:, 0] = 1
    for singular in (False, True):
        Hc = H.clone()
        if singular:
            Hc[3] = 0.0
        Hd, gd, pd, ud, dsd = (x.to(DEV).contiguous() for x in (Hc, rhs, p, u, ds))
        h = torch.empty(B, P, dtype=dt, device=DEV)
        pt = torch.empty(B, P, dtype=dt, device=DEV)
        ctrl = _ctrl()
        with _Force(force_direct):
            _lib.call('nlls_lm_solve', _lib.F64, _lib.ptr(gd), _lib.ptr(Hd), _lib.ptr(pd), _lib.ptr(ud), _lib.ptr(dsd), 0, B, P,
                      _lib.ptr(h), _lib.ptr(pt), _lib.ptr(ctrl), _lib.stream_ptr())
        torch.cuda.synchronize()
        keep = torch.ones(B, dtype=torch.bool)
        if singular:
            keep[3] = False
        want = -O.solve_fp64(Hc[keep], rhs[keep], dt)
        assert rel_err(h.cpu()[keep], want) <= 1e-9
        assert int(ctrl.view(torch.int32)[2]) == (1 if singular else 0)
        if singular:
            assert not torch.isfinite(h.cpu()[3]).all()        # propagates inf/NaN instead of trapping


@pytest.mark.parametrize('P', [4, 16])
def test_marquardt_scale_overflow_propagates_like_the_reference(P):
    """Q6: the compounding scale overflows to inf, its off-diagonal zeros turn NaN (0*inf); the kernel has to
    propagate that instead of trapping, and flag the solve."""
    dt = torch.float32
    B = 5
    H, rhs, p = _spd(B, P, 40 + P)
    H, rhs, p = H.to(dt), rhs.to(dt), p.to(dt)
    u = torch.full((B,), 1e-3, dtype=dt)
    ds = torch.zeros(B, 2, dtype=dt)
    ds[:, 0] = float('inf')
    Hd, gd, pd, ud, dsd = (x.to(DEV).contiguous() for x in (H, rhs, p, u, ds))
    h = torch.empty(B, P, dtype=dt, device=DEV)
    pt = torch.empty(B, P, dtype=dt, device=DEV)
    ctrl = _ctrl()
    _lib.call('nlls_lm_solve', _lib.F32, _lib.ptr(gd), _lib.ptr(Hd), _lib.ptr(pd), _lib.ptr(ud), _lib.ptr(dsd), 1, B, P,
              _lib.ptr(h), _lib.ptr(pt), _lib.ptr(ctrl), _lib.stream_ptr())
    torch.cuda.synchronize()
    out = dsd.cpu()
    assert torch.isinf(out[:, 0]).all() and torch.isnan(out[:, 1]).all()
    assert not torch.isfinite(h.cpu()).any()
    assert int(ctrl.view(torch.int32)[2]) == 1
