#!/usr/bin/env python
"""Secondary measurements (not the bench.py contract): every BASELINE.json config shape that fits
one GPU, per-kernel CUDA-event times and achieved algorithmic GB/s.  One JSON line per case.

    python scripts/bench_configs.py [case ...]     cases: cfg2-gn cfg2-lm cfg2-gd cfg2-gn-f32 cfg3-gd cfg3-lm cfg4-k1 cfg4-lm cfg4-gn
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                                        # noqa: E402
from bench import make_batch                                        # noqa: E402
from torchimize_b200 import _lib                                    # noqa: E402
from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel, gradient_descent_parallel   # noqa: E402
from torchimize_b200.models import EMGModel, SkewGaussModel, MultiEMGModel   # noqa: E402

DEV = 'cuda'
PEAK = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))).get('hbm_gbs', 6650.0)) if os.path.exists(
    os.path.join(ROOT, 'MEASURED_PEAKS.json')) else 6650.0
SOLVERS = {'gn': lsq_gna_parallel, 'lm': lsq_lma_parallel, 'gd': gradient_descent_parallel}


def skew_batch(B, N, P, dt, seed=3006):
    g = torch.Generator(device=DEV).manual_seed(seed)
    x = torch.linspace(-5, 5, N, device=DEV, dtype=torch.float64)
    gt = torch.tensor([10., -1., 2., 5.], dtype=torch.float64, device=DEV) * (
        1 + .1 * (torch.rand(B, 4, generator=g, dtype=torch.float64, device=DEV) - .5))
    p0 = torch.tensor([7.5, -.5, .5, 2.5], dtype=torch.float64, device=DEV) * (
        1 + .1 * (torch.rand(B, 4, generator=g, dtype=torch.float64, device=DEV) - .5))
    if P == 5:
        gt = torch.cat([gt, torch.zeros(B, 1, dtype=torch.float64, device=DEV)], 1)
        p0 = torch.cat([p0, torch.full((B, 1), .1, dtype=torch.float64, device=DEV)], 1)
    y = torch.empty(B, N, dtype=dt, device=DEV)
    for b0 in range(0, B, 16384):
        q = gt[b0:b0 + 16384]
        d = x - q[:, 1, None]
        m = q[:, 0, None] / (q[:, 2, None] * (2 * torch.pi) ** .5) * torch.exp(-d ** 2 / (2 * q[:, 2, None] ** 2)) * (
            1 + torch.erf(q[:, 3, None] * d / (q[:, 2, None] * 2 ** .5)))
        if P == 5:
            m = m + q[:, 4, None]
        y[b0:b0 + 16384] = (m + .01 * torch.randn(m.shape, generator=g, dtype=torch.float64, device=DEV)).to(dt)
    return p0.to(dt), x.to(dt), y


def timed_solve(solver, p0, model, args, w, max_iter, reps=3, **kw):
    tol = 1e-8
    p = torch.empty_like(p0)
    for _ in range(2):
        p.copy_(p0)
        out = solver(p, model.cost, model.jac, args=args, wvec=w, ftol=tol, ptol=tol, gtol=tol, max_iter=max_iter, **kw)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        p.copy_(p0)
        out = solver(p, model.cost, model.jac, args=args, wvec=w, ftol=tol, ptol=tol, gtol=tol, max_iter=max_iter, **kw)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    _lib.enable_timing(True)
    p.copy_(p0)
    n_it = len(solver(p, model.cost, model.jac, args=args, wvec=w, ftol=tol, ptol=tol, gtol=tol, max_iter=max_iter, **kw))
    recs = _lib.timing_records()
    _lib.enable_timing(False)
    per = {}
    for name, t in recs:
        per.setdefault(name, []).append(t)
    kern = {k: {'n': len(v), 'avg_ms': sum(v[:n_it]) / max(1, min(len(v), n_it))} for k, v in per.items()}
    return ms, len(out), kern


def report(case, B, ms, iters, kern, bytes_of):
    for k, v in kern.items():
        if k in bytes_of:
            v['alg_GB'] = bytes_of[k] / 1e9
            v['GBps'] = bytes_of[k] / (v['avg_ms'] * 1e-3) / 1e9
            v['frac_of_measured_hbm'] = v['GBps'] / PEAK
    print(json.dumps({'case': case, 'B': B, 'ms_per_solve': ms, 'iterations': iters, 'problems_per_s': B / (ms * 1e-3),
                      'problem_iters_per_s': B * iters / (ms * 1e-3), 'kernels': kern, 'hbm_peak_GBps': PEAK}), flush=True)


def case_cfg2(kind, dtname='f64'):
    B, C, N, P = 65536, 5, 1024, 4
    dt = torch.float64 if dtname == 'f64' else torch.float32
    s = 8 if dtname == 'f64' else 4
    p0, t, y = make_batch(B, C, N, 3008, torch.device(DEV), dt)
    w = torch.ones(C, dtype=dt, device=DEV)
    kw = dict(meth='lev') if kind == 'lm' else {}
    ms, iters, kern = timed_solve(SOLVERS[kind], p0, EMGModel(), (t, y), w, 40 if kind == 'lm' else 80, **kw)
    # DRAM traffic the kernels are designed to move: y once, m_prev read + m written ([B,N] each)
    bytes_of = {'nlls_model_normal_eq': s * N * (C + 2) * B, 'nlls_model_lm_trial': s * N * (C + 2) * B, 'nlls_model_cost_stat': s * C * N * B,
                'nlls_model_normal_eq_gda': (s * C * N + 2 * s * N) * B}
    report(f'cfg2-{kind}' + ('' if dtname == 'f64' else '-f32'), B, ms, iters, kern, bytes_of)


def case_cfg3(kind, P=5):
    B, C, N = 262144, 1, 1000
    dt, s = torch.float32, 4
    p0, x, y = skew_batch(B, N, P, dt)
    w = torch.ones(C, dtype=dt, device=DEV)
    kw = dict(meth='lev') if kind == 'lm' else {}
    ms, iters, kern = timed_solve(SOLVERS[kind], p0, SkewGaussModel(), (x, y), w, 40 if kind == 'lm' else 80, **kw)
    # DRAM traffic the kernels are designed to move: y once, m_prev read + m written ([B,N] each)
    bytes_of = {'nlls_model_normal_eq': s * N * (C + 2) * B, 'nlls_model_lm_trial': s * N * (C + 2) * B, 'nlls_model_cost_stat': s * C * N * B,
                'nlls_model_normal_eq_gda': (s * C * N + 2 * s * N) * B}
    report(f'cfg3-{kind}-P{P}', B, ms, iters, kern, bytes_of)


def multi_emg_batch(B, K, N, dt, seed=3010):
    """cfg-4 end to end: sum of K EMG peaks, centres spread over t, 16 parameters for K = 4."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    t = torch.linspace(-1e3, 1e3, N, device=DEV).double()
    centres = torch.linspace(-600, 600, K, dtype=torch.float64, device=DEV)
    base = torch.stack([torch.tensor([10., 0., 60., 3.], dtype=torch.float64, device=DEV) +
                        torch.tensor([0., 1., 0., 0.], dtype=torch.float64, device=DEV) * c for c in centres]).reshape(-1)
    scale = base.abs().clamp(min=20.)
    gt = base + scale * .05 * (torch.rand(B, 4 * K, generator=g, dtype=torch.float64, device=DEV) - .5)
    p0 = gt + scale * .04 * (torch.rand(B, 4 * K, generator=g, dtype=torch.float64, device=DEV) - .5)
    y = torch.empty(B, 1, N, dtype=dt, device=DEV)
    for b0 in range(0, B, 4096):
        q = gt[b0:b0 + 4096]
        m = 0
        for j in range(K):
            a, mu, sg, eta = (q[:, 4 * j + i, None] for i in range(4))
            d = t - mu
            m = m + a * torch.exp(-.5 * d ** 2 / sg ** 2) * (1 + torch.erf(eta * d / (sg * 2 ** .5)))
        y[b0:b0 + 4096, 0] = (m + .01 * torch.randn(m.shape, generator=g, dtype=torch.float64, device=DEV)).to(dt)
    return p0.to(dt), t.to(dt), y, gt


def case_cfg4(kind):
    """cfg-4 end to end: lsq_lma_parallel / lsq_gna_parallel on B=32768 four-peak EMG fits, C=1, N=4096, P=16, fp64,
    tensor path: nlls_model_eval materialises f and J, K1 (DMMA) reduces them, K2 (warp Cholesky) solves."""
    B, C, N, K = 32768, 1, 4096, 4
    P, dt, s = 4 * K, torch.float64, 8
    p0, t, y, gt = multi_emg_batch(B, K, N, dt)
    w = torch.ones(C, dtype=dt, device=DEV)
    kw = dict(meth='lev') if kind == 'lm' else {}
    ms, iters, kern = timed_solve(SOLVERS[kind], p0, MultiEMGModel(K), (t, y), w, 40, reps=2, **kw)
    bytes_of = {'nlls_normal_eq': s * C * N * (P + 2) * B, 'nlls_cost_stat': s * C * N * B}
    # nlls_model_eval is called for f alone, for J (+f) and, in LM, for the trial cost: report its average only
    report(f'cfg4-{kind}', B, ms, iters, kern, bytes_of)


def case_cfg4_k1():
    """Kernel-level: K1 (wide, P=16) + K2 on seeded N(0,1) tensors of the cfg-4 shape."""
    B, C, N, P = 32768, 1, 4096, 16
    dt, s = torch.float64, 8
    g = torch.Generator(device=DEV).manual_seed(4)
    J = torch.randn(B, C, N, P, generator=g, dtype=dt, device=DEV)
    f = torch.randn(B, C, N, generator=g, dtype=dt, device=DEV)
    fp = torch.randn(B, C, N, generator=g, dtype=dt, device=DEV)
    w = torch.ones(C, dtype=dt, device=DEV)
    gg = torch.empty(B, P, dtype=dt, device=DEV)
    H = torch.empty(B, P, P, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    p = torch.zeros(B, P, dtype=dt, device=DEV)
    u = torch.full((B,), 1e-3, dtype=dt, device=DEV)
    ds = torch.zeros(B, 2, dtype=dt, device=DEV)
    ds[:, 0] = 1
    h = torch.empty(B, P, dtype=dt, device=DEV)
    pt = torch.empty(B, P, dtype=dt, device=DEV)
    ctrl = _lib.new_ctrl(DEV)

    def k1():
        _lib.call('nlls_normal_eq', _lib.F64, _lib.ptr(J), _lib.ptr(f), _lib.ptr(fp), _lib.ptr(w), 1, B, C, N, P, _lib.ptr(gg),
                  _lib.ptr(H), _lib.ptr(st), None, _lib.stream_ptr())

    def k2():
        _lib.call('nlls_lm_solve', _lib.F64, _lib.ptr(gg), _lib.ptr(H), _lib.ptr(p), _lib.ptr(u), _lib.ptr(ds), 0, B, P,
                  _lib.ptr(h), _lib.ptr(pt), _lib.ptr(ctrl), _lib.stream_ptr())

    out = {}
    for name, fn, nbytes in (('nlls_normal_eq', k1, s * C * N * (P + 2) * B), ('nlls_lm_solve', k2, 8 * (P * P + 3 * P) * B)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        out[name] = {'avg_ms': ms, 'alg_GB': nbytes / 1e9, 'GBps': nbytes / (ms * 1e-3) / 1e9,
                     'frac_of_measured_hbm': nbytes / (ms * 1e-3) / 1e9 / PEAK}
    print(json.dumps({'case': 'cfg4-k1', 'B': B, 'shape': [B, C, N, P], 'kernels': out, 'hbm_peak_GBps': PEAK}), flush=True)


CASES = {'cfg2-gn': lambda: case_cfg2('gn'), 'cfg2-lm': lambda: case_cfg2('lm'), 'cfg2-gd': lambda: case_cfg2('gd'),
         'cfg2-gn-f32': lambda: case_cfg2('gn', 'f32'), 'cfg2-lm-f32': lambda: case_cfg2('lm', 'f32'),
         'cfg3-gd': lambda: case_cfg3('gd'), 'cfg3-lm': lambda: case_cfg3('lm'), 'cfg3-gn': lambda: case_cfg3('gn'),
         'cfg4-k1': case_cfg4_k1, 'cfg4-lm': lambda: case_cfg4('lm'), 'cfg4-gn': lambda: case_cfg4('gn')}

if __name__ == '__main__':
    assert _lib.load().nlls_device_check() == 0
    for c in (sys.argv[1:] or list(CASES)):
        try:
            CASES[c]()
        except Exception as e:                                   # noqa: BLE001
            print(json.dumps({'case': c, 'error': repr(e)}), flush=True)
        torch.cuda.empty_cache()
