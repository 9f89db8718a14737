"""How far does the REFERENCE ALGORITHM's own fp32 iteration count move under rounding-level changes of its arithmetic?

north_star: "identical convergence iteration counts except at tolerance boundaries, which are reported".  At the tests'
tolerances (1e-8, below fp32 resolution) the fp32 Levenberg-Marquardt stop decision of a whole batch is taken on rounding
noise: the reference needs EVERY problem to pass a stop test in the SAME iteration (lma:114-118, masks are not sticky).
This script runs the pinned oracle (bit-identical to the reference on these inputs: tests/test_oracle_golden.py) on every
fp32 LM golden case under legitimate re-roundings of the same algorithm and records the spread of the iteration count:

  exact_neq   normal equations accumulated in fp64 and rounded once (newton_parallel.py:40-43 sums in fp32 in bmm order)
  fp64_model  model and Jacobian evaluated in fp64 and rounded to fp32 (tests/emg.py in fp32 otherwise)
  both        exact_neq and fp64_model together (closest to what a fused kernel with its own exp/erf does)
  libm_K      model values and Jacobian entries off by up to K fp32 ulp (seeded, K = 2, 4; 2 seeds each): another math
              library (torch CPU vs CUDA expf/erff are specified to 1-2 ulp; the fused fp32 kernels' exp2-based exp and
              polynomial erfc are accurate to 4e-7 = ~4 ulp, csrc/models.cuh)
  ulp_p0      the initial parameters moved by one fp32 ulp (up / down / a seeded random direction per entry, 3 seeds)

tests/helpers.py::compare_plist allows an fp32 case to differ from the reference's count by at most max(1, spread of that
case) and still compares the true finals at 1e-4; every difference is written to profiles/r02_boundary_flips_*.json.

    python tests/golden/make_fp32_spread.py        -> tests/golden/fp32_count_spread.json
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
from oracle import nlls_oracle as O, models as M                  # noqa: E402

TOL = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8)
F32 = torch.float32


def variants(p0, fun, jac, args, w, kw):
    """iteration counts of the oracle's LM under the re-roundings above"""
    out = {}

    def run(p, f=fun, j=jac):
        try:
            return O.lma(p.clone(), f, j, args=args, wvec=w, **kw).iterations
        except RuntimeError:                                      # NaN iterate -> LAPACK argument error, as the reference (Q6)
            return kw['max_iter']

    out['reference'] = run(p0)
    orig = O.normal_equations
    O.normal_equations = lambda f, J, wv: tuple(t.to(f.dtype) for t in orig(f.double(), J.double(), wv.double()))
    try:
        out['exact_neq'] = run(p0)
    finally:
        O.normal_equations = orig
    f64 = lambda fn: (lambda p, *a: fn(p.double(), *[x.double() if torch.is_tensor(x) else x for x in a]).to(F32))
    out['fp64_model'] = run(p0, f64(fun), f64(jac))
    O.normal_equations = lambda f, J, wv: tuple(t.to(f.dtype) for t in orig(f.double(), J.double(), wv.double()))
    try:
        out['both'] = run(p0, f64(fun), f64(jac))
    finally:
        O.normal_equations = orig
    for K in (2, 4):
        for seed in (1, 2):
            gen = torch.Generator().manual_seed(100 * K + seed)

            def wob(fn, gen=gen, K=K):
                def g(p, *a):
                    v = fn(p, *a)
                    return v * (1 + K * 2.0 ** -24 * (2 * torch.rand(v.shape, generator=gen) - 1))
                return g
            out[f'libm_{K}ulp_s{seed}'] = run(p0, wob(fun), wob(jac))
    for seed in (1, 2, 3):
        sign = torch.where(torch.rand(p0.shape, generator=torch.Generator().manual_seed(seed)) < .5, -1., 1.)
        out[f'ulp_p0_rand{seed}'] = run(torch.nextafter(p0, sign * float('inf')))
    out['ulp_p0_up'] = run(torch.nextafter(p0, torch.full_like(p0, float('inf'))))
    out['ulp_p0_down'] = run(torch.nextafter(p0, torch.full_like(p0, -float('inf'))))
    return out


def main():
    torch.set_num_threads(8)
    res = {}
    load = lambda n: dict(np.load(os.path.join(HERE, n + '.npz')))
    fx, sy, sk, c5 = load('fixtures'), load('synthetic'), load('skewgauss'), load('cfg5')
    t32 = lambda a: torch.as_tensor(a).to(F32)
    emg2 = (lambda p, t, y: M.emg_cost(p, t, y, n_costs=2)), (lambda p, t, y: M.emg_jac(p, t, y, n_costs=2))
    lev = dict(meth='lev', max_iter=99, **TOL)
    res['lma_lev__f32'] = variants(t32(fx['p0']), *emg2, (t32(fx['t']), t32(fx['y'])), torch.ones(2), lev)
    res['lma_marq__f32'] = variants(t32(fx['p0']), *emg2, (t32(fx['t']), t32(fx['y'])), torch.ones(2), dict(meth='marq', max_iter=99, **TOL))
    res['synthetic:lma_lev__f32'] = variants(t32(sy['p0']), M.emg_cost, M.emg_jac, (t32(sy['t']), t32(sy['y'])), t32(sy['wvec']), lev)
    for P in (4, 5):
        pre = f'P{P}_'
        res[f'{pre}lma_lev__f32'] = variants(t32(sk[pre + 'p0']), M.skewgauss_cost, M.skewgauss_jac, (t32(sk[pre + 'x']), t32(sk[pre + 'y'])),
                                             torch.ones(1), lev)
    res['cfg5 lma_lev__f32'] = variants(t32(c5['p0']), M.emg_cost, M.emg_jac, (t32(c5['t']), t32(c5['y'])), t32(c5['wvec']),
                                        dict(meth='lev', max_iter=40, **TOL))
    for k, v in res.items():
        v['spread'] = max(abs(c - v['reference']) for c in v.values())
        print(k, v)
    with open(os.path.join(HERE, 'fp32_count_spread.json'), 'w') as f:
        json.dump(res, f, indent=1)


if __name__ == '__main__':
    main()
