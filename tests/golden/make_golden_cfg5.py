"""Golden vectors in the shape of BASELINE.json configs[4] (cfg-5: EMG fits, C=5, N=1024, P=4, fp32 and fp64) from the
UNMODIFIED reference (imported read-only from /root/reference): a seeded 16-problem batch that the GPU tests tile to
B >= 1 Mi problems (tests/test_gpu_fullsize.py::test_cfg5_*) and that the oracle is pinned against
(tests/test_oracle_golden.py).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_cfg5.py

Keys: t [N], y [B,C,N], p0 [B,P], gt [B,P], wvec [C]; ``<case>__<dtype>__plist`` [K,B,P] with the cases below
(max_iter as in scripts/bench_cfg5.py and README.rst:50-54,80-84: LM 40, GN 80)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, HERE)
from make_golden import emg_batch, models, lsq_lma_parallel, lsq_gna_parallel, TOL       # noqa: E402  (imports the reference)

CASES5 = {'lma_lev': (lsq_lma_parallel, dict(meth='lev', max_iter=40, **TOL)),
          'gna': (lsq_gna_parallel, dict(l=1., max_iter=80, **TOL))}

if __name__ == '__main__':
    torch.set_num_threads(8)
    B, C, N = 16, 5, 1024
    gt, p0, t, y = emg_batch(B, C, N, 3008)
    w = torch.ones(C, dtype=torch.float64)
    out = {'t': t.numpy(), 'y': y.numpy(), 'p0': p0.numpy(), 'gt': gt.numpy(), 'wvec': w.numpy()}
    for dt, dn in ((torch.float64, 'f64'), (torch.float32, 'f32')):
        for name, (solver, kw) in CASES5.items():
            p = p0.clone().to(dt)
            pl = solver(p=p, function=models.emg_cost, jac_function=models.emg_jac, args=(t.to(dt), y.to(dt)), wvec=w.to(dt), **kw)
            assert torch.equal(pl[-1], p)
            out[f'{name}__{dn}__plist'] = torch.stack(pl).numpy()
            print(f'cfg5 {name:8s} {dn}: {len(pl):3d} iterations, max |p - gt| = {float((pl[-1].double() - gt).abs().max()):.3e}')
    np.savez_compressed(os.path.join(HERE, 'cfg5.npz'), **out)
    print('cfg5', os.path.getsize(os.path.join(HERE, 'cfg5.npz')) // 1024, 'KiB')
