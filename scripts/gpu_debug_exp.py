"""debug: LM on the smoke fixtures under the current NLLS_LIB; prints the iterates and compares f at each snapshot with the oracle"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from torchimize_b200 import _lib
from torchimize_b200.functions import lsq_lma_parallel
from torchimize_b200.models import EMGModel
from oracle import nlls_oracle as O, models as M

gold = dict(np.load(os.path.join(ROOT, 'tests', 'golden', 'fixtures.npz')))
dt = torch.float64
t, y, p0 = (torch.as_tensor(gold[k]).to(dt) for k in ('t', 'y', 'p0'))
w = torch.ones(2, dtype=dt)
kw = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=99)
fun = lambda p, t, y: M.emg_cost(p, t, y, n_costs=2)
jac = lambda p, t, y: M.emg_jac(p, t, y, n_costs=2)
ref = O.lma(p0.clone(), fun, jac, args=(t, y), wvec=w, meth='lev', **kw).p_list
emg = EMGModel(n_costs=2)
dev = 'cuda'
for fused in (True, False):
    st = {}
    got = lsq_lma_parallel(p0.clone().to(dev), emg.cost, emg.jac, args=(t.to(dev), y.to(dev)), wvec=w.to(dev), fused=fused, meth='lev', stats=st, **kw)
    print('fused', fused, 'iterations', len(got), 'oracle', len(ref), {k: (v.tolist() if torch.is_tensor(v) else v) for k, v in st.items()})
    for i, (a, b) in enumerate(zip(got, ref)):
        err = ((a.cpu() - b).norm(dim=1) / b.norm(dim=1))
        print(' it', i, 'rel err per problem', ['%.1e' % e for e in err.tolist()])
        if float(err.max()) > 1e-6:
            # model values at the diverging iterate's predecessor, ours vs oracle
            pp = ref[i - 1] if i > 0 else p0
            f_ref = fun(pp, t, y)
            f_our = emg.cost(pp.to(dev), t.to(dev), y.to(dev)).cpu()
            d = (f_our - f_ref).abs()
            print('  f mismatch at predecessor: max abs', float(d.max()), 'at', np.unravel_index(int(d.argmax()), d.shape), 'f_ref there', float(f_ref.flatten()[int(d.argmax())]))
            J_ref = jac(pp, t, y)
            J_our = emg.jac(pp.to(dev), t.to(dev), y.to(dev)).cpu()
            dj = (J_our - J_ref).abs()
            print('  J mismatch: max abs', float(dj.max()), 'at', np.unravel_index(int(dj.argmax()), dj.shape), 'J_ref there', float(J_ref.flatten()[int(dj.argmax())]))
            print('  p there', pp[np.unravel_index(int(d.argmax()), d.shape)[0]].tolist())
            break
