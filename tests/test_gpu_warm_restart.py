"""lsq_lma_parallel(state=...): the damping state (u, v, Marquardt scale) of one call carried into the next
(SURVEY 8f rank 4).  k calls of max_iter = n walk the same iterates as one call of max_iter = k n, bit for bit,
on the fused (speculative) and on the tensor path."""
import pytest
import torch


@pytest.mark.gpu
@pytest.mark.parametrize('fused', [True, False])
@pytest.mark.parametrize('meth', ['lev', 'marq'])
def test_split_solve_walks_the_same_iterates(fused, meth):
    from torchimize_b200.functions import lsq_lma_parallel
    from torchimize_b200.models import EMGModel
    from bench import make_batch

    dev = torch.device('cuda:0')
    emg = EMGModel()
    B, C, N = 160, 3, 256
    p0, t, y = make_batch(B, C, N, 77, dev, torch.float64)
    w = torch.ones(C, dtype=torch.float64, device=dev)
    kw = dict(args=(t, y), wvec=w, ftol=0., ptol=0., gtol=0., meth=meth, fused=fused)      # never stops early

    p_full = p0.clone()
    full = lsq_lma_parallel(p_full, emg.cost, emg.jac, max_iter=12, **kw)
    assert len(full) == 12

    p_split, state, split = p0.clone(), {}, []
    for n in (5, 4, 3):
        split += lsq_lma_parallel(p_split, emg.cost, emg.jac, max_iter=n, state=state, **kw)
        assert state['iterations'] == n and state['meth'] == meth
        assert state['u'].shape == (B,) and state['v'].shape == (B,)
    assert len(split) == 12
    same = lambda a, b: torch.allclose(a, b, rtol=0., atol=0., equal_nan=True)      # Marquardt's scale overflows to NaN (lma:88)
    for a, b in zip(full, split):
        assert same(a, b)
    assert same(p_full, p_split)
    assert not torch.equal(state['u'], torch.full_like(state['u'], 1e-3))      # the damping did move

    # a cold restart (no state) from the same intermediate point differs: the state matters
    p_cold = p0.clone()
    cold = lsq_lma_parallel(p_cold, emg.cost, emg.jac, max_iter=5, **kw)
    cold += lsq_lma_parallel(p_cold, emg.cost, emg.jac, max_iter=7, **kw)
    if meth == 'lev':
        assert not same(p_cold, p_full)


@pytest.mark.gpu
def test_state_is_validated():
    from torchimize_b200._lib import NllsError
    from torchimize_b200.functions import lsq_lma_parallel
    from torchimize_b200.models import EMGModel
    from bench import make_batch

    dev = torch.device('cuda:0')
    emg = EMGModel()
    p0, t, y = make_batch(8, 2, 128, 5, dev, torch.float64)
    state = {}
    lsq_lma_parallel(p0.clone(), emg.cost, emg.jac, args=(t, y), max_iter=2, state=state)
    with pytest.raises(NllsError):
        lsq_lma_parallel(p0.clone(), emg.cost, emg.jac, args=(t, y), max_iter=2, meth='marq', state=dict(state))
    bad = dict(state, u=state['u'][:4])
    with pytest.raises(NllsError):
        lsq_lma_parallel(p0.clone(), emg.cost, emg.jac, args=(t, y), max_iter=2, state=bad)
