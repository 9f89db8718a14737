"""solve_host_batches: host batches through the device solvers with overlapped copies give exactly the
results of calling the drop-in solver directly on each batch (same kernels, same order)."""
import pytest
import torch

from torchimize_b200.pipeline import slot_schedule


def test_slot_schedule_orders_copy_before_solve_and_never_overwrites_a_live_slot():
    for n in (0, 1, 2, 5, 9):
        for depth in (1, 2, 3):
            ops = slot_schedule(n, depth)
            assert [b for op, b, _ in ops if op == 'solve'] == list(range(n))
            assert [b for op, b, _ in ops if op == 'copy'] == list(range(n))
            live = {}
            for op, b, s in ops:
                assert s == b % depth
                if op == 'copy':
                    assert s not in live, 'copy into a slot whose batch is not solved yet'
                    live[s] = b
                else:
                    assert live.pop(s) == b
            # the copy of batch k+depth-1 is enqueued before the solve of batch k
            for k in range(n):
                nxt = k + depth - 1
                if depth > 1 and nxt < n:
                    assert ops.index(('copy', nxt, nxt % depth)) < ops.index(('solve', k, k % depth))


@pytest.mark.gpu
@pytest.mark.parametrize('solver_name', ['gna', 'lma', 'gda'])
@pytest.mark.parametrize('depth', [1, 2, 3])
def test_host_batches_match_direct_solves(solver_name, depth):
    from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel, gradient_descent_parallel
    from torchimize_b200.models import EMGModel
    from torchimize_b200.pipeline import solve_host_batches
    from bench import make_batch

    solver = dict(gna=lsq_gna_parallel, lma=lsq_lma_parallel, gda=gradient_descent_parallel)[solver_name]
    dev = torch.device('cuda:0')
    emg = EMGModel()
    C = 3
    w = torch.ones(C, dtype=torch.float64)
    kw = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=30)
    batches = []
    for i, B in enumerate((96, 33, 96, 64, 7)):                 # ragged batch sizes: slots are re-fitted
        p0, t, y = make_batch(B, C, 256, 100 + i, 'cpu', torch.float64)
        batches.append((p0.pin_memory(), (t.pin_memory(), y.pin_memory())))
    direct = []
    for p0, (t, y) in batches:
        p = p0.to(dev)
        coeffs = solver(p, emg.cost, emg.jac, args=(t.to(dev), y.to(dev)), wvec=w.to(dev), **kw)
        direct.append((p.cpu(), len(coeffs)))
    got = [(out.clone(), n) for out, n in solve_host_batches(solver, iter(batches), emg.cost, emg.jac, wvec=w, depth=depth, **kw)]
    assert len(got) == len(direct)
    for (a, na), (b, nb) in zip(got, direct):
        assert na == nb
        assert torch.equal(a, b)
