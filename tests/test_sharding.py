"""Multi-GPU host logic exercised with world_size-2 process groups.

CPU part (gloo, runs everywhere): shard bounds, the one result all-gather, the 32-byte control-block
all-reduce hook and its installation/removal around a solver call.
GPU part (-m gpu): two ranks share cuda:0 over gloo and run the REAL kernels in mode='exact' on the
two halves of the reference's fixture batch; snapshots must equal the golden full-batch trajectory
of the reference (same global iteration count although each half alone would stop elsewhere)."""
import os
import socket
import sys
import traceback

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'tests')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from torchimize_b200 import sharding                                  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, fn, args, errq):
    try:
        os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group('gloo', rank=rank, world_size=world)
        fn(rank, world, *args)
        dist.barrier()
        dist.destroy_process_group()
    except Exception:                                                  # noqa: BLE001
        errq.put((rank, traceback.format_exc()))
        raise


def _spawn(fn, *args, world=2):
    ctx = mp.get_context('spawn')
    errq = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn, args, errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    assert not errs, '\n'.join(f'rank {r}:\n{tb}' for r, tb in errs)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]


# ------------------------------------------------------------------------------- pure
@pytest.mark.parametrize('B', [0, 1, 5, 6, 7, 65536, 4194304 + 3])
@pytest.mark.parametrize('world', [1, 2, 3, 4, 8])
def test_shard_bounds_partition(B, world):
    prev, sizes = 0, []
    for r in range(world):
        lo, hi = sharding.shard_bounds(B, world, r)
        assert lo == prev and hi >= lo
        prev = hi
        sizes.append(hi - lo)
    assert prev == B
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_shard_view_is_a_slice():
    t = torch.arange(14.).reshape(7, 2)
    a, b = sharding.shard(t, 2, 0), sharding.shard(t, 2, 1)
    assert torch.equal(torch.cat([a, b]), t) and a.shape[0] == 4 and b.shape[0] == 3
    assert a.data_ptr() == t.data_ptr()                                # view, no copy


# ------------------------------------------------------------------------- gloo, world 2
def _w_gather(rank, world, B):
    full = torch.arange(B * 4, dtype=torch.float64).reshape(B, 4) * 1.5
    mine = sharding.shard(full, world, rank).clone()
    out = sharding.gather_results(mine, B)
    assert torch.equal(out, full), (rank, out, full)


@pytest.mark.parametrize('B', [6, 7, 2, 1])
def test_gather_results_gloo(B):
    _spawn(_w_gather, B)


def _w_hook(rank, world):
    red = torch.tensor([1. + rank, 10. * rank, .25, 0.], dtype=torch.float64)
    sharding.make_reduce_hook()(red)
    assert torch.equal(red, torch.tensor([3., 10., .5, 0.], dtype=torch.float64))


def test_reduce_hook_sums_ctrl_block_gloo():
    _spawn(_w_hook)


def _w_solve_sharded(rank, world):
    from torchimize_b200.functions.parallel import _engine
    seen = {}

    def fake_solver(p, function, jac_function, args=(), wvec=None, **kw):
        hook = _engine._default_reduce_hook
        seen['hook'] = hook
        if hook is not None:
            red = torch.tensor([float(p.shape[0]), 0., 0., 0.], dtype=torch.float64)
            hook(red)
            seen['sum'] = float(red[0])
        p += 1.
        return [p.clone()]

    B = 7
    full = torch.zeros(B, 3)
    for mode in ('local', 'exact'):
        seen.clear()
        mine = sharding.shard(full, world, rank).clone()
        coeffs, out = sharding.solve_sharded(fake_solver, mine, None, mode=mode, B_total=B)
        assert (seen['hook'] is not None) == (mode == 'exact')
        if mode == 'exact':
            assert seen['sum'] == B                                    # all-reduced over both shards
        assert _engine._default_reduce_hook is None                    # hook removed afterwards
        assert out.shape == (B, 3) and torch.equal(out, torch.ones(B, 3)) and len(coeffs) == 1
    # B_total discovered by an all-reduce when not given
    mine = sharding.shard(full, world, rank).clone()
    _, out = sharding.solve_sharded(fake_solver, mine, None, mode='local')
    assert out.shape == (B, 3)
    with pytest.raises(ValueError):
        sharding.solve_sharded(fake_solver, mine, None, mode='bogus')
    # exact modes refuse a batch that leaves a rank without problems (it would never enter the exchange)
    for mode in ('exact', 'exact-peer'):
        with pytest.raises(ValueError):
            sharding.solve_sharded(fake_solver, sharding.shard(torch.zeros(1, 3), world, rank).clone(), None, mode=mode, B_total=1)
    _, out = sharding.solve_sharded(fake_solver, sharding.shard(torch.zeros(1, 3), world, rank).clone(), None, mode='local', B_total=1)
    assert out.shape == (1, 3)


def test_solve_sharded_host_logic_gloo():
    _spawn(_w_solve_sharded)


# ----------------------------------------------------------------- GPU: exact mode, real kernels
def _w_exact_gpu(rank, world, name, dn, path):
    import helpers as Hh
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel
    from torchimize_b200.functions.parallel import _engine
    from torchimize_b200.models import EMGModel
    from oracle import models as M
    gold = Hh.load('fixtures')
    dt = Hh.DT[dn]
    dev = torch.device('cuda:0')
    torch.cuda.set_device(0)
    solver_name, kw = Hh.CASES[name]
    solver = dict(lma=lsq_lma_parallel, gna=lsq_gna_parallel, gda=gradient_descent_parallel)[solver_name]
    t = torch.as_tensor(gold['t']).to(dt).to(dev)
    y = sharding.shard(torch.as_tensor(gold['y']).to(dt), world, rank).to(dev)
    p = sharding.shard(torch.as_tensor(gold['p0']).to(dt), world, rank).clone().to(dev)
    w = torch.ones(2, dtype=dt, device=dev)
    if path == 'fused':
        emg = EMGModel(n_costs=2)
        fun, jac = emg.cost, emg.jac
    else:
        fun = lambda q, tt, yy: M.emg_cost(q, tt, yy, n_costs=2)
        jac = lambda q, tt, yy: M.emg_jac(q, tt, yy, n_costs=2)

    def hook(red):                                                      # gloo: stage the 32 bytes through the host
        h = red.cpu()
        dist.all_reduce(h)
        red.copy_(h)

    old = _engine.set_default_reduce_hook(hook)
    try:
        got = solver(p, fun, jac, args=(t, y), wvec=w, **kw)
    finally:
        _engine.set_default_reduce_hook(old)
    want = torch.as_tensor(gold[f'{name}__{dn}__plist'])
    lo, hi = sharding.shard_bounds(want.shape[1], world, rank)
    Hh.compare_plist(got, want[:, lo:hi], dn, f'exact-sharded {name} {dn} {path} rank {rank}', allow_boundary=(dn == 'f32'),
                     spread_key=f'{name}__{dn}')


@pytest.mark.gpu
@pytest.mark.parametrize('path', ['fused', 'torch_callables'])
@pytest.mark.parametrize('name', ['lma_lev', 'gna', 'gda'])
def test_exact_mode_two_ranks_match_full_batch_reference(name, path):
    _spawn(_w_exact_gpu, name, 'f64', path)


@pytest.mark.gpu
def test_exact_peer_mode_matches_single_gpu_bit_for_bit_when_two_gpus_are_present():
    """mode='exact-peer' (nlls_iter_end_peer: the 32-byte exchange over NVLink peer memory by our own kernel) against
    mode='exact' (NCCL) and a single-GPU run: scripts/peer_exact_check.py under torchrun; needs >= 2 GPUs on the box."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs on one box')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
                        '127.0.0.1', '--master-port', '29533', os.path.join(root, 'scripts', 'peer_exact_check.py')],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and 'PEER EXACT CHECK OK' in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_solvers_run_on_the_device_of_p_not_on_the_current_device_when_two_gpus_are_present():
    """The reference accepts tensors on any device.  The library launches on the CURRENT device's stream, so every entry
    point switches to p's device first (_lib.on_device_of): tensors on cuda:1 while cuda:0 is current give the bits of the
    same solve on cuda:0 — fused and tensor path, the single-problem solver, newton_step_parallel and the host pipeline."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs on one box')
    from helpers import load
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel, newton_step_parallel, lsq_gna
    from torchimize_b200.models import EMGModel
    from torchimize_b200.pipeline import solve_host_batches
    gold = load('synthetic')
    dt = torch.float64
    emg = EMGModel()
    torch.cuda.set_device(0)
    out = {}
    for dev in ('cuda:0', 'cuda:1'):
        t, y, p0, w = (torch.as_tensor(gold[k]).to(dt).to(dev) for k in ('t', 'y', 'p0', 'wvec'))
        res = []
        for solver, kw in ((lsq_gna_parallel, {}), (lsq_lma_parallel, dict(meth='lev')), (gradient_descent_parallel, {})):
            for fused in (True, False):
                res.append(torch.stack(solver(p0.clone(), emg.cost, emg.jac, args=(t, y), wvec=w, max_iter=25, fused=fused, **kw)).cpu())
        _, f, g, H = newton_step_parallel(p0.clone(), lambda q: emg.cost(q, t, y), lambda q: emg.jac(q, t, y), w)
        res += [f.cpu(), g.cpu(), H.cpu()]
        one = lsq_gna(p0[0].clone(), lambda q: emg.cost(q[None], t, y[:1])[0, 0], lambda q: emg.jac(q[None], t, y[:1])[0, 0], max_iter=5)
        res.append(torch.stack([q.detach() for q in one]).cpu())
        hb = [(p0.cpu().pin_memory(), (t.cpu().pin_memory(), y.cpu().pin_memory()))]
        for p_fin, n_it in solve_host_batches(lsq_gna_parallel, hb, emg.cost, emg.jac, wvec=w, device=dev, max_iter=25):
            res.append(p_fin.clone())
        assert torch.cuda.current_device() == 0
        out[dev] = res
    for a, b in zip(out['cuda:0'], out['cuda:1']):
        assert torch.equal(a, b)
