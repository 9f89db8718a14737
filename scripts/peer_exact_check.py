#!/usr/bin/env python
"""torchrun check of mode='exact-peer' (nlls_iter_end_peer: the exact-mode exchange over NVLink peer memory):
every rank solves its shard in modes 'exact' (NCCL all-reduce) and 'exact-peer' and, on rank 0, the whole batch on one GPU;
the three have to agree bit for bit (same kernels, same global stop decisions), for GN, LM and GD.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/peer_exact_check.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                                        # noqa: E402
import torch.distributed as dist                                    # noqa: E402
from bench import make_batch                                        # noqa: E402
from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel, gradient_descent_parallel   # noqa: E402
from torchimize_b200.models import EMGModel                         # noqa: E402
from torchimize_b200.sharding import shard_bounds, solve_sharded, gather_results   # noqa: E402


def main():
    world, rank, local = int(os.environ['WORLD_SIZE']), int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    B, C, N = 4096 + 37, 3, 512
    p0, t, y = make_batch(B, C, N, 77, dev, torch.float64)          # same seed on every rank: identical full batch
    w = torch.ones(C, dtype=torch.float64, device=dev)
    emg = EMGModel()
    lo, hi = shard_bounds(B, world, rank)
    ok = True
    for name, solver, kw in (('gna', lsq_gna_parallel, dict(max_iter=60)), ('lma', lsq_lma_parallel, dict(max_iter=40, meth='lev')),
                             ('gda', gradient_descent_parallel, dict(max_iter=30))):
        res = {}
        for mode in ('exact', 'exact-peer', 'exact-peer'):          # twice: the mailbox is reused across solves (epochs)
            p = p0[lo:hi].clone()
            out, _ = solve_sharded(solver, p, emg.cost, emg.jac, args=(t, y[lo:hi]), wvec=w, mode=mode, gather=False, **kw)
            full = gather_results(p, B)
            res.setdefault(mode, []).append((len(out), full))
        single = p0.clone()
        n_single = len(solver(single, emg.cost, emg.jac, args=(t, y), wvec=w, **kw))
        for mode, runs in res.items():
            for n_it, full in runs:
                same = n_it == n_single and bool(torch.equal(full, single))
                ok = ok and same
                if rank == 0:
                    print(f'{name} {mode}: {n_it} iterations (single GPU {n_single}), bit-identical to the single-GPU run: {same}', flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print('PEER EXACT CHECK', 'OK' if int(flag.item()) == 1 else 'FAILED', flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == '__main__':
    main()
