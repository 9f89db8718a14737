#!/usr/bin/env python
"""BASELINE.json configs[4]: B (default 4 Mi) EMG fits (C=5, N=1024, P=4) sharded over the GPUs of one box,
fp32 and fp64, GN and LM through the public drop-in API (fused path).  One process per GPU:

    python scripts/bench_cfg5.py --B 1048576                                  (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_cfg5.py

Strong scaling: B is the TOTAL number of problems; rank r holds shard_bounds(B, world, r).  No per-iteration
collective (mode 'local'); one all-gather of the finals inside the timed region.  A shard that does not fit in HBM
(fp64 needs 57 KB per problem for y and the two model-value buffers) is solved in sequential chunks (--chunk).
One JSON line per (dtype, solver); times are CUDA-event times, max over ranks."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                                        # noqa: E402
import torch.distributed as dist                                    # noqa: E402
from bench import make_batch                                        # noqa: E402
from torchimize_b200 import _lib                                    # noqa: E402
from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel   # noqa: E402
from torchimize_b200.models import EMGModel                         # noqa: E402
from torchimize_b200.sharding import shard_bounds, gather_results, solve_sharded   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--B', type=int, default=4 * 1024 * 1024)
    ap.add_argument('--chunk', type=int, default=0, help='problems per sequential chunk on a rank (0 = fit to memory)')
    ap.add_argument('--dtypes', default='f32,f64')
    ap.add_argument('--solvers', default='gn,lm')
    ap.add_argument('--reps', type=int, default=2)
    ap.add_argument('--mode', default='local', choices=['local', 'exact', 'exact-peer'],
                    help="exact: all-reduce the 32-byte batch-global sums every iteration (sharding.solve_sharded)")
    args = ap.parse_args()
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    assert _lib.load().nlls_device_check() == 0
    C, N, P = 5, 1024, 4
    lo, hi = shard_bounds(args.B, world, rank)
    mine = hi - lo
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs'])
    except Exception:                                                # noqa: BLE001
        pass
    emg = EMGModel()
    for dn in args.dtypes.split(','):
        dt = torch.float64 if dn == 'f64' else torch.float32
        es = 8 if dn == 'f64' else 4
        free = torch.cuda.mem_get_info(dev)[0]
        per_problem = (C + 2) * N * es + 4096                        # y + the two model-value buffers (+ small per-problem state)
        chunk = args.chunk or min(mine, int(free * .85) // per_problem)
        n_chunks = (mine + chunk - 1) // chunk
        chunk = (mine + n_chunks - 1) // n_chunks
        w = torch.ones(C, dtype=dt, device=dev)
        for sv in args.solvers.split(','):
            solver = lsq_gna_parallel if sv == 'gn' else lsq_lma_parallel
            kw = dict(max_iter=80) if sv == 'gn' else dict(max_iter=40, meth='lev')
            finals = torch.empty(mine, P, dtype=dt, device=dev)
            total_ms, iters, launches = 0.0, [], 0
            for ci in range(n_chunks):
                c0, c1 = ci * chunk, min(mine, (ci + 1) * chunk)
                p0, t, y = make_batch(c1 - c0, C, N, 3008 + 7919 * (lo + c0), dev, dt)     # untimed: synthetic data
                best = None
                for _ in range(args.reps):
                    p = p0.clone()
                    torch.cuda.synchronize()
                    if world > 1:
                        dist.barrier()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    l0 = _lib.launches
                    a.record()
                    if args.mode != 'local' and world > 1:
                        out, _ = solve_sharded(solver, p, emg.cost, emg.jac, args=(t, y), wvec=w, mode=args.mode, gather=False,
                                               ftol=1e-8, ptol=1e-8, gtol=1e-8, **kw)
                    else:
                        out = solver(p, emg.cost, emg.jac, args=(t, y), wvec=w, ftol=1e-8, ptol=1e-8, gtol=1e-8, **kw)
                    b.record()
                    torch.cuda.synchronize()
                    ms = a.elapsed_time(b)
                    best = ms if best is None else min(best, ms)
                    n_it, n_l = len(out), _lib.launches - l0
                    del out
                finals[c0:c1] = p
                total_ms += best
                iters.append(n_it)
                launches += n_l
                del p0, t, y, p
                torch.cuda.empty_cache()
            if world > 1:
                gather_results(finals, args.B)                       # untimed warm-up: NCCL sets its communicator up lazily
                torch.cuda.synchronize()
                dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            full = gather_results(finals, args.B) if world > 1 else finals
            b.record()
            torch.cuda.synchronize()
            solve_ms, gather_ms = total_ms, a.elapsed_time(b)
            total_ms += gather_ms
            tm = torch.tensor([total_ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            ms = float(tm.item())
            ok = bool(torch.isfinite(full).all())
            if rank == 0:
                it = max(iters)
                passes = it + (1 if sv == 'lm' else 0)               # speculative LM: one model pass per iteration + the first
                alg = es * C * N * mine * passes                     # s*N_y per model evaluation
                print(json.dumps({'case': f'cfg5-{sv}-{dn}', 'B_total': args.B, 'n_gpus': world, 'B_per_gpu': mine,
                                  'chunks_per_gpu': n_chunks, 'iterations': it, 'ms': ms, 'solve_ms_rank0': solve_ms, 'gather_ms_rank0': gather_ms, 'problems_per_s': args.B / (ms * 1e-3),
                                  'problem_iters_per_s': args.B * it / (ms * 1e-3),
                                  'whole_solve_GBps_per_gpu': alg / (ms * 1e-3) / 1e9, 'frac_of_measured_hbm': alg / (ms * 1e-3) / 1e9 / peak,
                                  'gpu_launches_rank0': launches, 'finite': ok, 'scaling': 'strong', 'mode': args.mode}), flush=True)
            del finals, full
            torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
