#!/usr/bin/env python
"""bench.py — headline benchmark of the batched NLLS hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--path fused|tensor]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): lsq_gna_parallel EMG fit, B=65536 problems per GPU, C=5 costs,
N=1024 residuals, P=4 parameters, fp64, tol 1e-8, max_iter 80 (README.rst:80-84), synthetic data
(SURVEY.md §8d).  A "step" is ONE full solve of the batch through the public drop-in API
(torchimize_b200.functions.lsq_gna_parallel); metric = problems solved per second.

  value   whole-job problems/s with inputs resident in HBM (timed with CUDA events, max over ranks)
  e2e     the same solves through torchimize_b200.pipeline.solve_host_batches: every step copies its inputs (p0, t, y)
          from pinned host memory and reads the final parameters back to the host; the copy of step k+1 overlaps the
          solve of step k (copy stream, two device slots)
  roofline  dominant kernel of the timed steps, CUDA-event timed per launch INSIDE the timed region
          (sustained clocks); achieved = algorithmic bytes / average launch duration
  cpu_baseline  the oracle (CPU restatement of the reference's torch-op path) on a bounded
          sample of the same workload on the host cores (rank 0, N=1 only)

--impl reference times that oracle port alone on the host cores (the reference is pure Python over
torch CPU ops and cannot travel to the GPU box; the oracle is its pinned restatement).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch                                                     # noqa: E402

WORK = dict(B=65536, C=5, N=1024, P=4, dtype='f64', solver='gna', tol=1e-8, max_iter=80, seed=3008)
METRIC = 'LM/GN problems solved per second (EMG fits, B x C costs)'
UNIT = 'problems/s'


# ------------------------------------------------------------------------------- data
def emg_model_t(p, t):
    """Plain torch EMG model used ONLY to synthesise data (outside every timed region)."""
    a, mu, s, eta = (p[:, i, None] for i in range(4))
    d = t - mu
    return a * torch.exp(-.5 * d ** 2 / s ** 2) * (1 + torch.erf(eta * d / (s * 2 ** .5)))


def make_batch(B, C, N, seed, device, dtype):
    """Seeded synthetic EMG batch: gt = [10,-1,80,5] jittered +-5 %, init = gt
    jittered +-10 %, y = emg(gt) + 0.01 N(0,1), distinct noise per cost channel."""
    g = torch.Generator(device=device).manual_seed(seed)
    base = torch.tensor([10., -1., 80., 5.], dtype=torch.float64, device=device)
    gt = base * (1 + .1 * (torch.rand(B, 4, generator=g, dtype=torch.float64, device=device) - .5))
    p0 = gt * (1 + .2 * (torch.rand(B, 4, generator=g, dtype=torch.float64, device=device) - .5))
    t = torch.linspace(-1e3, 1e3, N, device=device).double()      # built in fp32 then cast (unit_test_parallel.py:42)
    y = torch.empty(B, C, N, dtype=dtype, device=device)
    clean = None
    chunk = 8192
    for b0 in range(0, B, chunk):
        clean = emg_model_t(gt[b0:b0 + chunk], t)
        for c in range(C):
            noise = torch.randn(clean.shape, generator=g, dtype=torch.float64, device=device)
            y[b0:b0 + chunk, c] = (clean + .01 * noise).to(dtype)
    return p0.to(dtype), t.to(dtype), y


# ----------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:                                        # noqa: BLE001
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: 'hw_slowdown',
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: 'hw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: 'sw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonSwPowerCap: 'sw_power_cap'}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:                                    # noqa: BLE001
                pass
            time.sleep(0.05)

    def summary(self):
        s = sorted(self.samples)
        return {'sm_mhz': (s[len(s) // 2] if s else None), 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
                'samples': len(s)}


# ------------------------------------------------------------------- reference (CPU) arm
def oracle_solve(B, work, threads):
    """One GN solve of a B-problem sample of the workload with the oracle on the host cores."""
    from oracle import nlls_oracle as O, models as M
    torch.set_num_threads(threads)
    dt = torch.float64 if work['dtype'] == 'f64' else torch.float32
    p0, t, y = make_batch(B, work['C'], work['N'], work['seed'], 'cpu', dt)
    w = torch.ones(work['C'], dtype=dt)
    t0 = time.perf_counter()
    tr = O.gna(p0.clone(), M.emg_cost, M.emg_jac, args=(t, y), wvec=w, ftol=work['tol'], ptol=work['tol'], gtol=work['tol'],
               max_iter=work['max_iter'])
    dt_s = time.perf_counter() - t0
    return dt_s, tr.iterations


def cpu_baseline(work, budget_s=20.0):
    threads = os.cpu_count() or 1
    probe_B = 64
    s, _ = oracle_solve(probe_B, work, threads)                  # warm-up + cost probe
    per_problem = s / probe_B
    B = int(max(64, min(4096, budget_s / max(per_problem, 1e-9))))
    s, iters = oracle_solve(B, work, threads)
    return {'value': B / s, 'unit': UNIT, 'cores': threads, 'kind': 'port',
            'sample': f'oracle GN on the first-seeded {B} of {work["B"]} problems (C={work["C"]},N={work["N"]},P=4,{work["dtype"]}), '
                      f'{iters} iterations, {s:.1f} s, torch {torch.__version__} CPU, {threads} threads',
            'iterations': iters}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    work = dict(WORK)
    s, _ = oracle_solve(64, work, threads)
    total_budget = 90.0
    B = int(max(64, min(2048, total_budget / (args.steps + args.warmup) / max(s / 64, 1e-9))))
    for _ in range(args.warmup):
        oracle_solve(B, work, threads)
    times, iters = [], 0
    for _ in range(args.steps):
        s, iters = oracle_solve(B, work, threads)
        times.append(s)
    tot = sum(times)
    val = B * args.steps / tot
    line = {'metric': METRIC, 'value': val, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': 1e3 * tot / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': work['dtype'], 'data': 'synthetic', 'impl': 'reference',
            'config': {'workload': f'lsq_gna_parallel EMG fit, C={work["C"]}, N={work["N"]}, P=4, {work["dtype"]}, tol 1e-8, '
                                   f'max_iter {work["max_iter"]}; bounded sample of {B} problems per step on host cores',
                       'iterations': iters},
            'cpu_baseline': {'value': val, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                             'sample': f'{B} problems per step, {args.steps} steps'},
            'e2e': {'value': val, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch.distributed as dist
    from torchimize_b200 import _lib
    from torchimize_b200.functions import lsq_gna_parallel
    from torchimize_b200.models import EMGModel
    from torchimize_b200.sharding import gather_results

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    work = dict(WORK)
    if args.B:
        work['B'] = args.B
    B, C, N, P = work['B'], work['C'], work['N'], work['P']
    dt = torch.float64 if work['dtype'] == 'f64' else torch.float32
    esz = 8 if dt == torch.float64 else 4
    assert _lib.load().nlls_device_check() == 0, _lib.load().nlls_last_error().decode()

    p0, t, y = make_batch(B, C, N, work['seed'] + 1000 * rank, dev, dt)      # weak scaling: B problems per GPU
    w = torch.ones(C, dtype=dt, device=dev)
    emg = EMGModel()
    fused = args.path == 'fused'
    tol = work['tol']

    def solve(p, tt, yy):
        return lsq_gna_parallel(p, emg.cost, emg.jac, args=(tt, yy), wvec=w, ftol=tol, ptol=tol, gtol=tol,
                                max_iter=work['max_iter'], fused=fused)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        evs = []
        barrier()
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = fn()
            b.record()
            evs.append((a, b))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        tms = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        return float(tms.item()), out

    # ---- resident-input steps ("value")
    p = torch.empty_like(p0)
    iters_seen = []

    def step_resident():
        p.copy_(p0)
        coeffs = solve(p, t, y)
        iters_seen.append(len(coeffs))
        if world > 1:
            gather_results(p, B * world)                         # the one collective of the path (finals)
        return coeffs

    for _ in range(args.warmup):
        step_resident()
    launches0 = _lib.launches
    sampler = ClockSampler(local)
    sampler.start()
    _lib.enable_timing(True)                                     # CUDA events around every launch of the TIMED steps
    ms_total, coeffs = timed(step_resident, args.steps)
    recs = _lib.timing_records()
    _lib.enable_timing(False)
    launches = _lib.launches - launches0
    iters = iters_seen[-1]

    # ---- the same batch through lsq_lma_parallel meth='lev' (BASELINE.json's metric is quoted for LM and GN): reported
    # beside the headline GN line, same timing rules (resident inputs, CUDA events, per-launch kernel times)
    from torchimize_b200.functions import lsq_lma_parallel
    lm_iters = []

    def step_lm():
        p.copy_(p0)
        coeffs = lsq_lma_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, ftol=tol, ptol=tol, gtol=tol, meth='lev',
                                  max_iter=40, fused=fused)
        lm_iters.append(len(coeffs))
        return coeffs

    lm_steps = max(1, min(args.steps, 3))
    for _ in range(2):
        step_lm()
    _lib.enable_timing(True)
    ms_lm, _ = timed(step_lm, lm_steps)
    lm_recs = _lib.timing_records()
    _lib.enable_timing(False)
    lm_per = {}
    for name, ms in lm_recs:
        lm_per.setdefault(name, []).append(ms)
    lm_top = max(lm_per, key=lambda k: sum(lm_per[k]))
    lm_live = [v for v in lm_per[lm_top] if v > .05 * max(lm_per[lm_top])]
    lm_avg = sum(lm_live) / len(lm_live)
    lm_bytes = (3 * esz * C * N if fused else esz * C * N * (P + 3)) * B    # fused: y + f_cur + f_trial; tensor: SURVEY 8d LMA
    also_lm = {'solver': "lsq_lma_parallel meth='lev', max_iter 40, same batch", 'ms_per_step': ms_lm / lm_steps, 'steps': lm_steps,
               'iterations': lm_iters[-1], 'kernel': lm_top, 'avg_launch_ms': lm_avg,
               'achieved_GBps': lm_bytes / (lm_avg * 1e-3) / 1e9}

    # ---- end-to-end steps: host buffers in, host result out, through the public host-batch call
    # (torchimize_b200.pipeline.solve_host_batches): every step copies ITS inputs (p0, t, y) from pinned host memory and
    # reads its finals back; the copy of step k+1 runs on a copy stream while step k solves (two device slots).  The
    # timed region holds exactly K copies, K solves and K result reads — the first copy is not hidden.
    from torchimize_b200.pipeline import solve_host_batches, host_memory_near
    with host_memory_near(local) as numa:                        # one process per GPU: pinned buffers on the GPU's NUMA node
        hp0, ht = p0.cpu().pin_memory(), t.cpu().pin_memory()
        hy = torch.empty(y.shape, dtype=y.dtype).pin_memory()    # first touched (by the copy below) under the policy
        hy.copy_(y)
    e2e_iters = []

    def run_e2e(steps):
        out = None
        for out, n_it in solve_host_batches(lsq_gna_parallel, ((hp0, (ht, hy)) for _ in range(steps)), emg.cost, emg.jac,
                                            wvec=w, ftol=tol, ptol=tol, gtol=tol, max_iter=work['max_iter'], fused=fused):
            e2e_iters.append(n_it)
        return out

    e2e_steps = max(1, args.steps)
    run_e2e(2)                                                   # warm-up: slot allocation, pinned result buffers
    ms_e2e, h_out = timed(lambda: run_e2e(e2e_steps), 1)
    assert e2e_iters[-1] == iters, f'e2e step ran {e2e_iters[-1]} iterations, resident step {iters}'
    assert bool(torch.isfinite(h_out).all())
    sampler.stop_flag = True
    sampler.join(timeout=2)
    h2d = (hp0.numel() + ht.numel() + hy.numel()) * esz
    d2h = h_out.numel() * esz

    # ---- per-kernel breakdown of the timed resident steps (sustained clocks, same launches as `value`)
    per = {}
    for name, ms in recs:
        per.setdefault(name, []).append(ms)
    tot = {k: sum(v) for k, v in per.items()}
    top_name = max(tot, key=tot.get)
    # launches enqueued after the batch converged return at once (ctrl->done): not part of the average
    live = [v for v in per[top_name] if v > .05 * max(per[top_name])]
    avg_ms = sum(live) / len(live)
    step_ms_profiled = sum(tot.values())
    share = sum(live) / step_ms_profiled if step_ms_profiled else None
    if top_name == 'nlls_model_normal_eq':
        bytes_per_problem = 3 * esz * C * N                      # y read + f_prev read + f write
        byte_note = 'y + f_prev + f per problem-iteration: 3*s*C*N (SURVEY §8d fused figure s*C*N + stored f_prev traffic)'
    elif top_name == 'nlls_normal_eq':
        bytes_per_problem = esz * C * N * (P + 2)                # GN s*C*N*(P+2)
        byte_note = 'J + f + f_prev per problem-iteration: s*C*N*(P+2) (SURVEY §8d)'
    else:                                                        # nlls_model_eval: y read, f + J written
        bytes_per_problem = esz * C * N * (P + 2)
        byte_note = 'y read + f, J written per problem-iteration: s*C*N*(P+2)'
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:                                            # noqa: BLE001
        pass
    peak = float(peaks.get('hbm_gbs', 6650.0))
    achieved = bytes_per_problem * B / (avg_ms * 1e-3) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json'))).get(top_name)
    except Exception:                                            # noqa: BLE001
        pass
    roof = {'bound': 'hbm', 'kernel': top_name, 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
            'traffic': traffic, 'peak_source': 'MEASURED_PEAKS.json hbm_gbs (measured)' if peaks else 'fallback 6650 GB/s',
            'avg_launch_ms': avg_ms, 'launches_timed': len(live), 'share_of_step': share, 'algorithmic_bytes_per_launch':
            bytes_per_problem * B, 'bytes_note': byte_note,
            'achieved_y_only': (esz * C * N * B / (avg_ms * 1e-3) / 1e9) if top_name == 'nlls_model_normal_eq' else None,
            'y_only_note': 'SURVEY 8d lower bound s*C*N (data read once, f_prev recomputed instead of stored): the kernel stores f '
                           'because a second model evaluation per sample costs more than the 2*s*C*N bytes it saves (DESIGN.md 4)',
            'kernels_ms_per_step': {k: sum(v) / args.steps for k, v in per.items()},
            'timing': 'CUDA events around every launch inside the timed steps (sustained clocks)'}

    if rank == 0:
        val = B * world * args.steps / (ms_total * 1e-3)
        e2e_val = B * world * e2e_steps / (ms_e2e * 1e-3)
        line = {'metric': METRIC, 'value': val, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
                'ms_per_step': ms_total / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
                'dtype': work['dtype'], 'data': 'synthetic',
                'config': {'workload': f'lsq_gna_parallel EMG fit B={B} per GPU, C={C}, N={N}, P={P}, {work["dtype"]}, tol 1e-8, '
                                       f'max_iter {work["max_iter"]} (BASELINE.json configs[1])', 'path': args.path,
                           'iterations': iters, 'problem_iters_per_s': val * iters,
                           'l2': f'inputs {y.numel() * esz / 1e9:.2f} GB per GPU >> 126 MB L2 (no flush needed)',
                           'parallelism': f'B sharded over {world} GPU(s), no per-iteration collective, one all-gather of finals'},
                'e2e': {'value': e2e_val, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                        'ms_per_step': ms_e2e / e2e_steps, 'steps': e2e_steps,
                        'host_numa_node': numa.node, 'host_numa_policy_set': numa.ok,
                        'api': 'torchimize_b200.pipeline.solve_host_batches (double-buffered H2D on a copy stream; K copies + '
                               'K solves + K result reads inside the timed region)'},
                'gpu_launches': launches, 'roofline': roof, 'clocks': sampler.summary()}
        also_lm['value'] = B * world * lm_steps / (ms_lm * 1e-3)
        also_lm['frac_of_hbm_peak'] = also_lm['achieved_GBps'] / peak
        line['also_lm'] = also_lm
        if world == 1 and not args.no_cpu_baseline:
            line['cpu_baseline'] = cpu_baseline(work)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--path', default='fused', choices=['fused', 'tensor'])
    ap.add_argument('--B', type=int, default=0, help='override problems per GPU (debug)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
