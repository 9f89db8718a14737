#!/usr/bin/env python
"""bench.py — headline benchmark of the batched NLLS hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--path fused|tensor]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): lsq_gna_parallel EMG fit, B=65536 problems per GPU, C=5 costs,
N=1024 residuals, P=4 parameters, fp64, tol 1e-8, max_iter 80 (README.rst:80-84), synthetic data
(SURVEY.md §8d).  A "step" is ONE full solve of the batch through the public drop-in API
(torchimize_b200.functions.lsq_gna_parallel); metric = problems solved per second.

  value     whole-job problems/s with inputs resident in HBM (timed with CUDA events, max over ranks)
  e2e       the same solves through torchimize_b200.pipeline.solve_host_batches: every step copies its inputs (p0, t, y)
            from pinned host memory and reads the final parameters back to the host; the copy of step k+1 overlaps the
            solve of step k.  e2e.h2d_ceiling_gbs = what plain concurrent cudaMemcpyAsync of the same buffers reaches on
            this box with all N ranks copying at once; e2e.frac_of_ceiling = the pipeline's H2D rate against it
  roofline  dominant kernel of the timed steps, CUDA-event timed per launch INSIDE the timed region (sustained clocks);
            achieved = SURVEY §8(d) algorithmic bytes / average launch duration (fused path: s*N_y per model evaluation)
  also_lm   lsq_lma_parallel meth='lev' on the same batch with its own roofline record (the metric names LM and GN)
  retire_mode  the opt-in mode='retire' on the same batch (a converged problem is frozen and never evaluated again): problems/s,
            mean iterations per problem, deviation of its finals from the exact mode — reported separately, not the headline
  tensor_path  K1 (nlls_normal_eq, the HBM-bound kernel of the tensor path) at the same shape: s*C*N*(P+2) bytes per launch
  exact_mode   (N >= 2) sharded solves in modes local / exact (NCCL all-reduce) / exact-peer (nlls_iter_end_peer over NVLink
            peer memory): bit identity with the single-GPU run and microseconds per iteration
  cfg5      (N >= 2) BASELINE.json configs[4]: 4 Mi EMG fits strong-scaled over the N ranks, LM and GN, fp32 and fp64,
            with a CPU sample of the same solves beside it
  cpu_baseline  the oracle (CPU restatement of the reference's torch-op path) on a bounded
            sample of the same workload on the host cores (rank 0, N=1 only)

--impl reference times that oracle port alone on the host cores (the reference is pure Python over
torch CPU ops; it cannot be pip-installed here — its setup.py imports docutils, which the image lacks — and copying its
sources is against the rules, so the arm runs the oracle, which tests/test_oracle_golden.py pins bit-for-bit to it).
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
CFG5_B = 4 * 1024 * 1024


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
def oracle_solve(B, work, threads, solver='gna', seed=None):
    """One solve of a B-problem sample of the workload with the oracle on the host cores."""
    from oracle import nlls_oracle as O, models as M
    torch.set_num_threads(threads)
    dt = torch.float64 if work['dtype'] == 'f64' else torch.float32
    p0, t, y = make_batch(B, work['C'], work['N'], work['seed'] if seed is None else seed, 'cpu', dt)
    w = torch.ones(work['C'], dtype=dt)
    tol = work['tol']
    t0 = time.perf_counter()
    if solver == 'lma':
        tr = O.lma(p0.clone(), M.emg_cost, M.emg_jac, args=(t, y), wvec=w, ftol=tol, ptol=tol, gtol=tol, meth='lev', max_iter=40)
    else:
        tr = O.gna(p0.clone(), M.emg_cost, M.emg_jac, args=(t, y), wvec=w, ftol=tol, ptol=tol, gtol=tol,
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
class Ctx:
    """process / device / timing plumbing shared by the legs of the benchmark"""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get('WORLD_SIZE', '1'))
        self.rank = int(os.environ.get('RANK', '0'))
        self.local = int(os.environ.get('LOCAL_RANK', '0'))
        torch.cuda.set_device(self.local)
        self.dev = torch.device('cuda', self.local)
        if self.world > 1:
            dist.init_process_group('nccl', device_id=self.dev)
        self.peaks = {}
        try:
            self.peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:                                        # noqa: BLE001
            pass
        self.hbm = float(self.peaks.get('hbm_gbs', 6650.0))
        self.peak_source = 'MEASURED_PEAKS.json hbm_gbs (measured)' if self.peaks else 'fallback 6650 GB/s'

    def barrier(self):
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, v):
        t = torch.tensor([v], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, v):
        t = torch.tensor([v], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def timed(self, fn, steps):
        """K calls of fn bracketed by barrier + synchronize; CUDA-event milliseconds, max over ranks"""
        evs = []
        out = None
        self.barrier()
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = fn()
            b.record()
            evs.append((a, b))
        self.barrier()
        return self.max_over_ranks(sum(a.elapsed_time(b) for a, b in evs)), out


def kernel_table(recs):
    per = {}
    for name, ms in recs:
        per.setdefault(name, []).append(ms)
    return per


def live_average(times):
    """launches enqueued after the batch converged return at once (ctrl->done): not part of the average"""
    live = [v for v in times if v > .05 * max(times)]
    return sum(live) / len(live), live


def ncu_record(kernel):
    """per-launch ncu counters of `kernel` from profiles/traffic.json — only when they were captured on THIS build
    (hash of csrc/ + the header); a stale capture is not a measurement of the benched binary"""
    from torchimize_b200 import _lib
    try:
        tr = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))
    except Exception:                                            # noqa: BLE001
        return None
    rec = tr.get(kernel)
    if not isinstance(rec, dict) or rec.get('source_hash') != _lib.source_hash():
        return None
    return rec


def roofline_record(ctx, kernel, avg_ms, n_live, share, B, C, N, P, esz, per_step):
    """SURVEY §8(d) roofline of one kernel; DRAM traffic and FP64-pipe use as named side fields"""
    fused = kernel in ('nlls_model_normal_eq', 'nlls_model_lm_trial')
    if fused:
        alg = esz * C * N * B
        note = 'SURVEY 8d fused path: s*N_y per model evaluation, N_y = C*N distinct data scalars, one evaluation per launch'
        dram_model = esz * N * (C + 2) * B
        dram_note = 'designed DRAM traffic: y once + m_prev read + m written = s*N*(C+2) per problem (DESIGN.md 4)'
    elif kernel == 'nlls_normal_eq':
        alg = esz * C * N * (P + 2) * B
        note = 'SURVEY 8d tensor path GN: J + f + f_prev = s*C*N*(P+2)'
        dram_model, dram_note = alg, 'same as the algorithmic bytes'
    else:                                                        # nlls_model_eval: y read, f + J written
        alg = esz * C * N * (P + 2) * B
        note = 'y read + f, J written per problem-iteration: s*C*N*(P+2)'
        dram_model, dram_note = alg, 'same as the algorithmic bytes'
    t = avg_ms * 1e-3
    rec = {'bound': 'hbm', 'kernel': kernel, 'achieved': alg / t / 1e9, 'peak': ctx.hbm, 'unit': 'GB/s',
           'frac': alg / t / 1e9 / ctx.hbm, 'traffic': None, 'peak_source': ctx.peak_source, 'avg_launch_ms': avg_ms,
           'launches_timed': n_live, 'share_of_step': share, 'algorithmic_bytes_per_launch': alg, 'bytes_note': note,
           'dram_model_bytes_per_launch': dram_model, 'dram_utilisation': dram_model / t / 1e9 / ctx.hbm, 'dram_note': dram_note,
           'kernels_ms_per_step': per_step,
           'timing': 'CUDA events around every launch inside the timed steps (sustained clocks; the event records are part of '
                     'the timed region, so `value` is conservative by their cost)'}
    ncu = ncu_record(kernel)
    if ncu:
        rec['traffic'] = ncu.get('dram_bytes')
        rec['traffic_source'] = ncu.get('source')
        if ncu.get('fp64_pipe_pct') is not None and ncu.get('duration_ms'):
            # pipe-busy cycles are a property of the instruction stream: rescale the ncu fraction to the live launch time
            scale = ncu['duration_ms'] / avg_ms
            rec['fp64_pipe_frac'] = ncu['fp64_pipe_pct'] / 100. * scale
            if ncu.get('issue_active_pct') is not None:
                rec['issue_slot_frac'] = ncu['issue_active_pct'] / 100. * scale
    else:
        rec['traffic_note'] = 'no ncu capture of this build in profiles/traffic.json (source hash differs): field dropped'
    if fused:
        rec['limiter'] = ('FP64 pipe and issue slots of the in-kernel model evaluation (2 exp + erf + 12 FP64 operations per '
                          'cost and sample), not HBM: DRAM moves s*N*(C+2) per problem (dram_utilisation)')
    return rec


def leg_retire(ctx, p0, t, y, w, emg, tol, work, exact_finals, fused):
    """mode='retire' (not the reference's semantics: reported separately): problems/s, iterations per problem, deviation of the
    finals from the exact mode on the same batch."""
    from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel
    out = {'note': "opt-in mode='retire': a problem whose own stop test fired is frozen at that iterate and never evaluated again "
                   '(device-resident active list, no host round-trip); NOT the reference semantics, hence not the headline'}
    B = p0.shape[0]
    for name, solver, kw in (('gna', lsq_gna_parallel, dict(max_iter=work['max_iter'])), ('lma', lsq_lma_parallel, dict(max_iter=40, meth='lev'))):
        p = torch.empty_like(p0)
        st = {}

        def step():
            p.copy_(p0)
            return solver(p, emg.cost, emg.jac, args=(t, y), wvec=w, ftol=tol, ptol=tol, gtol=tol, fused=fused, mode='retire', stats=st, **kw)
        for _ in range(2):
            step()
        steps = 3
        ms, coeffs = ctx.timed(step, steps)
        stop = st['stop_iteration']
        iters_pp = torch.where(stop >= 0, stop + 1, torch.full_like(stop, len(coeffs))).double()
        rec = {'value': B * ctx.world * steps / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms / steps, 'iterations': len(coeffs),
               'mean_iterations_per_problem': float(iters_pp.mean()), 'problems_never_stopped': int((stop < 0).sum())}
        ex = exact_finals.get(name)
        if ex is not None:
            rel = ((p - ex).norm(dim=1) / ex.norm(dim=1))
            rec['finals_vs_exact_mode'] = {'max_rel': float(rel.max()), 'median_rel': float(rel.median()),
                                           'note': 'exact mode keeps stepping every problem until all converge together'}
        out[name] = rec
        del coeffs
    return out


def leg_tensor_path(ctx, B, C, N, P, dt, esz):
    """K1 alone at the headline shape: the HBM-bound kernel of north_star item (1), driver-visible."""
    from torchimize_b200 import _lib
    dev = ctx.dev
    g = torch.Generator(device=dev).manual_seed(11)
    J = torch.randn(B, C, N, P, generator=g, dtype=dt, device=dev)
    f = torch.randn(B, C, N, generator=g, dtype=dt, device=dev)
    fp = torch.randn(B, C, N, generator=g, dtype=dt, device=dev)
    w = torch.ones(C, dtype=dt, device=dev)
    gg = torch.empty(B, P, dtype=dt, device=dev)
    H = torch.empty(B, P, P, dtype=dt, device=dev)
    st = torch.empty(B, 4, dtype=torch.float64, device=dev)
    code = _lib.dtype_code(J)

    def k1():
        _lib.call('nlls_normal_eq', code, _lib.ptr(J), _lib.ptr(f), _lib.ptr(fp), _lib.ptr(w), C, B, C, N, P, _lib.ptr(gg),
                  _lib.ptr(H), _lib.ptr(st), None, _lib.stream_ptr())
    for _ in range(3):
        k1()
    n = 10
    _lib.enable_timing(True)
    ms, _ = ctx.timed(k1, n)
    recs = _lib.timing_records()
    _lib.enable_timing(False)
    avg = sum(v for _, v in recs) / len(recs)
    alg = esz * C * N * (P + 2) * B
    rec = {'kernel': 'nlls_normal_eq', 'shape': [B, C, N, P], 'dtype': 'f64' if esz == 8 else 'f32', 'launches_timed': n,
           'avg_launch_ms': avg, 'algorithmic_bytes_per_launch': alg, 'bytes_note': 'SURVEY 8d: J + f + f_prev = s*C*N*(P+2)',
           'achieved': alg / (avg * 1e-3) / 1e9, 'unit': 'GB/s', 'peak': ctx.hbm, 'frac': alg / (avg * 1e-3) / 1e9 / ctx.hbm,
           'data': 'seeded N(0,1) J, f, f_prev resident in HBM (J is 10.7 GB >> L2)'}
    ncu = ncu_record('nlls_normal_eq')
    rec['traffic'] = ncu.get('dram_bytes') if ncu else None
    del J, f, fp
    torch.cuda.empty_cache()
    return rec


def leg_h2d_ceiling(ctx, hy, reps=3):
    """Plain pinned host->device copies of the e2e step's big buffer, one cudaMemcpyAsync per copy on one stream per rank,
    all ranks copying at the same time: what the box gives N concurrent readers of host memory."""
    dst = torch.empty(hy.shape, dtype=hy.dtype, device=ctx.dev)
    st = torch.cuda.Stream(ctx.dev)
    nbytes = hy.numel() * hy.element_size()
    best = None
    for i in range(reps + 1):
        ctx.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            a.record(st)
            dst.copy_(hy, non_blocking=True)
            b.record(st)
        st.synchronize()
        ms = ctx.max_over_ranks(a.elapsed_time(b))               # slowest rank: the copies overlap for that long
        if i > 0:
            best = ms if best is None else min(best, ms)
    del dst
    torch.cuda.empty_cache()
    return {'aggregate_gbs': nbytes * ctx.world / (best * 1e-3) / 1e9, 'per_gpu_gbs': nbytes / (best * 1e-3) / 1e9,
            'bytes_per_rank': nbytes, 'ms': best, 'ranks_copying': ctx.world}


def leg_exact_mode(ctx):
    """Sharded solves of ONE global batch in modes local / exact / exact-peer (latency regime: 2048 problems per GPU)."""
    from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel, gradient_descent_parallel
    from torchimize_b200.models import EMGModel
    from torchimize_b200.sharding import shard_bounds, solve_sharded, gather_results
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    per = 2048
    B, C, N = per * world + 37, 3, 512
    p0, t, y = make_batch(B, C, N, 77, dev, torch.float64)        # same seed on every rank: identical full batch
    w = torch.ones(C, dtype=torch.float64, device=dev)
    emg = EMGModel()
    lo, hi = shard_bounds(B, world, rank)
    tol = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8)
    out = {'problems_total': B, 'shape': [B, C, N, 4], 'dtype': 'f64', 'solvers': {}}
    all_ok = True
    for name, solver, kw in (('gna', lsq_gna_parallel, dict(max_iter=60)), ('lma', lsq_lma_parallel, dict(max_iter=40, meth='lev')),
                             ('gda', gradient_descent_parallel, dict(max_iter=30))):
        single = p0.clone()
        n_single = len(solver(single, emg.cost, emg.jac, args=(t, y), wvec=w, **tol, **kw))
        rec = {'iterations_single_gpu': n_single}
        for mode in ('local', 'exact', 'exact-peer'):
            res = None
            for rep in range(3):                                 # rep 0 warms the mode up (communicator / mailbox set-up)
                p = p0[lo:hi].clone()
                ctx.barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                coeffs, _ = solve_sharded(solver, p, emg.cost, emg.jac, args=(t, y[lo:hi]), wvec=w, mode=mode, gather=False, **tol, **kw)
                b.record()
                torch.cuda.synchronize()
                ms = ctx.max_over_ranks(a.elapsed_time(b))
                n_it = int(ctx.max_over_ranks(len(coeffs)))
                if rep > 0:
                    res = ms if res is None else min(res, ms)
            full = gather_results(p, B)
            same = bool(torch.equal(full, single)) and n_it == n_single
            m = {'iterations': n_it, 'ms_per_solve': res, 'us_per_iteration': 1e3 * res / max(1, n_it)}
            if mode != 'local':
                m['bit_identical_to_single_gpu'] = same
                all_ok = all_ok and same
            rec[mode] = m
        rec['exact_peer_overhead_us_per_iteration'] = rec['exact-peer']['us_per_iteration'] - rec['local']['us_per_iteration']
        rec['exact_nccl_overhead_us_per_iteration'] = rec['exact']['us_per_iteration'] - rec['local']['us_per_iteration']
        out['solvers'][name] = rec
    flag = torch.tensor([1 if all_ok else 0], device=dev)
    ctx.dist.all_reduce(flag, op=ctx.dist.ReduceOp.MIN)
    out['all_bit_identical'] = bool(int(flag.item()) == 1)
    out['note'] = ("'local': every shard stops on its own (no exchange); 'exact': 32-byte NCCL all-reduce per iteration; 'exact-peer': "
                   'nlls_iter_end_peer = iteration end + exchange in ONE kernel over NVLink peer memory (P2P stores + stamps)')
    del p0, t, y
    torch.cuda.empty_cache()
    return out


def leg_cfg5(ctx, B_total):
    """BASELINE.json configs[4]: B_total EMG fits (C=5, N=1024, P=4) strong-scaled over the ranks, LM and GN, fp32 and fp64."""
    from torchimize_b200 import _lib
    from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel
    from torchimize_b200.models import EMGModel
    from torchimize_b200.sharding import shard_bounds, gather_results
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    C, N, P = 5, 1024, 4
    lo, hi = shard_bounds(B_total, world, rank)
    mine = hi - lo
    emg = EMGModel()
    out = {'B_total': B_total, 'shape_per_problem': [C, N, P], 'scaling': 'strong', 'mode': 'local', 'cases': {},
           'note': 'rank r solves problems shard_bounds(B_total, N, r); no per-iteration collective; one all-gather of the finals '
                   'inside the timed region; a shard that does not fit in HBM is solved in sequential chunks'}
    threads = os.cpu_count() or 1
    for dn in ('f64', 'f32'):
        dt = torch.float64 if dn == 'f64' else torch.float32
        es = 8 if dn == 'f64' else 4
        free = torch.cuda.mem_get_info(dev)[0]
        per_problem = (C + 2) * N * es + 4096                    # y + the two model-value buffers (+ small per-problem state)
        chunk = min(mine, int(free * .8) // per_problem)
        n_chunks = int(ctx.max_over_ranks((mine + chunk - 1) // chunk))    # the same number of (barrier-separated) chunks on every rank
        chunk = (mine + n_chunks - 1) // n_chunks
        w = torch.ones(C, dtype=dt, device=dev)
        for sv in ('gn', 'lm'):
            solver = lsq_gna_parallel if sv == 'gn' else lsq_lma_parallel
            kw = dict(max_iter=80) if sv == 'gn' else dict(max_iter=40, meth='lev')
            # warm-up on a small batch: module load of this dtype's kernels, pinned flag ring
            wp, wt, wy = make_batch(1024, C, N, 5, dev, dt)
            solver(wp, emg.cost, emg.jac, args=(wt, wy), wvec=w, ftol=1e-8, ptol=1e-8, gtol=1e-8, **kw)
            del wp, wt, wy
            finals = torch.empty(mine, P, dtype=dt, device=dev)
            solve_ms, iters, launches = 0.0, [], 0
            for ci in range(n_chunks):
                c0, c1 = ci * chunk, min(mine, (ci + 1) * chunk)
                p, t, y = make_batch(c1 - c0, C, N, 3008 + 7919 * (lo + c0), dev, dt)      # untimed: synthetic data
                ctx.barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                l0 = _lib.launches
                a.record()
                coeffs = solver(p, emg.cost, emg.jac, args=(t, y), wvec=w, ftol=1e-8, ptol=1e-8, gtol=1e-8, **kw)
                b.record()
                torch.cuda.synchronize()
                solve_ms += a.elapsed_time(b)
                iters.append(len(coeffs))
                launches += _lib.launches - l0
                finals[c0:c1] = p
                del coeffs, p, t, y
                torch.cuda.empty_cache()
            gather_results(finals, B_total)                       # untimed warm-up of the collective
            ctx.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            full = gather_results(finals, B_total)
            b.record()
            torch.cuda.synchronize()
            gather_ms = a.elapsed_time(b)
            ms = ctx.max_over_ranks(solve_ms + gather_ms)
            finite = bool(torch.isfinite(full).all())
            it = int(ctx.max_over_ranks(max(iters)))
            passes = it + (1 if sv == 'lm' else 0)               # speculative LM: one model pass per iteration + the first
            alg = es * C * N * mine * passes                     # s*N_y per model evaluation
            dram = es * N * (C + 2) * mine * passes
            rec = {'ms': ms, 'iterations': it, 'problems_per_s': B_total / (ms * 1e-3), 'problem_iters_per_s': B_total * it / (ms * 1e-3),
                   'B_per_gpu': mine, 'chunks_per_gpu': n_chunks, 'gather_ms_rank0': gather_ms, 'finite': finite,
                   'gpu_launches_rank0': launches,
                   'roofline_per_gpu': {'achieved': alg / (ms * 1e-3) / 1e9, 'unit': 'GB/s', 'peak': ctx.hbm,
                                        'frac': alg / (ms * 1e-3) / 1e9 / ctx.hbm, 'bytes_note': 'whole solve: s*C*N per model evaluation (SURVEY 8d)',
                                        'dram_utilisation': dram / (ms * 1e-3) / 1e9 / ctx.hbm}}
            if rank == 0:                                        # the CPU figure beside it: the oracle on a small sample of the same solves
                wk = dict(WORK, dtype=dn)
                n_cpu = 64 if (sv, dn) == ('gn', 'f32') else 128      # fp32 GN runs to max_iter = 80 on the CPU as well
                s, cit = oracle_solve(n_cpu, wk, threads, solver='lma' if sv == 'lm' else 'gna', seed=3008 + 7919 * lo)
                rec['cpu_sample'] = {'value': n_cpu / s, 'unit': UNIT, 'cores': threads, 'kind': 'port', 'iterations': cit,
                                     'sample': f'oracle on the first {n_cpu} problems of rank 0\'s shard, {s:.1f} s'}
            out['cases'][f'{sv}-{dn}'] = rec
            del finals, full
            torch.cuda.empty_cache()
    return out


def run_ours(args):
    from torchimize_b200 import _lib
    from torchimize_b200.functions import lsq_gna_parallel, lsq_lma_parallel
    from torchimize_b200.models import EMGModel
    from torchimize_b200.sharding import gather_results

    ctx = Ctx()
    dist, world, rank, local, dev = ctx.dist, ctx.world, ctx.rank, ctx.local, ctx.dev
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
    ms_total, coeffs = ctx.timed(step_resident, args.steps)
    recs = _lib.timing_records()
    _lib.enable_timing(False)
    launches = _lib.launches - launches0
    iters = iters_seen[-1]
    del coeffs
    exact_finals = {'gna': p.clone()}

    # ---- the same batch through lsq_lma_parallel meth='lev' (BASELINE.json's metric is quoted for LM and GN): reported
    # beside the headline GN line, same timing rules (resident inputs, CUDA events, per-launch kernel times)
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
    ms_lm, _ = ctx.timed(step_lm, lm_steps)
    lm_per = kernel_table(_lib.timing_records())
    _lib.enable_timing(False)
    exact_finals['lma'] = p.clone()
    lm_top = max(lm_per, key=lambda k: sum(lm_per[k]))
    lm_avg, lm_live = live_average(lm_per[lm_top])
    lm_tot = sum(sum(v) for v in lm_per.values())
    also_lm = {'solver': "lsq_lma_parallel meth='lev', max_iter 40, same batch", 'ms_per_step': ms_lm / lm_steps, 'steps': lm_steps,
               'iterations': lm_iters[-1], 'value': B * world * lm_steps / (ms_lm * 1e-3), 'unit': UNIT,
               'roofline': roofline_record(ctx, lm_top, lm_avg, len(lm_live), sum(lm_live) / lm_tot if lm_tot else None, B, C, N, P, esz,
                                           {k: sum(v) / lm_steps for k, v in lm_per.items()})}

    retire_rec = leg_retire(ctx, p0, t, y, w, emg, tol, work, exact_finals, fused) if fused else None
    del exact_finals

    # ---- per-kernel breakdown of the timed resident steps (sustained clocks, same launches as `value`)
    per = kernel_table(recs)
    tot = {k: sum(v) for k, v in per.items()}
    top_name = max(tot, key=tot.get)
    avg_ms, live = live_average(per[top_name])
    step_ms_profiled = sum(tot.values())
    roof = roofline_record(ctx, top_name, avg_ms, len(live), sum(live) / step_ms_profiled if step_ms_profiled else None, B, C, N, P, esz,
                           {k: sum(v) / args.steps for k, v in per.items()})

    # ---- end-to-end steps: host buffers in, host result out, through the public host-batch call
    # (torchimize_b200.pipeline.solve_host_batches): every step copies ITS inputs (p0, t, y) from pinned host memory and
    # reads its finals back; the copy of step k+1 runs on a copy stream while step k solves (two device slots).  The
    # timed region holds exactly K copies, K solves and K result reads — the first copy is not hidden.
    from torchimize_b200.pipeline import solve_host_batches, host_memory_near
    with host_memory_near(local) as numa:                        # one process per GPU: pinned buffers on the GPU's NUMA node
        hp0, ht = p0.cpu().pin_memory(), t.cpu().pin_memory()
        hy = torch.empty(y.shape, dtype=y.dtype).pin_memory()    # first touched (by the copy below) under the policy
        hy.copy_(y)
    del y
    torch.cuda.empty_cache()
    ceiling = leg_h2d_ceiling(ctx, hy)
    e2e_iters = []

    def run_e2e(steps):
        out = None
        for out, n_it in solve_host_batches(lsq_gna_parallel, ((hp0, (ht, hy)) for _ in range(steps)), emg.cost, emg.jac,
                                            wvec=w, ftol=tol, ptol=tol, gtol=tol, max_iter=work['max_iter'], fused=fused):
            e2e_iters.append(n_it)
        return out

    e2e_steps = max(1, args.steps)
    run_e2e(2)                                                   # warm-up: slot allocation, pinned result buffers
    ms_e2e, h_out = ctx.timed(lambda: run_e2e(e2e_steps), 1)
    assert e2e_iters[-1] == iters, f'e2e step ran {e2e_iters[-1]} iterations, resident step {iters}'
    assert bool(torch.isfinite(h_out).all())
    sampler.stop_flag = True
    sampler.join(timeout=2)
    h2d = (hp0.numel() + ht.numel() + hy.numel()) * esz
    d2h = h_out.numel() * esz
    e2e_h2d_rate = h2d * world * e2e_steps / (ms_e2e * 1e-3) / 1e9
    del hy
    torch.cuda.empty_cache()

    # ---- K1 at the same shape (tensor path), multi-GPU records
    tensor_rec = leg_tensor_path(ctx, B, C, N, P, dt, esz)
    exact_rec = leg_exact_mode(ctx) if world > 1 else None
    cfg5_rec = None
    if world > 1 and not args.no_cfg5:
        del p0, t, p
        torch.cuda.empty_cache()
        cfg5_rec = leg_cfg5(ctx, args.cfg5_B)

    if rank == 0:
        val = B * world * args.steps / (ms_total * 1e-3)
        e2e_val = B * world * e2e_steps / (ms_e2e * 1e-3)
        line = {'metric': METRIC, 'value': val, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
                'ms_per_step': ms_total / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
                'dtype': work['dtype'], 'data': 'synthetic',
                'config': {'workload': f'lsq_gna_parallel EMG fit B={B} per GPU, C={C}, N={N}, P={P}, {work["dtype"]}, tol 1e-8, '
                                       f'max_iter {work["max_iter"]} (BASELINE.json configs[1])', 'path': args.path,
                           'iterations': iters, 'problem_iters_per_s': val * iters,
                           'l2': f'inputs {B * C * N * esz / 1e9:.2f} GB per GPU >> 126 MB L2 (no flush needed)',
                           'parallelism': f'B sharded over {world} GPU(s), no per-iteration collective, one all-gather of finals'},
                'e2e': {'value': e2e_val, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                        'ms_per_step': ms_e2e / e2e_steps, 'steps': e2e_steps,
                        'h2d_gbs': e2e_h2d_rate, 'h2d_ceiling_gbs': ceiling['aggregate_gbs'],
                        'h2d_ceiling_gbs_per_gpu': ceiling['per_gpu_gbs'], 'frac_of_ceiling': e2e_h2d_rate / ceiling['aggregate_gbs'],
                        'h2d_ceiling_note': f'{ceiling["ranks_copying"]} rank(s) each copying the same {ceiling["bytes_per_rank"] / 1e9:.2f} GB pinned '
                                            f'buffer with one cudaMemcpyAsync at the same time, best of 3: {ceiling["ms"]:.1f} ms',
                        'host_numa_node': numa.node, 'host_numa_policy_set': numa.ok,
                        'api': 'torchimize_b200.pipeline.solve_host_batches (double-buffered H2D on a copy stream; K copies + '
                               'K solves + K result reads inside the timed region)'},
                'gpu_launches': launches, 'roofline': roof, 'clocks': sampler.summary(), 'also_lm': also_lm,
                'tensor_path': tensor_rec}
        if retire_rec is not None:
            line['retire_mode'] = retire_rec
        if exact_rec is not None:
            line['exact_mode'] = exact_rec
        if cfg5_rec is not None:
            line['cfg5'] = cfg5_rec
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
    ap.add_argument('--no-cfg5', action='store_true', help='skip the cfg-5 record at N >= 2')
    ap.add_argument('--cfg5-B', type=int, default=CFG5_B, help='total problems of the cfg-5 record (default 4 Mi)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
