#!/usr/bin/env python
"""Where does the bench step spend the time that is not inside our kernels?  Prints the idle gaps between
consecutive C-ABI launches of one solve (CUDA events), and the step time with/without the workspace set-up."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import make_batch, WORK
from torchimize_b200 import _lib
from torchimize_b200.functions import lsq_gna_parallel
from torchimize_b200.models import EMGModel

dev = torch.device('cuda')
B, C, N = WORK['B'], WORK['C'], WORK['N']
p0, t, y = make_batch(B, C, N, 3008, dev, torch.float64)
w = torch.ones(C, dtype=torch.float64, device=dev)
emg = EMGModel()
p = torch.empty_like(p0)
def solve():
    p.copy_(p0)
    return lsq_gna_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=w, ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=80)
for _ in range(3):
    solve()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); solve(); b.record(); torch.cuda.synchronize()
print('step ms', a.elapsed_time(b))
_lib.enable_timing(True)
s0 = torch.cuda.Event(enable_timing=True); s0.record()
solve()
s1 = torch.cuda.Event(enable_timing=True); s1.record()
torch.cuda.synchronize()
recs = _lib._timing
print('first launch starts %.3f ms after step start' % s0.elapsed_time(recs[0][1]))
print('step end %.3f ms after last launch end' % recs[-1][2].elapsed_time(s1))
busy = sum(x.elapsed_time(y_) for _, x, y_ in recs)
gaps = [recs[i][2].elapsed_time(recs[i + 1][1]) for i in range(len(recs) - 1)]
print('launches', len(recs), 'busy ms %.3f' % busy, 'gaps total ms %.3f' % sum(gaps), 'max gap %.3f' % max(gaps))
big = sorted(((g, i, recs[i][0], recs[i + 1][0]) for i, g in enumerate(gaps)), reverse=True)[:8]
for g, i, n0, n1 in big:
    print('  gap %.3f ms after #%d %s -> %s' % (g, i, n0, n1))
per = {}
for n, x, y_ in recs:
    per.setdefault(n, []).append(x.elapsed_time(y_))
for n, v in per.items():
    print(' ', n, len(v), 'first %.3f' % v[0], 'avg %.4f' % (sum(v) / len(v)), 'last %.4f' % v[-1])
print('normal_eq per-launch ms:', ' '.join('%.3f' % v for v in per['nlls_model_normal_eq']))
import threading, time
try:
    import pynvml
    pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
    samples = []
    stop = False
    def samp():
        while not stop:
            samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.))
            time.sleep(0.01)
    th = threading.Thread(target=samp); th.start()
    _lib.enable_timing(False)
    for _ in range(20):
        solve()
    torch.cuda.synchronize()
    stop = True; th.join()
    print('clock/power samples during 20 solves:', samples[::8])
except Exception as e:
    print('nvml failed', e)
